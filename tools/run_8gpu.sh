#!/bin/bash
# The 8-GPU call of round 2 (charged x8: keep it short).  Order: what can fail first.
#   gpurun --gpus 8 --timeout 900 -- 'bash tools/run_8gpu.sh'
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
nvidia-smi topo -m > gpurun_out/topo_${N}gpu.txt 2>&1
timeout 300 python -m pytest tests -m gpu -q -rs -k "p2p or pod or nvml" > gpurun_out/pytest_gpu_${N}.log 2>&1; tail -5 gpurun_out/pytest_gpu_${N}.log
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --mount-iters 6 2> gpurun_out/bench${N}.err | tail -1 > gpurun_out/bench_${N}gpu.json; echo "bench rc=${PIPESTATUS[0]}"
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_${N}gpu.json"))
    print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms/call", round(d["e2e"]["ms_per_call"], 2), "cold", round(d["e2e"]["cold_call"]["ms_per_call"], 2))
    p = d.get("p2p", {})
    print("p2p ok", p.get("ok"), {k: p.get(k) for k in ("store_gbs", "load_gbs", "memcpy_peer_gbs", "bidir_gbs_per_gpu_min", "alltoall_gbs_per_gpu", "matrix_match_nvml", "mismatch_lanes", "unavailable")})
    m = d.get("mount", {}); u = m.get("mount_to_usable", {})
    print("mount_ok", d.get("mount_ok"), m.get("unavailable"), "cold in-pod", u.get("target"), "p50", u.get("p50_s"), "first", u.get("first_gpu_usable_p50_s"), u.get("median_phases_ms"))
    print("timeline", u.get("context_timeline_last_ms"))
    r = m.get("mount_to_usable_resident_worker", {}); print("resident", r.get("p50_s"), r.get("median_phases_ms"))
except Exception as e:
    print("bench summary failed:", e); print(open("gpurun_out/bench${N}.err").read()[-3000:])
PY
timeout 520 python tools/mount_latency.py --skip-mount-bench --churn-rounds 0 --churn-real-rounds 6 --churn-nodes-rounds 100 --usable-iters 5 --fanout --out gpurun_out/churn_real_${N}gpu.json > gpurun_out/churn.log 2>&1
python - <<PY
import json
d = json.load(open("gpurun_out/churn_real_${N}gpu.json"))
for k in ("mount_to_usable", "mount_to_usable_fanout"):
    for n, v in d.get(k, {}).items():
        print(k, n, "p50", round(v["p50_s"], 3), "p99", round(v["p99_s"], 3), "first", round(v["first_gpu_usable_p50_s"], 3), "cuInit", round(v["cuda_init_ms"]))
for k in ("churn_real", "churn_real_nodes_only"):
    c = d.get(k, {}); print(k, {x: c.get(x) for x in ("steps", "total_s", "mount_to_usable", "unmount_to_verified_gone", "leaked_fds", "leaked_gpu_processes", "devices_list_back_to_baseline", "device_free_bytes_delta", "zombies")})
PY
tail -c 600 gpurun_out/churn.log | grep -i -A5 "error\|Traceback" | head -20
