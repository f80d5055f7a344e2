#!/bin/bash
# One call on a 2-GPU box: peer-store sweep (bulk engine vs vector stores), the P2P tests, bench at N=2, start-up timeline.
#   gpurun --gpus 2 --timeout 900 -- 'bash tools/run_2gpu.sh'
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_2gpu.txt 2>&1
timeout 300 python tools/sweep_p2p.py --out gpurun_out/sweep_p2p.json 2>&1 | tail -16
timeout 400 python -m pytest tests -m gpu -q -k "p2p or pod or nvml" > gpurun_out/pytest_gpu_2.log 2>&1; tail -6 gpurun_out/pytest_gpu_2.log
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 2> gpurun_out/bench2.err | tail -1 > gpurun_out/bench_2gpu.json; echo "bench rc=${PIPESTATUS[0]}"
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_2gpu.json"))
    print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms/call", round(d["e2e"]["ms_per_call"], 2), "cold", round(d["e2e"]["cold_call"]["ms_per_call"], 2))
    print("p2p", json.dumps(d.get("p2p"))[:1200])
    m = d.get("mount", {}); u = m.get("mount_to_usable", {})
    print("mount_ok", d.get("mount_ok"), m.get("unavailable"), "cold in-pod", u.get("target"), u.get("p50_s"), u.get("median_phases_ms"))
    print("timeline", u.get("context_timeline_last_ms"))
except Exception as e:
    print("bench summary failed:", e); print(open("gpurun_out/bench2.err").read()[-3000:])
PY
timeout 300 python tools/ctx_timeline.py --holders none,init,ctx --sizes 1,2 --iters 3 --out gpurun_out/ctx_timeline_2gpu.json | head -8
