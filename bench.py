#!/usr/bin/env python
"""bench.py — the post-mount HBM probe on N B200s, one JSON line.

Workload (BASELINE.json configs[1], "AddGPU of 1 B200 ... then the post-mount HBM probe"):
one STEP is one probe pass over a pair of N-byte device buffers (N = 8 GiB, >> 126 MB L2):

    hbm_write_pattern(A)      N   algorithmic bytes
    hbm_read_verify(A)        N
    hbm_copy(A -> B)          2N
                              4N  per step

`value`   = 4N * K * world / (device time of K steps, max over ranks)      [GB/s], buffers resident in HBM,
            launched through the C ABI's device-pointer entry points on torch's current stream.
`e2e`     = the same 4N per call, divided by the WALL time of the reference-facing one-shot call
            gmp_probe_hbm() (device select, cudaMalloc of both buffers, the passes, a pinned-host DMA round trip
            of `dma` bytes each way, verdict read-back, cudaFree) — host buffers and copies inside the timed region.
`roofline`= the copy kernel (largest share of the step) against MEASURED_PEAKS.json's hbm_gbs.
`cpu_baseline` = the oracle's CPU port of the same three passes on the host cores (bounded sample).
`--impl reference` times that CPU port alone (the reference itself has no device-memory code).

Multi-GPU: the probe shards by device with no data-path collective (weak scaling); torch.distributed is used
only for the barrier and the max-over-ranks of the device time.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GIB = 1 << 30
METRIC = "probe HBM GB/s per job (write+read-verify+copy pass, algorithmic bytes)"
SEED = 0x9E3779B97F4A7C15


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, torch copy_ read+write)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(prefix="gmp_clocks_", suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu_index), f"--query-gpu={self.FIELDS}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        try:
            rows = [r.strip().split(", ") for r in open(self.path) if r.strip()]
            os.unlink(self.path)
        except Exception:
            return out
        sm, mx, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                power.append(float(r[3]))
            except ValueError:
                continue
            for name, v in zip(names, r[5:9]):
                if v.strip().lower() == "active":
                    reasons.add(name)
        if sm:
            # "under load" = samples in the upper half of the power range seen
            lo, hi = min(power), max(power)
            load = [s for s, p in zip(sm, power) if p >= lo + 0.5 * (hi - lo)] or sm
            out.update(sm_mhz=statistics.median(load), sm_max_mhz=max(mx), reasons=sorted(reasons),
                       samples=len(sm), power_w_max=hi)
        return out


def max_over_ranks(x: float, world: int, device: str = "cuda") -> float:
    """The slowest rank's time: every multi-GPU number is bracketed by barriers and reduced with MAX, never a wall clock
    on rank 0 alone.  Works on any backend (NCCL on the box, gloo in the CPU tests)."""
    if world <= 1:
        return float(x)
    import torch
    import torch.distributed as dist

    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def whole_job_gbs(step_bytes: int, steps: int, world: int, slowest_seconds: float) -> float:
    """Weak scaling: every rank moves step_bytes per step on its own GPU; the job is done when the slowest rank is."""
    return step_bytes * steps * world / slowest_seconds / 1e9


def host_cores() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def cpu_port_run(sample_bytes: int, threads: int, min_seconds: float, max_reps: int):
    """The oracle's CPU port of one step (fill, verify, copy) on `threads` host threads."""
    from oracle import gmo_ctypes as gmo  # the checker doubles as the timed CPU baseline, nothing else

    total_s, reps, per = 0.0, 0, [0.0, 0.0, 0.0]
    while reps < max_reps and (reps == 0 or total_s < min_seconds):
        secs, (bad, _first, x, s) = gmo.cpu_probe(sample_bytes, threads, SEED + reps)
        assert bad == 0 and (x, s) == gmo.pattern_checksum(sample_bytes, 0, SEED + reps)
        total_s += sum(secs)
        per = [a + b for a, b in zip(per, secs)]
        reps += 1
    gbs = 4.0 * sample_bytes * reps / total_s / 1e9
    return gbs, total_s / reps, reps, [p / reps for p in per]


def mount_section(n_gpus: int, iters: int):
    """Node-local mount latency (reference fork/exec path vs native) and mount-to-usable p50/p99 on n_gpus GPUs.
    Any failure is REPORTED ("mount_ok": false at the top level, the exception text here) and makes bench.py exit non-zero
    after the line is printed: a broken mount path must not hide behind a good kernel number."""
    try:
        from tools import mount_latency as ml
        from tools import standin_pod

        why = standin_pod.can_run()
        if why:
            return {"unavailable": why, "ok": False}
        nv = __import__("gpumounter_b200").probe.enumerate_nvml()
        gpus = [(nv[i % len(nv)].uuid, 100 + i) for i in range(n_gpus)]  # private /dev: minors shifted off the real nodes
        with standin_pod.Target() as t:
            ref_m, _ = ml.bench_reference(t, gpus, 30)
        with standin_pod.Target() as t:
            nat_m, _ = ml.bench_native(t, gpus, 30, batched=True)
        out = {"gpus": n_gpus, "host_cores": os.cpu_count(), "threads_used": 1,
               "reference_fork_exec_mount_ms": {"p50": ref_m["p50_ms"], "p99": ref_m["p99_ms"]},
               "native_mount_ms": {"p50": nat_m["p50_ms"], "p99": nat_m["p99_ms"]},
               "note": "slave-pod scheduling (ref: allocator.go:40-99) needs a cluster and is not included"}
        # cold: a fresh process in the pod starts CUDA on the mounted GPUs (what the tenant's first CUDA call pays too)
        # (gmp_probe_in_pod_fanout from 2 GPUs on: one probe process per GPU, each confined to its GPU; beside it the
        # single-process probe, which is also what a tenant's own multi-GPU process pays)
        out["mount_to_usable"] = ml.to_usable(iters, 1 << 30, n_gpus, fanout=n_gpus >= 2)
        if n_gpus >= 2:
            out["mount_to_usable_single_process"] = ml.to_usable(max(3, iters // 2), 1 << 30, n_gpus)
        # the two-stage hook of a resident worker (this process: cuInit and contexts already paid).  Contexts are kept
        # and peer access is left alone: the process still belongs to torch.  Kernels run from the HOST view.
        out["mount_to_usable_resident_worker"] = ml.to_usable_resident(iters, 1 << 30, n_gpus, keep_contexts=True,
                                                                      skip_p2p=True, release_at_end=False)
        out["ok"] = True
        return out
    except Exception as e:  # the bench line must still be printed; main() turns this into a non-zero exit
        import traceback

        return {"unavailable": f"{type(e).__name__}: {e}", "traceback": traceback.format_exc()[-1500:], "ok": False}


def p2p_section(uuids, bytes_per_pair: int, iters: int):
    """BASELINE's "P2P GB/s per GPU" at N >= 2: the whole P2P probe (gmp_probe_p2p through gpumounter-probe, a process of
    its own so that torch's contexts and peer mappings are not involved) over the job's GPUs: store / load / copy-engine
    matrices, bidirectional per GPU, all-to-all, and the cudaDeviceCanAccessPeer-vs-NVML matrix check."""
    try:
        from gpumounter_b200 import probe
        from tools import mount_latency as ml

        cmd = [probe.probe_executable(), "--host-view", "--skip-nodes", "--skip-hbm", "--p2p-bytes", str(bytes_per_pair),
               "--p2p-iters", str(iters)]
        for u in uuids:
            cmd += ["--uuid", u]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
        rep = json.loads(r.stdout.strip().splitlines()[-1])
        if "p2p" not in rep:
            return {"ok": False, "unavailable": f"probe exit {r.returncode}: {rep.get('detail')}"}
        out = ml.p2p_summary(rep["p2p"])
        out.update(iters=iters, status=rep["status"], detail=rep.get("detail", ""), ms_total=rep["p2p"]["ms_total"],
                   store_kernel="k_hbm_write_pattern on a peer pointer (32-byte vector stores; 713 GB/s is the fabric's ceiling for "
                                "128-byte write requests, profiles/r02_p2p_ncu.md); all-to-all: k_peer_alltoall_write, one launch per sender",
                   nvlink5_nominal_gbs_per_direction=900.0)
        # bound: NVLink.  Denominator: the pool's measured peer copy (B200_PROFILING.md: 770 GB/s per direction; nominal 900).
        # Algorithmic bytes = wire user bytes = bytes_per_pair per ordered pair (profiles/r02_p2p_ncu.md: equal on every path).
        out["roofline"] = {"bound": "nvlink", "peak": 770.0, "peak_source": "measured peer copy, /opt/skills/guides/B200_PROFILING.md",
                           "unit": "GB/s", "store": {"achieved": out["store_gbs"]["mean"], "frac": out["store_gbs"]["mean"] / 770.0},
                           "load_verify": {"achieved": out["load_gbs"]["mean"], "frac": out["load_gbs"]["mean"] / 770.0},
                           "alltoall_per_gpu": {"achieved": out["alltoall_gbs_per_gpu"], "frac": out["alltoall_gbs_per_gpu"] / 770.0},
                           "note": "stores are 128-byte NVLink requests with 24 bytes of protocol each: 713 GB/s of user bytes is 849 GB/s "
                                   "on the wire (profiles/r02_p2p_ncu.md)"}
        out["ok"] = bool(rep["status"] == 0 and out["matrix_match_nvml"] and out["mismatch_lanes"] == 0)
        return out
    except Exception as e:
        return {"ok": False, "unavailable": f"{type(e).__name__}: {e}"}


def run_reference(args, rank: int, world: int):
    if rank != 0:
        return
    threads = min(256, host_cores())
    sample = args.cpu_sample_gib * GIB
    # each "step" is one bounded sample of the workload
    for _ in range(min(args.warmup, 1)):
        cpu_port_run(sample, threads, 0.0, 1)
    t0 = time.perf_counter()
    per_step = []
    for _ in range(args.steps):
        gbs, sec, _reps, _ = cpu_port_run(sample, threads, 0.0, 1)
        per_step.append(sec)
    wall = time.perf_counter() - t0
    value = 4.0 * sample * args.steps / sum(per_step) / 1e9
    line = {
        "impl": "reference",
        "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sum(per_step) / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": f"CPU port of the probe step (fill+verify+copy) on host DRAM, "
                               f"{args.cpu_sample_gib} GiB per buffer sample of the {args.gib} GiB workload",
                   "bytes_per_buffer": sample, "algorithmic_bytes_per_step": 4 * sample},
        "cpu_baseline": {"value": value, "unit": "GB/s", "cores": threads, "kind": "port",
                         "sample": f"{args.cpu_sample_gib} GiB per buffer x {args.steps} steps; the reference has no "
                                   f"device-memory path, so this is oracle/gmo_pattern.c on pthreads"},
        "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": wall,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--gib", type=float, default=8.0, help="bytes per buffer in GiB (BASELINE: 8)")
    ap.add_argument("--dma-mib", type=int, default=16, help="pinned host<->device round trip inside the e2e call")
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 -> min(steps, 10)")
    ap.add_argument("--cpu-sample-gib", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-clocks", action="store_true")
    ap.add_argument("--no-mount", action="store_true", help="skip the mount-latency / mount-to-usable section")
    ap.add_argument("--mount-iters", type=int, default=8)
    ap.add_argument("--no-p2p", action="store_true", help="skip the P2P probe block (N >= 2)")
    ap.add_argument("--p2p-mib", type=int, default=1024, help="bytes per ordered pair (BASELINE.md: 1 GiB)")
    ap.add_argument("--p2p-iters", type=int, default=2)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return 0

    import torch
    import torch.distributed as dist

    from gpumounter_b200 import probe

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: gpumounter_b200 has no CPU fallback "
                         "(use --impl reference for the CPU port)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()

    nbytes = int(args.gib * GIB) & ~15
    a = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    b = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    acc = torch.zeros(4, dtype=torch.int64, device="cuda")
    stream = torch.cuda.current_stream()
    s = stream.cuda_stream
    pa, pb, pacc = a.data_ptr(), b.data_ptr(), acc.data_ptr()
    def step(i, ev=None):
        seed = SEED + i
        if ev:
            ev[0].record(stream)
        probe.hbm_write_pattern(pa, nbytes, probe.PAT_SPLITMIX, seed, s)
        if ev:
            ev[1].record(stream)
        probe.hbm_read_verify_async(pa, nbytes, probe.PAT_SPLITMIX, seed, pacc, s)
        if ev:
            ev[2].record(stream)
        probe.hbm_copy(pb, pa, nbytes, s)
        if ev:
            ev[3].record(stream)

    for i in range(args.warmup):
        step(i)
    r = probe.hbm_verify_fetch(pacc, s)
    assert r.ok, "warm-up verify failed"

    K = args.steps
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(K)]
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(local_rank)
    if not args.no_clocks and rank == 0:
        sampler.start()
        time.sleep(0.3)
    launches0 = probe.kernel_launches()
    barrier()
    torch.cuda.synchronize()
    t_start.record(stream)
    for i in range(K):
        step(args.warmup + i, evs[i])
    t_end.record(stream)
    torch.cuda.synchronize()
    barrier()
    dev_ms = t_start.elapsed_time(t_end)
    timed_launches = probe.kernel_launches() - launches0  # counted by the library at each launch site
    clocks = sampler.stop() if (not args.no_clocks and rank == 0) else None

    # the last step's verdicts: A verified in-step; verify the copy now (outside the timed region)
    r = probe.hbm_verify_fetch(pacc, s)
    last_seed = SEED + args.warmup + K - 1
    rb = probe.hbm_read_verify(pb, nbytes, probe.PAT_SPLITMIX, last_seed, s)
    assert r.ok and rb.ok and (r.xor64, r.sum64) == (rb.xor64, rb.sum64), "probe verification failed"

    max_ms = max_over_ranks(dev_ms, world)
    step_bytes = 4 * nbytes
    value = whole_job_gbs(step_bytes, K, world, max_ms * 1e-3)

    k_ms = {"write": [], "verify": [], "copy": []}
    for e in evs:
        k_ms["write"].append(e[0].elapsed_time(e[1]))
        k_ms["verify"].append(e[1].elapsed_time(e[2]))
        k_ms["copy"].append(e[2].elapsed_time(e[3]))
    avg = {k: sum(v) / len(v) for k, v in k_ms.items()}
    if world > 1:  # the slowest rank's average per kernel, not rank 0's
        avg = {k: max_over_ranks(avg[k], world) for k in ("write", "verify", "copy")}
    # the allocation-free verify (fixed scratch in the library, one CTA per 4 MiB segment) on the same 8 GiB buffer
    noalloc_ms = []
    if nbytes <= probe.NOALLOC_MAX_BYTES:
        last_seed_a = SEED + args.warmup + K - 1
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            rn = probe.hbm_read_verify_noalloc(pa, nbytes, probe.PAT_SPLITMIX, last_seed_a, s)
            e1.record(stream)
            e1.synchronize()
            noalloc_ms.append(e0.elapsed_time(e1))
            assert rn.ok and (rn.xor64, rn.sum64) == (r.xor64, r.sum64), "allocation-free verify disagrees"
        noalloc_ms = noalloc_ms[1:]
    algo = {"write": nbytes, "verify": nbytes, "copy": 2 * nbytes}
    gbs = {k: algo[k] / (avg[k] * 1e-3) / 1e9 for k in avg}
    peak, peak_src = load_peaks()

    # ---- e2e: the reference-facing one-shot call, wall clock, host buffers + copies inside ----
    e2e_steps = args.e2e_steps or min(K, 10)
    dma = args.dma_mib << 20
    del b  # the one-shot call allocates its own pair; keep A so that the device is not empty
    torch.cuda.empty_cache()

    my_uuid = probe.enumerate_cuda()[local_rank].uuid

    def one_shot(flags):
        return probe.probe_hbm(my_uuid, nbytes=nbytes, iters=1, seed=SEED, dma_bytes=dma, flags=flags)

    def timed_calls(flags, calls):
        for _ in range(2):
            one_shot(flags)
        barrier()
        torch.cuda.synchronize()
        w0 = time.perf_counter()
        last = None
        for _ in range(calls):
            last = one_shot(flags)
        torch.cuda.synchronize()
        w1 = time.perf_counter()
        barrier()
        assert last.status == 0 and last.mismatch_lanes == 0
        return max_over_ranks(w1 - w0, world), last

    def phases(last):
        return {k[3:]: getattr(last, k) / 1e6 for k in ("ns_select_device", "ns_context", "ns_alloc", "ns_first_kernel",
                                                        "ns_kernels", "ns_dma", "ns_teardown", "ns_total")}

    # headline: a resident worker's call (GMP_HBM_F_RETAIN: buffers, streams and the pinned staging block stay for the next
    # call on the device) - passes + DMA round trip + one synchronisation.  Beside it the cold call, which also pays
    # cudaMalloc/cudaFree of both buffers and the pinning of the staging block every time.
    e2e_seconds, last = timed_calls(probe.HBM_F_KEEP_DEVICE | probe.HBM_F_RETAIN, e2e_steps)
    probe.probe_hbm_release(my_uuid)
    cold_calls = max(2, e2e_steps // 2)
    cold_seconds, cold_last = timed_calls(probe.HBM_F_KEEP_DEVICE, cold_calls)
    e2e_value = whole_job_gbs(step_bytes, e2e_steps, world, e2e_seconds)
    e2e = {
        "value": e2e_value, "unit": "GB/s",
        "h2d_bytes_per_step": int(last.h2d_bytes), "d2h_bytes_per_step": int(last.d2h_bytes),
        "ms_per_call": 1e3 * e2e_seconds / e2e_steps, "calls": e2e_steps,
        "call": "gmp_probe_hbm(uuid, {bytes, iters=1, dma_bytes, flags=KEEP_DEVICE|RETAIN}) one-shot, context warm, "
                "buffers retained between calls (a resident worker probing again and again)",
        "phases_ms_last_call": phases(last),
        "in_call_gbs": {"write": last.write_gbs_best, "verify": last.read_gbs_best, "copy": last.copy_gbs_best,
                        "h2d": last.h2d_gbs, "d2h": last.d2h_gbs},
        "cold_call": {"value": whole_job_gbs(step_bytes, cold_calls, world, cold_seconds), "unit": "GB/s",
                      "ms_per_call": 1e3 * cold_seconds / cold_calls, "calls": cold_calls,
                      "call": "the same without RETAIN: cudaMalloc + cudaFree of both buffers and pinning inside every call",
                      "phases_ms_last_call": phases(cold_last)},
    }

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    line = {
        "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": K, "warmup": args.warmup,
        "ms_per_step": max_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": {"workload": "post-mount HBM probe step: write_pattern(A)+read_verify(A)+copy(A->B), "
                               "BASELINE configs[1]", "bytes_per_buffer": nbytes, "buffers": 2,
                   "algorithmic_bytes_per_step": step_bytes, "pattern": "splitmix",
                   "l2": "inputs larger than L2 (2 x %.1f GiB vs 126 MB), no flush needed" % (nbytes / GIB),
                   "parallelism": f"{world} independent per-GPU probes, no collective"},
        "gpu_launches": timed_launches,
        "kernels": {k: {"avg_ms": avg[k], "gbs": gbs[k], "frac_of_peak": gbs[k] / peak,
                        "frac_of_datasheet_8tbs": gbs[k] / 8000.0,  # secondary denominator, not the graded one
                        "share_of_step": avg[k] / sum(avg.values())} for k in avg},
        "kernels_over": "slowest rank's average per kernel" if world > 1 else "this GPU",
        "roofline": {"bound": "hbm", "kernel": "k_hbm_copy_bulk (gmp_hbm_copy: TMA bulk copy, the step's dominant kernel)", "achieved": gbs["copy"], "peak": peak, "unit": "GB/s",
                     "frac": gbs["copy"] / peak, "traffic": None, "traffic_source": None, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": 2 * nbytes},
        "e2e": e2e,
        "clocks": clocks,
    }
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            line["roofline"]["traffic"] = json.load(f).get("k_hbm_copy_dram_bytes_per_launch")
        line["roofline"]["traffic_source"] = ("profiles/ncu_traffic.json (one `ncu --set full` capture of this kernel, "
                                              "profiles/r02_ncu_summary.md) - a citation, not measured in this run")
    except Exception:
        pass
    if noalloc_ms:
        nm = sum(noalloc_ms) / len(noalloc_ms)
        line["kernels"]["verify_noalloc"] = {
            "avg_ms": nm, "gbs": nbytes / (nm * 1e-3) / 1e9, "frac_of_peak": nbytes / (nm * 1e-3) / 1e9 / peak,
            "note": "gmp_hbm_read_verify_noalloc: whole call (result-slot clear, k_hbm_read_verify_segments, k_noalloc_fold, "
                    "32-byte read-back), not part of the step"}

    if world == 1 and not args.no_cpu_baseline:
        threads = min(256, host_cores())
        sample = args.cpu_sample_gib * GIB
        cgbs, sec, reps, per = cpu_port_run(sample, threads, 8.0, 6)
        line["cpu_baseline"] = {
            "value": cgbs, "unit": "GB/s", "cores": threads, "kind": "port",
            "sample": f"oracle/gmo_pattern.c fill+verify+copy of {args.cpu_sample_gib} GiB per buffer on host DRAM, "
                      f"{reps} reps, {sec:.2f} s per rep (the reference has no device-memory code to time)",
            "seconds_per_pass": {"fill": per[0], "verify": per[1], "copy": per[2]},
        }
    del a, acc
    torch.cuda.empty_cache()
    failures = []
    if world >= 2 and not args.no_p2p:
        # BASELINE's "P2P GB/s per GPU": the whole P2P probe over the job's GPUs (rank 0 alone; the other ranks have left)
        uuids = [d.uuid for d in probe.enumerate_cuda()[:world]]
        line["p2p"] = p2p_section(uuids, args.p2p_mib << 20, args.p2p_iters)
        if not line["p2p"].get("ok"):
            failures.append("p2p")
    if not args.no_mount:
        # the other half of BASELINE's metric: mount latency next to the reference's own mount path on this host, and
        # mount-to-usable with the in-pod probe on all `world` GPUs (rank 0 alone; the other ranks have left)
        line["mount"] = mount_section(world, args.mount_iters)
        line["mount_ok"] = bool(line["mount"].get("ok"))
        if not line["mount_ok"]:
            failures.append("mount")
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if failures:
        print(f"bench.py: {', '.join(failures)} section failed (see the line above)", file=sys.stderr)
        return 3
    return 0


if __name__ == "__main__":
    sys.exit(main())
