#!/usr/bin/env python
"""Where the cold in-pod start-up goes when k GPUs are mounted (multi-GPU box), and what can overlap.

For k in --sizes, in a pod-like stand-in (tools/standin_pod.py) with exactly k GPUs mounted:
  single : ONE gpumounter-probe in the pod, one host thread per GPU (the product path): cuInit, and per GPU the absolute
           begin/end stamps of primary-context creation - do the intervals overlap or does the driver take them in turn?
  fanout : k gpumounter-probe processes started together, each in a CHILD of the pod's device cgroup that allows only
           its own GPU (a child can only narrow what the pod may open), so that each process's cuInit attaches one GPU
           and creates one context.  Does the driver run k such start-ups side by side?
Prints one JSON object; --out writes it too.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from gpumounter_b200 import device, probe, util  # noqa: E402
from tools import mount_latency as ml  # noqa: E402

PROBE_ARGS = ["--bytes", "256M", "--iters", "1", "--dma-bytes", "1M", "--skip-p2p"]


def child_cgroup(t, idx, minor, all_minors):
    """Created AFTER the mount: a v1 devices child starts with its parent's list (the mounted GPUs included) and may
    only be narrowed - here down to one GPU."""
    d = os.path.join(t.cgroup_dir, f"gmp-fan-{idx}")
    os.mkdir(d)
    for other in all_minors:
        if other != minor:
            with open(os.path.join(d, "devices.deny"), "w") as f:
                f.write(f"c 195:{other} rw")
    return d


BOOT = ("import os, sys\n"
        "open(sys.argv[1], 'w').write(str(os.getpid()) + '\\n')\n"
        "fd = os.open(sys.argv[2], os.O_RDONLY)\n"
        "os.setns(fd, os.CLONE_NEWNS)\n"
        "os.close(fd)\n"
        "os.execv(sys.argv[3], sys.argv[3:])\n")


def fanout_once(g, gpus, narrow=True, only_first=False, stagger_ms=0.0):
    """narrow: each process in a child cgroup that opens only its own GPU; else every process sees all k mounted GPUs
    (cuInit attaches all of them) and is told which one to probe (--host-view: more visible than named is fine)."""
    t = g.t
    recs = [device.New(m, u) for u, m in gpus]
    pod, env = ml.product_pod(), g.env()
    exe = probe.probe_executable()
    dirs = []
    try:
        t0 = time.monotonic_ns()
        util.MountGPUs(pod, recs, env)
        if narrow:
            dirs = [child_cgroup(t, i, m, [mm for _, mm in gpus]) for i, (_, m) in enumerate(gpus)]
        procs = []
        for i, (u, m) in enumerate(gpus[:1] if only_first else gpus):  # only_first: ONE narrowed process, alone
            d = dirs[i] if narrow else t.cgroup_dir
            argv = [sys.executable, "-c", BOOT, os.path.join(d, "cgroup.procs"), f"/proc/{t.pid}/ns/mnt", exe,
                    "--uuid", f"{u}:{m}"] + PROBE_ARGS + ([] if narrow else ["--host-view"])
            procs.append(subprocess.Popen(argv, stdout=subprocess.PIPE, text=True))
            if stagger_ms:  # per-GPU attach work is serialised node-wide: let each start-up finish its own before the next begins
                time.sleep(stagger_ms / 1e3)
        reps = [json.loads(p.communicate(timeout=120)[0].strip().splitlines()[-1]) for p in procs]
        util.UnmountGPUs(pod, recs, False, env)
    finally:
        for d in dirs:
            for _ in range(50):
                try:
                    os.rmdir(d)
                    break
                except OSError:
                    time.sleep(0.02)
    assert all(r["status"] == 0 and r["n_gpus"] == 1 for r in reps), [(r["status"], r["detail"]) for r in reps]
    usable = [(r["mono_usable_ns"] - t0) / 1e6 for r in reps]
    return {"all_usable_ms": max(usable), "first_usable_ms": min(usable),
            "python_boot_to_probe_start_ms": [round((r["mono_start_ns"] - t0) / 1e6, 1) for r in reps],
            "cuda_init_ms": [r["ms_cuda_init"] for r in reps],
            "context_ms": [r["hbm"][0]["ms_context"] for r in reps],
            "context_intervals_ms": [[round((r["hbm"][0]["mono_context_begin_ns"] - t0) / 1e6, 1),
                                      round((r["hbm"][0]["mono_context_end_ns"] - t0) / 1e6, 1)] for r in reps]}


def single_once(g, gpus):
    recs = [device.New(m, u) for u, m in gpus]
    pod, env = ml.product_pod(), g.env()
    t0 = time.monotonic_ns()
    util.MountGPUs(pod, recs, env)
    rc, rep = util.ProbeInPod(pod, recs, PROBE_ARGS, env, raise_on_error=False)
    assert rc == 0, (rc, rep)
    util.UnmountGPUs(pod, recs, False, env)
    return {"all_usable_ms": (rep["mono_usable_ns"] - t0) / 1e6,
            "first_usable_ms": (min(h["mono_first_verified_ns"] for h in rep["hbm"]) - t0) / 1e6,
            "probe_start_ms": (rep["mono_start_ns"] - t0) / 1e6, "cuda_init_ms": rep["ms_cuda_init"],
            "context_ms": [h["ms_context"] for h in rep["hbm"]],
            "context_intervals_ms": [[round((h["mono_context_begin_ns"] - t0) / 1e6, 1), round((h["mono_context_end_ns"] - t0) / 1e6, 1)]
                                     for h in rep["hbm"]]}


HOLDER = {
    # what else is alive on the node while the pod's first CUDA process starts
    "none": None,
    "init": "from cuda.bindings import driver as d\nimport time\nassert int(d.cuInit(0)[0]) == 0\nprint('up', flush=True)\ntime.sleep(900)\n",
    "ctx": ("from cuda.bindings import runtime as rt\nimport time\nn = rt.cudaGetDeviceCount()[1]\n"
            "for i in range(n):\n    rt.cudaSetDevice(i); rt.cudaFree(0)\nprint('up', flush=True)\ntime.sleep(900)\n"),
}


def start_holder(kind):
    """A process OUTSIDE the pod (the resident worker, or any other tenant of the node) that has initialised CUDA
    ("init") or holds a primary context on every GPU ("ctx")."""
    if HOLDER[kind] is None:
        return None
    p = subprocess.Popen([sys.executable, "-c", HOLDER[kind]], stdout=subprocess.PIPE, text=True)
    assert p.stdout.readline().strip() == "up"
    return p


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--holders", default="none,ctx", help="comma list of none,init,ctx")
    ap.add_argument("--sizes", default="1,2,4,8")
    ap.add_argument("--iters", type=int, default=4)
    ap.add_argument("--staggers", default="", help="comma list of ms between the starts of the narrowed per-GPU processes")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    os.environ.pop("CUDA_VISIBLE_DEVICES", None)
    all_gpus = [(d.uuid, d.minor) for d in probe.enumerate_nvml()]
    out = {"gpus_on_box": len(all_gpus), "probe_args": PROBE_ARGS, "iters": args.iters, "holders": {}}
    for kind in args.holders.split(","):
        holder = start_holder(kind)
        sizes = {}
        try:
            for k in [int(x) for x in args.sizes.split(",") if int(x) <= len(all_gpus)]:
                gpus = all_gpus[:k]
                row = {"single": [], "fanout": [], "fanout_shared": [], "one_narrowed_alone": []}
                for _ in range(args.iters):
                    with ml.GpuTarget(True) as g:
                        row["single"].append(single_once(g, gpus))
                    if k > 1:
                        with ml.GpuTarget(True) as g:
                            row["fanout"].append(fanout_once(g, gpus))
                        with ml.GpuTarget(True) as g:
                            row["fanout_shared"].append(fanout_once(g, gpus, narrow=False))
                        for st in [float(x) for x in args.staggers.split(",") if x]:
                            with ml.GpuTarget(True) as g:
                                row.setdefault(f"fanout_stagger_{st:g}ms", []).append(fanout_once(g, gpus, stagger_ms=st))
                        # does a node that EXISTS but is refused cost cuInit time, or is it the k start-ups sharing a lock?
                        with ml.GpuTarget(True) as g:
                            row["one_narrowed_alone"].append(fanout_once(g, gpus, only_first=True))
                sizes[str(k)] = {
                    mode: {"all_usable_p50_ms": statistics.median(r["all_usable_ms"] for r in runs),
                           "first_usable_p50_ms": statistics.median(r["first_usable_ms"] for r in runs),
                           "cuda_init_p50_ms": statistics.median(statistics.median(r["cuda_init_ms"]) if isinstance(r["cuda_init_ms"], list)
                                                                 else r["cuda_init_ms"] for r in runs),
                           "last_run": runs[-1]}
                    for mode, runs in row.items() if runs}
                print(kind, k, {m: round(v["all_usable_p50_ms"]) for m, v in sizes[str(k)].items()},
                      "cuInit", {m: round(v["cuda_init_p50_ms"]) for m, v in sizes[str(k)].items()}, flush=True)
        finally:
            if holder is not None:
                holder.kill()
                holder.wait()
        out["holders"][kind] = sizes
    text = json.dumps(out)
    if args.out:
        with open(args.out, "w") as f:
            f.write(text + "\n")
    print(text)


if __name__ == "__main__":
    main()
