#!/usr/bin/env python
"""Sweep the launch geometry of the three HBM kernels on a real B200 (benchmarking aid).

For every variant (THREADS, UNROLL, cache hint — see GMP_VARIANTS in gmp_kernels.cu) and CTAs-per-SM value,
time ITERS launches with CUDA events on an 8 GiB buffer pair and print GB/s (algorithmic bytes).
Also times the library baselines of SURVEY.md 2.3 (cudaMemsetAsync via torch zero_, D2D copy via torch copy_).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from gpumounter_b200 import probe

VARIANTS = {0: "256x4x16B default", 1: "256x4x16B .cs", 2: "256x8x16B .cs", 3: "512x4x16B .cs",
            4: "256x4x16B no_alloc", 5: "512x2x16B .cs", 6: "128x8x16B .cs", 7: "256x2x32B .cs", 8: "256x4x32B .cs",
            9: "512x2x32B .cs", 10: "256x2x32B default", 11: "1024x1x32B .cs", 12: "1024x2x16B .cs",
            13: "256x1x32B .cs", 14: "512x1x32B .cs", 15: "256x4x32B L2::evict_first"}
# copy only: the TMA engine (cp.async.bulk through shared memory), stages x chunk
BULK_VARIANTS = {16: "TMA bulk 6x16K", 17: "TMA bulk 3x32K", 18: "TMA bulk 12x8K", 19: "TMA bulk 3x64K", 20: "TMA bulk 2x96K",
                 21: "TMA bulk 2x112K", 22: "TMA bulk 4x48K", 23: "TMA bulk 7x32K"}


def timeit(fn, iters, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    best, tot = 1e30, 0.0
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        best = min(best, ms)
        tot += ms
    return best, tot / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gib", type=float, default=8.0)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--out", default="gpurun_out/sweep.json")
    ap.add_argument("--kernels", default="write,read,copy")
    ap.add_argument("--variants", default="", help="comma list; default all")
    ap.add_argument("--ctas", default="0,1,2,4,6,8,12,16,24,32")
    args = ap.parse_args()
    n = int(args.gib * (1 << 30))
    a = torch.empty(n, dtype=torch.uint8, device="cuda")
    b = torch.empty(n, dtype=torch.uint8, device="cuda")
    acc = torch.zeros(4, dtype=torch.int64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    rows = []
    ctas_list = [int(x) for x in args.ctas.split(",")]
    kinds = {
        "write": (lambda: probe.hbm_write_pattern(a.data_ptr(), n, 0, 1, s), n),
        "read": (lambda: probe.hbm_read_verify_async(a.data_ptr(), n, 0, 1, acc.data_ptr(), s), n),
        "copy": (lambda: probe.hbm_copy(b.data_ptr(), a.data_ptr(), n, s), 2 * n),
    }
    probe.hbm_write_pattern(a.data_ptr(), n, 0, 1, s)
    kinds = {k: v for k, v in kinds.items() if k in args.kernels.split(",")}
    bulk = BULK_VARIANTS
    for kind, (fn, nbytes) in kinds.items():
        variants = {**VARIANTS, **bulk} if kind == "copy" else VARIANTS
        if args.variants:
            variants = {v: variants[v] for v in variants if str(v) in args.variants.split(",")}
        for v in variants:
            for c in ctas_list:
                probe.tune(f"{kind}_variant", v)
                probe.tune(f"{kind}_ctas_per_sm", c)
                best, avg = timeit(fn, args.iters)
                row = {"kernel": kind, "variant": v, "desc": variants[v], "ctas_per_sm": c,
                       "best_gbs": nbytes / best / 1e6, "avg_gbs": nbytes / avg / 1e6}
                rows.append(row)
                print(json.dumps(row), flush=True)
    # library baselines
    for name, fn, nbytes in (("torch.zero_ (memset)", lambda: a.zero_(), n),
                             ("torch.copy_ (D2D)", lambda: b.copy_(a), 2 * n)):
        best, avg = timeit(fn, args.iters)
        row = {"kernel": "baseline", "desc": name, "best_gbs": nbytes / best / 1e6, "avg_gbs": nbytes / avg / 1e6}
        rows.append(row)
        print(json.dumps(row), flush=True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(rows, open(args.out, "w"), indent=1)
    for kind in kinds:
        top = sorted([r for r in rows if r["kernel"] == kind], key=lambda r: -r["avg_gbs"])[:5]
        print(kind, "top:", [(r["desc"], r["ctas_per_sm"], round(r["avg_gbs"])) for r in top])


if __name__ == "__main__":
    main()
