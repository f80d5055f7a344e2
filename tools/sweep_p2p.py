#!/usr/bin/env python
"""Sweep launch geometry of the kernels whose pointer is PEER memory (GPU0 -> GPU1 over NVLink), 2 GPUs.

write     : vector stores (k_hbm_write_pattern) and, variants 32.., the pattern staged in shared memory and pushed by the
            bulk-copy engine (k_hbm_write_pattern_bulk) - the candidate for beating the copy engines' 777 GB/s
read      : vector loads + verify (k_hbm_read_verify)
copy_push : local HBM -> peer through k_hbm_copy / k_hbm_copy_bulk (for comparison with the pattern push)
Baselines : torch copy_ both ways (cudaMemcpyPeer-class, copy engines).
Afterwards the winner is set through gmp_tune and the whole P2P probe (gmp_probe_p2p, 1 GiB per pair, verified) runs
once with it, so that the figure is also seen through the product entry point.  Writes gpurun_out/sweep_p2p.json.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from gpumounter_b200 import probe
from tools.sweep_hbm import BULK_VARIANTS, VARIANTS, timeit

BULK_WRITE_VARIANTS = {32: "bulk push 256thr 4x16K", 33: "bulk push 256thr 4x32K", 34: "bulk push 512thr 4x32K",
                       35: "bulk push 256thr 2x64K", 36: "bulk push 512thr 3x64K", 37: "bulk push 128thr 8x8K",
                       38: "bulk push 1024thr 2x96K", 39: "bulk push 256thr 6x32K"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kinds", default="write,read,copy_push")
    ap.add_argument("--iters", type=int, default=4)
    ap.add_argument("--out", default="gpurun_out/sweep_p2p.json")
    args = ap.parse_args()
    assert torch.cuda.device_count() >= 2
    n = 1 << 30
    remote = torch.empty(n, dtype=torch.uint8, device="cuda:1")
    local = torch.empty(n, dtype=torch.uint8, device="cuda:0")
    torch.cuda.set_device(0)
    probe.peer_access(0, 1, True)
    probe.peer_access(1, 0, True)
    acc = torch.zeros(4, dtype=torch.int64, device="cuda:0")
    s = torch.cuda.current_stream().cuda_stream
    rows = []
    probe.hbm_write_pattern(remote.data_ptr(), n, 0, 1, s)
    probe.hbm_write_pattern(local.data_ptr(), n, 0, 1, s)
    torch.cuda.synchronize()
    kinds = {"write": lambda: probe.hbm_write_pattern(remote.data_ptr(), n, 0, 1, s),
             "read": lambda: probe.hbm_read_verify_async(remote.data_ptr(), n, 0, 1, acc.data_ptr(), s),
             "copy_pull": lambda: probe.hbm_copy(local.data_ptr(), remote.data_ptr(), n, s),
             "copy_push": lambda: probe.hbm_copy(remote.data_ptr(), local.data_ptr(), n, s)}
    few = {1: VARIANTS[1], 11: VARIANTS[11], 12: VARIANTS[12]}  # the vector geometries that led round 1's sweep
    plan = {"write": ({**few, **BULK_WRITE_VARIANTS}, (1, 2, 3, 4, 6, 8, 16)),
            "read": ({1: VARIANTS[1], 8: VARIANTS[8], 10: VARIANTS[10]}, (4, 8, 16, 32)),
            "copy_pull": ({1: VARIANTS[1], **BULK_VARIANTS}, (1, 2, 4, 8, 32)),
            "copy_push": ({1: VARIANTS[1], **BULK_VARIANTS}, (1, 2, 4, 8, 32))}
    defaults = {k: probe.tune(k) for k in ("write_variant", "write_ctas_per_sm", "read_variant", "read_ctas_per_sm",
                                           "copy_variant", "copy_ctas_per_sm")}
    best_of = {}
    for kind in [k for k in args.kinds.split(",") if k]:
        fn = kinds[kind]
        knob = "copy" if kind.startswith("copy") else kind
        variants, ctas = plan[kind]
        for v, desc in variants.items():
            for c in ctas:
                probe.tune(f"{knob}_variant", v)
                probe.tune(f"{knob}_ctas_per_sm", c)
                best, avg = timeit(fn, args.iters)
                rows.append({"kernel": kind, "variant": v, "desc": desc, "ctas_per_sm": c, "best_gbs": n / best / 1e6,
                             "avg_gbs": n / avg / 1e6})
        mine = sorted([r for r in rows if r["kernel"] == kind], key=lambda r: -r["avg_gbs"])
        best_of[kind] = mine[0]
        print(kind, "top:", [(r["desc"], r["ctas_per_sm"], round(r["avg_gbs"])) for r in mine[:6]], flush=True)
        print(kind, "worst:", [(r["desc"], r["ctas_per_sm"], round(r["avg_gbs"])) for r in mine[-2:]], flush=True)
    for k, v in defaults.items():
        probe.tune(k, v)
    if "write" in best_of:
        # what was written through the winning geometry is what the pattern says (the sweep itself only times)
        w = best_of["write"]
        probe.tune("write_variant", w["variant"])
        probe.tune("write_ctas_per_sm", w["ctas_per_sm"])
        remote.zero_()
        torch.cuda.synchronize(1)  # zero_ runs on GPU 1's stream: it must be over before GPU 0 starts to write
        probe.hbm_write_pattern(remote.data_ptr(), n, 0, 77, s)
        torch.cuda.synchronize(0)
        probe.tune("write_variant", defaults["write_variant"])
        probe.tune("write_ctas_per_sm", defaults["write_ctas_per_sm"])
        torch.cuda.set_device(1)
        r = probe.hbm_read_verify(remote.data_ptr(), n, 0, 77, torch.cuda.current_stream().cuda_stream)
        torch.cuda.set_device(0)
        print("winner's bytes verified by the owner:", r.ok, r, flush=True)
        best_of["write"]["verified_by_owner"] = bool(r.ok)
    base = {}
    best, avg = timeit(lambda: local.copy_(remote), 5)
    base["torch_copy_pull_gbs"] = n / avg / 1e6
    best, avg = timeit(lambda: remote.copy_(local), 5)
    base["torch_copy_push_gbs"] = n / avg / 1e6
    print("copy engines:", {k: round(v) for k, v in base.items()}, flush=True)
    out = {"rows": rows, "baselines": base, "best": best_of}
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    json.dump(out, open(args.out, "w"), indent=1)
    del remote, local
    torch.cuda.empty_cache()
    probe.peer_access(0, 1, False)
    probe.peer_access(1, 0, False)
    if "write" in best_of:
        w = best_of["write"]
        probe.tune("peer_store_variant", w["variant"])
        probe.tune("peer_store_ctas_per_sm", w["ctas_per_sm"])
    if "read" in best_of:
        probe.tune("peer_load_variant", best_of["read"]["variant"])
        probe.tune("peer_load_ctas_per_sm", best_of["read"]["ctas_per_sm"])
    uuids = [d.uuid for d in probe.enumerate_cuda()[:2]]
    r = probe.probe_p2p(uuids, nbytes=1 << 30, iters=3, flags=probe.P2P_F_KEEP_DEVICE, raise_on_error=False)
    out["probe_with_winners"] = {"status": r.status, "store_gbs": r.matrix("write_gbs"), "load_gbs": r.matrix("read_gbs"),
                                 "memcpy_peer_gbs": r.matrix("memcpy_gbs"), "bidir_gbs": list(r.bidir_gbs[:2]),
                                 "alltoall_gbs": r.alltoall_gbs, "mismatch_lanes": r.mismatch_lanes,
                                 "matrix_match": r.matrix_match}
    print("gmp_probe_p2p with the winners:", json.dumps(out["probe_with_winners"]), flush=True)
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    json.dump(out, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
