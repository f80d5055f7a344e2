#!/usr/bin/env python
"""Smallest case that touches every kernel and every edge path (ragged tails, slow paths, peer-free), for
`compute-sanitizer --tool memcheck python tools/sanitize_case.py` (one tool per gpurun call)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from gpumounter_b200 import probe

torch.cuda.set_device(0)
s = torch.cuda.current_stream().cuda_stream
for nbytes in (0, 5, 16, 4096 + 9, (1 << 20) + 3, (5 << 20) + 21):
    a = torch.empty(nbytes, dtype=torch.uint8, device="cuda")       # exact size: any overrun is out of bounds
    b = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    for variant in (1, 10, 11, 12):
        for knob in ("write", "read", "copy"):
            probe.tune(f"{knob}_variant", variant)
        for pat in range(probe.PAT_COUNT):
            probe.hbm_write_pattern(a.data_ptr(), nbytes, pat, 7, s)
            probe.hbm_copy(b.data_ptr(), a.data_ptr(), nbytes, s)
            r = probe.hbm_read_verify(b.data_ptr(), nbytes, pat, 7, s)
            assert r.ok
            r2 = probe.hbm_read_verify_noalloc(b.data_ptr(), nbytes, pat, 7, s)
            assert (r2.mismatch_lanes, r2.xor64, r2.sum64) == (0, r.xor64, r.sum64)
    if nbytes:
        probe.hbm_corrupt(b.data_ptr(), nbytes, nbytes - 1, 1, s)
        assert probe.hbm_read_verify(b.data_ptr(), nbytes, probe.PAT_ADDRESS, 7, s).mismatch_lanes == 1
        assert probe.hbm_read_verify_noalloc(b.data_ptr(), nbytes, probe.PAT_ADDRESS, 7, s).mismatch_lanes == 1
res = probe.probe_hbm(nbytes=(8 << 20) + 16, iters=1, pattern_mask=31, dma_bytes=1 << 20, flags=probe.HBM_F_KEEP_DEVICE)
assert res.status == 0
print("sanitize case ok")
