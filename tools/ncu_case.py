#!/usr/bin/env python
"""One launch of every shipped kernel at N = 8 GiB, for `ncu --set full` (run once WITHOUT ncu first).

Order of launches (after one untimed warm-up of each): write_pattern, read_verify, copy_bulk (+ remainder copy),
read_verify_segments + noalloc_fold (gmp_hbm_read_verify_noalloc), write_pattern_bulk (the peer-store kernel, here on
local HBM: its SM side and its bulk stores, without NVLink behind them).
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from gpumounter_b200 import probe

n = 8 << 30
a = torch.empty(n, dtype=torch.uint8, device="cuda")
b = torch.empty(n, dtype=torch.uint8, device="cuda")
acc = torch.zeros(4, dtype=torch.int64, device="cuda")
s = torch.cuda.current_stream().cuda_stream
seed = 0x9E3779B97F4A7C15


def once(tag):
    probe.hbm_write_pattern(a.data_ptr(), n, 0, seed, s)
    probe.hbm_read_verify_async(a.data_ptr(), n, 0, seed, acc.data_ptr(), s)
    probe.hbm_copy(b.data_ptr(), a.data_ptr(), n, s)
    r = probe.hbm_read_verify_noalloc(b.data_ptr(), n, 0, seed, s)
    assert r.ok, tag
    wv, wc = probe.tune("write_variant"), probe.tune("write_ctas_per_sm")
    probe.tune("write_variant", probe.tune("peer_store_variant") if probe.tune("peer_store_variant") >= 32 else 33)
    probe.tune("write_ctas_per_sm", probe.tune("peer_store_ctas_per_sm"))
    probe.hbm_write_pattern(b.data_ptr(), n, 0, seed + 1, s)
    probe.tune("write_variant", wv)
    probe.tune("write_ctas_per_sm", wc)
    r = probe.hbm_read_verify(b.data_ptr(), n, 0, seed + 1, s)
    assert r.ok, tag + " bulk write"
    torch.cuda.synchronize()


once("warm-up")
once("captured")
print("ok")
