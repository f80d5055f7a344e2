#!/usr/bin/env python
"""GPU 0 -> GPU 1 over NVLink, one launch of each peer path at 1 GiB, for ncu's NVLink byte counters (2 GPUs, ONE process;
run once without ncu first): vector stores, the bulk-engine push, the load+verify pass."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from gpumounter_b200 import probe

n = 1 << 30
remote = torch.empty(n, dtype=torch.uint8, device="cuda:1")
torch.cuda.synchronize(1)
torch.cuda.set_device(0)
probe.peer_access(0, 1, True)
acc = torch.zeros(4, dtype=torch.int64, device="cuda:0")
s = torch.cuda.current_stream().cuda_stream


def once():
    for variant, ctas in ((11, 6), (33, 8)):  # 1024x1x32B .cs vector stores; bulk push 256thr 4x32K
        probe.tune("write_variant", variant)
        probe.tune("write_ctas_per_sm", ctas)
        probe.hbm_write_pattern(remote.data_ptr(), n, 0, 5, s)
    probe.tune("read_variant", 1)
    probe.tune("read_ctas_per_sm", 32)
    probe.hbm_read_verify_async(remote.data_ptr(), n, 0, 5, acc.data_ptr(), s)
    r = probe.hbm_verify_fetch(acc.data_ptr(), s)
    assert r.ok
    torch.cuda.synchronize()


once()
once()
print("ok")
