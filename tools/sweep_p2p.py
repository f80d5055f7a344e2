#!/usr/bin/env python
"""Sweep launch geometry of the write / read-verify kernels when the pointer is PEER memory (GPU0 -> GPU1 over NVLink)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from gpumounter_b200 import probe
from tools.sweep_hbm import BULK_VARIANTS, VARIANTS, timeit

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
torch.cuda.synchronize()
kinds = {"write": lambda: probe.hbm_write_pattern(remote.data_ptr(), n, 0, 1, s),
         "read": lambda: probe.hbm_read_verify_async(remote.data_ptr(), n, 0, 1, acc.data_ptr(), s),
         "copy_pull": lambda: probe.hbm_copy(local.data_ptr(), remote.data_ptr(), n, s),
         "copy_push": lambda: probe.hbm_copy(remote.data_ptr(), local.data_ptr(), n, s)}
for kind, fn in kinds.items():
    knob = "copy" if kind.startswith("copy") else kind
    # copies also through the TMA engine: peer stores from vector instructions plateau at 713 GB/s for every geometry
    # (profiles/r01_sweep_p2p.json) while the copy engines reach 777 - bulk stores to the peer have not been tried yet
    variants = {**VARIANTS, **BULK_VARIANTS} if knob == "copy" else VARIANTS
    for v in variants:
        for c in (0, 1, 2, 4, 8, 16, 32):
            probe.tune(f"{knob}_variant", v)
            probe.tune(f"{knob}_ctas_per_sm", c)
            best, avg = timeit(fn, 5)
            rows.append({"kernel": kind, "variant": v, "desc": variants[v], "ctas_per_sm": c, "best_gbs": n / best / 1e6, "avg_gbs": n / avg / 1e6})
    top = sorted([r for r in rows if r["kernel"] == kind], key=lambda r: -r["avg_gbs"])[:6]
    print(kind, "top:", [(r["desc"], r["ctas_per_sm"], round(r["avg_gbs"])) for r in top], flush=True)
    worst = sorted([r for r in rows if r["kernel"] == kind], key=lambda r: r["avg_gbs"])[:2]
    print(kind, "worst:", [(r["desc"], r["ctas_per_sm"], round(r["avg_gbs"])) for r in worst], flush=True)
best, avg = timeit(lambda: local.copy_(remote), 5)
print("torch copy_ pull (cudaMemcpyPeer-class):", round(n / avg / 1e6), flush=True)
best, avg = timeit(lambda: remote.copy_(local), 5)
print("torch copy_ push:", round(n / avg / 1e6), flush=True)
json.dump(rows, open("gpurun_out/sweep_p2p.json", "w"), indent=1)
