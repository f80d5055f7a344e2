#!/usr/bin/env python
"""Second opinion on the P2P verdict (optional): NCCL all-reduce and all-gather over the same GPUs.
Launch with torchrun, one rank per GPU.  Prints one JSON line from rank 0: algorithm and bus bandwidth in GB/s."""
import json
import os
import sys

import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 30
x = torch.ones(nbytes // 4, dtype=torch.float32, device="cuda")
out = {}
for name, fn, factor in (("all_reduce", lambda: dist.all_reduce(x), 2.0 * (world - 1) / world),):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record()
    e1.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / 10], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    alg = nbytes / (ms.item() * 1e-3) / 1e9
    out[name] = {"algbw_gbs": alg, "busbw_gbs": alg * factor, "ms": ms.item()}
ok = bool(torch.isfinite(x).all().item())
if rank == 0:
    print(json.dumps({"world": world, "bytes": nbytes, "finite": ok, **out}), flush=True)
dist.destroy_process_group()
