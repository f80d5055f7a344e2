"""Worker for tests/test_bench_dist.py: one rank of a world_size-2 gloo job exercising bench.py's multi-rank plumbing."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
dist.barrier()
# rank r pretends its K steps took (10 + 5 r) ms on the device
mine_ms = 10.0 + 5.0 * rank
slowest = bench.max_over_ranks(mine_ms, world, device="cpu")
value = bench.whole_job_gbs(step_bytes=4 << 30, steps=7, world=world, slowest_seconds=slowest * 1e-3)
dist.barrier()
os.write(1, (json.dumps({"rank": rank, "slowest_ms": slowest, "value": value}) + "\n").encode())  # one atomic write per rank
dist.destroy_process_group()
