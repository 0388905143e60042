#!/usr/bin/env python3
"""BASELINE config 4 (1500 queries x 200 000 map descriptors) with the map SHARDED over the ranks of a torchrun
job: every rank scans its shard (vslam_match_shard), the 16-byte-per-query parts are all-gathered
over NCCL (24 KB per rank -- the only exchange), rank 0 merges (vslam_match_merge_shards) and checks the result
against the unsharded call.  Prints one JSON line; device times are the max over ranks.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/c4_sharded.py
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import visnav_b200  # noqa: E402
from visnav_b200 import sharding, workloads as wl  # noqa: E402

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
dist = None
if world > 1:
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = visnav_b200.Context(local)
q, db = wl.reloc_query()
lo, hi = sharding.shard_range(len(db), rank, world)
shard = np.ascontiguousarray(db[lo:hi])
best_kernel, best_call = 1e9, 1e9
for rep in range(6):
    t0 = time.perf_counter()
    part = ctx.match_shard(q, shard, lo)
    best_call = min(best_call, time.perf_counter() - t0)
    best_kernel = min(best_kernel, ctx.last_match_ms())
sharding.gather_shard_parts(part, dist)           # first collective: NCCL communicator set-up, not timed
t0 = time.perf_counter()
parts = sharding.gather_shard_parts(part, dist)
t_gather = time.perf_counter() - t0
kernel_ms = sharding.max_over_ranks(best_kernel, dist)
call_ms = sharding.max_over_ranks(best_call * 1e3, dist)
if rank == 0:
    t0 = time.perf_counter()
    got = ctx.match_merge_shards(parts)
    t_merge = time.perf_counter() - t0
    ctx.match(q, db)
    want = ctx.match(q, db)
    full_ms = ctx.last_match_ms()
    ok = bool(np.array_equal(got, want) and len(got) == 1500)
    print(json.dumps({"workload": "config 4: 1500 x 200000, map sharded over %d GPU(s)" % world, "n_gpus": world,
                      "shard_rows": hi - lo, "shard_kernel_ms_max": kernel_ms, "shard_call_ms_max": call_ms,
                      "gather_ms_rank0": 1e3 * t_gather, "merge_call_ms": 1e3 * t_merge,
                      "unsharded_kernel_ms_one_gpu": full_ms, "dist_per_s_kernels": 1500 * 200000 / (kernel_ms * 1e-3),
                      "matches": int(len(got)), "equals_unsharded": ok, "gathered_bytes_per_rank": int(part.nbytes)}),
          flush=True)
    assert ok
if dist is not None:
    dist.barrier()
    dist.destroy_process_group()
