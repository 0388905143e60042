"""world_size-2 test of the N>1 path on CPU (gloo): pairs are split by rank, each rank runs its
shard independently (the CPU oracle stands in for the device here), results are gathered in rank
order and equal the unsharded run; the reported time is the max over ranks."""
import os
import socket

import numpy as np
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch.distributed as dist
    import oracle
    import visnav_b200  # noqa: F401
    from visnav_b200 import sharding, workloads as wl
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    o = oracle.load_oracle()
    kf = wl.keyframe_descriptors(5, 120, 25)
    jobs = wl.keyframe_pair_jobs(5)
    lo, hi = sharding.shard_range(len(jobs), rank, world)
    local = np.array([len(o.match(kf[i], kf[j], 70, 1.2)) for i, j in jobs[lo:hi]], np.int32)
    allc = sharding.gather_counts(local, dist)
    tmax = sharding.max_over_ranks(1.0 + rank, dist)
    if rank == 0:
        full = np.array([len(o.match(kf[i], kf[j], 70, 1.2)) for i, j in jobs], np.int32)
        q.put((np.array_equal(allc, full), tmax, int(allc.sum())))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_gathers_the_unsharded_result():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, tmax, total = q.get(timeout=240)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert ok and tmax == 2.0 and total > 0


def _worker_db(rank, world, port, q):
    """Config 4 with desc2 sharded by rank: each rank scans its shard (CPU restatement standing in for the
    device), the 16-byte-per-query parts are all-gathered, the merge equals the unsharded matchDescriptors."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch.distributed as dist
    import oracle
    import oracle.oracle as oo
    import visnav_b200  # noqa: F401
    from visnav_b200 import sharding, workloads as wl
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    o = oracle.load_oracle()
    rng = np.random.default_rng(3)
    db = wl.random_descriptors(701, 5)
    q_desc = wl.flip_bits(db[rng.integers(0, 701, 90)], 20, rng)
    db[400:420] = db[100:120]                      # equal best distances in different shards: lowest index must win
    lo, hi = sharding.shard_range(len(db), rank, world)
    results = []
    for thr, ratio in [(70, 1.2), (256, 1.0), (40, 1.5)]:
        part = oo.match_shard_part(o, q_desc, db[lo:hi], lo, thr, ratio)
        parts = sharding.gather_shard_parts(part, dist)
        got = oo.merge_shard_parts(parts, thr, ratio)
        rev = oo.merge_shard_parts(parts[::-1], thr, ratio)   # the merge does not depend on the order of the parts
        want = o.match(q_desc, db, thr, ratio)
        results.append(bool(np.array_equal(got, want) and np.array_equal(rev, want) and parts.shape == (world, 90, 4)))
    if rank == 0:
        q.put((all(results), len(want)))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_db_sharding_merges_to_the_unsharded_matches():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_db, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, n = q.get(timeout=240)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert ok and n > 0
