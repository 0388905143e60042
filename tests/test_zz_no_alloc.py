"""vslam_best2_noalloc (csrc/noalloc.cu): the allocation-free variant of vslam_best2 -- device pointers in and out, its
per-range partials in a fixed file-scope device array instead of cudaMalloc'd scratch.  No counterpart in the reference;
the row scan it computes is keypoints.h:331-354.

Checked through the C ABI against the existing entry point (vslam_best2) and the CPU oracle, bit-exact (integer work),
on two calls in a row with different shapes: the second call finds the first one's partials in the shared array and must
not see them.  Collected last (file name) so that a failure here hides no other test."""
import numpy as np
import pytest

import visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import VslamError

pytestmark = pytest.mark.gpu


def _noalloc(ctx, torch, q, db):
    """q / db: host uint32 [n][8] -> out3 [nq][3] through the device-pointer entry point (buffers owned by the test)."""
    dq = torch.from_numpy(np.ascontiguousarray(q).view(np.int32).reshape(-1, 8)).cuda()
    ddb = torch.from_numpy(np.ascontiguousarray(db).view(np.int32).reshape(-1, 8)).cuda()
    out = torch.full((max(len(q), 1), 3), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ctx.best2_noalloc(dq.data_ptr() if len(q) else 0, len(q), ddb.data_ptr() if len(db) else 0, len(db), out.data_ptr())
    ctx.sync()
    return out.cpu().numpy()[:len(q)]


def test_noalloc_best2_equals_best2_twice_in_a_row(ctx, oracle_port):
    import torch
    rng = np.random.default_rng(5)
    base = wl.random_descriptors(64, 21)
    # call 1: many ranges (64 x 512 columns), ragged row count, heavy ties (few base patterns, few flipped bits)
    q1 = wl.flip_bits(base[rng.integers(0, 64, 1337)], 5, rng)
    db1 = wl.flip_bits(base[rng.integers(0, 64, 32768 - 77)], 5, rng)
    # call 2: fewer rows, fewer and shorter ranges -- the array still holds call 1's partials at every slot
    q2 = wl.random_descriptors(300, 22)
    db2 = np.concatenate([wl.random_descriptors(700, 23), q2[::3]])
    for q, db in ((q1, db1), (q2, db2)):
        got = _noalloc(ctx, torch, q, db)
        want = ctx.best2(q, db)
        assert np.array_equal(got, want)
        assert np.array_equal(got, oracle_port.best2(q, db))
    # the same inputs again give the same answer (nothing accumulates in the array between calls)
    assert np.array_equal(_noalloc(ctx, torch, q1, db1), ctx.best2(q1, db1))


def test_noalloc_best2_edges_and_limits(ctx):
    import torch
    d = wl.random_descriptors(130, 4)
    # empty database: (256, 0, 256) for every query, as vslam_best2; single column; no queries
    assert np.array_equal(_noalloc(ctx, torch, d, d[:0]), ctx.best2(d, d[:0]))
    assert np.array_equal(_noalloc(ctx, torch, d, d[:1]), ctx.best2(d, d[:1]))
    assert _noalloc(ctx, torch, d[:0], d).shape == (0, 3)
    # beyond the fixed scratch: a clear error, nothing launched
    buf = torch.zeros((4097, 8), dtype=torch.int32, device="cuda")
    out = torch.zeros((4097, 3), dtype=torch.int32, device="cuda")
    with pytest.raises(VslamError, match="4096"):
        ctx.best2_noalloc(buf.data_ptr(), 4097, buf.data_ptr(), 10, out.data_ptr())
    with pytest.raises(VslamError, match="2\\^20"):
        ctx.best2_noalloc(buf.data_ptr(), 10, buf.data_ptr(), (1 << 20) + 1, out.data_ptr())
    with pytest.raises(VslamError, match="aligned"):
        ctx.best2_noalloc(buf.data_ptr() + 4, 10, buf.data_ptr(), 10, out.data_ptr())
    # the largest accepted query count runs
    q = wl.random_descriptors(4096, 6)
    assert np.array_equal(_noalloc(ctx, torch, q, d), ctx.best2(q, d))
