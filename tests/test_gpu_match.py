"""GPU parity of vslam_match / vslam_best2 / vslam_match_batch against the CPU oracle
(visnav::matchDescriptors, reference keypoints.h:323-369).  Bit-exact: integer work."""
import numpy as np
import pytest

import visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import DescStore, VSLAM_OPT_MATCH_CSA

pytestmark = pytest.mark.gpu


def _tied(n, seed, n_base=40, flips=6):
    """Descriptors with heavy distance ties: few base patterns, few flipped bits."""
    rng = np.random.default_rng(seed)
    base = wl.random_descriptors(n_base, seed + 1)
    d = base[rng.integers(0, n_base, n)]
    return wl.flip_bits(d, flips, rng) if n else d


CASES = [(1, 1), (1, 40), (40, 1), (31, 33), (64, 65), (257, 255), (1500, 1500), (1537, 1300), (3100, 700), (700, 3100)]


@pytest.mark.parametrize("csa", [1, 0])
@pytest.mark.parametrize("n1,n2", CASES)
def test_match_random_correlated(ctx, oracle_port, n1, n2, csa):
    ctx.set_option(VSLAM_OPT_MATCH_CSA, csa)
    rng = np.random.default_rng(n1 * 7919 + n2)
    d2 = wl.random_descriptors(n2, n1 + 3)
    d1 = wl.random_descriptors(n1, n2 + 5)
    k = min(n1, n2)
    perm = rng.permutation(n2)[:k]
    d1[:k] = wl.flip_bits(d2[perm], 20, rng)      # correlated part: produces matches
    for thr, ratio in [(70, 1.2), (256, 1.0), (300, 1.0), (40, 2.0), (0, 1.2)]:
        got = ctx.match(d1, d2, thr, ratio)
        want = oracle_port.match(d1, d2, thr, ratio)
        assert np.array_equal(got, want), (n1, n2, thr, ratio, len(got), len(want))
    ctx.set_option(VSLAM_OPT_MATCH_CSA, 1)


@pytest.mark.parametrize("seed", range(6))
def test_match_heavy_ties(ctx, oracle_port, seed):
    n1, n2 = 300 + 37 * seed, 280 + 53 * seed
    d1, d2 = _tied(n1, seed), _tied(n2, seed)   # same base patterns -> many equal distances
    for thr, ratio in [(70, 1.2), (256, 1.0), (20, 1.0)]:
        got = ctx.match(d1, d2, thr, ratio)
        want = oracle_port.match(d1, d2, thr, ratio)
        assert np.array_equal(got, want)
    assert np.array_equal(ctx.best2(d1, d2), oracle_port.best2(d1, d2))


def test_match_odd_parameters(ctx, oracle_port):
    """Thresholds and ratios away from (70, 1.2): ratios below 1 (never reject), negative, huge and infinite ones
    (reject everything but a zero best distance), thresholds above the largest distance and below zero; fused
    (one CTA per job) and column-split kernels; weakly related, strongly related and identical sets."""
    rng = np.random.default_rng(31)
    for n1, n2, flips in [(700, 650, 20), (1500, 1400, 45), (300, 5000, 30), (200, 180, 0), (64, 64, 100)]:
        d2 = wl.random_descriptors(n2, n1)
        d1 = wl.random_descriptors(n1, n2 + 1)
        k = min(n1, n2)
        d1[:k] = wl.flip_bits(d2[rng.permutation(n2)[:k]], flips, rng) if flips else d2[rng.permutation(n2)[:k]]
        for thr, ratio in [(70, 1.2), (70, 0.5), (70, 0.0), (70, -3.0), (70, 1e9), (70, float("inf")), (1, 1.2),
                           (50, 3.7), (256, 1.2), (257, 1.2), (1000, 20.0), (128, 2.0), (-5, 1.2)]:
            got = ctx.match(d1, d2, thr, ratio)
            want = oracle_port.match(d1, d2, thr, ratio)
            assert np.array_equal(got, want), (n1, n2, flips, thr, ratio, len(got), len(want))
        assert np.array_equal(ctx.best2(d1, d2), oracle_port.best2(d1, d2))


def test_match_extremes(ctx, oracle_port):
    z = np.zeros((5, 8), np.uint32)
    f = np.full((7, 8), 0xFFFFFFFF, np.uint32)
    # all distances 256: nothing updates (strict '<'), argmin stays 0
    assert np.array_equal(ctx.best2(z, f), oracle_port.best2(z, f))
    assert len(ctx.match(z, f, 300, 1.0)) == len(oracle_port.match(z, f, 300, 1.0))
    assert np.array_equal(ctx.match(z, f, 300, 1.0), oracle_port.match(z, f, 300, 1.0))
    # identical sets: every row ties at 0 with itself
    d = wl.random_descriptors(100, 9)
    assert np.array_equal(ctx.match(d, d, 70, 1.2), oracle_port.match(d, d, 70, 1.2))
    dd = np.concatenate([d, d])  # duplicates: second best == best -> ratio test rejects
    assert np.array_equal(ctx.match(dd, d, 70, 1.2), oracle_port.match(dd, d, 70, 1.2))
    # empty inputs: matches.clear() and nothing else
    e = np.zeros((0, 8), np.uint32)
    assert len(ctx.match(e, d)) == 0 and len(ctx.match(d, e)) == 0 and len(ctx.match(e, e)) == 0


def test_match_reference_library(ctx, oracle_ref):
    if oracle_ref is None:
        pytest.skip("oracle/_ref not shipped")
    rng = np.random.default_rng(5)
    d2 = wl.random_descriptors(900, 1)
    d1 = wl.flip_bits(d2[rng.permutation(900)[:800]], 25, rng)
    assert np.array_equal(ctx.match(d1, d2, 70, 1.2), oracle_ref.match(d1, d2, 70, 1.2))


def test_match_large_split_path(ctx, oracle_port):
    """Config 4 shape scaled down: queries vs a DB too big for one CTA (column ranges + merge)."""
    q, db = wl.reloc_query(300, 20000, 25)
    assert np.array_equal(ctx.best2(q, db), oracle_port.best2(q, db))
    got = ctx.match(q, db, 70, 1.2)
    want = oracle_port.match(q, db, 70, 1.2)
    assert np.array_equal(got, want) and len(got) > 100
    # and the transposed problem (many rows, few columns)
    got = ctx.match(db[:12000], q, 70, 1.2, cap=300)
    want = oracle_port.match(db[:12000], q, 70, 1.2)
    assert np.array_equal(got, want)


def test_match_full_size_properties(ctx):
    """Config 4 at full size through size-independent properties: planted matches are found,
    the predicate is symmetric under swapping the two sets."""
    q, db = wl.reloc_query(1500, 200000, 25)
    m = ctx.match(q, db, 70, 1.2)
    assert len(m) == 1500
    assert np.array_equal(m[:, 0], np.arange(1500)) and np.array_equal(m[:, 1], 133 * np.arange(1500))
    mt = ctx.match(db, q, 70, 1.2, cap=1500)
    assert np.array_equal(mt[np.argsort(mt[:, 1])][:, ::-1], m)


def test_match_batch_store(ctx, oracle_port):
    kf = wl.keyframe_descriptors(6, 700, 25)
    kf_list = [kf[0], kf[1][:650], kf[2], kf[3][:10], kf[4], kf[5]]
    store = DescStore(ctx, kf_list)
    jobs = wl.keyframe_pair_jobs(6)
    matches, counts = ctx.match_batch(store, jobs, 70, 1.2)
    for (i, j), m, c in zip(jobs, matches, counts):
        want = oracle_port.match(kf_list[i], kf_list[j], 70, 1.2)
        assert c == len(want) and np.array_equal(m[:c], want), (i, j)
    store.close()


@pytest.mark.parametrize("n_shards", [1, 2, 3, 8])
def test_match_db_sharded_equals_unsharded(ctx, oracle_port, n_shards):
    """Config 4 with desc2 split into shards (one per GPU in production; here one after the other on one GPU):
    vslam_match_shard + vslam_match_merge_shards == vslam_match on the whole DB == the oracle; equal best
    distances in different shards resolve to the lowest global index; an empty shard is harmless."""
    import oracle.oracle as oo
    from visnav_b200 import sharding, workloads as wl
    rng = np.random.default_rng(n_shards)
    db = wl.random_descriptors(20011, 8)
    q = wl.flip_bits(db[rng.integers(0, len(db), 700)], 22, rng)
    db[15000:15050] = db[3000:3050]
    for thr, ratio in [(70, 1.2), (256, 1.0)]:
        parts, cpu_parts = [], []
        for r in range(n_shards):
            lo, hi = sharding.shard_range(len(db), r, n_shards)
            parts.append(ctx.match_shard(q, db[lo:hi], lo, thr, ratio))
            cpu_parts.append(oo.match_shard_part(oracle_port, q, db[lo:hi], lo, thr, ratio))
        assert np.array_equal(np.stack(parts), np.stack(cpu_parts))
        parts.append(ctx.match_shard(q, db[:0], len(db), thr, ratio))      # empty shard
        want = oracle_port.match(q, db, thr, ratio)
        assert np.array_equal(ctx.match_merge_shards(np.stack(parts), thr, ratio), want)
        assert np.array_equal(ctx.match_merge_shards(np.stack(parts[::-1]), thr, ratio), want)
        assert np.array_equal(ctx.match(q, db, thr, ratio), want)


def test_match_db_sharded_config4_full_size(ctx):
    """BASELINE config 4 in full (1500 x 200 000) over 8 shards: every planted query finds its DB row."""
    from visnav_b200 import sharding, workloads as wl
    q, db = wl.reloc_query()
    parts = []
    for r in range(8):
        lo, hi = sharding.shard_range(len(db), r, 8)
        parts.append(ctx.match_shard(q, db[lo:hi], lo))
    got = ctx.match_merge_shards(np.stack(parts))
    assert np.array_equal(got, ctx.match(q, db))
    assert np.array_equal(got, np.stack([np.arange(1500), 133 * np.arange(1500)], 1))


@pytest.mark.parametrize("split", [1, 0])
@pytest.mark.parametrize("n_sets", [2, 5, 13])
def test_match_batch_small_batches_both_paths(ctx, oracle_port, split, n_sets):
    """Few jobs: the column-split path (every job spread over all SMs) and the one-CTA-per-job path give the
    oracle's lists -- ragged set sizes, an empty set, sets smaller than one stage, ties."""
    from visnav_b200.keypoints import VSLAM_OPT_MATCH_SPLIT_SMALL
    kf = wl.keyframe_descriptors(n_sets, 1400, 25)
    rng = np.random.default_rng(n_sets)
    sizes = [1400, 1337, 0, 63, 64, 65, 700, 1, 1400, 999, 130, 1400, 512][:n_sets]
    kf_list = [kf[i][:sizes[i]] for i in range(n_sets)]
    if n_sets > 4:
        kf_list[4] = kf_list[1][:64].copy()        # identical descriptors in two sets: distance 0 ties
    store = DescStore(ctx, kf_list)
    jobs = wl.keyframe_pair_jobs(n_sets)
    jobs = np.concatenate([jobs, jobs[:, ::-1][:3]])
    ctx.set_option(VSLAM_OPT_MATCH_SPLIT_SMALL, split)
    try:
        for thr, ratio in [(70, 1.2), (256, 1.0)]:
            matches, counts = ctx.match_batch(store, jobs, thr, ratio)
            for (i, j), m, c in zip(jobs, matches, counts):
                want = oracle_port.match(kf_list[i], kf_list[j], thr, ratio)
                assert c == len(want) and np.array_equal(m[:c], want), (split, i, j, thr)
    finally:
        ctx.set_option(VSLAM_OPT_MATCH_SPLIT_SMALL, 1)
    store.close()


def test_desc_store_append(ctx, oracle_port):
    """vslam_desc_store_append: sets added after creation (buffers grow and move) match like sets given at
    creation; earlier indices stay valid."""
    kf = wl.keyframe_descriptors(9, 900, 25)
    sets = [kf[0][:700], kf[1]]
    store = DescStore(ctx, sets)
    for k in range(2, 9):
        s = kf[k][:900 - 90 * (k % 4)] if k != 5 else kf[k][:0]
        assert store.append(s) == k
        sets.append(s)
        assert len(store) == k + 1
        jobs = np.array([(k, j) for j in range(k)] + [(0, k)], np.int32)
        matches, counts = ctx.match_batch(store, jobs, 70, 1.2)
        for (a, b), m, c in zip(jobs, matches, counts):
            want = oracle_port.match(sets[a], sets[b], 70, 1.2)
            assert c == len(want) and np.array_equal(m[:c], want), (k, a, b)
    store.close()


def test_config4_full_size_equals_oracle(ctx, oracle_port):
    """BASELINE config 4 at full size (1500 x 200 000) against the CPU oracle directly: the (best, index, second best)
    triple of every query and the match list -- bit-exact.  (The oracle needs ~1 s.)"""
    q, db = wl.reloc_query()
    assert np.array_equal(ctx.best2(q, db), oracle_port.best2(q, db))
    for thr, ratio in [(70, 1.2), (90, 1.05)]:
        assert np.array_equal(ctx.match(q, db, thr, ratio), oracle_port.match(q, db, thr, ratio)), (thr, ratio)


def test_config3_sampled_jobs_equal_oracle(ctx, oracle_port):
    """BASELINE config 3 shape (keyframes of 1500 descriptors, 25 flipped bits, all unordered pairs): 60 keyframes =
    1770 jobs through vslam_match_batch (tensor-core path: the batch fills the GPU), 240 of them compared with the
    oracle's match lists, the rest by count against the POPC kernels."""
    from visnav_b200.keypoints import VSLAM_OPT_MATCH_MMA
    kf = wl.keyframe_descriptors(60, 1500, 25)
    store = DescStore(ctx, kf)
    jobs = wl.keyframe_pair_jobs(60)
    try:
        matches, counts = ctx.match_batch(store, jobs, 70, 1.2)
        ctx.set_option(VSLAM_OPT_MATCH_MMA, 0)
        try:
            _, counts_popc = ctx.match_batch(store, jobs, 70, 1.2, want_matches=False)
        finally:
            ctx.set_option(VSLAM_OPT_MATCH_MMA, 1)
    finally:
        store.close()
    assert np.array_equal(counts, counts_popc) and counts.min() > 1000
    rng = np.random.default_rng(12)
    for k in rng.choice(len(jobs), 240, replace=False):
        i, j = jobs[k]
        want = oracle_port.match(kf[i], kf[j], 70, 1.2)
        assert counts[k] == len(want) and np.array_equal(matches[k][:counts[k]], want), (i, j)
    assert ctx.mma_timeouts() == 0
