"""GPU parity of the tensor-core matcher (csrc/match_mma.cu: tcgen05.mma kind::i8 on descriptors expanded to signed
bytes, dot = 256 - 2 * hamming) against the CPU oracle (visnav::matchDescriptors, reference keypoints.h:323-369) and
against the POPC kernels.  Bit-exact: integer work.  VSLAM_OPT_MATCH_MMA = 2 forces the path for any batch size."""
import numpy as np
import pytest

import visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import DescStore, VSLAM_OPT_MATCH_MMA

pytestmark = pytest.mark.gpu


@pytest.fixture()
def mma(ctx):
    ctx.set_option(VSLAM_OPT_MATCH_MMA, 2)
    yield ctx
    ctx.set_option(VSLAM_OPT_MATCH_MMA, 1)
    assert ctx.mma_timeouts() == 0, "a pipeline wait of the tensor-core matcher timed out"


def _check_jobs(ctx, oracle_port, sets, jobs, params):
    store = DescStore(ctx, sets)
    try:
        for thr, ratio in params:
            matches, counts = ctx.match_batch(store, jobs, thr, ratio)
            for (i, j), m, c in zip(jobs, matches, counts):
                want = oracle_port.match(sets[i], sets[j], thr, ratio)
                assert c == len(want) and np.array_equal(m[:c], want), (i, j, thr, ratio, c, len(want))
    finally:
        store.close()


def test_mma_ragged_sets(mma, oracle_port):
    """Set sizes around the 128-descriptor tile and the 32-column chunk, empty and single-descriptor sets, identical
    sets (distance-0 ties), both job orientations."""
    sizes = [1400, 1337, 0, 63, 64, 65, 127, 128, 129, 1, 255, 257, 700, 1500, 33, 31]
    kf = wl.keyframe_descriptors(len(sizes), 1500, 25)
    sets = [kf[i][:n] for i, n in enumerate(sizes)]
    sets[4] = sets[1][:64].copy()
    jobs = wl.keyframe_pair_jobs(len(sizes))
    jobs = np.concatenate([jobs, jobs[:, ::-1][::7], np.array([[0, 0], [3, 3], [9, 9]], np.int32)])
    _check_jobs(mma, oracle_port, sets, jobs, [(70, 1.2), (256, 1.0)])


def test_mma_heavy_ties_and_odd_parameters(mma, oracle_port):
    """Few base patterns with few flipped bits: many equal distances inside and across tiles (lowest index must win);
    thresholds / ratios far from (70, 1.2)."""
    rng = np.random.default_rng(3)
    base = wl.random_descriptors(40, 11)
    sets = []
    for k, n in enumerate([300, 517, 1100, 129, 1500]):
        sets.append(wl.flip_bits(base[rng.integers(0, 40, n)], 6, rng))
    jobs = np.array([(a, b) for a in range(5) for b in range(5)], np.int32)
    _check_jobs(mma, oracle_port, sets, jobs,
                [(70, 1.2), (256, 1.0), (20, 1.0), (70, 0.5), (70, -3.0), (70, float("inf")), (1, 1.2), (257, 1.2),
                 (1000, 20.0), (-5, 1.2)])


def test_mma_extremes(mma, oracle_port):
    """All distances 256 (nothing updates: strict '<'), all distances 0, complementary and duplicated sets."""
    z = np.zeros((150, 8), np.uint32)
    f = np.full((140, 8), 0xFFFFFFFF, np.uint32)
    d = wl.random_descriptors(200, 9)
    sets = [z, f, d, np.concatenate([d, d]), ~d, z[:1], f[:1]]
    jobs = np.array([(0, 1), (1, 0), (0, 0), (2, 2), (3, 2), (2, 3), (2, 4), (4, 2), (5, 6), (5, 5), (0, 2)], np.int32)
    _check_jobs(mma, oracle_port, sets, jobs, [(70, 1.2), (300, 1.0), (257, 0.0)])


def test_mma_equals_popc_on_a_full_batch(ctx, oracle_port):
    """A batch large enough to take the tensor-core path by itself (>= half the SM count): 24 keyframes, all 276
    unordered pairs; equal to the POPC kernels on every job and to the oracle on a sample."""
    kf = wl.keyframe_descriptors(24, 1500, 25)
    rng = np.random.default_rng(1)
    sets = [kf[i][:int(n)] for i, n in enumerate(rng.integers(900, 1501, 24))]
    jobs = wl.keyframe_pair_jobs(24)
    store = DescStore(ctx, sets)
    try:
        a_m, a_c = ctx.match_batch(store, jobs, 70, 1.2)
        ctx.set_option(VSLAM_OPT_MATCH_MMA, 0)
        try:
            b_m, b_c = ctx.match_batch(store, jobs, 70, 1.2)
        finally:
            ctx.set_option(VSLAM_OPT_MATCH_MMA, 1)
        assert np.array_equal(a_c, b_c) and (a_c > 0).all()
        for k in range(len(jobs)):
            assert np.array_equal(a_m[k][:a_c[k]], b_m[k][:b_c[k]]), jobs[k]
        for k in range(0, len(jobs), 23):
            i, j = jobs[k]
            want = oracle_port.match(sets[i], sets[j], 70, 1.2)
            assert a_c[k] == len(want) and np.array_equal(a_m[k][:a_c[k]], want)
    finally:
        store.close()
    assert ctx.mma_timeouts() == 0


def test_mma_store_append(mma, oracle_port):
    """Sets appended after the first tensor-core batch: the signed-byte copy of the store follows."""
    kf = wl.keyframe_descriptors(6, 800, 25)
    sets = [kf[0], kf[1][:650]]
    store = DescStore(mma, sets)
    try:
        for k in range(2, 6):
            s = kf[k][:800 - 100 * (k % 3)]
            assert store.append(s) == k
            sets.append(s)
            jobs = np.array([(k, j) for j in range(k)] + [(0, k)], np.int32)
            matches, counts = mma.match_batch(store, jobs, 70, 1.2)
            for (a, b), m, c in zip(jobs, matches, counts):
                want = oracle_port.match(sets[a], sets[b], 70, 1.2)
                assert c == len(want) and np.array_equal(m[:c], want), (k, a, b)
    finally:
        store.close()


def test_mma_stereo_batch_equals_popc_and_oracle(ctx, oracle_port):
    """The stereo pipeline with the matcher on the tensor cores (forced, 7 pairs of 320x240) and on the POPC pipe:
    identical match lists, equal to the oracle's on the pipeline's own descriptors."""
    imgs = wl.synth_stereo_batch(7, 320, 240, workers=1)
    ctx.set_option(VSLAM_OPT_MATCH_MMA, 2)
    try:
        a = ctx.stereo_batch(imgs, 600, True, 70, 1.2)
    finally:
        ctx.set_option(VSLAM_OPT_MATCH_MMA, 0)
    try:
        b = ctx.stereo_batch(imgs, 600, True, 70, 1.2)
    finally:
        ctx.set_option(VSLAM_OPT_MATCH_MMA, 1)
    assert np.array_equal(a["n_kp"], b["n_kp"]) and np.array_equal(a["n_matches"], b["n_matches"])
    for k in range(7):
        m = a["n_matches"][k]
        assert m > 20 and np.array_equal(a["matches"][k, :m], b["matches"][k, :m])
        dl = a["desc"][2 * k, :a["n_kp"][2 * k]]
        dr = a["desc"][2 * k + 1, :a["n_kp"][2 * k + 1]]
        assert np.array_equal(a["matches"][k, :m], oracle_port.match(dl, dr, 70, 1.2))
    assert ctx.mma_timeouts() == 0
