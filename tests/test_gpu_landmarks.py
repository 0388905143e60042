"""CUDA find_matches_landmarks (vslam_match_landmarks, csrc/landmarks.cu) against the CPU oracle,
the committed reference fixture and -- where oracle/_ref travelled -- the reference's own function
(vo_utils.h:83-167).  Bit-exact: the output is integer pairs in keypoint order."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
PARAMS = [(20.0, 70, 1.2), (35.0, 256, 1.0), (12.5, 40, 1.5), (20.0, 70, 0.9)]


@pytest.mark.parametrize("seed,n_kp,n_lm", [(0, 300, 250), (1, 1313, 900), (2, 64, 700), (3, 1500, 2500)])
def test_landmarks_vs_oracle(ctx, seed, n_kp, n_lm):
    import oracle.oracle as oo
    from visnav_b200 import workloads as wl
    sc = wl.landmark_scene(n_kp=n_kp, n_landmarks=n_lm, seed=seed)
    for r, thr, ratio in PARAMS:
        want = oo.match_landmarks_port(*sc, r, thr, ratio)
        got = ctx.match_landmarks(*sc, r, thr, ratio)
        assert np.array_equal(got, want), (seed, r, thr, ratio, len(got), len(want))
        ref = oo.match_landmarks_ref(*sc, r, thr, ratio)
        if ref is not None:
            assert np.array_equal(got, ref)


def test_landmarks_reference_fixture(ctx):
    g = np.load(os.path.join(GOLD, "landmarks_220x200.npz"))
    sc = (g["kp_xy"], g["kp_desc"], g["proj_xy"], g["obs_offsets"], g["obs_desc"])
    for k in [k for k in g.files if k.startswith("m_")]:
        _, r, thr, ratio = k.split("_")
        got = ctx.match_landmarks(*sc, float(r.replace("p", ".")), int(thr), float(ratio.replace("p", ".")))
        assert np.array_equal(got, g[k]), k


def test_landmarks_edge_cases(ctx):
    from visnav_b200 import workloads as wl
    from visnav_b200.keypoints import VslamError
    sc = wl.landmark_scene(n_kp=50, n_landmarks=40, seed=3)
    e2, e8 = np.zeros((0, 2)), np.zeros((0, 8), np.uint32)
    assert len(ctx.match_landmarks(e2, e8, sc[2], sc[3], sc[4])) == 0
    assert len(ctx.match_landmarks(sc[0], sc[1], e2, np.zeros(1, np.int32), e8)) == 0
    off0 = np.zeros(len(sc[2]) + 1, np.int32)
    assert len(ctx.match_landmarks(sc[0], sc[1], sc[2], off0, e8, 20.0, 256, 1.0)) == 0
    assert len(ctx.match_landmarks(*sc, 0.0, 70, 1.2)) == 0
    # a landmark with many observations (more than one warp step)
    rng = np.random.default_rng(5)
    kp_xy = np.array([[100.0, 100.0], [300.0, 200.0]])
    kp_desc = rng.integers(0, 2 ** 32, (2, 8), dtype=np.uint64).astype(np.uint32)
    obs = rng.integers(0, 2 ** 32, (100, 8), dtype=np.uint64).astype(np.uint32)
    obs[77] = kp_desc[1] ^ np.array([1, 0, 0, 0, 0, 0, 0, 3], np.uint32)   # distance 3 to keypoint 1
    got = ctx.match_landmarks(kp_xy, kp_desc, np.array([[301.0, 199.0]]), np.array([0, 100], np.int32), obs)
    assert np.array_equal(got, np.array([[1, 0]], np.int32))
    # capacity error
    with pytest.raises(VslamError):
        ctx.match_landmarks(*wl.landmark_scene(n_kp=200, n_landmarks=200, seed=9), 35.0, 256, 1.0, cap=3)
    # descending offsets are rejected
    with pytest.raises(VslamError):
        ctx.match_landmarks(sc[0], sc[1], sc[2][:2], np.array([0, 5, 3], np.int32), sc[4])
