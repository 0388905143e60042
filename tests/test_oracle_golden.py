"""The CPU oracle (oracle/visnav_oracle.c) against the committed golden vectors (tests/golden/,
made by tools/make_golden.py from cv2 and from the reference's own functions) -- this is what
pins the oracle.  Runs without a GPU."""
import os

import numpy as np
import pytest

import visnav_b200  # noqa: F401  (puts the package directory on the path)
from visnav_b200 import workloads as wl

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


def _libs(oracle_port, oracle_ref):
    return [oracle_port] + ([oracle_ref] if oracle_ref is not None else [])


@pytest.mark.parametrize("name", ["gftt_320x240", "gftt_161x123", "gftt_752x480"])
def test_gftt_matches_opencv(oracle_port, name):
    g = _load(name)
    img = g["img"]
    if img.size == 0:
        from visnav_b200 import workloads as wl
        img = wl.synth_image(int(g["seed"]), int(g["w"]), int(g["h"]))
    got = oracle_port.gftt(img, int(g["num_features"]))
    assert np.array_equal(got, g["corners_unopt"])      # ordered equality, scalar OpenCV
    assert np.array_equal(got, g["corners_default"])    # and the AVX2 dispatch on these images
    if "lam_unopt" in g.files:
        lam = oracle_port.response(img)
        assert np.array_equal(lam.view(np.uint32), g["lam_unopt"].view(np.uint32))   # bit-exact map


def test_detect_applies_border_filter(oracle_port):
    g = _load("gftt_320x240")
    xy = oracle_port.detect(g["img"], int(g["num_features"]))
    c = g["corners_unopt"]
    keep = (c[:, 0] >= 19) & (c[:, 0] < 320 - 19) & (c[:, 1] >= 19) & (c[:, 1] < 240 - 19)
    assert np.array_equal(xy, c[keep].astype(np.float64))


def test_angles_descriptors_match_reference(oracle_port, oracle_ref):
    g = _load("stereo_320x240")
    for lib in _libs(oracle_port, oracle_ref):
        for side in "lr":
            img = g["left" if side == "l" else "right"]
            xy = lib.detect(img, 400)
            assert np.array_equal(xy, g["xy_" + side])
            ang = lib.angles(img, xy, True)
            assert np.array_equal(ang, g["ang_" + side])      # same libm, same integer moments
            assert np.array_equal(lib.describe(img, xy, ang), g["desc_" + side])
        assert np.array_equal(lib.describe(g["left"], g["xy_l"], np.zeros(len(g["xy_l"]))), g["desc_l_norot"])
        assert np.array_equal(lib.angles(g["left"], g["xy_l"], False), np.zeros(len(g["xy_l"])))
        assert np.array_equal(lib.match(g["desc_l"], g["desc_r"], 70, 1.2), g["matches"])


def test_match_ties_golden(oracle_port, oracle_ref):
    g = _load("match_ties")
    for lib in _libs(oracle_port, oracle_ref):
        for key in [k for k in g.files if k.startswith("m_")]:
            _, thr, ratio = key.split("_")
            got = lib.match(g["d1"], g["d2"], int(thr), float(ratio.replace("p", ".")))
            assert np.array_equal(got, g[key]), key


def test_match_predicate_properties(oracle_port):
    """Size-independent properties of the predicate: symmetric under swapping the
    sets, ascending id1, one-to-one, empty inputs give nothing."""
    from visnav_b200 import workloads as wl
    rng = np.random.default_rng(4)
    d2 = wl.random_descriptors(400, 1)
    d1 = wl.flip_bits(d2[rng.permutation(400)[:350]], 22, rng)
    m = oracle_port.match(d1, d2, 70, 1.2)
    mt = oracle_port.match(d2, d1, 70, 1.2)
    assert len(m) > 300 and np.all(np.diff(m[:, 0]) > 0)
    assert len(set(m[:, 1])) == len(m)
    assert np.array_equal(mt[np.argsort(mt[:, 1])][:, ::-1], m)
    e = np.zeros((0, 8), np.uint32)
    assert len(oracle_port.match(e, d2)) == 0 and len(oracle_port.match(d1, e)) == 0
    b = oracle_port.best2(d1[:5], e)
    assert np.array_equal(b, np.tile([256, 0, 256], (5, 1)))


def test_contraction_modes_differ_only_rarely(oracle_port):
    g = _load("stereo_320x240")
    a = oracle_port.describe(g["left"], g["xy_l"], g["ang_l"], contract=True)
    b = oracle_port.describe(g["left"], g["xy_l"], g["ang_l"], contract=False)
    assert np.array_equal(a, g["desc_l"])                 # oracle/_ref was built with GCC's default contraction
    assert (a != b).any(1).mean() < 0.05                  # the two roundings almost never disagree


def _lm_params(key):
    _, r, thr, ratio = key.split("_")
    return float(r.replace("p", ".")), int(thr), float(ratio.replace("p", "."))


def test_landmark_matching_vs_reference_fixture():
    """vso_match_landmarks (incl. its restatement of libstdc++'s partial_sort tie order) against the
    outputs of the reference's own find_matches_landmarks (vo_utils.h:83-167)."""
    import oracle.oracle as oo
    g = _load("landmarks_220x200")
    sc = (g["kp_xy"], g["kp_desc"], g["proj_xy"], g["obs_offsets"], g["obs_desc"])
    keys = [k for k in g.files if k.startswith("m_")]
    assert len(keys) == 4
    tie_cases = 0
    for k in keys:
        r, thr, ratio = _lm_params(k)
        got = oo.match_landmarks_port(*sc, r, thr, ratio)
        assert np.array_equal(got, g[k]), k
        live = oo.match_landmarks_ref(*sc, r, thr, ratio)
        if live is not None:
            assert np.array_equal(live, g[k]), k
        tie_cases += len(g[k])
    assert tie_cases > 100


def test_landmark_matching_edge_cases():
    import oracle.oracle as oo
    from visnav_b200 import workloads as wl
    sc = wl.landmark_scene(n_kp=50, n_landmarks=40, seed=3)
    e2, e8 = np.zeros((0, 2)), np.zeros((0, 8), np.uint32)
    assert len(oo.match_landmarks_port(e2, e8, sc[2], sc[3], sc[4])) == 0          # no keypoints
    assert len(oo.match_landmarks_port(sc[0], sc[1], e2, np.zeros(1, np.int32), e8)) == 0   # no landmarks
    # landmarks without observations keep the initial 256 and can never pass a threshold <= 256
    off0 = np.zeros(len(sc[2]) + 1, np.int32)
    assert len(oo.match_landmarks_port(sc[0], sc[1], sc[2], off0, e8, 20.0, 256, 1.0)) == 0
    # radius 0: nothing is inside the gate
    assert len(oo.match_landmarks_port(*sc, 0.0, 70, 1.2)) == 0


def test_real_euroc_pair(oracle_port, oracle_ref):
    """Restated gftt == cv2 on a REAL image pair (scalar and default AVX2 dispatch agree there), and the
    restated angles / descriptors / matcher == the reference's own functions' committed outputs."""
    g = np.load(os.path.join(GOLD, "euroc_pair.npz"))
    descs = {}
    for side in ("l", "r"):
        img = g["img_" + side]
        assert np.array_equal(oracle_port.gftt(img, 1500), g["corners_unopt_" + side])
        assert np.array_equal(g["corners_unopt_" + side], g["corners_default_" + side])
        xy, ang, desc = oracle_port.detect_describe(img, 1500)
        assert np.array_equal(xy, g["xy_" + side])
        assert np.abs(ang - g["ang_" + side]).max() <= 1e-12
        assert np.array_equal(desc, g["desc_" + side])
        descs[side] = desc
    assert np.array_equal(oracle_port.match(descs["l"], descs["r"], 70, 1.2), g["matches"])
    if oracle_ref is not None:
        assert np.array_equal(oracle_ref.match(descs["l"], descs["r"], 70, 1.2), g["matches"])


def test_harris_option_matches_opencv(oracle_port):
    """goodFeaturesToTrack(useHarrisDetector=True, k=0.04).  Restatement == cv2 corners (scalar
    and default dispatch agree on the fixtures) and == cv2.cornerHarris bit for bit on the small image, whose pixel
    count is not a multiple of 4 (calcHarris' double-precision scalar tail, corner.cpp:144-150)."""
    g = np.load(os.path.join(GOLD, "harris.npz"))
    from visnav_b200 import workloads as wl
    oracle_port.set_harris(True, 0.04)
    try:
        for name in "abc":
            w, h, seed, nf = (int(v) for v in g["meta_" + name])
            img = g["img_" + name] if g["img_" + name].size else wl.synth_image(seed, w, h)
            assert np.array_equal(oracle_port.gftt(img, nf), g["corners_unopt_" + name]), name
            assert np.array_equal(g["corners_unopt_" + name], g["corners_default_" + name])
            if "harris_unopt_" + name in g.files:
                assert (w * h) % 4 != 0
                assert np.array_equal(oracle_port.response(img).view(np.uint32), g["harris_unopt_" + name].view(np.uint32))
    finally:
        oracle_port.set_harris(False)
    # off again: the min-eigenvalue fixture still holds
    g2 = np.load(os.path.join(GOLD, "gftt_161x123.npz"))
    assert np.array_equal(oracle_port.response(g2["img"]).view(np.uint32), g2["lam_unopt"].view(np.uint32))


def test_response_large_matches_vendored_opencv(oracle_port, oracle_ref):
    """tests/golden/response_large.npz: cv::cornerMinEigenVal of the reference's vendored OpenCV (scalar dispatch) at the
    benchmark sizes, as sha256 of the float bits, and its goodFeaturesToTrack corners.  The restatement (running column
    sum, box_filter.simd.hpp:206-262) reproduces the map bit for bit and the ordered corners; the direct three-row sum
    does not on the images the fixture marks."""
    import hashlib
    g = np.load(os.path.join(GOLD, "response_large.npz"))
    n_direct = 0
    for key in g["names"]:
        w, h, seed = int(key.split("x")[0]), int(key.split("x")[1].split("_")[0]), int(key.split("_")[1])
        img = wl.synth_image(seed, w, h)
        if hashlib.sha256(img.tobytes()).digest() != g[key + "_img_sha"].tobytes():
            pytest.skip("the synthetic image generator no longer reproduces the fixture images")
        lam = oracle_port.response(img)
        assert hashlib.sha256(lam.tobytes()).digest() == g[key + "_lam_sha"].tobytes(), key
        nf = 1500 if w == 752 else 4000
        assert np.array_equal(oracle_port.gftt(img, nf), g[key + "_corners_scalar"]), key
        # default (AVX2/FMA) dispatch: the same corner SET; the order may differ where responses tie in the last bit
        assert set(map(tuple, g[key + "_corners_default"])) == set(map(tuple, g[key + "_corners_scalar"]))
        if int(g[key + "_n_direct_sum_differs"]):
            oracle_port.set_box_direct(1)
            try:
                d = oracle_port.response(img)
            finally:
                oracle_port.set_box_direct(0)
            nd = int((d.view(np.uint32) != lam.view(np.uint32)).sum())
            assert nd == int(g[key + "_n_direct_sum_differs"])
            n_direct += nd
        if oracle_ref is not None and oracle_ref.has_opencv and h <= 480:     # live, where the vendored build exists
            oracle_ref.set_use_optimized(False)
            try:
                assert np.array_equal(oracle_ref.cv_response(img).view(np.uint32), lam.view(np.uint32))
            finally:
                oracle_ref.set_use_optimized(True)
    assert n_direct > 0
