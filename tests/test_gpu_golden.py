"""The CUDA path against the committed golden vectors (cv2 / the reference's own functions)."""
import os

import numpy as np
import pytest

import visnav_b200  # noqa: F401  (puts the package directory on the path)
from visnav_b200 import workloads as wl

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


@pytest.mark.parametrize("name", ["gftt_320x240", "gftt_161x123", "gftt_752x480"])
def test_corners_and_response_vs_opencv(ctx, name):
    g = _load(name)
    img = g["img"]
    if img.size == 0:
        from visnav_b200 import workloads as wl
        img = wl.synth_image(int(g["seed"]), int(g["w"]), int(g["h"]))
    h, w = img.shape
    c = g["corners_unopt"]
    keep = (c[:, 0] >= 19) & (c[:, 0] < w - 19) & (c[:, 1] >= 19) & (c[:, 1] < h - 19)
    assert np.array_equal(ctx.detect(img, int(g["num_features"])), c[keep].astype(np.float64))
    if "lam_unopt" in g.files:
        assert np.array_equal(ctx.response(img).view(np.uint32), g["lam_unopt"].view(np.uint32))


def test_stereo_vs_reference_functions(ctx):
    g = _load("stereo_320x240")
    for side, img in (("l", g["left"]), ("r", g["right"])):
        xy, ang, desc = ctx.detect_describe(img, 400, True)
        assert np.array_equal(xy, g["xy_" + side])
        assert np.abs(ang - g["ang_" + side]).max() <= 1e-5
        assert np.array_equal(desc, g["desc_" + side])
        assert np.array_equal(ctx.describe(img, g["xy_" + side], g["ang_" + side]), g["desc_" + side])
    assert np.array_equal(ctx.describe(g["left"], g["xy_l"], np.zeros(len(g["xy_l"]))), g["desc_l_norot"])
    assert np.array_equal(ctx.match(g["desc_l"], g["desc_r"], 70, 1.2), g["matches"])


def test_match_ties_vs_reference(ctx):
    g = _load("match_ties")
    for key in [k for k in g.files if k.startswith("m_")]:
        _, thr, ratio = key.split("_")
        assert np.array_equal(ctx.match(g["d1"], g["d2"], int(thr), float(ratio.replace("p", "."))), g[key]), key


def test_real_euroc_pair_vs_opencv_and_reference(ctx):
    """A real EuRoC stereo pair (bundled with the reference, data/euroc_V1): corners == cv2.goodFeaturesToTrack
    (+ the reference's 19-px filter), angles <= 1e-5 rad, descriptors and stereo matches bit-exact vs the
    reference's own functions -- slam.cpp:1125-1143 end to end, single calls and the batched call."""
    g = _load("euroc_pair")
    descs = {}
    for side in ("l", "r"):
        img = g["img_" + side]
        c = g["corners_unopt_" + side]
        keep = (c[:, 0] >= 19) & (c[:, 0] < 752 - 19) & (c[:, 1] >= 19) & (c[:, 1] < 480 - 19)
        xy, ang, desc = ctx.detect_describe(img, 1500, True)
        assert np.array_equal(xy, c[keep].astype(np.float64)) and np.array_equal(xy, g["xy_" + side])
        assert np.abs(ang - g["ang_" + side]).max() <= 1e-5
        assert np.array_equal(desc, g["desc_" + side])
        descs[side] = desc
    assert np.array_equal(ctx.match(descs["l"], descs["r"], 70, 1.2), g["matches"])
    out = ctx.stereo_batch(np.ascontiguousarray(np.stack([g["img_l"], g["img_r"]])[None]), 1500, True, 70, 1.2)
    m = int(out["n_matches"][0])
    assert np.array_equal(out["matches"][0, :m], g["matches"])
    assert np.array_equal(out["desc"][0, :out["n_kp"][0]], g["desc_l"])


def test_harris_option_vs_opencv(ctx, oracle_port):
    """vslam_set_harris(1, 0.04): corners == cv2.goodFeaturesToTrack(useHarrisDetector=True) (+ the 19-px filter),
    response == cv2.cornerHarris bit for bit on the small fixture (incl. calcHarris' double-precision tail), and ==
    the oracle on other sizes; switching it off restores the min-eigenvalue detector."""
    from visnav_b200 import workloads as wl
    g = _load("harris")
    ctx.set_harris(True, 0.04)
    oracle_port.set_harris(True, 0.04)
    try:
        for name in "abc":
            w, h, seed, nf = (int(v) for v in g["meta_" + name])
            img = g["img_" + name] if g["img_" + name].size else wl.synth_image(seed, w, h)
            c = g["corners_unopt_" + name]
            keep = (c[:, 0] >= 19) & (c[:, 0] < w - 19) & (c[:, 1] >= 19) & (c[:, 1] < h - 19)
            assert np.array_equal(ctx.detect(img, nf), c[keep].astype(np.float64)), name
            if "harris_unopt_" + name in g.files:
                assert np.array_equal(ctx.response(img).view(np.uint32), g["harris_unopt_" + name].view(np.uint32))
        for w, h, seed in [(203, 77, 3), (333, 201, 4), (752, 480, 8)]:
            img = wl.synth_image(seed, w, h)
            assert np.array_equal(ctx.response(img).view(np.uint32), oracle_port.response(img).view(np.uint32)), (w, h)
            assert np.array_equal(ctx.detect(img, 500), oracle_port.detect(img, 500))
        edges = np.zeros((120, 160), np.uint8)
        edges[:, 80:] = 200                                   # a pure edge: every Harris response <= 0 -> no corner
        assert len(ctx.detect(edges, 100)) == 0 == len(oracle_port.detect(edges, 100))
    finally:
        ctx.set_harris(False)
        oracle_port.set_harris(False)
    g2 = _load("gftt_161x123")
    assert np.array_equal(ctx.response(g2["img"]).view(np.uint32), g2["lam_unopt"].view(np.uint32))


def test_response_large_fixture(ctx):
    """Benchmark-size response maps against the reference's vendored OpenCV (tests/golden/response_large.npz, scalar
    dispatch): the full-height kernel (what batches that fill the GPU run) reproduces the map bit for bit -- compared by
    sha256 -- and both kernel modes return OpenCV's ordered corners."""
    import hashlib
    from visnav_b200.keypoints import VSLAM_OPT_RESPONSE_BAND
    g = np.load(os.path.join(GOLD, "response_large.npz"))
    for key in g["names"]:
        w, h, seed = int(key.split("x")[0]), int(key.split("x")[1].split("_")[0]), int(key.split("_")[1])
        img = wl.synth_image(seed, w, h)
        if hashlib.sha256(img.tobytes()).digest() != g[key + "_img_sha"].tobytes():
            pytest.skip("the synthetic image generator no longer reproduces the fixture images")
        nf = 1500 if w == 752 else 4000
        border = lambda c: c[(c[:, 0] >= 19) & (c[:, 0] < w - 19) & (c[:, 1] >= 19) & (c[:, 1] < h - 19)]
        want = border(g[key + "_corners_scalar"]).astype(np.float64)
        assert np.array_equal(ctx.detect(img, nf), want), key          # banded launch (single image)
        ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 65535)
        try:
            lam = ctx.response(img)
            assert hashlib.sha256(lam.tobytes()).digest() == g[key + "_lam_sha"].tobytes(), key
            assert np.array_equal(ctx.detect(img, nf), want), key
        finally:
            ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 0)


def test_equal_strength_corners_keep_opencv_order(ctx, oracle_port):
    """The image the fuzz sweep found (tools/fuzz_parity.py, seed 23, iteration 6026; 379 x 626, quantised content): two
    corners of bit-equal strength, which cv::goodFeaturesToTrack orders by address (featureselect.cpp:55-60).  With row
    bands a last-bit difference of the directly summed box filter swapped them; the default paths (the small-count path
    for one image, full-height boxes for a batch) carry OpenCV's running column sum and keep the order."""
    import os
    from visnav_b200.keypoints import VSLAM_OPT_RESPONSE_BAND
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "tie_order_case.npz"))
    img, nf = np.ascontiguousarray(d["img"]), int(d["nf"])
    want = oracle_port.detect(img, nf)
    assert np.array_equal(ctx.detect(img, nf), want)                                    # one image: small-count path
    xy, _, _, n = ctx.detect_describe_batch(np.stack([img] * 20), nf)                   # 20 images: full-height boxes
    for k in (0, 19):
        assert n[k] == len(want) and np.array_equal(xy[k, :n[k]], want)
    ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 65535)                                      # the streaming kernel on one image
    try:
        assert np.array_equal(ctx.detect(img, nf), want)
        ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 1)                                      # bands: same corners, order may differ
        got = ctx.detect(img, nf)
        assert len(got) == len(want) and set(map(tuple, got)) == set(map(tuple, want))
    finally:
        ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 0)
