"""The CUDA path against the committed golden vectors (cv2 / the reference's own functions)."""
import os

import numpy as np
import pytest

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
