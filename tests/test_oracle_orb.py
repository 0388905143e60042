"""CPU tests of the cv::ORB oracle (ORB half): the C restatement (oracle/visnav_oracle_orb.c) against the
committed fixture tests/golden/orb.npz -- produced by cv2.ORB_create(n, 1.2, 8, 19, 0, 2, FAST_SCORE) itself
(tools/make_golden.py orb) -- and, where cv2 is importable, against cv2 live on more images and on the single stages
(resize, FAST, blur, fastAtan2).  Bit-exact: keypoint set, angle, response, size and all 256 descriptor bits."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import oracle as oo
from visnav_b200 import workloads as wl

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

try:
    import cv2
except Exception:  # pragma: no cover
    cv2 = None


def canonical(kp, desc):
    """Sorted by (octave, y, x): the order inside a level is unspecified in the reference (std::nth_element)."""
    order = np.lexsort((kp[:, 0], kp[:, 1], kp[:, 5]))
    return kp[order], desc[order]


def images():
    return {"synth": wl.synth_image(1234, 752, 480), "euroc": np.load(os.path.join(GOLD, "euroc_pair.npz"))["img_l"]}


@pytest.mark.parametrize("name", ["synth", "euroc"])
@pytest.mark.parametrize("nf", [1500, 300])
def test_port_matches_golden_cv2_orb(name, nf):
    g = np.load(os.path.join(GOLD, "orb.npz"))
    kp, desc = canonical(*oo.orb_detect_compute(images()[name], nf))
    assert np.array_equal(kp.view(np.uint32), g["%s_%d_kp" % (name, nf)].view(np.uint32))
    assert np.array_equal(desc, g["%s_%d_desc" % (name, nf)])
    assert len(kp) >= nf     # ties at the cut-off score all stay (KeyPointsFilter::retainBest)


@pytest.mark.parametrize("case", range(6))
def test_port_matches_golden_cv2_orb_other_parameters(case):
    """tests/golden/orb_params.npz (tools/make_golden.py orb_params): cv2.ORB at scale factors 1.1 ... 2.0, 1 ... 8
    levels, FAST thresholds 5 ... 40, edge thresholds 19 / 25 / 31 on odd-sized images -- away from the reference's one
    parameter set, which the entry point also accepts."""
    g = np.load(os.path.join(GOLD, "orb_params.npz"))
    seed, w, h, nf, sf, nl, et, ft = g["cases"][case]
    img = wl.synth_image(int(seed), int(w), int(h))
    kp, desc = canonical(*oo.orb_detect_compute(img, int(nf), int(nl), float(sf), int(et), int(ft)))
    assert np.array_equal(kp.view(np.uint32), g["kp%d" % case].view(np.uint32))
    assert np.array_equal(desc, g["desc%d" % case])


def test_gaussian_kernel_is_cv2s():
    g = np.load(os.path.join(GOLD, "orb.npz"))
    assert np.array_equal(oo.gaussian_kernel7().view(np.uint32), g["k7"].view(np.uint32))


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
def test_stages_against_cv2_live():
    lib = oo.load_oracle().lib
    P = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    img = wl.synth_image(7, 752, 480)
    for dw, dh in [(627, 400), (522, 333), (210, 134), (751, 479), (100, 37), (900, 500)]:
        out = np.zeros((dh, dw), np.uint8)
        lib.vso_resize_linear_exact(P(img), 752, 480, 752, P(out), dw, dh, dw)
        assert np.array_equal(out, cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR_EXACT)), (dw, dh)
    score = np.zeros((480, 752), np.uint8)
    lib.vso_fast_score_map(P(img), 752, 480, 752, 20, P(score))
    s = np.pad(score.astype(np.int32), 1)
    nb = np.maximum.reduce([s[1 + dy:481 + dy, 1 + dx:753 + dx] for dy in (-1, 0, 1) for dx in (-1, 0, 1) if (dy, dx) != (0, 0)])
    mask = (score > 0) & (score > nb)
    got = {(int(x), int(y), int(score[y, x])) for y, x in zip(*np.nonzero(mask))}
    kps = cv2.FastFeatureDetector_create(20, True).detect(img, None)
    assert got == {(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in kps} and len(got) > 1000
    k7 = oo.gaussian_kernel7()
    out = np.zeros_like(img)
    lib.vso_gauss7(P(img), 752, 480, 752, P(k7), 1, P(out), 752)
    assert np.array_equal(out, cv2.sepFilter2D(img, cv2.CV_8U, k7, k7, borderType=cv2.BORDER_REFLECT_101))
    lib.vso_fast_atan2.restype = C.c_float
    lib.vso_fast_atan2.argtypes = [C.c_float, C.c_float]
    rng = np.random.default_rng(0)
    for y, x in rng.integers(-300000, 300000, (2000, 2)).astype(np.float32):
        assert lib.vso_fast_atan2(float(y), float(x)) == cv2.fastAtan2(float(y), float(x))


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
@pytest.mark.parametrize("seed,w,h,nf", [(7, 752, 480, 1500), (21, 640, 480, 800), (33, 321, 243, 200), (5, 1000, 700, 2000)])
def test_port_matches_cv2_live(seed, w, h, nf):
    img = wl.synth_image(seed, w, h)
    kps, d = cv2.ORB_create(nf, 1.2, 8, 19, 0, 2, cv2.ORB_FAST_SCORE).detectAndCompute(img, None)
    want = np.array([[k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave] for k in kps], np.float32).reshape(-1, 6)
    want, wd = canonical(want, d if d is not None else np.zeros((0, 32), np.uint8))
    kp, desc = canonical(*oo.orb_detect_compute(img, nf))
    assert np.array_equal(kp.view(np.uint32), want.view(np.uint32)) and np.array_equal(desc, wd)
