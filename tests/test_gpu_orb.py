"""GPU parity of the cv::ORB kernels (ORB half: vslam_orb_detect_compute_batch) and of
visnav::compute_bow_vector as a whole (vslam_compute_bow_vector_batch) through the C ABI, against the fixture cv2.ORB
itself produced (tests/golden/orb.npz) and against the CPU oracle on seeded images.  Keypoint set, response, size
and angle bit-exact; descriptors bit-exact up to the libm difference the test states (cosf / sinf of the angle)."""
import os

import numpy as np
import pytest

from oracle import oracle as oo
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import BowVocabulary, OrbParams, VslamError

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def canonical(kp, desc):
    order = np.lexsort((kp[:, 0], kp[:, 1], kp[:, 5]))
    return kp[order], desc[order]


def check_against(kp, desc, want_kp, want_desc, what):
    """Same keypoints (bit-exact floats, raster order per level); descriptor bits: the rotation's cos / sin come from
    CUDA's double-precision functions rounded to float, the reference's from glibc's cosf / sinf -- both correctly
    rounded except in rare cases, and a one-ulp difference only matters when a rotated sample lands within 1e-6 of a
    rounding boundary.  Allowed: at most 2 differing bits per 1000 keypoints; observed: 0."""
    assert kp.shape == want_kp.shape, (what, kp.shape, want_kp.shape)
    assert np.array_equal(kp.view(np.uint32), want_kp.view(np.uint32)), what
    bits = int(np.unpackbits(desc ^ want_desc).sum())
    assert bits <= 2 * (1 + len(kp) // 1000), (what, bits)
    return bits


def test_golden_cv2_orb(ctx):
    g = np.load(os.path.join(GOLD, "orb.npz"))
    imgs = {"synth": wl.synth_image(1234, 752, 480), "euroc": np.load(os.path.join(GOLD, "euroc_pair.npz"))["img_l"]}
    total_bits = 0
    for nf in (1500, 300):
        batch = np.stack([imgs["synth"], imgs["euroc"]])
        kp, desc, n = ctx.orb_detect_compute_batch(batch, nf)
        for i, name in enumerate(("synth", "euroc")):
            # the CUDA order is already canonical: level by level, raster order inside a level
            k, d = kp[i, :n[i]], desc[i, :n[i]]
            ck, cd = canonical(k, d)
            assert np.array_equal(k, ck)
            total_bits += check_against(k, d, g["%s_%d_kp" % (name, nf)], g["%s_%d_desc" % (name, nf)], (name, nf))
    assert total_bits == 0   # observed; check_against states what would be tolerated


@pytest.mark.parametrize("case", range(6))
def test_golden_cv2_orb_other_parameters(ctx, case):
    """cv2.ORB's own output at other scale factors, level counts, FAST and edge thresholds (orb_params.npz)."""
    g = np.load(os.path.join(GOLD, "orb_params.npz"))
    seed, w, h, nf, sf, nl, et, ft = g["cases"][case]
    img = wl.synth_image(int(seed), int(w), int(h))
    kp, desc, n = ctx.orb_detect_compute_batch(img[None], params=OrbParams(int(nf), float(sf), int(nl), int(et), int(ft)))
    check_against(kp[0, :n[0]], desc[0, :n[0]], g["kp%d" % case], g["desc%d" % case], ("case", case))


@pytest.mark.parametrize("w,h,nf,count", [(752, 480, 1500, 5), (640, 480, 700, 3), (321, 243, 200, 4), (1000, 700, 2000, 2),
                                          (130, 100, 500, 3), (60, 50, 100, 2)])
def test_against_oracle(ctx, w, h, nf, count):
    imgs = np.stack([wl.synth_image(100 + 7 * i + w, w, h) for i in range(count)])
    kp, desc, n = ctx.orb_detect_compute_batch(imgs, nf)
    for i in range(count):
        okp, odesc = oo.orb_detect_compute(imgs[i], nf)
        check_against(kp[i, :n[i]], desc[i, :n[i]], okp, odesc, (w, h, nf, i))


def test_parameters_and_degenerate_images(ctx):
    flat = np.full((2, 200, 300), 90, np.uint8)               # no corners at all
    kp, desc, n = ctx.orb_detect_compute_batch(flat, 500)
    assert list(n) == [0, 0]
    rng = np.random.default_rng(4)
    noise = rng.integers(0, 256, (1, 240, 320), dtype=np.uint8)   # thousands of FAST corners, heavy score ties
    for prm in (OrbParams(400, 1.2, 8, 19, 20), OrbParams(400, 1.5, 4, 31, 10), OrbParams(0, 1.2, 8, 19, 20),
                OrbParams(100000, 1.2, 3, 19, 40)):
        kp, desc, n = ctx.orb_detect_compute_batch(noise, params=prm, cap=8192)
        okp, odesc = oo.orb_detect_compute(noise[0], prm.nfeatures, prm.nlevels, prm.scale_factor, prm.edge_threshold,
                                           prm.fast_threshold)
        check_against(kp[0, :n[0]], desc[0, :n[0]], okp, odesc, (prm.nfeatures, prm.nlevels))
    with pytest.raises(VslamError):                            # ties make n exceed a tight capacity
        ctx.orb_detect_compute_batch(noise, 400, cap=100)
    with pytest.raises(VslamError):
        ctx.orb_detect_compute_batch(noise, params=OrbParams(400, 1.2, 9, 19, 20))
    with pytest.raises(VslamError):                            # the sampling pattern reaches 18 px
        ctx.orb_detect_compute_batch(noise, params=OrbParams(400, 1.2, 8, 18, 20))


def test_compute_bow_vector_is_orb_then_transform(ctx):
    """visnav::compute_bow_vector: the fused call equals cv::ORB followed by the vocabulary transform -- against the
    CPU oracles of both halves, BowVector doubles bit for bit."""
    voc = wl.synth_vocabulary(10, 4, 21, prune=0.05, stop=0.05)
    o = oo.BowVocab.from_arrays(10, 4, 0, 0, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    v = BowVocabulary.from_arrays(ctx, 10, 4, 0, 0, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    imgs = np.stack([wl.synth_image(900 + i, 752, 480) for i in range(4)])
    vecs, extra = v.compute_bow_vector_batch(imgs, 1500, levelsup=2)
    for i in range(4):
        okp, odesc = oo.orb_detect_compute(imgs[i], 1500)
        n = extra["n_kp"][i]
        bits = check_against(extra["kp"][i, :n], extra["desc"][i, :n], okp, odesc, i)
        ids, vals, fw, fn, fwt = o.transform(odesc, 2)
        if bits == 0:
            assert np.array_equal(vecs[i][0], ids) and np.array_equal(vecs[i][1].view(np.uint64), vals.view(np.uint64))
            assert np.array_equal(extra["feat_node"][i, :n], fn)
            assert np.array_equal(extra["feat_weight"][i, :n].view(np.uint64), fwt.view(np.uint64))
        assert abs(vecs[i][1].sum() - 1.0) < 1e-12
    # the same scene scores higher against itself (shifted) than against another scene
    shifted = np.roll(imgs[0], -12, axis=1)[None]
    (vs,), _ = v.compute_bow_vector_batch(shifted, 1500)
    assert v.score(vecs[0], vs) > 1.3 * v.score(vecs[0], vecs[1])
    v.close()
    o.close()


def test_row_pitch_and_image_stride_are_honoured(ctx):
    """Images handed over with padding after every row and between images (pitch > w, stride > pitch * h)."""
    import ctypes as C
    w, h, pitch, count, nf, cap = 333, 250, 360, 3, 400, 1024
    stride = pitch * h + 128
    imgs = np.stack([wl.synth_image(40 + i, w, h) for i in range(count)])
    buf = np.full(stride * count, 255, np.uint8)
    for i in range(count):
        buf[i * stride:i * stride + pitch * h].reshape(h, pitch)[:, :w] = imgs[i]
    kp = np.zeros((count, cap, 6), np.float32)
    desc = np.zeros((count, cap, 32), np.uint8)
    n = np.zeros(count, np.int32)
    prm = OrbParams(nf)
    P = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    ctx._check(ctx.lib.vslam_orb_detect_compute_batch(ctx.h, P(buf), count, w, h, pitch, stride, C.byref(prm), P(kp), P(desc),
                                                      P(n), cap))
    for i in range(count):
        okp, odesc = oo.orb_detect_compute(imgs[i], nf)
        check_against(kp[i, :n[i]], desc[i, :n[i]], okp, odesc, i)
