"""GPU parity of detect / angles / describe and the batched front ends against the CPU oracle.
Corners: ordered equality (integer coordinates).  Response map: bit-exact.  Angles: <= 1e-5 rad
(north_star tolerance; fp64 atan2 of CUDA vs glibc differ by ulps).  Descriptors, given identical
corners and angles: bit-exact."""
import numpy as np
import pytest

import visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import (KeypointsData, VSLAM_OPT_FP_CONTRACT, VSLAM_OPT_RESPONSE_BAND,
                                   VSLAM_OPT_RESPONSE_MULTI, VSLAM_OPT_TWO_STREAMS, StereoPlan)

pytestmark = pytest.mark.gpu

ANGLE_TOL = 1e-5

SIZES = [(752, 480), (640, 481), (121, 97), (39, 39), (250, 64), (1920, 1080)]


def _image(w, h, seed):
    return wl.synth_image(seed, w, h)


# The 3x3 box filter behind the response: OpenCV keeps a RUNNING column sum in double (box_filter.simd.hpp:206-262).
# A warp that walks a whole image column reproduces it bit for bit (boxes as tall as the image: the default).  Row-banded
# launches (VSLAM_OPT_RESPONSE_BAND = 1 or >= 8: an opt-in for latency) sum the three rows directly; the two differ only
# where a double addition was inexact.  Stated tolerance of the banded mode against OpenCV's order on these images: at
# most MAX_DIFF_PER_MPX pixels per million differ, each by at most MAX_DIFF_ULP float ulps, below the candidate threshold
# (0.01 x the image maximum), and the corner list is identical.  (In general a differing pixel can lie above the
# threshold and swap two corners of equal strength -- tools/fuzz_parity.py found one such image in 19 000 -- which is
# why bands are not the default.)
MAX_DIFF_PER_MPX = 16
MAX_DIFF_ULP = 16


def _default_band_is_full_height(h):
    return True          # boxes span the image height unless VSLAM_OPT_RESPONSE_BAND asks for bands


def _check_response(oracle_port, img, got, full_height):
    want = oracle_port.response(img)                      # OpenCV's running column sum
    assert got.shape == want.shape
    if full_height:
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), float((got != want).mean())
        return 0
    oracle_port.set_box_direct(1)
    try:
        direct = oracle_port.response(img)
    finally:
        oracle_port.set_box_direct(0)
    assert np.array_equal(got.view(np.uint32), direct.view(np.uint32)), float((got != direct).mean())
    diff = got.view(np.uint32) != want.view(np.uint32)
    n = int(diff.sum())
    assert n <= max(2, MAX_DIFF_PER_MPX * got.size / 1e6), n
    if n:
        ulp = np.abs(got.view(np.int32)[diff].astype(np.int64) - want.view(np.int32)[diff].astype(np.int64))
        assert ulp.max() <= MAX_DIFF_ULP, int(ulp.max())
        assert got[diff].max() < 0.01 * want.max() and want[diff].max() < 0.01 * want.max()
    return n


@pytest.mark.parametrize("w,h", SIZES)
def test_response_bit_exact(ctx, oracle_port, w, h):
    img = _image(w, h, 7 + w)
    _check_response(oracle_port, img, ctx.response(img), _default_band_is_full_height(h))


@pytest.mark.parametrize("band", [8, 26, 29, 61, 100, 65535])
@pytest.mark.parametrize("w,h", [(752, 480), (121, 97), (333, 131)])
def test_response_box_heights(ctx, oracle_port, w, h, band):
    """The response kernel streams a box through a two-chunk (2 x 32 rows) shared-memory ring; the box height is chosen
    from the batch size (whole image height for a batch that fills the SMs).  Chunk boundaries at all phases of the row
    pipeline, reflected rows read from the previous chunk, one chunk only.  Banded heights give the direct-sum bits,
    the full height gives OpenCV's bits (see _check_response); the corners are the oracle's at every height."""
    img = _image(w, h, 5 + w + band)
    want_xy = oracle_port.detect(img, 1500)
    ctx.set_option(VSLAM_OPT_RESPONSE_BAND, band)
    try:
        got_map, got_xy = ctx.response(img), ctx.detect(img, 1500)
    finally:
        ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 0)
    _check_response(oracle_port, img, got_map, band >= h)
    assert got_xy.shape == want_xy.shape and np.array_equal(got_xy, want_xy)


def test_response_running_sum_matters(ctx, oracle_port):
    """The difference the full-height mode exists for is real: over a set of 1080p images the direct sum and OpenCV's
    running sum disagree somewhere, the full-height kernel follows OpenCV on every pixel, the banded kernel stays inside
    the stated tolerance and yields the same corners."""
    n_diff = 0
    for seed in range(4):
        img = _image(1920, 1080, 300 + seed)
        want_xy = oracle_port.detect(img, 4000)
        ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 1)      # bands chosen by the host (latency mode)
        try:
            n_diff += _check_response(oracle_port, img, ctx.response(img), False)
            assert np.array_equal(ctx.detect(img, 4000), want_xy)
        finally:
            ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 0)  # default: full-height boxes
        _check_response(oracle_port, img, ctx.response(img), True)
        assert np.array_equal(ctx.detect(img, 4000), want_xy)
    assert n_diff > 0, "direct and running box sums never differed on this set: the test images no longer probe it"


@pytest.mark.parametrize("w,h", SIZES)
@pytest.mark.parametrize("nf", [1500, 100, 0])
def test_detect_ordered_equal(ctx, oracle_port, w, h, nf):
    img = _image(w, h, 11 + h)
    got = ctx.detect(img, nf)
    want = oracle_port.detect(img, nf)
    assert got.shape == want.shape and np.array_equal(got, want)


@pytest.mark.parametrize("w,h", [(752, 480), (640, 200), (296, 130), (130, 70), (121, 97), (177, 64)])
def test_narrow_last_strip_shared_by_images(ctx, oracle_port, w, h):
    """The narrow last column strip of two or three images in one warp (response_multi_kernel; 752 = 6 x 120 + 32 gives
    three groups of ten lanes, 640 two of twelve, 296 two of sixteen, 130 / 121 three of five / three; 177 has a strip of
    57 columns and keeps one warp per strip): corners identical to one warp per strip for every image count modulo the
    group count, in row bands and in full-height boxes, and equal to the oracle's."""
    imgs = np.stack([_image(w, h, 4000 + 13 * k + w) for k in range(7)])
    want = [oracle_port.detect(im, 400) for im in imgs]
    for band in (1, 0):      # bands chosen by the host (opt-in latency mode), full-height boxes (default)
        ctx.set_option(VSLAM_OPT_RESPONSE_BAND, band)
        try:
            for count in (2, 3, 4, 7):
                ctx.set_option(VSLAM_OPT_RESPONSE_MULTI, 2)
                a_xy, _, a_desc, a_n = ctx.detect_describe_batch(imgs[:count], 400)
                ctx.set_option(VSLAM_OPT_RESPONSE_MULTI, 0)
                b_xy, _, b_desc, b_n = ctx.detect_describe_batch(imgs[:count], 400)
                assert np.array_equal(a_n, b_n), (band, count)
                for k in range(count):
                    n = a_n[k]
                    assert np.array_equal(a_xy[k, :n], b_xy[k, :n]) and np.array_equal(a_desc[k, :n], b_desc[k, :n])
                    assert n == len(want[k]) and np.array_equal(a_xy[k, :n], want[k]), (band, count, k)
        finally:
            ctx.set_option(VSLAM_OPT_RESPONSE_MULTI, 1)
            ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 0)


def test_response_ring_bottom_alignments(ctx, oracle_port):
    """Every position of the last image row inside a 32-row chunk: the reflected rows h-2, h-3 that rows h, h+1 read
    sit in the same chunk or in the previous one (the other half of the ring)."""
    ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 65535)
    try:
        for h in list(range(56, 72)) + list(range(88, 100)):
            img = _image(150, h, 900 + h)
            _check_response(oracle_port, img, ctx.response(img), True)
            assert np.array_equal(ctx.detect(img, 500), oracle_port.detect(img, 500)), h
    finally:
        ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 0)


def test_response_full_height_1080p(ctx, oracle_port):
    """A 1080-row box: 34 chunks through the ring, 16 boxes across (config 5 geometry)."""
    img = _image(1920, 1080, 123)
    ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 65535)
    try:
        got_map, got_xy = ctx.response(img), ctx.detect(img, 4000)
    finally:
        ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 0)
    want_map = oracle_port.response(img)
    assert np.array_equal(got_map.view(np.uint32), want_map.view(np.uint32))
    assert np.array_equal(got_xy, oracle_port.detect(img, 4000))


def test_detect_strongest_subset_paths(ctx, oracle_port):
    """1080p images have ~24 k candidates above the threshold: the select kernel decides on the strongest
    8192 in shared memory when they yield num_features corners (4000) and must fall back to the complete
    set on the global workspace when they do not (9000)."""
    img = _image(1920, 1080, 77)
    for nf in (4000, 9000):
        got = ctx.detect(img, nf)
        want = oracle_port.detect(img, nf)
        assert got.shape == want.shape and np.array_equal(got, want), nf
    rng = np.random.default_rng(8)
    noise = rng.integers(0, 256, (600, 800), dtype=np.uint8)   # very dense candidates, many equal responses
    for nf in (300, 3000, 6000):
        assert np.array_equal(ctx.detect(noise, nf), oracle_port.detect(noise, nf)), nf


def test_detect_4k_image(ctx, oracle_port):
    """3840 x 2160: 129 600 minDistance cells, more than the select kernel's shared-memory table holds -- the whole
    selection runs on the global workspace.  goodFeaturesToTrack has no image-size limit."""
    img = _image(3840, 2160, 4)
    for nf in (1500, 20000):
        got, want = ctx.detect(img, nf), oracle_port.detect(img, nf)
        assert got.shape == want.shape and np.array_equal(got, want), nf


def test_detect_special_images(ctx, oracle_port):
    flat = np.full((120, 160), 77, np.uint8)              # constant: no response anywhere
    assert len(ctx.detect(flat, 100)) == 0 == len(oracle_port.detect(flat, 100))
    chk = ((np.indices((200, 240)).sum(0) // 8) % 2 * 255).astype(np.uint8)   # plateaus and exact ties
    assert np.array_equal(ctx.detect(chk, 0), oracle_port.detect(chk, 0))
    rng = np.random.default_rng(3)
    noise = rng.integers(0, 256, (300, 333), dtype=np.uint8)                 # dense candidates
    assert np.array_equal(ctx.detect(noise, 0), oracle_port.detect(noise, 0))
    tiny = rng.integers(0, 256, (20, 30), dtype=np.uint8)                    # nothing survives the 19 px border
    assert len(ctx.detect(tiny, 10)) == 0


def test_detect_pitched_input(ctx, oracle_port):
    big = _image(800, 500, 5)
    view = big[10:490, 20:772]            # 752 x 480 view, pitch 800
    assert np.array_equal(ctx.detect(view, 1500), oracle_port.detect(np.ascontiguousarray(view), 1500))


@pytest.mark.parametrize("rotate", [True, False])
def test_angles_and_descriptors(ctx, oracle_port, oracle_ref, rotate):
    img = _image(752, 480, 21)
    xy = oracle_port.detect(img, 1500)
    ang_o = oracle_port.angles(img, xy, rotate)
    ang_g = ctx.angles(img, xy, rotate)
    assert np.abs(ang_g - ang_o).max() <= ANGLE_TOL
    for contract in (1, 0):
        ctx.set_option(VSLAM_OPT_FP_CONTRACT, contract)
        got = ctx.describe(img, xy, ang_o)          # identical keypoints AND angles -> bit-exact
        want = oracle_port.describe(img, xy, ang_o, contract=bool(contract))
        assert np.array_equal(got, want), int((got != want).any(1).sum())
    ctx.set_option(VSLAM_OPT_FP_CONTRACT, 1)
    if oracle_ref is not None:
        assert np.array_equal(ctx.describe(img, xy, ang_o), oracle_ref.describe(img, xy, ang_o))
        assert np.abs(ctx.angles(img, xy, rotate) - oracle_ref.angles(img, xy, rotate)).max() <= ANGLE_TOL


def test_describe_border_keypoints(ctx, oracle_port):
    """Keypoints nearer than 19 px to the edge (never produced by detect): edge replication, like the oracle."""
    img = _image(160, 120, 2)
    xy = np.array([[0, 0], [159, 119], [5, 60], [80, 3], [150, 110], [80, 60]], np.float64)
    ang = oracle_port.angles(img, xy, True)
    assert np.abs(ctx.angles(img, xy, True) - ang).max() <= ANGLE_TOL
    assert np.array_equal(ctx.describe(img, xy, ang), oracle_port.describe(img, xy, ang))


def test_reference_api_mirror(ctx, oracle_port):
    """The reference's call sequence (slam.cpp:1125-1143) through the mirrored free functions."""
    left, right = wl.synth_stereo_pair(3)
    kdl, kdr = KeypointsData(), KeypointsData()
    visnav_b200.detectKeypointsAndDescriptors(left, kdl, 1500, True)
    kd2 = KeypointsData()
    visnav_b200.detectKeypoints(right, kd2, 1500)
    visnav_b200.computeAngles(right, kd2, True)
    visnav_b200.computeDescriptors(right, kd2)
    visnav_b200.detectKeypointsAndDescriptors(right, kdr, 1500, True)
    assert np.array_equal(kd2.corners, kdr.corners) and np.array_equal(kd2.corner_descriptors, kdr.corner_descriptors)
    matches = [(-1, -1)]
    visnav_b200.matchDescriptors(kdl.corner_descriptors, kdr.corner_descriptors, matches, 70, 1.2)
    xl, al, dl = oracle_port.detect_describe(left, 1500)
    xr, ar, dr = oracle_port.detect_describe(right, 1500)
    assert np.array_equal(kdl.corners, xl) and np.array_equal(kdr.corners, xr)
    assert np.abs(kdl.corner_angles - al).max() <= ANGLE_TOL
    assert np.array_equal(kdl.corner_descriptors, dl) and np.array_equal(kdr.corner_descriptors, dr)
    want = oracle_port.match(dl, dr, 70, 1.2)
    assert matches == [(int(a), int(b)) for a, b in want] and len(matches) > 900


def test_stereo_batch_matches_oracle(ctx, oracle_port):
    n = 5
    imgs = wl.synth_stereo_batch(n, workers=1)
    out = ctx.stereo_batch(imgs, 1500, True, 70, 1.2)
    xy2, ang2, desc2, nkp2 = ctx.detect_describe_batch(imgs.reshape(2 * n, 480, 752), 1500, True)
    for k in range(n):
        ds = []
        for side in range(2):
            i = 2 * k + side
            x, a, d = oracle_port.detect_describe(imgs[k, side], 1500)
            c = out["n_kp"][i]
            assert c == len(x) == nkp2[i]
            assert np.array_equal(out["xy"][i, :c], x) and np.array_equal(xy2[i, :c], x)
            assert np.abs(out["angles"][i, :c] - a).max() <= ANGLE_TOL
            assert np.array_equal(out["desc"][i, :c], d) and np.array_equal(desc2[i, :c], d)
            ds.append(d)
        want = oracle_port.match(ds[0], ds[1], 70, 1.2)
        m = out["n_matches"][k]
        assert m == len(want) and np.array_equal(out["matches"][k, :m], want)


def test_stereo_batch_chunked_pipeline(ctx):
    """More pairs than one pipeline chunk (128): results must not depend on the chunking."""
    base = wl.synth_stereo_batch(3, workers=1)
    n = 150
    imgs = np.ascontiguousarray(base[np.arange(n) % 3])
    out = ctx.stereo_batch(imgs, 1500)
    for k in range(n):
        r = k % 3
        for side in range(2):
            c = out["n_kp"][2 * k + side]
            assert c == out["n_kp"][2 * r + side]
            assert np.array_equal(out["desc"][2 * k + side, :c], out["desc"][2 * r + side, :c])
        m = out["n_matches"][k]
        assert m == out["n_matches"][r] and np.array_equal(out["matches"][k, :m], out["matches"][r, :m])


@pytest.mark.parametrize("n", [1, 3, 40])
def test_stereo_batch_small_counts_split_matcher(ctx, n):
    """Small batches take the column-split matcher; the lists equal those of the one-CTA-per-job matcher."""
    from visnav_b200.keypoints import VSLAM_OPT_MATCH_SPLIT_SMALL
    base = wl.synth_stereo_batch(3, workers=1)
    imgs = np.ascontiguousarray(base[np.arange(n) % 3])
    a = ctx.stereo_batch(imgs, 1500)
    ctx.set_option(VSLAM_OPT_MATCH_SPLIT_SMALL, 0)
    try:
        b = ctx.stereo_batch(imgs, 1500)
    finally:
        ctx.set_option(VSLAM_OPT_MATCH_SPLIT_SMALL, 1)
    assert np.array_equal(a["n_matches"], b["n_matches"]) and (a["n_matches"] > 0).all()
    for k in range(n):
        m = a["n_matches"][k]
        assert np.array_equal(a["matches"][k, :m], b["matches"][k, :m])


def test_stereo_batch_single_pair_graph_replay(ctx, oracle_port):
    """Config 1 the way slam.cpp:1125-1143 issues it: one pair per call, again and again.  From the third call with the
    same arguments the pipeline plan replays its launches as one CUDA graph and the results travel in one copy; every
    call returns the oracle's result, also when the arguments or the image change in between, and with VSLAM_OPT_GRAPH off."""
    from visnav_b200.keypoints import VSLAM_OPT_GRAPH
    pairs = wl.synth_stereo_batch(3, workers=1)
    want = {}
    for k in range(3):
        dl = oracle_port.detect_describe(pairs[k, 0], 1500)
        dr = oracle_port.detect_describe(pairs[k, 1], 1500)
        want[k] = (dl, dr, oracle_port.match(dl[2], dr[2], 70, 1.2), oracle_port.match(dl[2], dr[2], 50, 1.5))

    def check(out, k, m_idx):
        for side, d in enumerate(want[k][:2]):
            n = out["n_kp"][side]
            assert n == len(d[0]) and np.array_equal(out["xy"][side, :n], d[0]) and np.array_equal(out["desc"][side, :n], d[2])
            assert np.abs(out["angles"][side, :n] - d[1]).max() <= ANGLE_TOL
        m = out["n_matches"][0]
        assert m == len(want[k][m_idx]) and np.array_equal(out["matches"][0, :m], want[k][m_idx])

    for graph in (1, 0, 1):
        ctx.set_option(VSLAM_OPT_GRAPH, graph)
        try:
            for rep in range(5):                      # same arguments: plain, captured, replayed, replayed, replayed
                k = rep % 3                           # ... on changing images
                check(ctx.stereo_batch(pairs[k:k + 1], 1500, True, 70, 1.2), k, 2)
            for rep in range(4):                      # other arguments: the graph is rebuilt
                check(ctx.stereo_batch(pairs[1:2], 1500, True, 50, 1.5), 1, 3)
            check(ctx.stereo_batch(pairs[2:3], 1500, True, 70, 1.2), 2, 2)
        finally:
            ctx.set_option(VSLAM_OPT_GRAPH, 1)


def test_stereo_plan_two_streams_equal_one_stream(ctx, oracle_port):
    """Large resident batches run as two halves on two streams (vslam_stereo_plan_run, VSLAM_OPT_TWO_STREAMS): every
    output equal to the one-stream run, for an odd pair count (halves of 296 and 297 pairs), and equal to the oracle on
    a pair of each half."""
    n = 593
    base = wl.synth_stereo_batch(12, 320, 240, workers=1)
    imgs = np.ascontiguousarray(base[np.arange(n) % 12])
    plan = StereoPlan(ctx, 320, 240, n, 600)
    try:
        plan.upload(imgs)
        outs = []
        for two in (1, 0, 1):
            ctx.set_option(VSLAM_OPT_TWO_STREAMS, two)
            plan.run(n, 600, True, 70, 1.2)
            outs.append(plan.download(n))
        a, b, c = outs
        for key in ("n_kp", "n_matches"):
            assert np.array_equal(a[key], b[key]) and np.array_equal(a[key], c[key]), key
        for k in range(n):
            m = a["n_matches"][k]
            assert np.array_equal(a["matches"][k, :m], b["matches"][k, :m])
            for side in (2 * k, 2 * k + 1):
                q = a["n_kp"][side]
                assert np.array_equal(a["xy"][side, :q], b["xy"][side, :q]) and np.array_equal(a["desc"][side, :q], b["desc"][side, :q])
                assert np.array_equal(a["angles"][side, :q], b["angles"][side, :q])
        for k in (5, 400):
            dl, dr = [oracle_port.detect_describe(imgs[k, s], 600) for s in (0, 1)]
            assert np.array_equal(a["xy"][2 * k, :a["n_kp"][2 * k]], dl[0]) and np.array_equal(a["desc"][2 * k + 1, :a["n_kp"][2 * k + 1]], dr[2])
            m = a["n_matches"][k]
            assert np.array_equal(a["matches"][k, :m], oracle_port.match(dl[2], dr[2], 70, 1.2))
    finally:
        ctx.set_option(VSLAM_OPT_TWO_STREAMS, 1)
        plan.close()
