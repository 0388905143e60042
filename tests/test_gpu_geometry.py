"""CUDA camera models / findInliersEssential / project_landmarks (csrc/geometry.cu through the C ABI) against the
CPU oracle, the committed reference fixture and -- where oracle/_ref travelled -- the reference's own code
(camera_models.h, matching_utils.h:63-88, vo_utils.h:48-81).

fp64 tolerances: project <= 1e-9 px, unproject <= 1e-12, p0^T E p1 <= 1e-12.  Inlier and visibility lists must be
EQUAL (ordered) whenever no tested quantity lies within the tolerance of its threshold; borderline items (counted
and bounded) may differ."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
MODELS = ["pinhole", "eucm", "ds", "kb4"]
TOL_PX, TOL_UNIT = 1e-9, 1e-12


def _cam(m, P, w=752, h=480):
    from visnav_b200.keypoints import Camera
    return Camera.from_data(m, P, w, h)


@pytest.mark.parametrize("m", MODELS)
def test_camera_models(ctx, m):
    import oracle.oracle as oo
    from visnav_b200 import workloads as wl
    geo = oo.load_geometry()
    P = wl.CAMERAS[m]
    rng = np.random.default_rng(4)
    xyz = rng.normal(size=(20000, 3)) * [3, 2, 3] + [0, 0, 5]
    xyz[0] = [0, 0, 2.0]
    uv = rng.uniform([0, 0], [752, 480], size=(20000, 2))
    uv[0] = P[2:4]
    assert np.abs(ctx.cam_project(_cam(m, P), xyz) - geo.project(m, P, xyz)).max() <= TOL_PX
    got = ctx.cam_unproject(_cam(m, P), uv)
    assert np.abs(got - geo.unproject(m, P, uv)).max() <= TOL_UNIT
    assert np.abs(ctx.cam_project(_cam(m, P), got) - uv).max() <= 1e-7       # round trip on the device
    g = np.load(os.path.join(GOLD, "geometry.npz"))
    assert np.abs(ctx.cam_project(_cam(m, g[m + "_param"]), g[m + "_xyz"]) - g[m + "_project"]).max() <= TOL_PX
    assert np.abs(ctx.cam_unproject(_cam(m, g[m + "_param"]), g[m + "_uv"]) - g[m + "_unproject"]).max() <= TOL_UNIT


@pytest.mark.parametrize("m", MODELS)
@pytest.mark.parametrize("n", [40, 1300, 5000])
def test_find_inliers_essential(ctx, m, n):
    import oracle.oracle as oo
    from visnav_b200 import workloads as wl
    geo, ref = oo.load_geometry(), oo.load_geometry_ref()
    P = wl.CAMERAS[m]
    xy1, xy2, mt, pose, E = wl.epipolar_scene(m, n, seed=n)
    for thr in (5e-3, 1e-3, 0.0, 10.0):
        want, werr = geo.find_inliers_essential(xy1, xy2, (m, P), (m, P), E, thr, mt)
        got, gerr = ctx.find_inliers_essential(xy1, xy2, _cam(m, P), _cam(m, P), E, thr, mt, want_err=True)
        assert np.abs(gerr - werr).max() <= TOL_UNIT
        if (np.abs(np.abs(werr) - thr) > TOL_UNIT).all():
            assert np.array_equal(got, want), (m, n, thr, len(got), len(want))
            if ref is not None:
                assert np.array_equal(got, ref.find_inliers_essential(xy1, xy2, (m, P), (m, P), E, thr, mt)[0])
        else:  # only borderline matches may differ
            sure = mt[np.abs(np.abs(werr) - thr) > TOL_UNIT]
            a = {tuple(r) for r in got} & {tuple(r) for r in sure}
            b = {tuple(r) for r in want} & {tuple(r) for r in sure}
            assert a == b


@pytest.mark.parametrize("m", MODELS)
def test_inliers_and_projection_reference_fixture(ctx, m):
    g = np.load(os.path.join(GOLD, "geometry.npz"))
    cam = _cam(m, g[m + "_param"])
    for thr in (5e-3, 1e-3):
        got, err = ctx.find_inliers_essential(g[m + "_xy1"], g[m + "_xy2"], cam, cam, g[m + "_E"], thr,
                                              g[m + "_matches"], want_err=True)
        assert np.abs(err - g[m + "_err"]).max() <= TOL_UNIT
        assert np.array_equal(got, g[m + "_inliers_%g" % thr])
    xy, idx = ctx.project_landmarks(g[m + "_lm_pose"], cam, g[m + "_lm_xyz"], 0.1)
    assert np.array_equal(idx, g[m + "_proj_idx"])
    assert np.abs(xy - g[m + "_proj_xy"]).max() <= TOL_PX


@pytest.mark.parametrize("m", MODELS)
@pytest.mark.parametrize("n", [1, 1000, 1025, 200000])
def test_project_landmarks(ctx, m, n):
    import oracle.oracle as oo
    from visnav_b200 import workloads as wl
    geo = oo.load_geometry()
    P = wl.CAMERAS[m]
    pose, pts = wl.landmark_cloud(min(n, 20000), seed=n)
    if n > len(pts):
        pts = np.tile(pts, (n // len(pts), 1))
    wxy, widx = geo.project_landmarks(pose, (m, P), 752, 480, pts, 0.1)
    gxy, gidx = ctx.project_landmarks(pose, _cam(m, P), pts, 0.1)
    if np.array_equal(gidx, widx):
        assert len(gxy) == 0 or np.abs(gxy - wxy).max() <= TOL_PX
    else:
        # items whose depth or pixel lies within the tolerance of a bound may flip (e.g. z == threshold exactly,
        # computed through a rotation): everything else must agree
        diff = np.setxor1d(gidx, widx)
        assert len(diff) <= max(2, n // 40), len(diff)
        both = np.intersect1d(gidx, widx)
        assert np.abs(gxy[np.isin(gidx, both)] - wxy[np.isin(widx, both)]).max() <= TOL_PX
        assert len(both) >= len(widx) - len(diff)


def test_geometry_edge_cases(ctx):
    from visnav_b200 import workloads as wl
    from visnav_b200.keypoints import Camera, VslamError
    cam = _cam("ds", wl.CAMERAS["ds"])
    e2 = np.zeros((0, 2))
    assert len(ctx.find_inliers_essential(e2, e2, cam, cam, np.eye(3), 1e-3, np.zeros((0, 2), np.int32))) == 0
    xy, idx = ctx.project_landmarks(np.array([0, 0, 0, 1.0, 0, 0, 0]), cam, np.zeros((0, 3)), 0.1)
    assert len(xy) == 0 and len(idx) == 0
    assert ctx.cam_project(cam, np.zeros((0, 3))).shape == (0, 2)
    xy1, xy2, mt, pose, E = wl.epipolar_scene("ds", 200, seed=1)
    with pytest.raises(VslamError):                                   # capacity
        ctx.find_inliers_essential(xy1, xy2, cam, cam, E, 10.0, mt, cap=3)
    bad = mt.copy()
    bad[5, 1] = len(xy2)
    with pytest.raises(VslamError):                                   # id outside the corner list
        ctx.find_inliers_essential(xy1, xy2, cam, cam, E, 1e-3, bad)
    with pytest.raises(VslamError):                                   # unknown model
        ctx.cam_project(Camera(7, 752, 480, 0), np.zeros((1, 3)))
    with pytest.raises(VslamError):
        Camera.from_data("fisheye", [0] * 8)
    # behind the camera / outside the image: nothing survives
    pose = np.array([0, 0, 0, 1.0, 0, 0, 0])
    xy, idx = ctx.project_landmarks(pose, cam, np.array([[0, 0, -1.0], [50.0, 0, 1.0], [0, 0, 0.05]]), 0.1)
    assert len(idx) == 0
    xy, idx = ctx.project_landmarks(pose, cam, np.array([[0, 0, -1.0], [0.1, 0, 1.0]]), 0.1)
    assert list(idx) == [1]


def test_stereo_batch_fused_epipolar(ctx):
    """vslam_stereo_batch_epipolar: the inlier lists equal findInliersEssential applied to the batch's own corners
    and matches (oracle), and the other outputs are unchanged by the extra stage."""
    import oracle.oracle as oo
    from visnav_b200 import workloads as wl
    from visnav_b200.keypoints import Epipolar
    geo = oo.load_geometry()
    imgs = wl.synth_stereo_batch(5)
    P = wl.CAMERAS["ds"]
    pose, E = wl.stereo_geometry(0)
    thr = 0.02
    ep = Epipolar.make(_cam("ds", P), _cam("ds", P), E, thr)
    base = ctx.stereo_batch(imgs, 1500, True, 70, 1.2)
    out = ctx.stereo_batch(imgs, 1500, True, 70, 1.2, epipolar=ep)
    assert np.array_equal(base["n_kp"], out["n_kp"]) and np.array_equal(base["n_matches"], out["n_matches"])
    total = 0
    for k in range(len(imgs)):
        nl, nr, nm = out["n_kp"][2 * k], out["n_kp"][2 * k + 1], out["n_matches"][k]
        assert np.array_equal(out["matches"][k, :nm], base["matches"][k, :nm])
        want, err = geo.find_inliers_essential(out["xy"][2 * k, :nl], out["xy"][2 * k + 1, :nr], ("ds", P), ("ds", P),
                                               E, thr, out["matches"][k, :nm])
        assert (np.abs(np.abs(err) - thr) > TOL_UNIT).all()
        ni = out["n_inliers"][k]
        assert ni == len(want) and np.array_equal(out["inliers"][k, :ni], want), k
        total += ni
    assert total > 0
