"""CPU: the geometry restatement (oracle/visnav_oracle_geo.c) against the committed reference fixture
(tests/golden/geometry.npz, made by the reference's own camera_models.h / findInliersEssential /
project_landmarks) and, where oracle/_ref was built, against the reference's code itself.

Tolerances (fp64; the reference's -O3 build fuses multiply-adds, the restatement does not):
  project <= 1e-9 px, unproject <= 1e-12, p0^T E p1 <= 1e-12; inlier / visibility decisions equal
  except where the tested quantity lies within that tolerance of its threshold (none in the fixture)."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
MODELS = ["pinhole", "eucm", "ds", "kb4"]
TOL_PX, TOL_UNIT = 1e-9, 1e-12


@pytest.fixture(scope="module")
def geo():
    import oracle.oracle as oo
    return oo.load_geometry()


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "geometry.npz"))


@pytest.mark.parametrize("m", MODELS)
def test_camera_models_vs_fixture(geo, gold, m):
    P = gold[m + "_param"]
    assert np.abs(geo.project(m, P, gold[m + "_xyz"]) - gold[m + "_project"]).max() <= TOL_PX
    assert np.abs(geo.unproject(m, P, gold[m + "_uv"]) - gold[m + "_unproject"]).max() <= TOL_UNIT
    # project(unproject(uv)) == uv: the models are inverses of each other on the image
    uv = gold[m + "_uv"]
    assert np.abs(geo.project(m, P, geo.unproject(m, P, uv)) - uv).max() <= 1e-7


@pytest.mark.parametrize("m", MODELS)
def test_inliers_and_projection_vs_fixture(geo, gold, m):
    P = gold[m + "_param"]
    cam = (m, P)
    for thr in (5e-3, 1e-3):
        inl, err = geo.find_inliers_essential(gold[m + "_xy1"], gold[m + "_xy2"], cam, cam, gold[m + "_E"], thr,
                                              gold[m + "_matches"])
        assert np.abs(err - gold[m + "_err"]).max() <= TOL_UNIT
        assert np.abs(np.abs(gold[m + "_err"]) - thr).min() > TOL_UNIT       # no borderline case in the fixture
        assert np.array_equal(inl, gold[m + "_inliers_%g" % thr])
    xy, idx = geo.project_landmarks(gold[m + "_lm_pose"], cam, 752, 480, gold[m + "_lm_xyz"], 0.1)
    assert np.array_equal(idx, gold[m + "_proj_idx"])
    assert np.abs(xy - gold[m + "_proj_xy"]).max() <= TOL_PX


@pytest.mark.parametrize("m", MODELS)
def test_port_vs_reference_library(geo, m):
    import oracle.oracle as oo
    from visnav_b200 import workloads as wl
    ref = oo.load_geometry_ref()
    if ref is None:
        pytest.skip("oracle/_ref not built here")
    P = wl.CAMERAS[m]
    for seed in range(3):
        xy1, xy2, mt, pose, E = wl.epipolar_scene(m, 900, seed=seed)
        assert np.abs(ref.compute_essential(pose) - E).max() <= 1e-15
        a, ea = geo.find_inliers_essential(xy1, xy2, (m, P), (m, P), E, 5e-3, mt)
        b, eb = ref.find_inliers_essential(xy1, xy2, (m, P), (m, P), E, 5e-3, mt)
        assert np.abs(ea - eb).max() <= TOL_UNIT
        sure = np.abs(np.abs(eb) - 5e-3) > TOL_UNIT
        assert sure.all() and np.array_equal(a, b)
        pose_w, pts = wl.landmark_cloud(3000, seed)
        xa, ia = geo.project_landmarks(pose_w, (m, P), 752, 480, pts, 0.1)
        xb, ib = ref.project_landmarks(pose_w, (m, P), 752, 480, pts, 0.1)
        assert np.array_equal(ia, ib) and np.abs(xa - xb).max() <= TOL_PX


def test_empty_inputs(geo):
    from visnav_b200 import workloads as wl
    P = wl.CAMERAS["ds"]
    inl, err = geo.find_inliers_essential(np.zeros((0, 2)), np.zeros((0, 2)), ("ds", P), ("ds", P), np.eye(3), 1e-3,
                                          np.zeros((0, 2), np.int32))
    assert len(inl) == 0 and len(err) == 0
    xy, idx = geo.project_landmarks(np.array([0, 0, 0, 1.0, 0, 0, 0]), ("ds", P), 752, 480, np.zeros((0, 3)), 0.1)
    assert len(xy) == 0 and len(idx) == 0
