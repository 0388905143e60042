"""ctypes bindings for oracle/liboracle.so (C restatement, "port") and
oracle/_ref/libvisnav_ref.so (the reference's own functions, "reference").

TEST INFRASTRUCTURE ONLY -- see oracle/visnav_oracle.c.  Both libraries expose the
same operations; `Oracle` hides the small signature differences so tests can run the
same checks against either.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_u8p = C.POINTER(C.c_uint8)
_f64p = C.POINTER(C.c_double)
_f32p = C.POINTER(C.c_float)
_u32p = C.POINTER(C.c_uint32)
_i32p = C.POINTER(C.c_int32)


def build(quiet=True):
    """make -C oracle (liboracle.so always; _ref only where /root/reference exists)."""
    subprocess.run(["make", "-C", _HERE], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


def _p(a, t):
    return a.ctypes.data_as(t)


def _img(img):
    img = np.asarray(img)
    assert img.dtype == np.uint8 and img.ndim == 2
    if img.strides[1] != 1:
        img = np.ascontiguousarray(img)
    return img, img.shape[1], img.shape[0], img.strides[0]


class Oracle:
    """kind == "port": C restatement.  kind == "reference": reference's own code."""

    def __init__(self, lib, kind):
        self.lib, self.kind = lib, kind
        pre = "vso_" if kind == "port" else "ref_"
        self._pre = pre
        f = getattr(lib, pre + "detect")
        f.restype = C.c_int
        f.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, C.c_int, _f64p, C.c_int]
        f = getattr(lib, pre + "angles")
        f.restype = None
        f.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, _f64p, C.c_int, C.c_int, _f64p]
        f = getattr(lib, pre + "describe")
        f.restype = None
        if kind == "port":
            f.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, _f64p, _f64p, C.c_int, C.c_int, _u32p]
        else:
            f.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, _f64p, _f64p, C.c_int, _u32p]
        f = getattr(lib, pre + "match")
        f.restype = C.c_int
        f.argtypes = [_u32p, C.c_int, _u32p, C.c_int, C.c_int, C.c_double, _i32p, C.c_int]
        f = getattr(lib, pre + "detect_describe")
        f.restype = C.c_int
        if kind == "port":
            f.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f64p, _f64p, _u32p, C.c_int]
        else:
            f.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f64p, _f64p, _u32p, C.c_int]
        if kind == "port":
            lib.vso_response.restype = C.c_int
            lib.vso_response.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, _f32p]
            lib.vso_gftt.restype = C.c_int
            lib.vso_gftt.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, _f32p, C.c_int]
            lib.vso_set_harris.restype = None
            lib.vso_set_harris.argtypes = [C.c_int, C.c_double]
            lib.vso_set_box_direct.restype = None
            lib.vso_set_box_direct.argtypes = [C.c_int]
            lib.vso_best2.restype = None
            lib.vso_best2.argtypes = [_u32p, C.c_int, _u32p, C.c_int, _i32p]

        self.gftt_impl = "port (C restatement vso_gftt)"
        if kind == "reference" and hasattr(lib, "ref_gftt_impl"):
            lib.ref_gftt_impl.restype = C.c_char_p
            self.gftt_impl = lib.ref_gftt_impl().decode()
        self.has_opencv = kind == "reference" and hasattr(lib, "ref_gftt")
        if self.has_opencv:
            lib.ref_set_use_optimized.restype = None
            lib.ref_set_use_optimized.argtypes = [C.c_int]
            lib.ref_set_num_threads.restype = None
            lib.ref_set_num_threads.argtypes = [C.c_int]
            lib.ref_gftt.restype = C.c_int
            lib.ref_gftt.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_int, C.c_int,
                                     C.c_double, _f32p, C.c_int]
            lib.ref_corner_response.restype = None
            lib.ref_corner_response.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, _f32p]

    # --- the reference's vendored OpenCV (only in oracle/_ref/libvisnav_ref.so built with it) -----------
    def set_use_optimized(self, on):
        """cv::setUseOptimized of the vendored build: False = scalar dispatch (the bit-exact pin)."""
        self.lib.ref_set_use_optimized(int(bool(on)))

    def set_num_threads(self, n):
        self.lib.ref_set_num_threads(int(n))

    def cv_gftt(self, img, max_corners, quality=0.01, min_distance=8.0, block_size=3, harris=False, k=0.04):
        """cv::goodFeaturesToTrack of the vendored OpenCV (featureselect.cpp:382-548)."""
        img, w, h, pitch = _img(img)
        cap = max_corners if max_corners > 0 else w * h
        xy = np.empty((cap, 2), np.float32)
        n = self.lib.ref_gftt(_p(img, _u8p), w, h, pitch, max_corners, quality, min_distance, block_size, int(harris), k,
                              _p(xy, _f32p), cap)
        return xy[:min(n, cap)].copy()

    def cv_response(self, img, block_size=3, ksize=3, harris=False, k=0.04):
        """cv::cornerMinEigenVal / cv::cornerHarris of the vendored OpenCV."""
        img, w, h, pitch = _img(img)
        out = np.empty((h, w), np.float32)
        self.lib.ref_corner_response(_p(img, _u8p), w, h, pitch, block_size, ksize, int(harris), k, _p(out, _f32p))
        return out

    def set_box_direct(self, on):
        """0: OpenCV's running column sum in the box filter (default); 1: direct three-row sum (port only)."""
        self.lib.vso_set_box_direct(int(bool(on)))

    def set_harris(self, on, k=0.04):
        """goodFeaturesToTrack(useHarrisDetector, k) for the restatement's detection functions (port only)."""
        self.lib.vso_set_harris(int(bool(on)), float(k))

    # --- detection -------------------------------------------------------------------------
    def response(self, img):
        img, w, h, pitch = _img(img)
        lam = np.empty((h, w), np.float32)
        rc = self.lib.vso_response(_p(img, _u8p), w, h, pitch, _p(lam, _f32p))
        assert rc == 0, rc
        return lam

    def gftt(self, img, max_corners, quality=0.01, min_distance=8.0):
        img, w, h, pitch = _img(img)
        cap = max_corners if max_corners > 0 else w * h
        xy = np.empty((cap, 2), np.float32)
        n = self.lib.vso_gftt(_p(img, _u8p), w, h, pitch, max_corners, quality, min_distance,
                              _p(xy, _f32p), cap)
        assert n >= 0, n
        return xy[:n].copy()

    def detect(self, img, num_features):
        img, w, h, pitch = _img(img)
        cap = num_features if num_features > 0 else w * h
        xy = np.empty((cap, 2), np.float64)
        n = getattr(self.lib, self._pre + "detect")(_p(img, _u8p), w, h, pitch, num_features,
                                                    _p(xy, _f64p), cap)
        assert 0 <= n <= cap, n
        return xy[:n].copy()

    # --- orientation / description -----------------------------------------------------------
    def angles(self, img, xy, rotate=True):
        img, w, h, pitch = _img(img)
        xy = np.ascontiguousarray(xy, np.float64)
        out = np.empty(len(xy), np.float64)
        getattr(self.lib, self._pre + "angles")(_p(img, _u8p), w, h, pitch, _p(xy, _f64p), len(xy),
                                                int(rotate), _p(out, _f64p))
        return out

    def describe(self, img, xy, angles, contract=True):
        img, w, h, pitch = _img(img)
        xy = np.ascontiguousarray(xy, np.float64)
        angles = np.ascontiguousarray(angles, np.float64)
        out = np.empty((len(xy), 8), np.uint32)
        f = getattr(self.lib, self._pre + "describe")
        if self.kind == "port":
            f(_p(img, _u8p), w, h, pitch, _p(xy, _f64p), _p(angles, _f64p), len(xy), int(contract), _p(out, _u32p))
        else:
            f(_p(img, _u8p), w, h, pitch, _p(xy, _f64p), _p(angles, _f64p), len(xy), _p(out, _u32p))
        return out

    def detect_describe(self, img, num_features, rotate=True, contract=True):
        img, w, h, pitch = _img(img)
        cap = num_features if num_features > 0 else w * h
        xy = np.empty((cap, 2), np.float64)
        ang = np.empty(cap, np.float64)
        desc = np.empty((cap, 8), np.uint32)
        f = getattr(self.lib, self._pre + "detect_describe")
        if self.kind == "port":
            n = f(_p(img, _u8p), w, h, pitch, num_features, int(rotate), int(contract),
                  _p(xy, _f64p), _p(ang, _f64p), _p(desc, _u32p), cap)
        else:
            n = f(_p(img, _u8p), w, h, pitch, num_features, int(rotate),
                  _p(xy, _f64p), _p(ang, _f64p), _p(desc, _u32p), cap)
        assert n >= 0, n
        return xy[:n].copy(), ang[:n].copy(), desc[:n].copy()

    # --- matching ----------------------------------------------------------------------------
    def match(self, d1, d2, threshold=70, ratio=1.2):
        d1 = np.ascontiguousarray(d1, np.uint32).reshape(-1, 8)
        d2 = np.ascontiguousarray(d2, np.uint32).reshape(-1, 8)
        cap = max(1, min(len(d1), len(d2)))
        pairs = np.empty((cap, 2), np.int32)
        n = getattr(self.lib, self._pre + "match")(_p(d1, _u32p), len(d1), _p(d2, _u32p), len(d2),
                                                   int(threshold), float(ratio), _p(pairs, _i32p), cap)
        assert 0 <= n <= cap, n
        return pairs[:n].copy()

    def best2(self, q, db):
        q = np.ascontiguousarray(q, np.uint32).reshape(-1, 8)
        db = np.ascontiguousarray(db, np.uint32).reshape(-1, 8)
        out = np.empty((len(q), 3), np.int32)
        self.lib.vso_best2(_p(q, _u32p), len(q), _p(db, _u32p), len(db), _p(out, _i32p))
        return out


def _lm_args(kp_xy, kp_desc, proj_xy, obs_offsets, obs_desc):
    kp_xy = np.ascontiguousarray(kp_xy, np.float64).reshape(-1, 2)
    kp_desc = np.ascontiguousarray(kp_desc, np.uint32).reshape(-1, 8)
    proj_xy = np.ascontiguousarray(proj_xy, np.float64).reshape(-1, 2)
    obs_offsets = np.ascontiguousarray(obs_offsets, np.int32)
    obs_desc = np.ascontiguousarray(obs_desc, np.uint32).reshape(-1, 8)
    assert len(kp_xy) == len(kp_desc) and len(obs_offsets) == len(proj_xy) + 1
    return kp_xy, kp_desc, proj_xy, obs_offsets, obs_desc


_LM_ARGTYPES = [_f64p, _u32p, C.c_int, _f64p, C.c_int, _i32p, _u32p, C.c_double, C.c_int, C.c_double, _i32p, C.c_int]


def _match_landmarks(fn, kp_xy, kp_desc, proj_xy, obs_offsets, obs_desc, max_dist_2d, threshold, ratio):
    kp_xy, kp_desc, proj_xy, obs_offsets, obs_desc = _lm_args(kp_xy, kp_desc, proj_xy, obs_offsets, obs_desc)
    cap = max(1, len(kp_xy))
    pairs = np.empty((cap, 2), np.int32)
    fn.restype = C.c_int
    fn.argtypes = _LM_ARGTYPES
    n = fn(_p(kp_xy, _f64p), _p(kp_desc, _u32p), len(kp_xy), _p(proj_xy, _f64p), len(proj_xy),
           _p(obs_offsets, _i32p), _p(obs_desc, _u32p), float(max_dist_2d), int(threshold), float(ratio),
           _p(pairs, _i32p), cap)
    assert 0 <= n <= cap, n
    return pairs[:n].copy()


def match_landmarks_port(kp_xy, kp_desc, proj_xy, obs_offsets, obs_desc, max_dist_2d=20.0, threshold=70, ratio=1.2):
    """C restatement of find_matches_landmarks (vo_utils.h:83-167): [(feature id, projected index)]."""
    return _match_landmarks(load_oracle().lib.vso_match_landmarks, kp_xy, kp_desc, proj_xy, obs_offsets, obs_desc,
                            max_dist_2d, threshold, ratio)


def match_landmarks_ref(kp_xy, kp_desc, proj_xy, obs_offsets, obs_desc, max_dist_2d=20.0, threshold=70, ratio=1.2):
    """The reference's own find_matches_landmarks (oracle/_ref/libvisnav_ref_lm.so), or None if not built."""
    path = os.path.join(_HERE, "_ref", "libvisnav_ref_lm.so")
    if not os.path.exists(path):
        return None
    return _match_landmarks(C.CDLL(path).ref_match_landmarks, kp_xy, kp_desc, proj_xy, obs_offsets, obs_desc,
                            max_dist_2d, threshold, ratio)


# --- geometry after matching: camera models, findInliersEssential, project_landmarks ------
CAM_MODELS = {"pinhole": 0, "eucm": 1, "ds": 2, "kb4": 3}


class Geometry:
    """kind == "port": oracle/visnav_oracle_geo.c.  kind == "reference": the reference's own camera_models.h,
    matching_utils.h:63-88 and vo_utils.h:48-81 (oracle/_ref/libvisnav_ref_geo.so)."""

    def __init__(self, lib, kind):
        self.lib, self.kind = lib, kind
        pre = "vso_" if kind == "port" else "ref_"
        self._project = getattr(lib, pre + "cam_project")
        self._project.restype = None
        self._project.argtypes = [C.c_int, _f64p, _f64p, C.c_int, _f64p]
        self._unproject = getattr(lib, pre + "cam_unproject")
        self._unproject.restype = None
        self._unproject.argtypes = [C.c_int, _f64p, _f64p, C.c_int, _f64p]
        self._inl = getattr(lib, pre + "find_inliers_essential")
        self._inl.restype = C.c_int
        self._inl.argtypes = [_f64p, C.c_int, _f64p, C.c_int, C.c_int, _f64p, C.c_int, _f64p, _f64p, C.c_double,
                              _i32p, C.c_int, _i32p, _f64p]
        self._pl = getattr(lib, pre + "project_landmarks")
        self._pl.restype = C.c_int
        self._pl.argtypes = [_f64p, C.c_int, _f64p, C.c_int, C.c_int, _f64p, C.c_int, C.c_double, _f64p, _i32p]
        if kind == "reference":
            lib.ref_compute_essential.restype = None
            lib.ref_compute_essential.argtypes = [_f64p, _f64p]

    @staticmethod
    def _cam(model, param):
        model = CAM_MODELS[model] if isinstance(model, str) else int(model)
        param = np.ascontiguousarray(param, np.float64)
        assert param.shape == (8,)
        return model, param

    def project(self, model, param, xyz):
        model, param = self._cam(model, param)
        xyz = np.ascontiguousarray(xyz, np.float64).reshape(-1, 3)
        uv = np.empty((len(xyz), 2), np.float64)
        self._project(model, _p(param, _f64p), _p(xyz, _f64p), len(xyz), _p(uv, _f64p))
        return uv

    def unproject(self, model, param, uv):
        model, param = self._cam(model, param)
        uv = np.ascontiguousarray(uv, np.float64).reshape(-1, 2)
        xyz = np.empty((len(uv), 3), np.float64)
        self._unproject(model, _p(param, _f64p), _p(uv, _f64p), len(uv), _p(xyz, _f64p))
        return xyz

    def compute_essential(self, pose_0_1):
        """matching_utils.h:55-61 (reference library only): E = skew(t/|t|) R, row-major 3x3."""
        pose = np.ascontiguousarray(pose_0_1, np.float64)
        E = np.empty((3, 3), np.float64)
        self.lib.ref_compute_essential(_p(pose, _f64p), _p(E, _f64p))
        return E

    def find_inliers_essential(self, xy1, xy2, cam1, cam2, E, thr, matches):
        """-> (inliers [m][2] int32, err [n_matches] = p0^T E p1)."""
        xy1 = np.ascontiguousarray(xy1, np.float64).reshape(-1, 2)
        xy2 = np.ascontiguousarray(xy2, np.float64).reshape(-1, 2)
        m1, p1 = self._cam(*cam1)
        m2, p2 = self._cam(*cam2)
        E = np.ascontiguousarray(E, np.float64).reshape(3, 3)
        matches = np.ascontiguousarray(matches, np.int32).reshape(-1, 2)
        inl = np.empty((max(1, len(matches)), 2), np.int32)
        err = np.empty(max(1, len(matches)), np.float64)
        n = self._inl(_p(xy1, _f64p), len(xy1), _p(xy2, _f64p), len(xy2), m1, _p(p1, _f64p), m2, _p(p2, _f64p),
                      _p(E, _f64p), float(thr), _p(matches, _i32p), len(matches), _p(inl, _i32p), _p(err, _f64p))
        return inl[:n].copy(), err[:len(matches)].copy()

    def project_landmarks(self, pose_w_c, cam, width, height, xyz, cam_z_threshold):
        """-> (projected_points [m][2], landmark indices [m]) in ascending landmark index."""
        pose = np.ascontiguousarray(pose_w_c, np.float64)
        assert pose.shape == (7,)
        model, param = self._cam(*cam)
        xyz = np.ascontiguousarray(xyz, np.float64).reshape(-1, 3)
        xy = np.empty((max(1, len(xyz)), 2), np.float64)
        idx = np.empty(max(1, len(xyz)), np.int32)
        n = self._pl(_p(pose, _f64p), model, _p(param, _f64p), int(width), int(height), _p(xyz, _f64p), len(xyz),
                     float(cam_z_threshold), _p(xy, _f64p), _p(idx, _i32p))
        return xy[:n].copy(), idx[:n].copy()


def load_geometry():
    return Geometry(load_oracle().lib, "port")


def load_geometry_ref():
    path = os.path.join(_HERE, "_ref", "libvisnav_ref_geo.so")
    if not os.path.exists(path):
        return None
    return Geometry(C.CDLL(path), "reference")


# --- DB-sharded matchDescriptors (config 4 across GPUs): CPU statement of the part / merge pair -------------
def _passes(m1, m2, threshold, ratio):
    """keypoints.h:355-357: best1 < threshold and not (best2 < best1 * dist_2_best) (int * double)."""
    return (m1 < threshold) & ~(m2.astype(np.float64) < m1.astype(np.float64) * ratio)


def match_shard_part(o, d1, d2_shard, index_offset, threshold=70, ratio=1.2):
    """[n1][4] int32 (m1, global a1, m2, mutual flag) of one desc2 shard, from the restated scan
    (keypoints.h:331-354 for the rows, :278-313 isPQiffQP for the flag)."""
    d1 = np.ascontiguousarray(d1, np.uint32).reshape(-1, 8)
    d2 = np.ascontiguousarray(d2_shard, np.uint32).reshape(-1, 8)
    part = np.empty((len(d1), 4), np.int32)
    if len(d2) == 0:
        part[:] = [256, index_offset, 256, 0]
        return part
    rows = o.best2(d1, d2)
    cols = o.best2(d2, d1)
    c = cols[rows[:, 1]]
    flag = _passes(c[:, 0], c[:, 2], threshold, ratio) & (c[:, 1] == np.arange(len(d1)))
    part[:, 0], part[:, 1], part[:, 2], part[:, 3] = rows[:, 0], rows[:, 1] + index_offset, rows[:, 2], flag
    return part


def merge_shard_parts(parts, threshold=70, ratio=1.2):
    """Associative merge of the row triples (lowest global index wins ties), then threshold / ratio / mutual flag."""
    parts = np.asarray(parts, np.int32)
    best = parts[0].copy()
    for v in parts[1:]:
        m2 = np.minimum(np.minimum(best[:, 2], v[:, 2]), np.maximum(best[:, 0], v[:, 0]))
        take = (v[:, 0] < best[:, 0]) | ((v[:, 0] == best[:, 0]) & (v[:, 1] < best[:, 1]))
        best[take] = v[take]
        best[:, 2] = m2
    ok = _passes(best[:, 0], best[:, 2], threshold, ratio) & (best[:, 3] != 0)
    idx = np.nonzero(ok)[0]
    return np.stack([idx, best[idx, 1]], 1).astype(np.int32)


def load_oracle():
    path = os.path.join(_HERE, "liboracle.so")
    if not os.path.exists(path):
        build()
    return Oracle(C.CDLL(path), "port")


def load_ref():
    """The reference's own functions, or None where oracle/_ref was not built/shipped."""
    path = os.path.join(_HERE, "_ref", "libvisnav_ref.so")
    if not os.path.exists(path):
        return None
    return Oracle(C.CDLL(path), "reference")


# --- bag of words (DBoW2 half) ------------------------------------------------------------------
def _as_u8(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def _as_u32(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32))


def _as_f64(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _desc_bytes(desc):
    """[n][32] u8 view of descriptors given as [n][8] u32 (little endian) or [n][32] u8."""
    d = np.ascontiguousarray(desc)
    if d.dtype == np.uint32:
        d = d.view(np.uint8)
    return np.ascontiguousarray(d, np.uint8).reshape(-1, 32)


class BowVocab:
    """The C restatement (visnav_oracle_bow.c) of TemplatedVocabulary<FORB>: transform and score."""
    kind = "port"

    def __init__(self, lib, handle):
        self.lib, self.h = lib, handle
        info = np.zeros(6, np.int32)
        lib.vso_bow_vocab_info(C.c_void_p(handle), info.ctypes.data_as(C.POINTER(C.c_int32)))
        self.k, self.L, self.scoring, self.weighting, self.n_nodes, self.n_words = (int(x) for x in info)

    @classmethod
    def _lib(cls):
        lib = load_oracle().lib
        lib.vso_bow_vocab_create.restype = C.c_void_p
        lib.vso_bow_vocab_load_text.restype = C.c_void_p
        lib.vso_bow_score.restype = C.c_double
        return lib

    @classmethod
    def from_text(cls, path):
        lib = cls._lib()
        h = lib.vso_bow_vocab_load_text(path.encode())
        if not h:
            raise ValueError("not a vocabulary text file: %s" % path)
        return cls(lib, h)

    @classmethod
    def from_arrays(cls, k, L, scoring, weighting, parent, is_leaf, desc, weight):
        lib = cls._lib()
        parent = np.ascontiguousarray(parent, np.int32)
        is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = _desc_bytes(desc)
        weight = np.ascontiguousarray(weight, np.float64)
        h = lib.vso_bow_vocab_create(k, L, scoring, weighting, len(parent), parent.ctypes.data_as(C.POINTER(C.c_int32)),
                                     _as_u8(is_leaf), _as_u8(desc), _as_f64(weight))
        if not h:
            raise ValueError("bad vocabulary arrays")
        return cls(lib, h)

    def close(self):
        if self.h:
            self.lib.vso_bow_vocab_free(C.c_void_p(self.h))
            self.h = None

    def transform(self, desc, levelsup=4):
        """-> ids [m] u32 ascending, vals [m] f64, feat_word [n], feat_node [n], feat_weight [n]."""
        d = _desc_bytes(desc)
        n = len(d)
        ids = np.zeros(max(n, 1), np.uint32)
        vals = np.zeros(max(n, 1), np.float64)
        fw = np.zeros(max(n, 1), np.uint32)
        fn = np.zeros(max(n, 1), np.uint32)
        fwt = np.zeros(max(n, 1), np.float64)
        m = self.lib.vso_bow_transform(C.c_void_p(self.h), _as_u8(d), n, levelsup, _as_u32(ids), _as_f64(vals), _as_u32(fw), _as_u32(fn),
                                       _as_f64(fwt))
        return ids[:m].copy(), vals[:m].copy(), fw[:n].copy(), fn[:n].copy(), fwt[:n].copy()

    def score(self, v1, v2):
        i1, x1 = np.ascontiguousarray(v1[0], np.uint32), np.ascontiguousarray(v1[1], np.float64)
        i2, x2 = np.ascontiguousarray(v2[0], np.uint32), np.ascontiguousarray(v2[1], np.float64)
        return float(self.lib.vso_bow_score(self.scoring, _as_u32(i1), _as_f64(x1), len(i1), _as_u32(i2), _as_f64(x2), len(i2)))


class BowVocabRef:
    """The reference's own DBoW2 (oracle/_ref/libvisnav_ref_bow.so)."""
    kind = "reference"

    def __init__(self, lib, handle):
        self.lib, self.h = lib, handle
        info = np.zeros(6, np.int32)
        lib.vref_bow_info(C.c_void_p(handle), info.ctypes.data_as(C.POINTER(C.c_int32)))
        self.k, self.L, self.scoring, self.weighting, _, self.n_words = (int(x) for x in info)

    @staticmethod
    def lib():
        path = os.path.join(_HERE, "_ref", "libvisnav_ref_bow.so")
        if not os.path.exists(path):
            return None
        lib = C.CDLL(path)
        lib.vref_bow_load_text.restype = C.c_void_p
        lib.vref_bow_score.restype = C.c_double
        return lib

    @classmethod
    def from_text(cls, path):
        lib = cls.lib()
        if lib is None:
            return None
        h = lib.vref_bow_load_text(path.encode())
        if not h:
            raise ValueError("the reference rejected %s" % path)
        return cls(lib, h)

    @classmethod
    def create_save(cls, training_sets, k, L, weighting, scoring, path):
        """TemplatedVocabulary::create (k-means++ tree) on the training descriptor sets + saveToTextFile."""
        lib = cls.lib()
        sets = [_desc_bytes(s) for s in training_sets]
        begin = np.zeros(len(sets) + 1, np.int32)
        begin[1:] = np.cumsum([len(s) for s in sets])
        d = np.ascontiguousarray(np.concatenate(sets))
        return int(lib.vref_bow_create_save(_as_u8(d), begin.ctypes.data_as(C.POINTER(C.c_int32)), len(sets), k, L, weighting,
                                            scoring, path.encode()))

    def close(self):
        if self.h:
            self.lib.vref_bow_free(C.c_void_p(self.h))
            self.h = None

    def transform(self, desc, levelsup=4):
        """-> ids, vals, fv_node [n] (0xFFFFFFFF where the feature was stopped)."""
        d = _desc_bytes(desc)
        n = len(d)
        ids = np.zeros(max(n, 1), np.uint32)
        vals = np.zeros(max(n, 1), np.float64)
        fv = np.zeros(max(n, 1), np.uint32)
        m = self.lib.vref_bow_transform(C.c_void_p(self.h), _as_u8(d), n, levelsup, _as_u32(ids), _as_f64(vals), _as_u32(fv))
        return ids[:m].copy(), vals[:m].copy(), fv[:n].copy()

    def words(self, desc):
        d = _desc_bytes(desc)
        w = np.zeros(max(len(d), 1), np.uint32)
        self.lib.vref_bow_words(C.c_void_p(self.h), _as_u8(d), len(d), _as_u32(w))
        return w[:len(d)].copy()

    def score(self, v1, v2):
        i1, x1 = np.ascontiguousarray(v1[0], np.uint32), np.ascontiguousarray(v1[1], np.float64)
        i2, x2 = np.ascontiguousarray(v2[0], np.uint32), np.ascontiguousarray(v2[1], np.float64)
        return float(self.lib.vref_bow_score(C.c_void_p(self.h), _as_u32(i1), _as_f64(x1), len(i1), _as_u32(i2), _as_f64(x2), len(i2)))


# --- cv::ORB as compute_bow_vector configures it (ORB half) -----------------------------------------------
def gaussian_kernel7():
    """getGaussianKernel(7, 2, CV_32F): exp(-x^2 / (2 sigma^2)) normalised in double, rounded to float."""
    x = np.arange(-3, 4, dtype=np.float64)
    k = np.exp(-(x * x) / 8.0)
    return (k / k.sum()).astype(np.float32)


def orb_detect_compute(img, nfeatures=1500, nlevels=8, scale_factor=1.2, edge_threshold=19, fast_threshold=20, fused=1):
    """The C restatement (visnav_oracle_orb.c) of cv::ORB::create(nfeatures, 1.2f, 8, 19, 0, 2, FAST_SCORE)
    ->detectAndCompute.  -> kp [n, 6] float32 (x, y, size, angle, response, octave), desc [n, 32] uint8; level by
    level, raster order inside a level."""
    lib = load_oracle().lib
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    f = lib.vso_orb_detect_compute
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_void_p,
                  C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    f.restype = C.c_int
    k7 = gaussian_kernel7()
    cap = nfeatures * 2 + 4096
    while True:
        kp = np.zeros((cap, 6), np.float32)
        desc = np.zeros((cap, 32), np.uint8)
        n = f(img.ctypes.data, w, h, w, nfeatures, nlevels, float(np.float32(scale_factor)), edge_threshold,
              fast_threshold, k7.ctypes.data, fused, kp.ctypes.data, desc.ctypes.data, cap, None)
        if n <= cap:
            return kp[:n].copy(), desc[:n].copy()
        cap = n
