"""Host-side mirror of the reference's operator interface for the hot path, over the C ABI.

Same names and argument meaning as /root/reference/include/visnav/keypoints.h:
  detectKeypoints(img, kd, num_features)                       :133-150
  computeAngles(img, kd, rotate_features)                      :152-189
  computeDescriptors(img, kd)                                  :191-221
  detectKeypointsAndDescriptors(img, kd, num_features, rotate) :223-229
  matchDescriptors(desc1, desc2, matches, threshold, dist_2_best) :323-369
`KeypointsData` mirrors include/visnav/common_types.h:111-122 with numpy arrays in place of the
std::vectors (corners N x 2 float64, corner_angles N float64, corner_descriptors N x 8 uint32 =
std::bitset<256> words).  `matches` is a python list that is cleared and refilled with (id1, id2)
tuples, like the reference's std::vector<std::pair<int,int>>.

Everything here calls libvisnav_b200.so through ctypes; there is no CPU implementation.  A missing
library or GPU raises VslamError.
"""
import ctypes as C
import os

import numpy as np

PATCH_SIZE = 31          # keypoints.h:49
HALF_PATCH_SIZE = 15     # keypoints.h:50
EDGE_THRESHOLD = 19      # keypoints.h:51

_HERE = os.path.dirname(os.path.abspath(__file__))

ABI_SYMBOLS = [
    "vslam_create", "vslam_destroy", "vslam_last_error", "vslam_set_option", "vslam_stream", "vslam_sync",
    "vslam_launch_count", "vslam_host_alloc", "vslam_host_free", "vslam_detect", "vslam_angles",
    "vslam_describe", "vslam_detect_describe", "vslam_match", "vslam_best2", "vslam_best2_noalloc", "vslam_response",
    "vslam_match_landmarks", "vslam_pyramid_detect_describe_batch",
    "vslam_detect_describe_batch", "vslam_stereo_batch", "vslam_stereo_plan_create",
    "vslam_stereo_plan_destroy", "vslam_stereo_plan_upload", "vslam_stereo_plan_run",
    "vslam_stereo_plan_download", "vslam_stereo_plan_last_ms", "vslam_desc_store_create",
    "vslam_desc_store_destroy", "vslam_match_batch", "vslam_last_match_ms", "vslam_popc_peak",
    "vslam_event_record", "vslam_event_elapsed_ms",
    "vslam_cam_project", "vslam_cam_unproject", "vslam_find_inliers_essential", "vslam_project_landmarks",
    "vslam_stereo_batch_epipolar", "vslam_stereo_plan_set_epipolar", "vslam_stereo_plan_download_inliers",
    "vslam_stereo_plan_last_epipolar_ms", "vslam_match_shard", "vslam_match_merge_shards", "vslam_set_harris",
    "vslam_desc_store_append", "vslam_desc_store_size",
    "vslam_bow_vocab_create", "vslam_bow_vocab_load_text", "vslam_bow_vocab_destroy", "vslam_bow_vocab_info",
    "vslam_bow_transform_batch", "vslam_bow_score", "vslam_bow_db_create", "vslam_bow_db_destroy", "vslam_bow_db_add",
    "vslam_bow_db_size", "vslam_bow_db_query", "vslam_orb_detect_compute_batch", "vslam_compute_bow_vector_batch",
    "vslam_debug_mma_timeouts",
]

VSLAM_OPT_FP_CONTRACT = 1
VSLAM_OPT_MATCH_CSA = 2
VSLAM_OPT_BATCH_CHUNK = 3
VSLAM_OPT_BATCH_DEPTH = 4
VSLAM_OPT_MATCH_SPLIT_SMALL = 5
VSLAM_OPT_RESPONSE_BAND = 6
VSLAM_OPT_MATCH_MMA = 7
VSLAM_OPT_GRAPH = 8
VSLAM_OPT_RESPONSE_MULTI = 9
VSLAM_OPT_TWO_STREAMS = 10


class VslamError(RuntimeError):
    pass


class StereoParams(C.Structure):
    _fields_ = [("w", C.c_int), ("h", C.c_int), ("pitch", C.c_int), ("image_stride", C.c_size_t),
                ("num_features", C.c_int), ("rotate_features", C.c_int), ("match_threshold", C.c_int),
                ("match_dist_2_best", C.c_double)]


CAM_MODELS = {"pinhole": 0, "eucm": 1, "ds": 2, "kb4": 3}  # camera_models.h getName() -> VSLAM_CAM_*


class Camera(C.Structure):
    """vslam_camera: one of the reference's camera models (camera_models.h) with its 8-double parameter vector."""
    _fields_ = [("model", C.c_int), ("width", C.c_int), ("height", C.c_int), ("reserved", C.c_int),
                ("param", C.c_double * 8)]

    @classmethod
    def from_data(cls, name, param, width=0, height=0):
        """AbstractCamera::from_data (camera_models.h:475-498) + width()/height()."""
        if name not in CAM_MODELS:
            raise VslamError("Camera model %s is not implemented." % name)
        param = [float(v) for v in param]
        if len(param) != 8:
            raise VslamError("camera parameter vector must have 8 entries")
        return cls(CAM_MODELS[name], int(width), int(height), 0, (C.c_double * 8)(*param))


class Epipolar(C.Structure):
    _fields_ = [("cam1", Camera), ("cam2", Camera), ("E", C.c_double * 9), ("epipolar_error_threshold", C.c_double)]

    @classmethod
    def make(cls, cam1, cam2, E, thr):
        E = np.ascontiguousarray(E, np.float64).reshape(9)
        return cls(cam1, cam2, (C.c_double * 9)(*E.tolist()), float(thr))


def lib_path():
    # VSLAM_LIB: another build of the same library (profiling variants of tools/probes); never a different implementation
    return os.environ.get("VSLAM_LIB") or os.path.join(_HERE, "csrc", "libvisnav_b200.so")


_lib = None


def load_library():
    """dlopen the C-ABI library and declare its prototypes.  No GPU needed for this."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise VslamError("libvisnav_b200.so is not built (run __graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(path)
    vp, i, d, sz = C.c_void_p, C.c_int, C.c_double, C.c_size_t
    L.vslam_create.argtypes = [i, C.POINTER(vp)]
    L.vslam_destroy.argtypes = [vp]
    L.vslam_destroy.restype = None
    L.vslam_last_error.argtypes = [vp]
    L.vslam_last_error.restype = C.c_char_p
    L.vslam_set_option.argtypes = [vp, i, i]
    L.vslam_stream.argtypes = [vp]
    L.vslam_stream.restype = vp
    L.vslam_sync.argtypes = [vp]
    L.vslam_launch_count.argtypes = [vp]
    L.vslam_launch_count.restype = C.c_longlong
    L.vslam_host_alloc.argtypes = [sz]
    L.vslam_host_alloc.restype = vp
    L.vslam_host_free.argtypes = [vp]
    L.vslam_host_free.restype = None
    L.vslam_detect.argtypes = [vp, vp, i, i, i, i, vp, i, C.POINTER(i)]
    L.vslam_angles.argtypes = [vp, vp, i, i, i, vp, i, i, vp]
    L.vslam_describe.argtypes = [vp, vp, i, i, i, vp, vp, i, vp]
    L.vslam_detect_describe.argtypes = [vp, vp, i, i, i, i, i, vp, vp, vp, i, C.POINTER(i)]
    L.vslam_match.argtypes = [vp, vp, i, vp, i, i, d, vp, i, C.POINTER(i)]
    L.vslam_best2.argtypes = [vp, vp, i, vp, i, vp]
    L.vslam_best2_noalloc.argtypes = [vp, vp, i, vp, i, vp, vp]
    L.vslam_match_landmarks.argtypes = [vp, vp, vp, i, vp, i, vp, vp, d, i, d, vp, i, C.POINTER(i)]
    L.vslam_response.argtypes = [vp, vp, i, i, i, vp]
    L.vslam_detect_describe_batch.argtypes = [vp, vp, i, i, i, i, sz, i, i, vp, vp, vp, vp, i]
    L.vslam_pyramid_detect_describe_batch.argtypes = [vp, vp, i, i, i, i, sz, i, vp, i, vp, vp, vp, vp, i]
    L.vslam_stereo_batch.argtypes = [vp, vp, i, C.POINTER(StereoParams), vp, vp, vp, vp, vp, vp, i]
    L.vslam_stereo_plan_create.argtypes = [vp, i, i, i, i, C.POINTER(vp)]
    L.vslam_stereo_plan_destroy.argtypes = [vp]
    L.vslam_stereo_plan_destroy.restype = None
    L.vslam_stereo_plan_upload.argtypes = [vp, vp, i, i, sz]
    L.vslam_stereo_plan_run.argtypes = [vp, i, i, i, i, d]
    L.vslam_stereo_plan_download.argtypes = [vp, i, vp, vp, vp, vp, vp, vp]
    L.vslam_stereo_plan_last_ms.argtypes = [vp, C.POINTER(C.c_float)]
    L.vslam_desc_store_create.argtypes = [vp, vp, vp, vp, i, C.POINTER(vp)]
    L.vslam_desc_store_destroy.argtypes = [vp]
    L.vslam_desc_store_destroy.restype = None
    L.vslam_match_batch.argtypes = [vp, vp, vp, i, i, d, vp, vp, i]
    L.vslam_last_match_ms.argtypes = [vp]
    L.vslam_last_match_ms.restype = C.c_float
    L.vslam_popc_peak.argtypes = [vp, C.POINTER(d)]
    L.vslam_debug_mma_timeouts.argtypes = [vp]
    L.vslam_event_record.argtypes = [vp, i]
    L.vslam_event_elapsed_ms.argtypes = [vp, i, i, C.POINTER(C.c_float)]
    L.vslam_set_harris.argtypes = [vp, i, d]
    L.vslam_desc_store_append.argtypes = [vp, vp, i, C.POINTER(i)]
    L.vslam_desc_store_size.argtypes = [vp]
    L.vslam_match_shard.argtypes = [vp, vp, i, vp, i, i, i, d, vp]
    L.vslam_match_merge_shards.argtypes = [vp, vp, i, i, i, d, vp, i, C.POINTER(i)]
    cam = C.POINTER(Camera)
    L.vslam_cam_project.argtypes = [vp, cam, vp, i, vp]
    L.vslam_cam_unproject.argtypes = [vp, cam, vp, i, vp]
    L.vslam_find_inliers_essential.argtypes = [vp, vp, i, vp, i, cam, cam, vp, d, vp, i, vp, i, C.POINTER(i), vp]
    L.vslam_project_landmarks.argtypes = [vp, vp, cam, vp, i, d, vp, vp, i, C.POINTER(i)]
    L.vslam_stereo_batch_epipolar.argtypes = [vp, vp, i, C.POINTER(StereoParams), C.POINTER(Epipolar), vp, vp, vp, vp,
                                              vp, vp, vp, vp, i]
    L.vslam_stereo_plan_set_epipolar.argtypes = [vp, C.POINTER(Epipolar)]
    L.vslam_stereo_plan_download_inliers.argtypes = [vp, i, vp, vp]
    L.vslam_stereo_plan_last_epipolar_ms.argtypes = [vp, C.POINTER(C.c_float)]
    L.vslam_bow_vocab_create.argtypes = [vp, i, i, i, i, i, vp, vp, vp, vp, C.POINTER(vp)]
    L.vslam_bow_vocab_load_text.argtypes = [vp, C.c_char_p, C.POINTER(vp)]
    L.vslam_bow_vocab_destroy.argtypes = [vp]
    L.vslam_bow_vocab_destroy.restype = None
    L.vslam_bow_vocab_info.argtypes = [vp, vp]
    L.vslam_bow_transform_batch.argtypes = [vp, vp, vp, vp, i, i, vp, vp, vp, i, vp, vp, vp]
    L.vslam_bow_score.argtypes = [vp, vp, vp, vp, i, vp, vp, i, C.POINTER(d)]
    L.vslam_bow_db_create.argtypes = [vp, vp, C.POINTER(vp)]
    L.vslam_bow_db_destroy.argtypes = [vp]
    L.vslam_bow_db_destroy.restype = None
    L.vslam_bow_db_add.argtypes = [vp, vp, vp, i, C.POINTER(i)]
    L.vslam_bow_db_size.argtypes = [vp]
    L.vslam_bow_db_query.argtypes = [vp, vp, vp, vp, i, vp, vp]
    L.vslam_orb_detect_compute_batch.argtypes = [vp, vp, i, i, i, i, sz, C.POINTER(OrbParams), vp, vp, vp, i]
    L.vslam_compute_bow_vector_batch.argtypes = [vp, vp, vp, i, i, i, i, sz, C.POINTER(OrbParams), i, vp, vp, vp, i, vp, vp,
                                                 vp, vp, vp, i]
    _lib = L
    return L


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _img(img):
    """(array kept alive, w, h, pitch) of a 2-D uint8 image; rows may be strided (pitch)."""
    img = np.asarray(img)
    if img.dtype != np.uint8 or img.ndim != 2:
        raise VslamError("image must be a 2-D uint8 array")
    if img.strides[1] != 1 or img.strides[0] < img.shape[1]:
        img = np.ascontiguousarray(img)
    return img, img.shape[1], img.shape[0], img.strides[0]


class KeypointsData:
    """include/visnav/common_types.h:111-122."""

    def __init__(self):
        self.corners = np.zeros((0, 2), np.float64)
        self.corner_angles = np.zeros(0, np.float64)
        self.corner_descriptors = np.zeros((0, 8), np.uint32)


class Context:
    """One vslam_ctx (one GPU, one host thread)."""

    def __init__(self, device=0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.vslam_create(int(device), C.byref(h))
        if rc != 0 or not h:
            raise VslamError("vslam_create(device=%d) failed (%d): no usable sm_100 GPU; there is no CPU fallback"
                             % (device, rc))
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.lib.vslam_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise VslamError("vslam error %d: %s" % (rc, self.lib.vslam_last_error(self.h).decode()))

    def set_option(self, opt, value):
        self._check(self.lib.vslam_set_option(self.h, opt, int(value)))

    def set_harris(self, use_harris_detector, k=0.04):
        """goodFeaturesToTrack(useHarrisDetector, k) for this context's detection calls (reference: always False)."""
        self._check(self.lib.vslam_set_harris(self.h, int(bool(use_harris_detector)), float(k)))

    def stream(self):
        return self.lib.vslam_stream(self.h)

    def sync(self):
        self._check(self.lib.vslam_sync(self.h))

    def event_record(self, slot):
        self._check(self.lib.vslam_event_record(self.h, slot))

    def event_elapsed_ms(self, a, b):
        ms = C.c_float(0)
        self._check(self.lib.vslam_event_elapsed_ms(self.h, a, b, C.byref(ms)))
        return ms.value

    def host_array(self, shape, dtype):
        """numpy array over pinned host memory (cudaMallocHost); freed with the returned handle's .free()."""
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = self.lib.vslam_host_alloc(max(n, 1))
        if not p:
            raise VslamError("vslam_host_alloc(%d) failed" % n)
        buf = (C.c_uint8 * max(n, 1)).from_address(p)
        arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
        self._pinned = getattr(self, "_pinned", [])
        self._pinned.append(p)
        return arr

    def orb_detect_compute_batch(self, images, num_features=1500, params=None, cap=None):
        """cv::ORB::create(num_features, 1.2, 8, 19, 0, 2, FAST_SCORE)->detectAndCompute for [count, h, w] images
        (keypoints.h:249-250).  -> kp [count, cap, 6] (x, y, size, angle, response, octave), desc [count, cap, 32], n_kp."""
        images = _image_batch(images)
        count, h, w = images.shape
        prm = params or OrbParams(num_features)
        cap = cap or prm.nfeatures + prm.nfeatures // 4 + 256
        kp = np.zeros((max(count, 1), cap, 6), np.float32)
        desc = np.zeros((max(count, 1), cap, 32), np.uint8)
        n_kp = np.zeros(max(count, 1), np.int32)
        self._check(self.lib.vslam_orb_detect_compute_batch(self.h, _ptr(images), count, w, h, w, w * h, C.byref(prm),
                                                            _ptr(kp), _ptr(desc), _ptr(n_kp), cap))
        return kp[:count], desc[:count], n_kp[:count]

    def launch_count(self):
        return int(self.lib.vslam_launch_count(self.h))

    # ---- single image ----------------------------------------------------------------------
    def detect(self, img, num_features):
        img, w, h, pitch = _img(img)
        cap = num_features if num_features > 0 else w * h
        xy = np.empty((max(cap, 1), 2), np.float64)
        n = C.c_int(0)
        self._check(self.lib.vslam_detect(self.h, _ptr(img), w, h, pitch, int(num_features), _ptr(xy), cap, C.byref(n)))
        return xy[:n.value].copy()

    def angles(self, img, xy, rotate=True):
        img, w, h, pitch = _img(img)
        xy = np.ascontiguousarray(xy, np.float64).reshape(-1, 2)
        out = np.empty(len(xy), np.float64)
        self._check(self.lib.vslam_angles(self.h, _ptr(img), w, h, pitch, _ptr(xy), len(xy), int(bool(rotate)), _ptr(out)))
        return out

    def describe(self, img, xy, angles):
        img, w, h, pitch = _img(img)
        xy = np.ascontiguousarray(xy, np.float64).reshape(-1, 2)
        angles = np.ascontiguousarray(angles, np.float64)
        if len(angles) != len(xy):
            raise VslamError("corners and corner_angles differ in length")
        out = np.empty((len(xy), 8), np.uint32)
        self._check(self.lib.vslam_describe(self.h, _ptr(img), w, h, pitch, _ptr(xy), _ptr(angles), len(xy), _ptr(out)))
        return out

    def detect_describe(self, img, num_features, rotate=True):
        img, w, h, pitch = _img(img)
        cap = num_features if num_features > 0 else w * h
        xy = np.empty((max(cap, 1), 2), np.float64)
        ang = np.empty(max(cap, 1), np.float64)
        desc = np.empty((max(cap, 1), 8), np.uint32)
        n = C.c_int(0)
        self._check(self.lib.vslam_detect_describe(self.h, _ptr(img), w, h, pitch, int(num_features), int(bool(rotate)),
                                                   _ptr(xy), _ptr(ang), _ptr(desc), cap, C.byref(n)))
        n = n.value
        return xy[:n].copy(), ang[:n].copy(), desc[:n].copy()

    def response(self, img):
        img, w, h, pitch = _img(img)
        lam = np.empty((h, w), np.float32)
        self._check(self.lib.vslam_response(self.h, _ptr(img), w, h, pitch, _ptr(lam)))
        return lam

    def match(self, d1, d2, threshold=70, dist_2_best=1.2, cap=None):
        d1 = np.ascontiguousarray(d1, np.uint32).reshape(-1, 8)
        d2 = np.ascontiguousarray(d2, np.uint32).reshape(-1, 8)
        if cap is None:
            cap = max(1, min(len(d1), len(d2)))
        pairs = np.empty((cap, 2), np.int32)
        n = C.c_int(0)
        self._check(self.lib.vslam_match(self.h, _ptr(d1), len(d1), _ptr(d2), len(d2), int(threshold),
                                         float(dist_2_best), _ptr(pairs), cap, C.byref(n)))
        return pairs[:n.value].copy()

    def match_landmarks(self, kp_xy, kp_desc, proj_xy, obs_offsets, obs_desc, match_max_dist_2d=20.0, threshold=70,
                        dist_2_best=1.2, cap=None):
        """find_matches_landmarks (vo_utils.h:83-167): returns [(feature id, projected index)] as int32 [m, 2]."""
        kp_xy = np.ascontiguousarray(kp_xy, np.float64).reshape(-1, 2)
        kp_desc = np.ascontiguousarray(kp_desc, np.uint32).reshape(-1, 8)
        proj_xy = np.ascontiguousarray(proj_xy, np.float64).reshape(-1, 2)
        obs_offsets = np.ascontiguousarray(obs_offsets, np.int32)
        obs_desc = np.ascontiguousarray(obs_desc, np.uint32).reshape(-1, 8)
        assert len(kp_xy) == len(kp_desc) and len(obs_offsets) == len(proj_xy) + 1
        cap = len(kp_xy) if cap is None else cap
        pairs = np.empty((max(cap, 1), 2), np.int32)
        n = C.c_int(0)
        self._check(self.lib.vslam_match_landmarks(self.h, _ptr(kp_xy), _ptr(kp_desc), len(kp_xy), _ptr(proj_xy),
                                                   len(proj_xy), _ptr(obs_offsets), _ptr(obs_desc),
                                                   C.c_double(match_max_dist_2d), int(threshold),
                                                   C.c_double(dist_2_best), _ptr(pairs), cap, C.byref(n)))
        return pairs[:n.value].copy()

    # --- geometry after matching (camera_models.h, matching_utils.h:63-88, vo_utils.h:48-81) ---
    def cam_project(self, cam, xyz):
        xyz = np.ascontiguousarray(xyz, np.float64).reshape(-1, 3)
        uv = np.empty((len(xyz), 2), np.float64)
        self._check(self.lib.vslam_cam_project(self.h, C.byref(cam), _ptr(xyz), len(xyz), _ptr(uv)))
        return uv

    def cam_unproject(self, cam, uv):
        uv = np.ascontiguousarray(uv, np.float64).reshape(-1, 2)
        xyz = np.empty((len(uv), 3), np.float64)
        self._check(self.lib.vslam_cam_unproject(self.h, C.byref(cam), _ptr(uv), len(uv), _ptr(xyz)))
        return xyz

    def find_inliers_essential(self, xy1, xy2, cam1, cam2, E, thr, matches, want_err=False, cap=None):
        """findInliersEssential: -> inliers [m][2] (and p0^T E p1 per match when want_err)."""
        xy1 = np.ascontiguousarray(xy1, np.float64).reshape(-1, 2)
        xy2 = np.ascontiguousarray(xy2, np.float64).reshape(-1, 2)
        matches = np.ascontiguousarray(matches, np.int32).reshape(-1, 2)
        E = np.ascontiguousarray(E, np.float64).reshape(9)
        cap = len(matches) if cap is None else cap
        inl = np.empty((max(cap, 1), 2), np.int32)
        err = np.empty(max(len(matches), 1), np.float64) if want_err else None
        n = C.c_int(0)
        self._check(self.lib.vslam_find_inliers_essential(self.h, _ptr(xy1), len(xy1), _ptr(xy2), len(xy2),
                                                          C.byref(cam1), C.byref(cam2), _ptr(E), float(thr),
                                                          _ptr(matches), len(matches), _ptr(inl), cap, C.byref(n),
                                                          _ptr(err)))
        inl = inl[:n.value].copy()
        return (inl, err[:len(matches)]) if want_err else inl

    def project_landmarks(self, pose_w_c, cam, xyz, cam_z_threshold, cap=None):
        """project_landmarks: -> (projected_points [m][2], landmark indices [m])."""
        pose = np.ascontiguousarray(pose_w_c, np.float64)
        if pose.shape != (7,):
            raise VslamError("pose must be Sophus::SE3d::data(): qx qy qz qw tx ty tz")
        xyz = np.ascontiguousarray(xyz, np.float64).reshape(-1, 3)
        cap = len(xyz) if cap is None else cap
        xy = np.empty((max(cap, 1), 2), np.float64)
        idx = np.empty(max(cap, 1), np.int32)
        n = C.c_int(0)
        self._check(self.lib.vslam_project_landmarks(self.h, _ptr(pose), C.byref(cam), _ptr(xyz), len(xyz),
                                                     float(cam_z_threshold), _ptr(xy), _ptr(idx), cap, C.byref(n)))
        return xy[:n.value].copy(), idx[:n.value].copy()

    def match_shard(self, d1, d2_shard, index_offset, threshold=70, dist_2_best=1.2):
        """One DB shard of a sharded matchDescriptors: -> part [n1][4] int32 (m1, global a1, m2, mutual flag)."""
        d1 = np.ascontiguousarray(d1, np.uint32).reshape(-1, 8)
        d2 = np.ascontiguousarray(d2_shard, np.uint32).reshape(-1, 8)
        part = np.empty((len(d1), 4), np.int32)
        self._check(self.lib.vslam_match_shard(self.h, _ptr(d1), len(d1), _ptr(d2), len(d2), int(index_offset),
                                               int(threshold), float(dist_2_best), _ptr(part)))
        return part

    def match_merge_shards(self, parts, threshold=70, dist_2_best=1.2, cap=None):
        """parts: [n_parts][n1][4] int32 from match_shard (any order) -> matches [m][2], as match() on the whole DB."""
        parts = np.ascontiguousarray(parts, np.int32)
        assert parts.ndim == 3 and parts.shape[2] == 4
        n_parts, n1 = parts.shape[:2]
        cap = n1 if cap is None else cap
        pairs = np.empty((max(cap, 1), 2), np.int32)
        n = C.c_int(0)
        self._check(self.lib.vslam_match_merge_shards(self.h, _ptr(parts), n_parts, n1, int(threshold),
                                                      float(dist_2_best), _ptr(pairs), cap, C.byref(n)))
        return pairs[:n.value].copy()

    def best2(self, q, db):
        q = np.ascontiguousarray(q, np.uint32).reshape(-1, 8)
        db = np.ascontiguousarray(db, np.uint32).reshape(-1, 8)
        out = np.empty((len(q), 3), np.int32)
        self._check(self.lib.vslam_best2(self.h, _ptr(q), len(q), _ptr(db), len(db), _ptr(out)))
        return out

    def best2_noalloc(self, d_q, nq, d_db, ndb, d_out3, stream=None):
        """vslam_best2 on DEVICE buffers (integer addresses) with no allocation inside the call: see the header.  Not
        synchronised; issued on `stream` (a cudaStream_t value) or the context's own stream."""
        self._check(self.lib.vslam_best2_noalloc(self.h, C.c_void_p(d_q), int(nq), C.c_void_p(d_db), int(ndb),
                                                 C.c_void_p(d_out3), C.c_void_p(stream)))

    def last_match_ms(self):
        return float(self.lib.vslam_last_match_ms(self.h))

    def mma_timeouts(self):
        """1 if the tensor-core matcher ever hit a pipeline timeout in this process (a bug), else 0."""
        return int(self.lib.vslam_debug_mma_timeouts(self.h))

    def popc_peak(self):
        v = C.c_double(0)
        self._check(self.lib.vslam_popc_peak(self.h, C.byref(v)))
        return v.value

    # ---- batches ----------------------------------------------------------------------------
    def detect_describe_batch(self, images, num_features, rotate=True, cap=None):
        images = np.ascontiguousarray(images, np.uint8)
        count, h, w = images.shape
        cap = cap or num_features
        xy = np.empty((count, cap, 2), np.float64)
        ang = np.empty((count, cap), np.float64)
        desc = np.empty((count, cap, 8), np.uint32)
        nkp = np.empty(count, np.int32)
        self._check(self.lib.vslam_detect_describe_batch(self.h, _ptr(images), count, w, h, w, w * h, int(num_features),
                                                         int(bool(rotate)), _ptr(xy), _ptr(ang), _ptr(desc), _ptr(nkp), cap))
        return xy, ang, desc, nkp

    def pyramid_detect_describe_batch(self, images, num_features, rotate=True):
        """Multi-scale extension (config 5): num_features = per-level caps.  Returns per image and level
        lists via (xy, angles, desc, n_kp[count][levels], level offsets)."""
        images = np.ascontiguousarray(images, np.uint8)
        count, h, w = images.shape
        nf = np.ascontiguousarray(num_features, np.int32)
        cap = int(nf.sum())
        xy = np.empty((count, cap, 2), np.float64)
        ang = np.empty((count, cap), np.float64)
        desc = np.empty((count, cap, 8), np.uint32)
        nkp = np.empty((count, len(nf)), np.int32)
        self._check(self.lib.vslam_pyramid_detect_describe_batch(self.h, _ptr(images), count, w, h, w, w * h, len(nf),
                                                                 _ptr(nf), int(bool(rotate)), _ptr(xy), _ptr(ang),
                                                                 _ptr(desc), _ptr(nkp), cap))
        off = np.concatenate([[0], np.cumsum(nf)[:-1]]).astype(np.int64)
        return xy, ang, desc, nkp, off

    def stereo_batch(self, images, num_features=1500, rotate=True, threshold=70, dist_2_best=1.2, cap=None,
                     out=None, epipolar=None):
        """images: [count, 2, h, w] uint8 (C-contiguous; pinned memory makes the copies asynchronous).
        Returns dict(xy, angles, desc, n_kp, matches, n_matches) of fixed-stride slabs; with `epipolar`
        (an Epipolar) also inliers / n_inliers (findInliersEssential fused behind the matcher)."""
        images = np.asarray(images)
        if images.dtype != np.uint8 or images.ndim != 4 or images.shape[1] != 2 or not images.flags.c_contiguous:
            raise VslamError("images must be a C-contiguous [count, 2, h, w] uint8 array")
        count, _, h, w = images.shape
        cap = cap or num_features
        if out is None:
            out = dict(xy=np.empty((count * 2, cap, 2), np.float64), angles=np.empty((count * 2, cap), np.float64),
                       desc=np.empty((count * 2, cap, 8), np.uint32), n_kp=np.empty(count * 2, np.int32),
                       matches=np.empty((count, cap, 2), np.int32), n_matches=np.empty(count, np.int32))
        p = StereoParams(w, h, w, w * h, int(num_features), int(bool(rotate)), int(threshold), float(dist_2_best))
        if epipolar is not None:
            out.setdefault("inliers", np.empty((count, cap, 2), np.int32))
            out.setdefault("n_inliers", np.empty(count, np.int32))
            self._check(self.lib.vslam_stereo_batch_epipolar(
                self.h, _ptr(images), count, C.byref(p), C.byref(epipolar), _ptr(out.get("xy")), _ptr(out.get("angles")),
                _ptr(out.get("desc")), _ptr(out.get("n_kp")), _ptr(out.get("matches")), _ptr(out.get("n_matches")),
                _ptr(out["inliers"]), _ptr(out["n_inliers"]), cap))
            return out
        self._check(self.lib.vslam_stereo_batch(self.h, _ptr(images), count, C.byref(p), _ptr(out.get("xy")),
                                                _ptr(out.get("angles")), _ptr(out.get("desc")), _ptr(out.get("n_kp")),
                                                _ptr(out.get("matches")), _ptr(out.get("n_matches")), cap))
        return out

    def match_batch(self, store, jobs, threshold=70, dist_2_best=1.2, cap=None, want_matches=True):
        jobs = np.ascontiguousarray(jobs, np.int32).reshape(-1, 2)
        cap = cap or store.max_count
        n = len(jobs)
        matches = np.empty((n, cap, 2), np.int32) if want_matches else None
        counts = np.empty(n, np.int32)
        self._check(self.lib.vslam_match_batch(self.h, store.h, _ptr(jobs), n, int(threshold), float(dist_2_best),
                                               _ptr(matches), _ptr(counts), cap))
        return matches, counts


class DescStore:
    """Device-resident descriptor sets (keyframes) for Context.match_batch."""

    def __init__(self, ctx, sets):
        """sets: [n_sets, n, 8] uint32 array, or a list of [n_k, 8] arrays."""
        self.ctx = ctx
        if isinstance(sets, np.ndarray) and sets.ndim == 3:
            flat = np.ascontiguousarray(sets, np.uint32).reshape(-1, 8)
            counts = np.full(sets.shape[0], sets.shape[1], np.int32)
        else:
            sets = [np.ascontiguousarray(s, np.uint32).reshape(-1, 8) for s in sets]
            counts = np.array([len(s) for s in sets], np.int32)
            flat = np.concatenate(sets) if len(sets) else np.zeros((0, 8), np.uint32)
        offsets = np.concatenate([[0], np.cumsum(counts[:-1], dtype=np.int64)]).astype(np.int64)
        self.max_count = int(counts.max()) if len(counts) else 0
        h = C.c_void_p()
        ctx._check(ctx.lib.vslam_desc_store_create(ctx.h, _ptr(flat), _ptr(offsets), _ptr(counts), len(counts), C.byref(h)))
        self.h = h

    def append(self, desc):
        """Adds one descriptor set (a new keyframe); returns its set index.  Earlier indices stay valid."""
        d = np.ascontiguousarray(desc, np.uint32).reshape(-1, 8)
        idx = C.c_int(-1)
        self.ctx._check(self.ctx.lib.vslam_desc_store_append(self.h, _ptr(d), len(d), C.byref(idx)))
        self.max_count = max(self.max_count, len(d))
        return idx.value

    def __len__(self):
        return int(self.ctx.lib.vslam_desc_store_size(self.h))

    def close(self):
        if getattr(self, "h", None):
            self.ctx.lib.vslam_desc_store_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class OrbParams(C.Structure):
    """vslam_orb_params: cv::ORB::create(nfeatures, scale_factor, nlevels, edge_threshold, 0, 2, FAST_SCORE, 31, fast_threshold);
    the defaults are the reference's call (keypoints.h:249)."""
    _fields_ = [("nfeatures", C.c_int), ("scale_factor", C.c_float), ("nlevels", C.c_int), ("edge_threshold", C.c_int),
                ("fast_threshold", C.c_int)]

    def __init__(self, nfeatures=1500, scale_factor=1.2, nlevels=8, edge_threshold=19, fast_threshold=20):
        super().__init__(int(nfeatures), float(scale_factor), int(nlevels), int(edge_threshold), int(fast_threshold))


def _image_batch(images):
    images = np.ascontiguousarray(images)
    if images.dtype != np.uint8 or images.ndim != 3:
        raise VslamError("images must be a [count, h, w] uint8 array")
    return images


BOW_SCORING = {"L1_NORM": 0, "L2_NORM": 1, "CHI_SQUARE": 2, "KL": 3, "BHATTACHARYYA": 4, "DOT_PRODUCT": 5}
BOW_WEIGHTING = {"TF_IDF": 0, "TF": 1, "IDF": 2, "BINARY": 3}


def _desc_bytes(desc):
    """[n][32] u8 view of descriptors given as [n][8] u32 (little endian) or [n][32] u8 (cv::ORB rows)."""
    d = np.ascontiguousarray(desc)
    if d.dtype == np.uint32:
        d = d.view(np.uint8)
    return np.ascontiguousarray(d, np.uint8).reshape(-1, 32)


class BowVocabulary:
    """visnav::ORBVocabulary (ORBVocabulary.h:13 = DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>) on the
    device: transform() and score() as the reference calls them (keypoints.h:253, loop_closure_utils.h:119)."""

    def __init__(self, ctx, handle):
        self.ctx, self.h = ctx, handle
        info = np.zeros(6, np.int32)
        ctx.lib.vslam_bow_vocab_info(handle, _ptr(info))
        self.k, self.L, self.scoring, self.weighting, self.n_nodes, self.n_words = (int(x) for x in info)

    @classmethod
    def load_text(cls, ctx, path):
        """TemplatedVocabulary::loadFromTextFile (slam.cpp:373)."""
        h = C.c_void_p()
        ctx._check(ctx.lib.vslam_bow_vocab_load_text(ctx.h, str(path).encode(), C.byref(h)))
        return cls(ctx, h)

    @classmethod
    def from_arrays(cls, ctx, k, L, scoring, weighting, parent, is_leaf, desc, weight):
        parent = np.ascontiguousarray(parent, np.int32)
        is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = _desc_bytes(desc)
        weight = np.ascontiguousarray(weight, np.float64)
        assert len(is_leaf) == len(parent) == len(desc) == len(weight)
        h = C.c_void_p()
        ctx._check(ctx.lib.vslam_bow_vocab_create(ctx.h, int(k), int(L), int(scoring), int(weighting), len(parent),
                                                  _ptr(parent), _ptr(is_leaf), _ptr(desc), _ptr(weight), C.byref(h)))
        return cls(ctx, h)

    def transform_batch(self, desc_sets, levelsup=4, cap=None):
        """voc->transform(features, bow_vector, feature_vector, levelsup) for a list of descriptor sets.
        -> list of (ids u32 ascending, vals f64) per image, and feat_word, feat_node, feat_weight (flat, per feature;
        feature f is in FeatureVector[feat_node[f]] iff feat_weight[f] > 0) with set_begin."""
        sets = [_desc_bytes(s) for s in desc_sets]
        begin = np.zeros(len(sets) + 1, np.int32)
        if sets:
            begin[1:] = np.cumsum([len(s) for s in sets])
        total = int(begin[-1])
        flat = np.ascontiguousarray(np.concatenate(sets)) if total else np.zeros((0, 32), np.uint8)
        if cap is None:
            cap = max(1, max((len(s) for s in sets), default=1))
        n = len(sets)
        ids = np.zeros((max(n, 1), cap), np.uint32)
        vals = np.zeros((max(n, 1), cap), np.float64)
        n_bow = np.zeros(max(n, 1), np.int32)
        fw = np.zeros(max(total, 1), np.uint32)
        fn = np.zeros(max(total, 1), np.uint32)
        fwt = np.zeros(max(total, 1), np.float64)
        self.ctx._check(self.ctx.lib.vslam_bow_transform_batch(self.ctx.h, self.h, _ptr(flat), _ptr(begin), n, int(levelsup),
                                                               _ptr(ids), _ptr(vals), _ptr(n_bow), cap, _ptr(fw), _ptr(fn),
                                                               _ptr(fwt)))
        vecs = [(ids[i, :n_bow[i]].copy(), vals[i, :n_bow[i]].copy()) for i in range(n)]
        return vecs, fw[:total], fn[:total], fwt[:total], begin

    def transform(self, desc, levelsup=4):
        """One image: (ids, vals), FeatureVector as {node id: [feature indices]} (FeatureVector.cpp:28-42)."""
        vecs, fw, fn, fwt, _ = self.transform_batch([desc], levelsup)
        fv = {}
        for f in np.nonzero(fwt > 0)[0]:
            fv.setdefault(int(fn[f]), []).append(int(f))
        return vecs[0], dict(sorted(fv.items()))

    def compute_bow_vector_batch(self, images, num_features=1500, levelsup=4, params=None, cap=None):
        """visnav::compute_bow_vector (keypoints.h:243-254) for [count, h, w] images: cv::ORB, then transform; the
        descriptors stay on the device.  -> list of (ids, vals), and a dict with kp [count, cap, 6], desc [count, cap,
        32], n_kp, feat_node, feat_weight."""
        images = _image_batch(images)
        count, h, w = images.shape
        prm = params or OrbParams(num_features)
        cap = cap or min(8192, prm.nfeatures + prm.nfeatures // 4 + 256)
        ids = np.zeros((max(count, 1), cap), np.uint32)
        vals = np.zeros((max(count, 1), cap), np.float64)
        n_bow = np.zeros(max(count, 1), np.int32)
        kp = np.zeros((max(count, 1), cap, 6), np.float32)
        desc = np.zeros((max(count, 1), cap, 32), np.uint8)
        n_kp = np.zeros(max(count, 1), np.int32)
        fn = np.zeros((max(count, 1), cap), np.uint32)
        fwt = np.zeros((max(count, 1), cap), np.float64)
        self.ctx._check(self.ctx.lib.vslam_compute_bow_vector_batch(
            self.ctx.h, self.h, _ptr(images), count, w, h, w, w * h, C.byref(prm), int(levelsup), _ptr(ids), _ptr(vals),
            _ptr(n_bow), cap, _ptr(kp), _ptr(desc), _ptr(n_kp), _ptr(fn), _ptr(fwt), cap))
        vecs = [(ids[i, :n_bow[i]].copy(), vals[i, :n_bow[i]].copy()) for i in range(count)]
        return vecs, dict(kp=kp[:count], desc=desc[:count], n_kp=n_kp[:count], feat_node=fn[:count], feat_weight=fwt[:count])

    def score(self, v1, v2):
        i1, x1 = np.ascontiguousarray(v1[0], np.uint32), np.ascontiguousarray(v1[1], np.float64)
        i2, x2 = np.ascontiguousarray(v2[0], np.uint32), np.ascontiguousarray(v2[1], np.float64)
        out = C.c_double(0.0)
        self.ctx._check(self.ctx.lib.vslam_bow_score(self.ctx.h, self.h, _ptr(i1), _ptr(x1), len(i1), _ptr(i2), _ptr(x2),
                                                     len(i2), C.byref(out)))
        return out.value

    def close(self):
        if getattr(self, "h", None):
            self.ctx.lib.vslam_bow_vocab_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class BowDatabase:
    """The keyframes' BowVectors on the device; query() = voc->score(query, entry) for every entry plus the number
    of shared words (loop_closure_utils.h:152-201, tracking.h:176-208)."""

    def __init__(self, ctx, voc):
        self.ctx, self.voc = ctx, voc
        h = C.c_void_p()
        ctx._check(ctx.lib.vslam_bow_db_create(ctx.h, voc.h, C.byref(h)))
        self.h = h

    def add(self, vec):
        ids, vals = np.ascontiguousarray(vec[0], np.uint32), np.ascontiguousarray(vec[1], np.float64)
        idx = C.c_int(-1)
        self.ctx._check(self.ctx.lib.vslam_bow_db_add(self.h, _ptr(ids), _ptr(vals), len(ids), C.byref(idx)))
        return idx.value

    def __len__(self):
        return int(self.ctx.lib.vslam_bow_db_size(self.h))

    def query(self, vec):
        ids, vals = np.ascontiguousarray(vec[0], np.uint32), np.ascontiguousarray(vec[1], np.float64)
        n = len(self)
        scores = np.zeros(max(n, 1), np.float64)
        common = np.zeros(max(n, 1), np.int32)
        self.ctx._check(self.ctx.lib.vslam_bow_db_query(self.ctx.h, self.h, _ptr(ids), _ptr(vals), len(ids), _ptr(scores),
                                                        _ptr(common)))
        return scores[:n], common[:n]

    def close(self):
        if getattr(self, "h", None):
            self.ctx.lib.vslam_bow_db_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class StereoPlan:
    """Upload / run / download split of the stereo front end (bench.py times run() alone)."""

    def __init__(self, ctx, w, h, max_pairs, cap):
        self.ctx, self.w, self.h, self.max_pairs, self.cap = ctx, w, h, max_pairs, cap
        h_ = C.c_void_p()
        ctx._check(ctx.lib.vslam_stereo_plan_create(ctx.h, w, h, max_pairs, cap, C.byref(h_)))
        self.p = h_

    def upload(self, images):
        images = np.asarray(images)
        assert images.dtype == np.uint8 and images.flags.c_contiguous and images.shape[1:] == (2, self.h, self.w)
        self.ctx._check(self.ctx.lib.vslam_stereo_plan_upload(self.p, _ptr(images), images.shape[0], self.w, self.w * self.h))

    def run(self, count, num_features=1500, rotate=True, threshold=70, dist_2_best=1.2):
        self.ctx._check(self.ctx.lib.vslam_stereo_plan_run(self.p, count, int(num_features), int(bool(rotate)),
                                                           int(threshold), float(dist_2_best)))

    def download(self, count):
        cap = self.cap
        out = dict(xy=np.empty((count * 2, cap, 2), np.float64), angles=np.empty((count * 2, cap), np.float64),
                   desc=np.empty((count * 2, cap, 8), np.uint32), n_kp=np.empty(count * 2, np.int32),
                   matches=np.empty((count, cap, 2), np.int32), n_matches=np.empty(count, np.int32))
        self.ctx._check(self.ctx.lib.vslam_stereo_plan_download(self.p, count, _ptr(out["xy"]), _ptr(out["angles"]),
                                                                _ptr(out["desc"]), _ptr(out["n_kp"]), _ptr(out["matches"]),
                                                                _ptr(out["n_matches"])))
        return out

    def set_epipolar(self, epipolar):
        """Fuse findInliersEssential behind the matcher (None switches it off)."""
        self.ctx._check(self.ctx.lib.vslam_stereo_plan_set_epipolar(self.p, C.byref(epipolar) if epipolar is not None else None))

    def download_inliers(self, count):
        inl = np.empty((count, self.cap, 2), np.int32)
        n = np.empty(count, np.int32)
        self.ctx._check(self.ctx.lib.vslam_stereo_plan_download_inliers(self.p, count, _ptr(inl), _ptr(n)))
        return inl, n

    def last_epipolar_ms(self):
        ms = C.c_float(0)
        self.ctx._check(self.ctx.lib.vslam_stereo_plan_last_epipolar_ms(self.p, C.byref(ms)))
        return ms.value

    def last_ms(self):
        ms = (C.c_float * 5)()
        self.ctx._check(self.ctx.lib.vslam_stereo_plan_last_ms(self.p, ms))
        return list(ms)

    def close(self):
        if getattr(self, "p", None):
            self.ctx.lib.vslam_stereo_plan_destroy(self.p)
            self.p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default = None


def default_context():
    global _default
    if _default is None:
        _default = Context(int(os.environ.get("LOCAL_RANK", "0")))
    return _default


# ---- the reference's free functions ------------------------------------------------------------

def detectKeypoints(img_raw, kd, num_features):
    """keypoints.h:133-150: clears kd, fills kd.corners."""
    kd.corners = default_context().detect(img_raw, num_features)
    kd.corner_angles = np.zeros(0, np.float64)
    kd.corner_descriptors = np.zeros((0, 8), np.uint32)


def computeAngles(img_raw, kd, rotate_features):
    """keypoints.h:152-189."""
    kd.corner_angles = default_context().angles(img_raw, kd.corners, rotate_features)


def computeDescriptors(img_raw, kd):
    """keypoints.h:191-221."""
    kd.corner_descriptors = default_context().describe(img_raw, kd.corners, kd.corner_angles)


def detectKeypointsAndDescriptors(img_raw, kd, num_features, rotate_features):
    """keypoints.h:223-229 -- one device round trip instead of three."""
    kd.corners, kd.corner_angles, kd.corner_descriptors = default_context().detect_describe(
        img_raw, num_features, rotate_features)


def matchDescriptors(corner_descriptors_1, corner_descriptors_2, matches, threshold, dist_2_best):
    """keypoints.h:323-369: matches.clear(), then (id1, id2) in ascending id1."""
    del matches[:]
    pairs = default_context().match(corner_descriptors_1, corner_descriptors_2, threshold, dist_2_best)
    matches.extend((int(a), int(b)) for a, b in pairs)


def findInliersEssential(kd1, kd2, cam1, cam2, E, epipolar_error_threshold, md):
    """matching_utils.h:63-88: md.inliers.clear(), then the matches whose |p0^T E p1| is not above the threshold.
    kd1/kd2: KeypointsData; cam1/cam2: Camera; md: object with .matches / .inliers lists of (id1, id2)."""
    del md.inliers[:]
    inl = default_context().find_inliers_essential(kd1.corners, kd2.corners, cam1, cam2, E, epipolar_error_threshold,
                                                   np.asarray(md.matches, np.int32).reshape(-1, 2))
    md.inliers.extend((int(a), int(b)) for a, b in inl)


def project_landmarks(current_pose, cam, landmarks, cam_z_threshold, projected_points, projected_track_ids):
    """vo_utils.h:48-81.  current_pose: 7 doubles (Sophus::SE3d::data()); landmarks: dict TrackId -> 3-vector
    (Landmark::p).  Clears and fills projected_points / projected_track_ids (in the dict's iteration order)."""
    del projected_points[:]
    del projected_track_ids[:]
    ids = list(landmarks.keys())
    xyz = np.array([landmarks[k] for k in ids], np.float64).reshape(-1, 3)
    xy, idx = default_context().project_landmarks(current_pose, cam, xyz, cam_z_threshold)
    projected_points.extend(xy[k].copy() for k in range(len(xy)))
    projected_track_ids.extend(ids[k] for k in idx)
