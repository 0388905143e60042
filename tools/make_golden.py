#!/usr/bin/env python3
"""Generates tests/golden/*.npz -- golden vectors produced by the REFERENCE side, in the build
container only (needs cv2 and /root/reference; neither exists on the GPU box):

  * cv2 (OpenCV, python wheel): goodFeaturesToTrack / cornerMinEigenVal, with
    cv2.setUseOptimized(False) (scalar operation order = the oracle's and the CUDA kernel's) and
    with the default AVX2 dispatch (what a stock visnav build runs).  SURVEY.md section 8c: the
    python wheel returns the same ordered corners as the vendored OpenCV 4.7.0-dev.
  * oracle/_ref/libvisnav_ref.so: the reference's own computeAngles / computeDescriptors /
    matchDescriptors compiled from /root/reference/include/visnav/keypoints.h (oracle/Makefile).
  * oracle/_ref/libvisnav_ref_lm.so: the reference's own find_matches_landmarks
    (/root/reference/include/visnav/vo_utils.h:83-167, with its vendored Eigen).
  * oracle/_ref/libvisnav_ref_geo.so: the reference's own camera models (camera_models.h),
    findInliersEssential (matching_utils.h:63-88) and project_landmarks (vo_utils.h:48-81).

The inputs are seeded synthetic images (visual-slam_b200/workloads.py); they are stored too, so the
fixtures do not depend on numpy's generator staying bit-stable.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cv2  # noqa: E402
import oracle  # noqa: E402
import visnav_b200  # noqa: E402,F401
from visnav_b200 import workloads as wl  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def gftt_cv2(img, n, optimized):
    cv2.setUseOptimized(optimized)
    pts = cv2.goodFeaturesToTrack(img, n, 0.01, 8, None, None, 3, False)
    cv2.setUseOptimized(True)
    return np.zeros((0, 2), np.float32) if pts is None else pts.reshape(-1, 2).astype(np.float32)


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = oracle.load_ref()
    assert ref is not None, "build oracle/_ref first (make -C oracle)"
    # ---- detection: corners (both dispatches) and the response map ----
    for name, (w, h, seed, nf) in {"gftt_320x240": (320, 240, 5, 300), "gftt_161x123": (161, 123, 9, 0),
                                  "gftt_752x480": (752, 480, 1234, 1500)}.items():
        img = wl.synth_image(seed, w, h)
        rec = dict(img=img if w < 700 else np.zeros((0, 0), np.uint8), seed=seed, w=w, h=h, num_features=nf,
                   corners_unopt=gftt_cv2(img, nf, False), corners_default=gftt_cv2(img, nf, True))
        if w < 200:
            cv2.setUseOptimized(False)
            rec["lam_unopt"] = cv2.cornerMinEigenVal(img, 3, ksize=3)
            cv2.setUseOptimized(True)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
        print(name, len(rec["corners_unopt"]), "corners; default dispatch identical:",
              np.array_equal(rec["corners_unopt"], rec["corners_default"]))
    # ---- orientation / description / stereo matching by the reference's own functions ----
    left, right = wl.synth_stereo_pair(7, 320, 240)
    xl = ref.detect(left, 400)
    xr = ref.detect(right, 400)
    al, ar = ref.angles(left, xl, True), ref.angles(right, xr, True)
    dl, dr = ref.describe(left, xl, al), ref.describe(right, xr, ar)
    m = ref.match(dl, dr, 70, 1.2)
    np.savez_compressed(os.path.join(OUT, "stereo_320x240.npz"), left=left, right=right, xy_l=xl, xy_r=xr, ang_l=al,
                        ang_r=ar, desc_l=dl, desc_r=dr, matches=m, desc_l_norot=ref.describe(left, xl, np.zeros(len(xl))))
    print("stereo", len(xl), len(xr), len(m))
    # ---- matching with heavy ties ----
    rng = np.random.default_rng(77)
    base = wl.random_descriptors(24, 3)
    d1 = wl.flip_bits(base[rng.integers(0, 24, 220)], 5, rng)
    d2 = wl.flip_bits(base[rng.integers(0, 24, 190)], 5, rng)
    rec = dict(d1=d1, d2=d2)
    for thr, ratio in [(70, 1.2), (256, 1.0), (30, 1.0), (300, 0.5)]:
        rec["m_%d_%s" % (thr, str(ratio).replace(".", "p"))] = ref.match(d1, d2, thr, ratio)
    np.savez_compressed(os.path.join(OUT, "match_ties.npz"), **rec)
    print("match_ties", {k: len(v) for k, v in rec.items() if k.startswith("m_")})


def landmarks():
    """find_matches_landmarks by the reference's own function (oracle/_ref/libvisnav_ref_lm.so)."""
    import oracle.oracle as oo
    sc = wl.landmark_scene(n_kp=220, n_landmarks=200, seed=11)
    rec = dict(kp_xy=sc[0], kp_desc=sc[1], proj_xy=sc[2], obs_offsets=sc[3], obs_desc=sc[4])
    for r, thr, ratio in [(20.0, 70, 1.2), (35.0, 256, 1.0), (12.5, 40, 1.5), (20.0, 70, 0.9)]:
        m = oo.match_landmarks_ref(*sc, r, thr, ratio)
        assert m is not None, "build oracle/_ref first (make -C oracle)"
        rec["m_%s_%d_%s" % (str(r).replace(".", "p"), thr, str(ratio).replace(".", "p"))] = m
    np.savez_compressed(os.path.join(OUT, "landmarks_220x200.npz"), **rec)
    print("landmarks", {k: len(v) for k, v in rec.items() if k.startswith("m_")})


def geometry():
    """Camera models, findInliersEssential and project_landmarks by the reference's own code
    (oracle/_ref/libvisnav_ref_geo.so: camera_models.h, matching_utils.h:50-88, vo_utils.h:48-81)."""
    import oracle.oracle as oo
    ref = oo.load_geometry_ref()
    assert ref is not None, "build oracle/_ref first (make -C oracle)"
    rng = np.random.default_rng(21)
    rec = {}
    for m, P in wl.CAMERAS.items():
        xyz = rng.normal(size=(160, 3)) * [3, 2, 3] + [0, 0, 5]
        xyz[0] = [0, 0, 2.0]                                   # on the optical axis (kb4 r == 0 branch)
        uv = rng.uniform([0, 0], [752, 480], size=(160, 2))
        uv[0] = P[2:4]                                         # principal point (kb4 ru == 0 branch)
        xy1, xy2, mt, pose, _ = wl.epipolar_scene(m, 300, seed=3)
        E = ref.compute_essential(pose)
        lp, lx = wl.landmark_cloud(600, 5)
        rec.update({m + "_param": np.array(P, np.float64), m + "_xyz": xyz, m + "_uv": uv,
                    m + "_project": ref.project(m, P, xyz), m + "_unproject": ref.unproject(m, P, uv),
                    m + "_xy1": xy1, m + "_xy2": xy2, m + "_matches": mt, m + "_pose01": pose, m + "_E": E,
                    m + "_lm_pose": lp, m + "_lm_xyz": lx})
        for thr in (5e-3, 1e-3):
            inl, err = ref.find_inliers_essential(xy1, xy2, (m, P), (m, P), E, thr, mt)
            rec[m + "_inliers_%g" % thr] = inl
            rec[m + "_err"] = err
        pxy, pidx = ref.project_landmarks(lp, (m, P), 752, 480, lx, 0.1)
        rec[m + "_proj_xy"], rec[m + "_proj_idx"] = pxy, pidx
        print("geometry", m, {k: len(v) for k, v in rec.items() if k.startswith(m + "_inliers")}, len(pidx))
    np.savez_compressed(os.path.join(OUT, "geometry.npz"), **rec)


def harris():
    """goodFeaturesToTrack(useHarrisDetector=True, k=0.04) and cornerHarris by cv2 (scalar and default dispatch):
    SURVEY 8a row a14 (the reference itself always passes false, keypoints.h:138-139)."""
    rec = {}
    for name, (w, h, seed, nf) in {"a": (161, 123, 9, 0), "b": (320, 240, 5, 300), "c": (752, 480, 1234, 1500)}.items():
        img = wl.synth_image(seed, w, h)
        rec["img_" + name] = img if w < 700 else np.zeros((0, 0), np.uint8)
        rec["meta_" + name] = np.array([w, h, seed, nf])
        for tag, opt in (("unopt", False), ("default", True)):
            cv2.setUseOptimized(opt)
            pts = cv2.goodFeaturesToTrack(img, nf, 0.01, 8, None, None, 3, True, 0.04)
            rec["corners_%s_%s" % (tag, name)] = (np.zeros((0, 2), np.float32) if pts is None else
                                                  pts.reshape(-1, 2).astype(np.float32))
            if w < 200 and not opt:
                rec["harris_unopt_" + name] = cv2.cornerHarris(img, 3, 3, 0.04)
            cv2.setUseOptimized(True)
        print("harris", name, len(rec["corners_unopt_" + name]), "default dispatch identical:",
              np.array_equal(rec["corners_unopt_" + name], rec["corners_default_" + name]))
    np.savez_compressed(os.path.join(OUT, "harris.npz"), **rec)


def euroc():
    """One REAL EuRoC stereo pair bundled with the reference (data/euroc_V1, 752x480): decoded here with cv2
    (the decoded bytes are stored, so the fixture does not depend on a JPEG decoder), corners by cv2's
    goodFeaturesToTrack (scalar and default dispatch), angles / descriptors / stereo matches by the reference's
    own functions -- the call sequence of slam.cpp:1125-1143 with num_features = 1500."""
    ref = oracle.load_ref()
    assert ref is not None
    d = "/root/reference/data/euroc_V1"
    stem = "1403715273262142976"
    rec = {}
    kd = {}
    for side, cam in (("l", 0), ("r", 1)):
        img = cv2.imread(os.path.join(d, "%s_%d.jpg" % (stem, cam)), cv2.IMREAD_GRAYSCALE)
        assert img is not None and img.shape == (480, 752)
        c0, c1 = gftt_cv2(img, 1500, False), gftt_cv2(img, 1500, True)
        xy = ref.detect(img, 1500)
        keep = (c0[:, 0] >= 19) & (c0[:, 0] < 752 - 19) & (c0[:, 1] >= 19) & (c0[:, 1] < 480 - 19)
        assert np.array_equal(xy, c0[keep].astype(np.float64)), "restated gftt differs from cv2 on a real image"
        ang = ref.angles(img, xy, True)
        desc = ref.describe(img, xy, ang)
        kd[side] = desc
        rec.update({"img_" + side: img, "corners_unopt_" + side: c0, "corners_default_" + side: c1, "xy_" + side: xy,
                    "ang_" + side: ang, "desc_" + side: desc})
    rec["matches"] = ref.match(kd["l"], kd["r"], 70, 1.2)
    np.savez_compressed(os.path.join(OUT, "euroc_pair.npz"), **rec)
    print("euroc", len(rec["xy_l"]), len(rec["xy_r"]), len(rec["matches"]), "default dispatch identical:",
          np.array_equal(rec["corners_unopt_l"], rec["corners_default_l"]),
          np.array_equal(rec["corners_unopt_r"], rec["corners_default_r"]))


def bow():
    """Bag of words by the reference's own DBoW2 (oracle/_ref/libvisnav_ref_bow.so): a vocabulary built by
    TemplatedVocabulary::create (hierarchical k-means++, k = 5, L = 3) on synthetic training descriptors and written
    by saveToTextFile (its trailing newline removed -- see vslam_bow_vocab_load_text), then, per scoring type, the
    BowVectors / FeatureVector nodes of four feature sets (levelsup 2 and 4) and all pairwise scores."""
    import oracle.oracle as oo
    path = os.path.join(OUT, "bow_vocab.txt")
    base = wl.random_descriptors(3000, 3)
    rng = np.random.default_rng(4)
    training = [wl.flip_bits(base[rng.integers(0, 3000, 80)], 30, rng) for _ in range(60)]
    n_words = oo.BowVocabRef.create_save(training, 5, 3, 0, 0, path)
    txt = open(path).read().rstrip("\n")
    open(path, "w").write(txt)
    sets = [wl.flip_bits(base[rng.integers(0, 3000, n)], 34, rng) for n in (500, 431, 64, 1)]
    rec = {"n_words": n_words, "set_sizes": np.array([len(s) for s in sets])}
    for i, s_ in enumerate(sets):
        rec["set%d" % i] = s_
    lines = txt.split("\n")
    for scoring in (0, 1, 2, 4, 5):
        for weighting in (0, 1, 2, 3):
            tmp = os.path.join("/tmp", "bow_vocab_%d_%d.txt" % (scoring, weighting))
            k_, l_ = lines[0].split()[:2]
            open(tmp, "w").write("\n".join(["%s %s %d %d" % (k_, l_, scoring, weighting)] + lines[1:]))
            r = oo.BowVocabRef.from_text(tmp)
            vecs = []
            for i, s_ in enumerate(sets):
                for lu in (2, 4):
                    ids, vals, fv = r.transform(s_, lu)
                    key = "s%d_w%d_set%d_l%d_" % (scoring, weighting, i, lu)
                    rec[key + "ids"], rec[key + "vals"], rec[key + "fv"] = ids, vals, fv
                vecs.append((ids, vals))
                if scoring == 0 and weighting == 0:
                    rec["words_set%d" % i] = r.words(s_)
            rec["s%d_w%d_scores" % (scoring, weighting)] = np.array([[r.score(a, b) for b in vecs] for a in vecs])
            r.close()
    np.savez_compressed(os.path.join(OUT, "bow.npz"), **rec)
    print("bow: words", n_words, "vector sizes", [len(rec["s0_w0_set%d_l4_ids" % i]) for i in range(4)])


def orb():
    """cv::ORB as visnav::compute_bow_vector configures it (keypoints.h:249), by cv2 itself: keypoints and descriptors
    of a synthetic image (workloads.synth_image(1234), not stored) and of the real EuRoC image already in
    euroc_pair.npz, for 1500 and 300 features; sorted by (octave, y, x) -- the reference's order inside a level is
    std::nth_element's.  Also the Gaussian kernel and one blurred image (the float separable filter)."""
    rec = {"k7": cv2.getGaussianKernel(7, 2, cv2.CV_32F).ravel().astype(np.float32)}
    imgs = {"synth": wl.synth_image(1234, 752, 480), "euroc": np.load(os.path.join(OUT, "euroc_pair.npz"))["img_l"]}
    for name, img in imgs.items():
        for nf in (1500, 300):
            o = cv2.ORB_create(nf, 1.2, 8, 19, 0, 2, cv2.ORB_FAST_SCORE)
            kps, d = o.detectAndCompute(img, None)
            order = sorted(range(len(kps)), key=lambda i: (kps[i].octave, kps[i].pt[1], kps[i].pt[0]))
            rec["%s_%d_kp" % (name, nf)] = np.array([[kps[i].pt[0], kps[i].pt[1], kps[i].size, kps[i].angle, kps[i].response,
                                                      kps[i].octave] for i in order], np.float32)
            rec["%s_%d_desc" % (name, nf)] = d[order]
            print("orb", name, nf, len(kps))
    np.savez_compressed(os.path.join(OUT, "orb.npz"), **rec)


ORB_PARAM_CASES = [  # (seed, w, h, nfeatures, scale, nlevels, edge, fast)
    (2001, 333, 251, 300, 1.5, 4, 19, 10), (2002, 640, 200, 1000, 1.1, 8, 25, 20), (2003, 161, 123, 50, 2.0, 3, 19, 5),
    (2004, 480, 360, 2500, 1.3, 6, 31, 40), (2005, 257, 511, 600, 1.2, 1, 19, 20), (2006, 97, 89, 300, 1.2, 8, 19, 7)]


def orb_params():
    """cv2.ORB away from the reference's parameter set (other scale factors, level counts, FAST thresholds, edge
    thresholds, odd sizes, levels that come out empty): the entry point takes them, so the restatement is pinned there
    too.  Images are workloads.synth_image(seed, w, h), not stored."""
    rec = {"cases": np.array(ORB_PARAM_CASES, np.float64)}
    for i, (seed, w, h, nf, sf, nl, et, ft) in enumerate(ORB_PARAM_CASES):
        img = wl.synth_image(seed, w, h)
        o = cv2.ORB_create(nf, sf, nl, et, 0, 2, cv2.ORB_FAST_SCORE, 31, ft)
        kps, d = o.detectAndCompute(img, None)
        d = np.zeros((0, 32), np.uint8) if d is None else d
        order = sorted(range(len(kps)), key=lambda i: (kps[i].octave, kps[i].pt[1], kps[i].pt[0]))
        rec["kp%d" % i] = np.array([[kps[j].pt[0], kps[j].pt[1], kps[j].size, kps[j].angle, kps[j].response,
                                     kps[j].octave] for j in order], np.float32).reshape(-1, 6)
        rec["desc%d" % i] = d[order]
        print("orb_params", i, len(kps))
    np.savez_compressed(os.path.join(OUT, "orb_params.npz"), **rec)


def response_large():
    """Response maps at the benchmark sizes (752x480, 1920x1080) from the reference's OWN vendored OpenCV
    (oracle/_ref/libvisnav_ref.so links thirdparty/opencv core+imgproc): cv::cornerMinEigenVal under the scalar
    dispatch (cv::setUseOptimized(false)) as sha256 of the float32 bits -- the maps themselves would be 1.4 / 8 MB
    each -- plus the ordered goodFeaturesToTrack corners under both dispatches.  The images are the seeded synthetic
    ones; their sha256 is stored so that a drifting generator is told apart from a drifting kernel.  The 1080p seeds
    are ones where OpenCV's running column sum and a direct three-row sum differ (what the fixture exists to pin)."""
    import hashlib
    ref = oracle.load_ref()
    assert ref is not None and ref.has_opencv, "oracle/_ref must be built against the vendored OpenCV"
    port = oracle.load_oracle()
    ref.set_num_threads(1)
    rec = {}
    names = []
    for (w, h, seed, nf) in [(752, 480, 1234, 1500), (752, 480, 1, 1500), (1920, 1080, 300, 4000), (1920, 1080, 301, 4000),
                             (1920, 1080, 302, 4000), (1920, 1080, 303, 4000)]:
        img = wl.synth_image(seed, w, h)
        key = "%dx%d_%d" % (w, h, seed)
        names.append(key)
        ref.set_use_optimized(False)
        lam = ref.cv_response(img)
        c_scalar = ref.cv_gftt(img, nf)
        ref.set_use_optimized(True)
        lam_def = ref.cv_response(img)
        c_default = ref.cv_gftt(img, nf)
        port.set_box_direct(1)
        lam_direct = port.response(img)
        port.set_box_direct(0)
        rec[key + "_img_sha"] = np.frombuffer(hashlib.sha256(img.tobytes()).digest(), np.uint8)
        rec[key + "_lam_sha"] = np.frombuffer(hashlib.sha256(lam.tobytes()).digest(), np.uint8)
        rec[key + "_lam_max"] = np.float32(lam.max())
        rec[key + "_corners_scalar"] = c_scalar
        rec[key + "_corners_default"] = c_default
        rec[key + "_n_direct_sum_differs"] = np.int32((lam.view(np.uint32) != lam_direct.view(np.uint32)).sum())
        rec[key + "_n_default_dispatch_differs"] = np.int32((lam.view(np.uint32) != lam_def.view(np.uint32)).sum())
        print(key, "direct-sum pixels differing:", int(rec[key + "_n_direct_sum_differs"]), " default-dispatch pixels differing:",
              int(rec[key + "_n_default_dispatch_differs"]), " corners equal across dispatches:",
              np.array_equal(c_scalar, c_default), " set-equal:",
              set(map(tuple, c_scalar)) == set(map(tuple, c_default)))
    rec["names"] = np.array(names)
    rec["opencv"] = np.array(ref.gftt_impl)
    np.savez_compressed(os.path.join(OUT, "response_large.npz"), **rec)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "response_large":
        response_large()
    elif len(sys.argv) > 1 and sys.argv[1] == "orb":
        orb()
    elif len(sys.argv) > 1 and sys.argv[1] == "orb_params":
        orb_params()
    elif len(sys.argv) > 1 and sys.argv[1] == "bow":
        bow()
    elif len(sys.argv) > 1 and sys.argv[1] == "landmarks":
        landmarks()
    elif len(sys.argv) > 1 and sys.argv[1] == "geometry":
        geometry()
    elif len(sys.argv) > 1 and sys.argv[1] == "euroc":
        euroc()
    elif len(sys.argv) > 1 and sys.argv[1] == "harris":
        harris()
    else:
        main()
        response_large()
        landmarks()
        geometry()
        euroc()
        harris()
        bow()
        orb()
        orb_params()
