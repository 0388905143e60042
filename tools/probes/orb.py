#!/usr/bin/env python3
"""Timing probe of visnav::compute_bow_vector on the GPU: cv::ORB (1500 features, 8 levels) + vocabulary
transform (k = 10, L = 6 tree) for a batch of 752x480 images through the C ABI, pinned host buffers, copies inside the
timed region.  CPU beside it: cv2.ORB (the reference's detector) + the C restatement of the DBoW2 transform, one
thread and all host threads.  Writes one JSON line."""
import argparse
import ctypes as C
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np  # noqa: E402
import visnav_b200  # noqa: E402
from oracle import oracle as oo  # noqa: E402
from visnav_b200 import workloads as wl  # noqa: E402
from visnav_b200.keypoints import BowVocabulary, OrbParams  # noqa: E402

try:
    import cv2
except Exception:  # pragma: no cover
    cv2 = None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=256)
    ap.add_argument("--levels", type=int, default=6)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    ctx = visnav_b200.Context(0)
    voc = wl.synth_vocabulary(10, a.levels, 31, flips=(110, 80, 56, 40, 28, 20))
    v = BowVocabulary.from_arrays(ctx, 10, a.levels, 0, 0, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    w, h, nf, cap = 752, 480, 1500, 2048
    pairs = wl.synth_stereo_batch((a.images + 1) // 2, w, h, first=5000)
    imgs = ctx.host_array((a.images, h, w), np.uint8)
    imgs[:] = pairs.reshape(-1, h, w)[:a.images]
    ids, vals = ctx.host_array((a.images, cap), np.uint32), ctx.host_array((a.images, cap), np.float64)
    n_bow, n_kp = ctx.host_array((a.images,), np.int32), ctx.host_array((a.images,), np.int32)
    kp, desc = ctx.host_array((a.images, cap, 6), np.float32), ctx.host_array((a.images, cap, 32), np.uint8)
    prm = OrbParams(nf)
    P = lambda x: None if x is None else x.ctypes.data_as(C.c_void_p)  # noqa: E731

    def bow_call(count, with_kp):
        ctx._check(ctx.lib.vslam_compute_bow_vector_batch(ctx.h, v.h, P(imgs), count, w, h, w, w * h, C.byref(prm), 4, P(ids),
                                                          P(vals), P(n_bow), cap, P(kp) if with_kp else None,
                                                          P(desc) if with_kp else None, P(n_kp), None, None, cap))

    def orb_call(count):
        ctx._check(ctx.lib.vslam_orb_detect_compute_batch(ctx.h, P(imgs), count, w, h, w, w * h, C.byref(prm), P(kp), P(desc),
                                                          P(n_kp), cap))

    def timed(f, *args):
        f(*args)
        ts = []
        for _ in range(a.reps):
            t = time.perf_counter()
            f(*args)
            ts.append(time.perf_counter() - t)
        return float(np.median(ts))

    out = dict(workload="compute_bow_vector: cv::ORB(1500, 1.2, 8, 19, 0, 2, FAST_SCORE) + transform(levelsup 4), 752x480",
               images=a.images, tree=dict(k=10, L=a.levels, words=v.n_words))
    t = timed(bow_call, a.images, False)
    out["bow_vector_images_per_s"] = a.images / t
    out["bow_vector_ms_per_call"] = 1e3 * t
    t = timed(orb_call, a.images)
    out["orb_images_per_s_keypoints_and_descriptors_downloaded"] = a.images / t
    out["one_image_bow_vector_ms"] = 1e3 * timed(bow_call, 1, False)
    out["keypoints_per_image"] = float(n_kp[:a.images].mean())
    out["h2d_bytes_per_image"] = w * h
    # CPU: the reference's detector (cv2.ORB) + the restated DBoW2 transform
    o = oo.BowVocab.from_arrays(10, a.levels, 0, 0, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    bow_call(a.images, True)
    if cv2 is not None:
        cv2.setNumThreads(1)

        def cpu_one(i):
            orb = cv2.ORB_create(nf, 1.2, 8, 19, 0, 2, cv2.ORB_FAST_SCORE)
            _, d = orb.detectAndCompute(imgs[i], None)
            return d
        n1 = min(a.images, 16)
        t = time.perf_counter()
        ds = [cpu_one(i) for i in range(n1)]
        t_orb = time.perf_counter() - t
        t = time.perf_counter()
        bv = [o.transform(d)[:2] for d in ds]
        t_tr = time.perf_counter() - t
        out["cpu_1_thread"] = dict(kind="reference (cv2.ORB) + port (DBoW2 transform)", images_per_s=n1 / (t_orb + t_tr),
                                   orb_ms_per_image=1e3 * t_orb / n1, transform_ms_per_image=1e3 * t_tr / n1)
        cores = os.cpu_count() or 1
        nall = min(a.images, 8 * cores)
        t = time.perf_counter()
        with ThreadPoolExecutor(cores) as ex:
            list(ex.map(cpu_one, range(nall)))
        out["cpu_all_threads"] = dict(kind="reference (cv2.ORB only)", cores=cores, images_per_s=nall / (time.perf_counter() - t))
        # parity on this workload: BowVectors identical where the descriptor sets are
        same = 0
        for i in range(n1):
            g = (ids[i, :n_bow[i]], vals[i, :n_bow[i]])
            same += int(np.array_equal(g[0], bv[i][0]) and np.array_equal(g[1].view(np.uint64), bv[i][1].view(np.uint64)))
        out["bow_vectors_bit_identical_to_cpu"] = "%d of %d" % (same, n1)
    os.makedirs("gpurun_out", exist_ok=True)
    print(json.dumps(out))
    open("gpurun_out/orb_probe.json", "w").write(json.dumps(out) + "\n")


if __name__ == "__main__":
    main()
