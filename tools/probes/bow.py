#!/usr/bin/env python3
"""Timing probe of the bag-of-words kernels (DBoW2 half) on a tree of ORBvoc's shape (k = 10, L = 6).
  transform: `images` x 1500 features per call through the C ABI (host buffers, copies included)
  query:     one BowVector against `keyframes` stored vectors
CPU beside it: the reference's own DBoW2 (oracle/_ref) where it travelled, else the C restatement; one thread.
Writes one JSON line."""
import argparse
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np  # noqa: E402
import visnav_b200  # noqa: E402
from oracle import oracle as oo  # noqa: E402
from visnav_b200 import workloads as wl  # noqa: E402
from visnav_b200.keypoints import BowDatabase, BowVocabulary  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=256)
    ap.add_argument("--keyframes", type=int, default=4000)
    ap.add_argument("--levels", type=int, default=6)
    a = ap.parse_args()
    ctx = visnav_b200.Context(0)
    voc = wl.synth_vocabulary(10, a.levels, 31, flips=(110, 80, 56, 40, 28, 20))
    v = BowVocabulary.from_arrays(ctx, 10, a.levels, 0, 0, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    rng = np.random.default_rng(1)
    words = np.nonzero(voc["is_leaf"])[0]
    scene = words[rng.integers(0, len(words), 6000)]          # the words a "place" shows
    sets = [wl.flip_bits(np.ascontiguousarray(voc["desc"][scene[rng.integers(0, 6000, 1500)]]), 8, rng)
            for _ in range(a.images)]
    out = dict(tree=dict(k=10, L=a.levels, nodes=v.n_nodes, words=v.n_words), images=a.images, features=1500)
    vecs, *_ = v.transform_batch(sets)
    # the C call alone, pinned host buffers on both sides (copies inside the timed region)
    total, cap = 1500 * a.images, 1500
    flat = ctx.host_array((total, 32), np.uint8)
    flat[:] = np.concatenate(sets).view(np.uint8).reshape(-1, 32)
    begin = (1500 * np.arange(a.images + 1)).astype(np.int32)
    ids, vals = ctx.host_array((a.images, cap), np.uint32), ctx.host_array((a.images, cap), np.float64)
    n_bow = ctx.host_array((a.images,), np.int32)
    fw, fn = ctx.host_array((total,), np.uint32), ctx.host_array((total,), np.uint32)
    P = lambda x: x.ctypes.data_as(C.c_void_p)  # noqa: E731

    def call(count):
        ctx._check(ctx.lib.vslam_bow_transform_batch(ctx.h, v.h, P(flat), P(begin), count, 4, P(ids), P(vals), P(n_bow), cap,
                                                     P(fw), P(fn), None))
    ts = []
    for _ in range(7):
        t = time.perf_counter()
        call(a.images)
        ts.append(time.perf_counter() - t)
    assert all(np.array_equal(ids[i, :n_bow[i]], vecs[i][0]) for i in range(a.images))
    out["transform_ms_per_call"] = 1e3 * float(np.median(ts))
    out["transform_images_per_s"] = a.images / float(np.median(ts))
    ts = []
    for _ in range(7):
        t = time.perf_counter()
        call(1)
        ts.append(time.perf_counter() - t)
    out["transform_one_image_ms"] = 1e3 * float(np.median(ts))
    db = BowDatabase(ctx, v)
    for i in range(a.keyframes):
        db.add(vecs[i % len(vecs)])
    db.query(vecs[0])
    ts = []
    for i in range(10):
        t = time.perf_counter()
        scores, common = db.query(vecs[i % len(vecs)])
        ts.append(time.perf_counter() - t)
    out["query_ms"] = 1e3 * float(np.median(ts))
    out["query_scores_per_s"] = a.keyframes / float(np.median(ts))
    # CPU: one thread
    path = "/tmp/bow_probe_vocab.txt"
    r = None
    if oo.BowVocabRef.lib() is not None and a.levels <= 4:
        wl.write_vocabulary_text(path, voc)
        r = oo.BowVocabRef.from_text(path)
    o = r if r is not None else oo.BowVocab.from_arrays(10, a.levels, 0, 0, voc["parent"], voc["is_leaf"], voc["desc"],
                                                       voc["weight"])
    n_cpu = min(a.images, 32)
    t = time.perf_counter()
    cpu_vecs = [o.transform(s)[:2] for s in sets[:n_cpu]]
    dt = time.perf_counter() - t
    out["cpu"] = dict(kind=o.kind, threads=1, transform_images_per_s=n_cpu / dt)
    assert all(np.array_equal(c[0], g[0]) and np.array_equal(c[1].view(np.uint64), g[1].view(np.uint64))
               for c, g in zip(cpu_vecs, vecs))
    t = time.perf_counter()
    cpu_scores = [o.score(vecs[0], vecs[i % len(vecs)]) for i in range(min(a.keyframes, 2000))]
    dt = time.perf_counter() - t
    out["cpu"]["scores_per_s"] = min(a.keyframes, 2000) / dt
    s0, _ = db.query(vecs[0])
    assert np.array_equal(np.array(cpu_scores).view(np.uint64), s0[:len(cpu_scores)].view(np.uint64))
    out["parity"] = "transform and scores bit-identical to the CPU %s on this workload" % o.kind
    os.makedirs("gpurun_out", exist_ok=True)
    print(json.dumps(out))
    open("gpurun_out/bow_probe.json", "w").write(json.dumps(out) + "\n")


if __name__ == "__main__":
    main()
