#!/usr/bin/env python3
"""Randomised parity sweep of the cv::ORB kernels and the vocabulary transform against the CPU oracles:
random image sizes, contents (synthetic scenes, noise, quantised plateaus, flat), feature budgets, pyramid depths,
scale factors, FAST thresholds.  Every image: keypoint set + fields bit-exact, descriptor bits counted, BowVector
compared bit for bit.  Writes gpurun_out/fuzz_orb.json.   python tools/fuzz_orb.py [--seconds 120]"""
import argparse
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import visnav_b200  # noqa: E402
from oracle import oracle as oo  # noqa: E402
from visnav_b200 import workloads as wl  # noqa: E402
from visnav_b200.keypoints import BowVocabulary, OrbParams  # noqa: E402


def make_image(rng, w, h):
    kind = rng.integers(0, 5)
    if kind == 0:
        return wl.synth_image(int(rng.integers(0, 1 << 30)), w, h)
    if kind == 1:
        return rng.integers(0, 256, (h, w), dtype=np.uint8)
    if kind == 2:   # few grey levels: plateaus, massive score ties
        return (rng.integers(0, 4, (h // 6 + 1, w // 6 + 1), dtype=np.uint8).repeat(6, 0).repeat(6, 1)[:h, :w] * 70).astype(np.uint8)
    if kind == 3:
        img = wl.synth_image(int(rng.integers(0, 1 << 30)), w, h)
        return (img // 32 * 32).astype(np.uint8)
    return np.full((h, w), int(rng.integers(0, 256)), np.uint8)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=120)
    a = ap.parse_args()
    ctx = visnav_b200.Context(0)
    voc = wl.synth_vocabulary(9, 4, 77, prune=0.1, stop=0.1)
    v = BowVocabulary.from_arrays(ctx, 9, 4, 0, 0, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    o = oo.BowVocab.from_arrays(9, 4, 0, 0, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
    rng = np.random.default_rng(2024)
    rep = dict(images=0, keypoints=0, keypoint_set_or_field_mismatches=0, descriptor_bits_differing=0, descriptors=0,
               bow_vectors=0, bow_vectors_differing=0, limit_errors=0, failures=[])
    t0 = time.time()
    while time.time() - t0 < a.seconds:
        w, h = int(rng.integers(40, 1300)), int(rng.integers(40, 900))
        count = int(rng.integers(1, 5))
        prm = OrbParams(int(rng.choice([0, 50, 300, 1500, 4000])), float(rng.choice([1.2, 1.2, 1.1, 1.5, 2.0])),
                        int(rng.integers(1, 9)), int(rng.choice([19, 19, 24, 31])), int(rng.choice([20, 20, 5, 40, 100])))
        imgs = np.stack([make_image(rng, w, h) for _ in range(count)])
        with ThreadPoolExecutor(4) as ex:
            want = list(ex.map(lambda im: oo.orb_detect_compute(im, prm.nfeatures, prm.nlevels, prm.scale_factor,
                                                                prm.edge_threshold, prm.fast_threshold), imgs))
        cap = min(8192, max(len(k) for k, _ in want) + 8)
        try:
            vecs, extra = v.compute_bow_vector_batch(imgs, params=prm, cap=cap, levelsup=2)
        except visnav_b200.VslamError as e:
            if max(len(k) for k, _ in want) > 8192 or "same size" in str(e) or "8192" in str(e) or "too small" in str(e):
                rep["limit_errors"] += 1    # documented limits (more than 8192 keypoints, two levels of one size)
                continue
            rep["failures"].append(dict(w=w, h=h, nf=prm.nfeatures, levels=prm.nlevels, scale=prm.scale_factor,
                                        edge=prm.edge_threshold, fast=prm.fast_threshold, error=str(e)))
            break
        for i in range(count):
            okp, odesc = want[i]
            n = int(extra["n_kp"][i])
            kp, desc = extra["kp"][i, :n], extra["desc"][i, :n]
            rep["images"] += 1
            rep["keypoints"] += len(okp)
            if kp.shape != okp.shape or not np.array_equal(kp.view(np.uint32), okp.view(np.uint32)):
                rep["keypoint_set_or_field_mismatches"] += 1
                rep["failures"].append(dict(w=w, h=h, nf=prm.nfeatures, levels=prm.nlevels, scale=prm.scale_factor,
                                            edge=prm.edge_threshold, fast=prm.fast_threshold, got=int(n), want=len(okp)))
                continue
            bits = int(np.unpackbits(desc ^ odesc).sum())
            rep["descriptor_bits_differing"] += bits
            rep["descriptors"] += len(okp)
            if bits == 0:
                ids, vals, *_ = o.transform(odesc, 2)
                rep["bow_vectors"] += 1
                if not (np.array_equal(vecs[i][0], ids) and np.array_equal(vecs[i][1].view(np.uint64), vals.view(np.uint64))):
                    rep["bow_vectors_differing"] += 1
    rep["failures"] = rep["failures"][:20]
    rep["seconds"] = time.time() - t0
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rep, open("gpurun_out/fuzz_orb.json", "w"), indent=1)
    print(json.dumps(rep))
    sys.exit(1 if rep["keypoint_set_or_field_mismatches"] or rep["bow_vectors_differing"] or rep["failures"] else 0)


if __name__ == "__main__":
    main()
