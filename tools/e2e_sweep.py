#!/usr/bin/env python3
"""Sweeps the chunk size / pipeline depth of vslam_stereo_batch (pinned host buffers in and out)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import VSLAM_OPT_BATCH_CHUNK, VSLAM_OPT_BATCH_DEPTH

n_pairs, cap, W, H = int(os.environ.get("SWEEP_PAIRS", "2000")), 1500, 752, 480
ctx = visnav_b200.Context(0)
base = wl.synth_stereo_batch(64, W, H)
imgs = ctx.host_array((n_pairs, 2, H, W), np.uint8)
for k in range(0, n_pairs, 64):
    m = min(64, n_pairs - k)
    imgs[k:k + m] = base[:m]
out = dict(xy=ctx.host_array((n_pairs * 2, cap, 2), np.float64), angles=ctx.host_array((n_pairs * 2, cap), np.float64),
           desc=ctx.host_array((n_pairs * 2, cap, 8), np.uint32), n_kp=ctx.host_array((n_pairs * 2,), np.int32),
           matches=ctx.host_array((n_pairs, cap, 2), np.int32), n_matches=ctx.host_array((n_pairs,), np.int32))
for depth, chunk in [(4, 128), (4, 64), (4, 96), (4, 192), (4, 256)]:
    ctx.set_option(VSLAM_OPT_BATCH_CHUNK, chunk)
    ctx.set_option(VSLAM_OPT_BATCH_DEPTH, depth)
    ctx.stereo_batch(imgs, 1500, True, 70, 1.2, cap, out=out)
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter()
        ctx.stereo_batch(imgs, 1500, True, 70, 1.2, cap, out=out)
        best = min(best, time.perf_counter() - t0)
    print("depth %d chunk %4d: %.1f ms  %.0f pairs/s  (H2D alone would be %.1f ms at 55 GB/s)" %
          (depth, chunk, best * 1e3, n_pairs / best, imgs.nbytes / 55e9 * 1e3), flush=True)
