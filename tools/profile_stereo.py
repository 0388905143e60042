#!/usr/bin/env python3
"""Short stereo-plan run for ncu: PROF_PAIRS pairs (default 256) resident in HBM, PROF_REPS runs
(default 3) of response -> select -> orient_describe -> match (5 launches per run incl. the init
kernel).  Also prints pinned H2D / D2H bandwidth (plain run only; set PROF_PCIE=1)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import StereoPlan

npairs = int(os.environ.get("PROF_PAIRS", "256"))
reps = int(os.environ.get("PROF_REPS", "3"))
w, h = int(os.environ.get("PROF_W", "752")), int(os.environ.get("PROF_H", "480"))
nfeat = int(os.environ.get("PROF_FEAT", "1500"))
ctx = visnav_b200.Context(0)
distinct = min(npairs, 64)
base = wl.synth_stereo_batch(distinct, w, h)
imgs = ctx.host_array((npairs, 2, h, w), np.uint8)
for k in range(0, npairs, distinct):
    m = min(distinct, npairs - k)
    imgs[k:k + m] = base[:m]
plan = StereoPlan(ctx, w, h, npairs, nfeat)
plan.upload(imgs)
for _ in range(reps):
    plan.run(npairs, nfeat)
    ms = plan.last_ms()
print("pairs %d  total %.3f ms  response %.3f  select %.3f  describe %.3f  match %.3f  -> %.0f pairs/s" %
      (npairs, ms[0], ms[1], ms[2], ms[3], ms[4], npairs / (ms[0] * 1e-3)))
res = plan.download(npairs)
print("mean kp %.1f  mean matches %.1f" % (res["n_kp"].mean(), res["n_matches"].mean()))
if os.environ.get("PROF_PCIE"):
    for _ in range(3):
        t0 = time.perf_counter(); plan.upload(imgs); ctx.sync(); t1 = time.perf_counter()
    print("H2D pinned: %.1f GB/s (%.1f MB)" % (imgs.nbytes / (t1 - t0) / 1e9, imgs.nbytes / 1e6))
    for _ in range(3):
        t0 = time.perf_counter(); res = plan.download(npairs); t1 = time.perf_counter()
    nb = sum(v.nbytes for v in res.values())
    print("D2H (pageable numpy targets): %.1f GB/s (%.1f MB)" % (nb / (t1 - t0) / 1e9, nb / 1e6))
