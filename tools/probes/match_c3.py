"""Matcher throughput on a config-3 sample (780 jobs of 1500 x 1500) and on 2000 stereo-shaped jobs."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import DescStore
ctx = visnav_b200.Context(0)
kf = wl.keyframe_descriptors(40, 1500, 25)
store = DescStore(ctx, kf)
jobs = wl.keyframe_pair_jobs(40)
best = 1e9
for rep in range(5):
    _, counts = ctx.match_batch(store, jobs, 70, 1.2, want_matches=False)
    best = min(best, ctx.last_match_ms())
print("c3 sample: %.3f ms  %.3e dist/s  mean matches %.1f" % (best, len(jobs) * 2.25e6 / (best * 1e-3), counts.mean()))
kf2 = [k[:1312] for k in kf]
store2 = DescStore(ctx, kf2)
jobs2 = np.stack([np.arange(2000) % 40, (np.arange(2000) * 7 + 1) % 40], 1).astype(np.int32)
best = 1e9
for rep in range(5):
    _, counts = ctx.match_batch(store2, jobs2, 70, 1.2, want_matches=False)
    best = min(best, ctx.last_match_ms())
print("2000 jobs of 1312^2: %.3f ms  %.3e dist/s" % (best, 2000 * 1312.0 ** 2 / (best * 1e-3)))
