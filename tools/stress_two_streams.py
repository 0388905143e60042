#!/usr/bin/env python3
"""Stress of the two-stream stereo plan (vslam_stereo_plan_run, VSLAM_OPT_TWO_STREAMS): random batch sizes at and above the
switch-over, the same batch on two streams and on one, every output compared.  Prints one JSON line; exit 1 on a difference."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import StereoPlan, VSLAM_OPT_TWO_STREAMS

iters = int(os.environ.get("STRESS_ITERS", "30"))
rng = np.random.default_rng(int(os.environ.get("STRESS_SEED", "3")))
ctx = visnav_b200.Context(0)
base = wl.synth_stereo_batch(24, 320, 240, workers=1)
bad, runs = [], 0
plan = StereoPlan(ctx, 320, 240, 900, 600)
for it in range(iters):
    n = int(rng.integers(4 * 148, 901))
    imgs = np.ascontiguousarray(base[rng.integers(0, 24, n)])
    plan.upload(imgs)
    outs = []
    for two in (1, 0):
        ctx.set_option(VSLAM_OPT_TWO_STREAMS, two)
        for _ in range(2):          # back-to-back runs on the same buffers as well
            plan.run(n, 600, True, 70, 1.2)
        outs.append(plan.download(n))
    runs += 1
    a, b = outs
    ok = np.array_equal(a["n_kp"], b["n_kp"]) and np.array_equal(a["n_matches"], b["n_matches"])
    if ok:
        for k in range(n):
            m = a["n_matches"][k]
            ok = ok and np.array_equal(a["matches"][k, :m], b["matches"][k, :m])
            for s in (2 * k, 2 * k + 1):
                q = a["n_kp"][s]
                ok = ok and np.array_equal(a["xy"][s, :q], b["xy"][s, :q]) and np.array_equal(a["desc"][s, :q], b["desc"][s, :q])
    if not ok:
        bad.append(dict(iteration=it, pairs=n))
ctx.set_option(VSLAM_OPT_TWO_STREAMS, 1)
plan.close()
print(json.dumps(dict(iterations=runs, failures=bad, mma_timeouts=ctx.mma_timeouts())))
sys.exit(1 if bad or ctx.mma_timeouts() else 0)
