"""One stereo pair (BASELINE config 1, the way slam.cpp:1125-1143 issues the path): device stage times and the wall time of
one vslam_stereo_batch call with pageable and with pinned host buffers."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import StereoPlan
ctx = visnav_b200.Context(0)
if os.environ.get("RESPONSE_BAND"):      # e.g. 65535: full-height boxes (OpenCV's running column sum, bit-exact)
    from visnav_b200.keypoints import VSLAM_OPT_RESPONSE_BAND
    ctx.set_option(VSLAM_OPT_RESPONSE_BAND, int(os.environ["RESPONSE_BAND"]))
pl, pr = wl.synth_stereo_pair(0)
one = np.ascontiguousarray(np.stack([pl, pr])[None])
plan = StereoPlan(ctx, 752, 480, 1, 1500)
plan.upload(one)
for _ in range(5):
    plan.run(1)
    ms = plan.last_ms()
plan.close()
out = {"device_ms": {"total": float(ms[0]), "response": float(ms[1]), "select": float(ms[2]), "describe": float(ms[3]), "match": float(ms[4])}}


def timed(fn, n=200):
    for _ in range(20):
        fn()
    t = []
    for _ in range(n):
        t0 = time.perf_counter(); fn(); t.append(time.perf_counter() - t0)
    return 1e3 * float(np.median(t)), 1e3 * float(np.percentile(t, 10))


out["call_ms_pageable"] = timed(lambda: ctx.stereo_batch(one, 1500))
cap = 1500
pin = ctx.host_array((1, 2, 480, 752), np.uint8)
pin[:] = one
bufs = dict(xy=ctx.host_array((2, cap, 2), np.float64), angles=ctx.host_array((2, cap), np.float64),
            desc=ctx.host_array((2, cap, 8), np.uint32), n_kp=ctx.host_array((2,), np.int32),
            matches=ctx.host_array((1, cap, 2), np.int32), n_matches=ctx.host_array((1,), np.int32))
out["call_ms_pinned"] = timed(lambda: ctx.stereo_batch(pin, 1500, True, 70, 1.2, cap, out=bufs))
# the reference's call sequence: two detectKeypointsAndDescriptors + one matchDescriptors
def three():
    _, _, dl = ctx.detect_describe(pl, 1500)
    _, _, dr = ctx.detect_describe(pr, 1500)
    ctx.match(dl, dr, 70, 1.2)
out["three_calls_ms_pageable"] = timed(three, 100)
out["launches_per_call"] = None
l0 = ctx.launch_count(); ctx.stereo_batch(pin, 1500, True, 70, 1.2, cap, out=bufs); out["launches_per_call"] = ctx.launch_count() - l0
print(json.dumps(out))
