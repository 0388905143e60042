"""Stage times of a resident plan of n pairs: default response path (small-count path up to 16 images, full-height boxes
beyond), the streaming kernel forced (65535) and row bands (1)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import StereoPlan, VSLAM_OPT_RESPONSE_BAND, VSLAM_OPT_GRAPH
ctx = visnav_b200.Context(0)
ctx.set_option(VSLAM_OPT_GRAPH, 0)
base = wl.synth_stereo_batch(8, workers=1)
for n in [int(v) for v in os.environ.get("PAIRS", "1,2,4,6,8,12,24,48,120").split(",")]:
    imgs = np.ascontiguousarray(base[np.arange(n) % 8])
    plan = StereoPlan(ctx, 752, 480, n, 1500); plan.upload(imgs)
    for band in (0, 65535, 1):
        ctx.set_option(VSLAM_OPT_RESPONSE_BAND, band)
        for _ in range(4): plan.run(n, 1500, True, 70, 1.2); ms = plan.last_ms()
        print(n, "pairs, band", band, ["%.3f" % v for v in ms[:5]])
    ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 0)
    plan.close()
