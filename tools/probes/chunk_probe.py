import sys, os
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo"))
import numpy as np, visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import StereoPlan, VSLAM_OPT_RESPONSE_BAND
ctx = visnav_b200.Context(0)
base = wl.synth_stereo_batch(8, workers=1)
for n in (24, 48, 96, 192):
    imgs = np.ascontiguousarray(base[np.arange(n) % 8])
    plan = StereoPlan(ctx, 752, 480, n, 1500); plan.upload(imgs)
    for band in (0, 1):
        ctx.set_option(VSLAM_OPT_RESPONSE_BAND, band)
        for _ in range(4): plan.run(n, 1500, True, 70, 1.2); ms = plan.last_ms()
        print(n, "band", band, ["%.3f" % v for v in ms[:5]])
    plan.close()
