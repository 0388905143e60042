import sys; sys.path.insert(0,'/root/repo')
import numpy as np, time, visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import StereoPlan
ctx=visnav_b200.Context(0)
for n in (1,2,4,8,16):
    imgs=wl.synth_stereo_batch(n)
    plan=StereoPlan(ctx,752,480,n,1500); plan.upload(imgs)
    for r in range(5): plan.run(n); ms=plan.last_ms()
    print(n, ["%.3f"%m for m in ms])
    plan.close()
