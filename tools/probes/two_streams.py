"""Does splitting the resident batch over two streams (two contexts) let kernels of different stages share SMs?"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import StereoPlan
N = 2000
base = wl.synth_stereo_batch(100)
imgs = np.ascontiguousarray(base[np.arange(N) % 100])
def run(nctx):
    ctxs = [visnav_b200.Context(0) for _ in range(nctx)]
    per = N // nctx
    plans = []
    for k, c in enumerate(ctxs):
        p = StereoPlan(c, 752, 480, per, 1500); p.upload(imgs[k * per:(k + 1) * per]); plans.append(p)
    for _ in range(3):
        for p in plans: p.run(per)
    for c in ctxs: c.sync()
    t0 = time.perf_counter()
    K = 10
    for _ in range(K):
        for p in plans: p.run(per)
    for c in ctxs: c.sync()
    dt = (time.perf_counter() - t0) / K
    print("%d stream(s): %.3f ms per %d pairs -> %.0f pairs/s" % (nctx, dt * 1e3, N, N / dt))
    for p in plans: p.close()
    for c in ctxs: c.close()
run(1); run(2); run(4)
