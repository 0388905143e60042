"""End-to-end vslam_stereo_batch (pinned host buffers, 2000 pairs) with full-height boxes (default) and with row bands."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import VSLAM_OPT_RESPONSE_BAND, VSLAM_OPT_BATCH_CHUNK, VSLAM_OPT_BATCH_DEPTH
ctx = visnav_b200.Context(0)
N, cap = 2000, 1500
base = wl.synth_stereo_batch(100)
host = ctx.host_array((N, 2, 480, 752), np.uint8)
host[:] = base[np.arange(N) % 100]
out = dict(xy=ctx.host_array((N * 2, cap, 2), np.float64), angles=ctx.host_array((N * 2, cap), np.float64),
           desc=ctx.host_array((N * 2, cap, 8), np.uint32), n_kp=ctx.host_array((N * 2,), np.int32),
           matches=ctx.host_array((N, cap, 2), np.int32), n_matches=ctx.host_array((N,), np.int32))
def run(tag):
    ctx.stereo_batch(host, 1500, True, 70, 1.2, cap, out=out)
    ts = []
    for _ in range(4):
        t0 = time.perf_counter(); ctx.stereo_batch(host, 1500, True, 70, 1.2, cap, out=out); ts.append(time.perf_counter() - t0)
    print(tag, "%.2f ms -> %.0f pairs/s" % (1e3 * min(ts), N / min(ts)))
for band in (0, 1):
    ctx.set_option(VSLAM_OPT_RESPONSE_BAND, band); run("band %d chunk 96" % band)
ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 0)
for chunk in [int(c) for c in os.environ.get("CHUNKS", "64,128,192").split(",")]:
    ctx.set_option(VSLAM_OPT_BATCH_CHUNK, chunk); run("band 0 chunk %d" % chunk)
