"""Cycle counters of CTA 0 of match_mma_kernel (library built with VSLAM_NVCC_EXTRA=-DVSLAM_MMA_PROFILE and loaded
through VSLAM_LIB): where the MMA issuer and the epilogue warps wait."""
import sys, os, ctypes as C, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import StereoPlan
ctx = visnav_b200.Context(0)
L = ctx.lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 592
base = wl.synth_stereo_batch(8, workers=1)
imgs = np.ascontiguousarray(base[np.arange(n) % 8])
plan = StereoPlan(ctx, 752, 480, n, 1500)
plan.upload(imgs)
for _ in range(2):
    plan.run(n, 1500, True, 70, 1.2)
clk = (C.c_longlong * 32)()
assert L.vslam_debug_mma_clocks(clk, 1) == 0
plan.run(n, 1500, True, 70, 1.2)
ms = plan.last_ms()
assert L.vslam_debug_mma_clocks(clk, 0) == 0
c = list(clk)
res = plan.download(n)
jobs0 = (n + 147) // 148
out = {"pairs": n, "jobs_of_cta0": jobs0, "match_ms": float(ms[4]),
       "mma_thread": {"wait_a_full": c[1], "wait_b_full": c[2], "wait_t_empty": c[3], "issue": c[4]},
       "epilogue_d1_warp0": {"wait_t_full": c[9], "fold_and_merge": c[10], "finalize": c[11]},
       "epilogue_d2_warp4": {"wait_t_full": c[17], "fold_and_merge": c[18]},
       "mean_kp": float(res["n_kp"].mean())}
print(json.dumps(out))
