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
prof = hasattr(L, "vslam_debug_mma_clocks")
if not prof:
    L = type("NoProf", (), {"vslam_debug_mma_clocks": staticmethod(lambda *a: 0)})()
ctx.sync()
plan.download(n)          # synchronises the plan's stream: the counters below belong to ONE run
assert L.vslam_debug_mma_clocks(clk, 1) == 0
plan.run(n, 1500, True, 70, 1.2)
ms = plan.last_ms()
res = plan.download(n)
assert L.vslam_debug_mma_clocks(clk, 0) == 0
c = list(clk)
jobs0 = (n + 147) // 148
out = {"pairs": n, "jobs_of_cta0": jobs0, "match_ms": float(ms[4]), "match_clk_at_1965": float(ms[4]) * 1.965e6,
       "tile_pairs_of_cta0": c[12],
       "mma_thread": {"wait_a_full": c[1], "wait_b_full": c[2], "wait_t_empty": c[3], "issue": c[4]},
       "epilogue_d1_warp0": {"wait_t_full": c[9], "fold_and_merge": c[10], "finalize": c[11]},
       "epilogue_d2_warp4": {"wait_t_full": c[17], "fold_and_merge": c[18]},
       "mean_kp": float(res["n_kp"].mean())}
print(json.dumps(out))
if prof and hasattr(L, "vslam_debug_mma_trace"):
    tr = (C.c_longlong * 1536)()
    assert L.vslam_debug_mma_trace(tr) == 0
    t = np.array(list(tr), np.int64).reshape(128, 12)
    t0 = t[0, 0]
    print("pair | pre_b_full loop_top  t_empty_ok  issue_start  issued  issued_d2 committed | w0_full w0_loaded w0_done w4_done   (cycles since the first issue)")
    for k in list(range(0, 30)) + list(range(118, 126)):
        r = t[k] - t0
        print("%4d | %8d %8d %8d %8d %8d %8d %8d | %8d %8d %8d %8d" % (k, r[8], r[4], r[5], r[0], r[6], r[9], r[1], r[10], r[7], r[2], r[3]))
