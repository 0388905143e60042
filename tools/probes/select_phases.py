"""Phase clocks of the select kernel's CTA 0 (library built with VSLAM_NVCC_EXTRA=-DVSLAM_SEL_PROFILE)."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, visnav_b200
from visnav_b200 import workloads as wl
ctx = visnav_b200.Context(0)
L = ctx.lib
img = wl.synth_stereo_pair(0)[0]
for rep in range(3):
    ctx.detect(img, 1500)
clk = (C.c_longlong * 16)()
assert L.vslam_debug_select_clocks(clk) == 0
t = list(clk)[:7]
names = ["filter", "binning", "fixed point", "compact", "sort", "cap+border+write"]
tot = t[6] - t[0]
for k, nm in enumerate(names):
    print("%-18s %8d cycles %5.1f%%" % (nm, t[k + 1] - t[k], 100.0 * (t[k + 1] - t[k]) / tot))
print("total %d cycles = %.1f us at 1.965 GHz" % (tot, tot / 1965.0))
full = list(clk)
print("fixed-point sweeps:", full[7], " undecided after sweep 1:", full[15], " sweep end clocks relative to loop start:",
      [full[8 + k] - t[2] for k in range(min(7, full[7]))])
