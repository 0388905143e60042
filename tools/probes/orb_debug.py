import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import visnav_b200
from oracle import oracle as oo
from visnav_b200 import workloads as wl
ctx = visnav_b200.Context(0)
img = wl.synth_image(1234, 752, 480)
kp, desc, n = ctx.orb_detect_compute_batch(img[None], 1500)
k = kp[0, :n[0]]
okp, od = oo.orb_detect_compute(img, 1500)
a = {tuple(r[[5, 0, 1]]): r for r in k}
b = {tuple(r[[5, 0, 1]]): r for r in okp}
print("gpu", len(a), "cpu", len(b))
for key in sorted(set(a) - set(b)):
    print("extra", key, a[key])
for key in sorted(set(b) - set(a)):
    print("missing", key, b[key])
for l in range(8):
    ga = [r for r in k if r[5] == l]; gb = [r for r in okp if r[5] == l]
    print(l, len(ga), len(gb), min((r[4] for r in ga), default=0), min((r[4] for r in gb), default=0))
print("=== level 0 detail")
ga = {(r[0], r[1]): r[4] for r in k if r[5] == 0}; gb = {(r[0], r[1]): r[4] for r in okp if r[5] == 0}
print("common", len(set(ga) & set(gb)), "only gpu", len(set(ga) - set(gb)), "only cpu", len(set(gb) - set(ga)))
for key in list(set(ga) & set(gb))[:5]:
    print(key, ga[key], gb[key])
for key in sorted(set(ga) - set(gb))[:8]:
    print("gpu only", key, ga[key])
for key in sorted(set(gb) - set(ga))[:8]:
    print("cpu only", key, gb[key])
