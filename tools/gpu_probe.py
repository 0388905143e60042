#!/usr/bin/env python3
"""Quick on-GPU probe: POPC peak, matcher throughput (configs 1/3/4 shapes), stereo plan stage times.
Writes gpurun_out/probe.json.  Not a benchmark of record (bench.py is)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import DescStore, StereoPlan, VSLAM_OPT_MATCH_CSA

out = {}
ctx = visnav_b200.Context(0)
out["popc_per_s"] = ctx.popc_peak()
print("popc/s", out["popc_per_s"], flush=True)

# config 3 sample: 40 keyframes -> 780 jobs of 1500 x 1500
kf = wl.keyframe_descriptors(40, 1500, 25)
store = DescStore(ctx, kf)
jobs = wl.keyframe_pair_jobs(40)
for csa in (1, 0):
    ctx.set_option(VSLAM_OPT_MATCH_CSA, csa)
    best = 1e9
    for rep in range(4):
        _, counts = ctx.match_batch(store, jobs, 70, 1.2, want_matches=False)
        best = min(best, ctx.last_match_ms())
    dist = len(jobs) * 1500.0 * 1500.0
    out["c3_csa%d" % csa] = dict(ms=best, dist_per_s=dist / (best * 1e-3), popc_equiv_per_s=8 * dist / (best * 1e-3),
                                 mean_matches=float(counts.mean()))
    print("c3 csa", csa, out["c3_csa%d" % csa], flush=True)
ctx.set_option(VSLAM_OPT_MATCH_CSA, 1)
# config 4: 1500 x 200000
q, db = wl.reloc_query()
best = 1e9
for rep in range(4):
    m = ctx.match(q, db, 70, 1.2)
    best = min(best, ctx.last_match_ms())
out["c4"] = dict(ms=best, dist_per_s=1500 * 200000 / (best * 1e-3), matches=len(m))
print("c4", out["c4"], flush=True)
# config 1 single call latency
d1, d2 = kf[0], kf[1]
t = []
for rep in range(20):
    t0 = time.perf_counter(); ctx.match(d1, d2); t.append(time.perf_counter() - t0)
out["c1_match_call_ms"] = 1e3 * float(np.median(t)); out["c1_match_kernel_ms"] = ctx.last_match_ms()
print("c1", out["c1_match_call_ms"], out["c1_match_kernel_ms"], flush=True)

# config 1: ONE stereo pair the way slam.cpp:1125-1143 issues it (two detectKeypointsAndDescriptors calls + one
# matchDescriptors call, host buffers), and the same pair through one vslam_stereo_batch call
pl, pr = wl.synth_stereo_pair(0)
one = np.ascontiguousarray(np.stack([pl, pr])[None])
t3, t1 = [], []
for rep in range(30):
    t0 = time.perf_counter()
    _, _, dl = ctx.detect_describe(pl, 1500)
    _, _, dr = ctx.detect_describe(pr, 1500)
    m = ctx.match(dl, dr, 70, 1.2)
    t3.append(time.perf_counter() - t0)
    t0 = time.perf_counter(); o1 = ctx.stereo_batch(one, 1500); t1.append(time.perf_counter() - t0)
out["c1_single_pair"] = dict(three_calls_ms=1e3 * float(np.median(t3)), one_batch_call_ms=1e3 * float(np.median(t1)),
                             keypoints=[len(dl), len(dr)], matches=len(m))
print("c1 pair", out["c1_single_pair"], flush=True)

# stereo plan: stage times
npairs = int(os.environ.get("PROBE_PAIRS", "256"))
imgs = wl.synth_stereo_batch(npairs)
plan = StereoPlan(ctx, 752, 480, npairs, 1500)
plan.upload(imgs)
for rep in range(4):
    plan.run(npairs)
    ms = plan.last_ms()
res = plan.download(npairs)
out["stereo"] = dict(pairs=npairs, ms_total=ms[0], ms_response=ms[1], ms_select=ms[2], ms_describe=ms[3], ms_match=ms[4],
                     pairs_per_s=npairs / (ms[0] * 1e-3), mean_kp=float(res["n_kp"].mean()),
                     mean_matches=float(res["n_matches"].mean()), min_kp=int(res["n_kp"].min()))
print("stereo", out["stereo"], flush=True)
t0 = time.perf_counter(); ctx.stereo_batch(imgs); t1 = time.perf_counter()
out["stereo_e2e_pageable_pairs_per_s"] = npairs / (t1 - t0)
print("e2e pageable", out["stereo_e2e_pageable_pairs_per_s"])

# config 3 in full: 500 keyframes x 1500, all 124 750 unordered pairs (counts only come back)
if os.environ.get("PROBE_C3_FULL", "1") == "1":
    kf = wl.keyframe_descriptors(500, 1500, 25)
    store = DescStore(ctx, kf)
    jobs = wl.keyframe_pair_jobs(500)
    t0 = time.perf_counter()
    _, counts = ctx.match_batch(store, jobs, 70, 1.2, want_matches=False)
    wall = time.perf_counter() - t0
    ms = ctx.last_match_ms()
    dist = len(jobs) * 1500.0 * 1500.0
    out["c3_full"] = dict(jobs=len(jobs), kernel_ms=ms, wall_s=wall, dist_per_s=dist / (ms * 1e-3),
                          popc_equiv_per_s=8 * dist / (ms * 1e-3), frac_of_popc_peak=8 * dist / (ms * 1e-3) / out["popc_per_s"],
                          mean_matches=float(counts.mean()), min_matches=int(counts.min()))
    print("c3 full", out["c3_full"], flush=True)
    store.close()

# config 5: 1920x1080, 4000 features -- single-scale stereo plan and the 4-level pyramid extension
n5 = int(os.environ.get("PROBE_C5_PAIRS", "296"))
base5 = wl.synth_stereo_batch(min(n5, 8), 1920, 1080)
imgs5 = np.concatenate([base5] * ((n5 + len(base5) - 1) // len(base5)))[:n5]
plan5 = StereoPlan(ctx, 1920, 1080, n5, 4000)
plan5.upload(imgs5)
for rep in range(3):
    plan5.run(n5, 4000)
    ms5 = plan5.last_ms()
r5 = plan5.download(n5)
out["c5_single_scale"] = dict(pairs=n5, ms_total=ms5[0], ms_response=ms5[1], ms_select=ms5[2], ms_describe=ms5[3],
                              ms_match=ms5[4], pairs_per_s=n5 / (ms5[0] * 1e-3), mean_kp=float(r5["n_kp"].mean()),
                              mean_matches=float(r5["n_matches"].mean()))
print("c5 single-scale", out["c5_single_scale"], flush=True)
plan5.close()
flat = np.ascontiguousarray(imgs5.reshape(-1, 1080, 1920))
caps = [3012, 753, 188, 47]
ctx.pyramid_detect_describe_batch(flat[:4], caps)
t0 = time.perf_counter()
res = ctx.pyramid_detect_describe_batch(flat, caps)
dt = time.perf_counter() - t0
out["c5_pyramid_e2e"] = dict(images=len(flat), wall_s=dt, images_per_s=len(flat) / dt,
                             mean_kp_per_level=[float(v) for v in res[3].mean(axis=0)])
print("c5 pyramid (pageable host arrays, H2D + D2H included)", out["c5_pyramid_e2e"], flush=True)

# find_matches_landmarks: one frame against the projected map (slam.cpp:1159)
sc = wl.landmark_scene(n_kp=1313, n_landmarks=2000, seed=1)
t = []
for rep in range(20):
    t0 = time.perf_counter(); m = ctx.match_landmarks(*sc); t.append(time.perf_counter() - t0)
out["landmarks"] = dict(n_kp=1313, n_landmarks=2000, call_ms=1e3 * float(np.median(t)), kernel_ms=ctx.last_match_ms(),
                        matches=len(m))
print("landmarks", out["landmarks"], flush=True)

# geometry after matching: single calls with host buffers, and the inlier filter fused into the plan
from visnav_b200.keypoints import Camera, Epipolar
P = wl.CAMERAS["ds"]
cam = Camera.from_data("ds", P, 752, 480)
xy1, xy2, mt, pose01, E = wl.epipolar_scene("ds", 1300, seed=2)
t = []
for rep in range(20):
    t0 = time.perf_counter(); inl = ctx.find_inliers_essential(xy1, xy2, cam, cam, E, 5e-3, mt); t.append(time.perf_counter() - t0)
out["inliers_essential"] = dict(matches=len(mt), inliers=len(inl), call_ms=1e3 * float(np.median(t)))
pose_w, pts = wl.landmark_cloud(20000, 1)
pts = np.tile(pts, (5, 1))
t = []
for rep in range(20):
    t0 = time.perf_counter(); pxy, pidx = ctx.project_landmarks(pose_w, cam, pts, 0.1); t.append(time.perf_counter() - t0)
out["project_landmarks"] = dict(landmarks=len(pts), projected=len(pidx), call_ms=1e3 * float(np.median(t)))
plan.set_epipolar(Epipolar.make(cam, cam, wl.stereo_geometry(0)[1], 0.02))
for rep in range(3):
    plan.run(npairs)
inl, ninl = plan.download_inliers(npairs)
out["stereo_fused_epipolar"] = dict(pairs=npairs, epipolar_ms=plan.last_epipolar_ms(), total_ms=plan.last_ms()[0],
                                    mean_inliers=float(ninl.mean()))
plan.set_epipolar(None)
print("geometry", out["inliers_essential"], out["project_landmarks"], out["stereo_fused_epipolar"], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/probe.json", "w"), indent=1)
