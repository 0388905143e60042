#!/usr/bin/env python3
"""bench.py -- stereo front-end throughput (BASELINE.json metric) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W]            our arm (CUDA, through the C ABI)
  python bench.py --impl reference [--gpus N] [--steps K] ...    the reference's CPU path (oracle/_ref)

Workload (config.workload): BASELINE.json configs[1] -- a batch of 2000 synthetic 752x480 stereo
pairs PER GPU (weak scaling: pairs are independent, no data-path collective), each step =
detectKeypointsAndDescriptors on both images (1500 features, rotate) + matchDescriptors(70, 1.2).
  value : pairs/s, images resident in HBM, CUDA events on the launching stream, max over ranks
  e2e   : the same through vslam_stereo_batch with pinned HOST buffers in and out (H2D + D2H timed)
One JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

W_IMG, H_IMG, NUM_FEATURES, THRESH, RATIO = 752, 480, 1500, 70, 1.2
PAIRS_PER_GPU = 2000
DISTINCT_PAIRS = 500       # distinct synthetic pairs per rank, repeated cyclically to fill the batch
METRIC = "stereo frames/s (752x480, 1500 ORB) detect+describe+match"
UNIT = "stereo pairs/s"
# algorithmic bytes per image of the dominant kernel (response): the u8 image read once + the
# 1500 x 8 B corner list it ultimately yields
RESP_BYTES_PER_IMAGE = W_IMG * H_IMG + NUM_FEATURES * 8
# DRAM traffic of the response kernel per image, from the ncu --set full capture of the same kernel
# at the bench's own launch size (profiles/r1x_response_raw.txt: dram__bytes_read.sum 1.444258 GB +
# dram__bytes_write.sum 136.05 MB for 4000 images): the image is read once (1.00x its size), the rest is the
# candidate list.
RESP_DRAM_BYTES_PER_IMAGE = (1.444258e9 + 136.047616e6) / 4000
CPU_GFTT_IMPL = "?"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pairs", type=int, default=PAIRS_PER_GPU, help="pairs per GPU per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=0, help="0: min(steps, 5)")
    return ap.parse_args()


def peaks():
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(pk["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.lines, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---- the reference's CPU path (test infrastructure; timed as the baseline only) ----------------
def cpu_pairs_per_s(pairs, n_threads, budget_s=20.0):
    """Runs the reference's detect+describe (both images) + match on `pairs` ([n,2,h,w] u8) with
    n_threads host threads (ctypes releases the GIL), each thread on its own pairs.  Stops handing
    out pairs after budget_s.  Returns (pairs/s, pairs done, kind)."""
    import oracle
    from concurrent.futures import ThreadPoolExecutor
    lib = oracle.load_ref()
    kind = "reference"
    if lib is None:
        lib, kind = oracle.load_oracle(), "port"
    elif not lib.has_opencv:
        kind = "reference+port"      # reference functions, goodFeaturesToTrack = C restatement (OpenCV not built)
    else:
        lib.set_num_threads(1)       # one OpenCV thread per worker: the workers are the parallelism
    global CPU_GFTT_IMPL
    CPU_GFTT_IMPL = lib.gftt_impl

    t0 = time.perf_counter()
    deadline = t0 + budget_s

    def one(k):
        if time.perf_counter() > deadline:
            return 0
        _, _, dl = lib.detect_describe(pairs[k, 0], NUM_FEATURES)
        _, _, dr = lib.detect_describe(pairs[k, 1], NUM_FEATURES)
        lib.match(dl, dr, THRESH, RATIO)
        return 1

    with ThreadPoolExecutor(n_threads) as ex:
        done = sum(ex.map(one, range(len(pairs))))
    dt = time.perf_counter() - t0
    return done / dt, done, kind


def make_pairs(n_distinct, first):
    from visnav_b200 import workloads as wl
    return wl.synth_stereo_batch(n_distinct, W_IMG, H_IMG, first=first)


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def bind_to_gpu_numa_node(local):
    """Pin this rank's host threads to the CPU cores next to its GPU (NVML affinity mask) BEFORE the pinned
    buffers are allocated, so that first-touch places them on the GPU's NUMA node: at N > 1 every rank then
    streams its 1.44 GB per step over its own memory controller and PCIe root.  Host-side policy only."""
    try:
        import pynvml
        pynvml.nvmlInit()
        idx = local
        cvd = os.environ.get("CUDA_VISIBLE_DEVICES", "")
        if cvd and all(t.strip().isdigit() for t in cvd.split(",")):
            idx = int(cvd.split(",")[local])
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * w + b for w, v in enumerate(words) for b in range(64) if (v >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return 0


def run_reference(args):
    rank, world, local = dist_env()
    if rank != 0:
        return  # rank 0 alone times the CPU path; the others exit 0 without work
    cores = os.cpu_count() or 1
    sample = min(4 * cores, 256)
    pairs = make_pairs(sample, 0)
    for _ in range(max(args.warmup, 0)):
        cpu_pairs_per_s(pairs[:max(cores, 1)], cores, budget_s=5.0)
    t_tot, n_tot, kind = 0.0, 0, "reference"
    for _ in range(max(args.steps, 1)):
        t0 = time.perf_counter()
        _, done, kind = cpu_pairs_per_s(pairs, cores, budget_s=15.0)
        t_tot += time.perf_counter() - t0
        n_tot += done
    v = n_tot / t_tot
    sample_txt = ("%d synthetic 752x480 stereo pairs per step on %d host threads (one pair per thread at a time, "
                  "cv::setNumThreads(1)); detectKeypoints/computeAngles/computeDescriptors/matchDescriptors are the "
                  "reference's own functions compiled from keypoints.h (oracle/_ref); cv::goodFeaturesToTrack = %s"
                  % (sample, cores, CPU_GFTT_IMPL))
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / max(args.steps, 1), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32/f64/u32 (cpu)", "data": "synthetic",
            "config": {"workload": "configs[1] bounded sample: %d pairs/step, CPU" % sample},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample_txt},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def run_ours(args):
    rank, world, local = dist_env()
    import visnav_b200
    from visnav_b200.keypoints import StereoPlan
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    numa_cpus = bind_to_gpu_numa_node(local) if world > 1 else 0
    ctx = visnav_b200.Context(local)          # raises without an sm_100 GPU: no CPU fallback
    n_pairs, cap = args.pairs, NUM_FEATURES

    # ---- synthetic batch in pinned host memory (generation is not timed) ----
    n_distinct = min(DISTINCT_PAIRS, n_pairs)
    distinct = make_pairs(n_distinct, first=rank * n_distinct)
    host_imgs = ctx.host_array((n_pairs, 2, H_IMG, W_IMG), np.uint8)
    for k in range(0, n_pairs, n_distinct):
        m = min(n_distinct, n_pairs - k)
        host_imgs[k:k + m] = distinct[:m]

    def barrier():
        ctx.sync()
        if dist is not None:
            import torch
            torch.cuda.synchronize()
            dist.barrier()

    # ---- value: device-resident batch, K timed steps, CUDA events on the ctx stream ----
    plan = StereoPlan(ctx, W_IMG, H_IMG, n_pairs, cap)
    plan.upload(host_imgs)
    for _ in range(max(args.warmup, 3)):
        plan.run(n_pairs, NUM_FEATURES, True, THRESH, RATIO)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = ctx.launch_count()
    stage_ms = np.zeros(5)
    ctx.event_record(0)
    for _ in range(args.steps):
        plan.run(n_pairs, NUM_FEATURES, True, THRESH, RATIO)
    ctx.event_record(1)
    t_ms = ctx.event_elapsed_ms(0, 1)
    barrier()
    launches = ctx.launch_count() - launches0
    stage_ms[:] = plan.last_ms()              # stages of the last timed step
    res = plan.download(n_pairs)
    assert (res["n_kp"] >= 0).all(), "a frame overflowed an internal limit"

    # ---- e2e: pinned host buffers in and out through vslam_stereo_batch ----
    out = dict(xy=ctx.host_array((n_pairs * 2, cap, 2), np.float64), angles=ctx.host_array((n_pairs * 2, cap), np.float64),
               desc=ctx.host_array((n_pairs * 2, cap, 8), np.uint32), n_kp=ctx.host_array((n_pairs * 2,), np.int32),
               matches=ctx.host_array((n_pairs, cap, 2), np.int32), n_matches=ctx.host_array((n_pairs,), np.int32))
    e2e_steps = args.e2e_steps or min(args.steps, 5)
    ctx.stereo_batch(host_imgs, NUM_FEATURES, True, THRESH, RATIO, cap, out=out)   # warm-up (allocates the pipeline)
    barrier()
    # vslam_stereo_batch is synchronous at return, so events recorded on the (idle) ctx stream around the loop
    # carry device timestamps of the whole region: copies in, kernels, copies out
    t0 = time.perf_counter()
    ctx.event_record(2)
    for _ in range(e2e_steps):
        ctx.stereo_batch(host_imgs, NUM_FEATURES, True, THRESH, RATIO, cap, out=out)
    ctx.event_record(3)
    t_e2e = ctx.event_elapsed_ms(2, 3) * 1e-3
    t_e2e_wall = time.perf_counter() - t0
    barrier()
    clocks = sampler.stop()
    assert np.array_equal(out["n_matches"], res["n_matches"]) and np.array_equal(out["n_kp"], res["n_kp"])
    h2d = int(host_imgs.nbytes)
    d2h = int(sum(v.nbytes for v in out.values()))

    # ---- max over ranks ----
    if dist is not None:
        import torch
        t = torch.tensor([t_ms, t_e2e], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_ms, t_e2e = float(t[0]), float(t[1])
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    total_pairs = world * n_pairs
    value = total_pairs * args.steps / (t_ms * 1e-3)
    e2e_value = total_pairs * e2e_steps / t_e2e
    peak, peak_src = peaks()
    resp_ms = float(stage_ms[1])
    achieved = RESP_BYTES_PER_IMAGE * 2 * n_pairs / (resp_ms * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": t_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8 -> f32/f64 (response), f64 (rBRIEF rotation), u32 popc (matching)", "data": "synthetic",
        "config": {"workload": "BASELINE configs[1]: %d synthetic 752x480 stereo pairs per GPU per step (%d distinct, "
                               "repeated), detect+describe both images (1500 features) + stereo match (70, 1.2)"
                               % (n_pairs, n_distinct),
                   "pairs_per_gpu": n_pairs, "l2": "inputs (1.47 GB per step) larger than L2, no flush needed",
                   "sharding": "pairs split by rank, no collective on the data path"},
        "stage_ms": {"response": stage_ms[1], "select": stage_ms[2], "describe": stage_ms[3], "match": stage_ms[4]},
        "roofline": {"kernel": "response_kernel (Sobel+cov+box+min-eig+local-max, TMA tiles)", "bound": "hbm",
                     "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": RESP_DRAM_BYTES_PER_IMAGE * 2 * n_pairs,
                     "traffic_source": "ncu --set full, profiles/r1x_response_raw.txt (per image x images per launch)",
                     "ms_per_launch": resp_ms, "algorithmic_bytes_per_launch": RESP_BYTES_PER_IMAGE * 2 * n_pairs},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                "timing": "CUDA events around the synchronous C-ABI calls (rank-0 host wall clock: %.1f ms/step)"
                          % (1e3 * t_e2e_wall / e2e_steps),
                "host_threads_bound_to_gpu_numa_node": numa_cpus},
        "gpu_launches": int(launches), "clocks": clocks,
        "results": {"mean_keypoints": float(res["n_kp"].mean()), "mean_matches": float(res["n_matches"].mean())},
    }
    # The matcher against the integer-popcount roofline north_star names (8 popc.b32 per 256-bit distance;
    # the kernel's carry-save adders turn them into 4 real POPCs, so the fraction may exceed 1).  Peak = the
    # register-only POPC microbenchmark of the library, run now on this GPU.
    kp = res["n_kp"].astype(np.int64).reshape(-1, 2)
    n_dist = float((kp[:, 0] * kp[:, 1]).sum())
    popc_peak = ctx.popc_peak()
    match_ms = float(stage_ms[4])
    line["roofline_match"] = {"kernel": "match_fused_kernel (all-pairs Hamming + best/second + ratio + mutual check)",
                              "bound": "popc", "unit": "popc.b32/s", "achieved": 8.0 * n_dist / (match_ms * 1e-3),
                              "peak": popc_peak, "peak_source": "vslam_popc_peak microbenchmark, this run",
                              "frac": 8.0 * n_dist / (match_ms * 1e-3) / popc_peak,
                              "distances_per_s": n_dist / (match_ms * 1e-3), "ms_per_launch": match_ms,
                              "distances_per_launch": n_dist}
    if not args.no_cpu_baseline and world == 1:
        cores = os.cpu_count() or 1
        sample = min(16 * cores, n_distinct)      # ~15-25 s of CPU work (73 ms per pair and core)
        v, done, kind = cpu_pairs_per_s(distinct[:sample], cores, budget_s=20.0)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                                "sample": "%d of the same synthetic pairs on %d host threads (reference functions from "
                                          "keypoints.h in oracle/_ref; cv::goodFeaturesToTrack = %s)" % (done, cores, CPU_GFTT_IMPL)}
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


_REAL_STDOUT = None


def quiet_stdout():
    """Libraries under us (NCCL prints its version banner) write to fd 1; the contract is ONE JSON line on stdout.
    Everything else goes to stderr: fd 1 is pointed at fd 2 and the JSON line is written to the saved descriptor."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _REAL_STDOUT is None:
        os.write(1, data)
    else:
        os.write(_REAL_STDOUT, data)


def main():
    args = parse()
    quiet_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
