#!/usr/bin/env python3
"""Parity report (BASELINE.json north_star: "keypoint sets match the reference's OpenCV corners, with the mismatch
rate reported ... descriptors and match lists bit-exact ... angles within 1e-5 rad").

Runs the CUDA path through the C ABI on N synthetic stereo pairs of each size and compares every output with the
CPU oracle on the same inputs (test infrastructure: oracle/liboracle.so, the restatement pinned to cv2 and to the
reference's own functions; oracle/_ref where it travelled).  Where cv2 is importable the corners are ALSO compared
with cv2.goodFeaturesToTrack itself, under the scalar dispatch and under the default (AVX2) dispatch -- the latter
is what a stock visnav build runs, and the only place where floating-point ties may legitimately differ.
Writes gpurun_out/parity.json.

  python tools/parity_report.py [--pairs 64] [--pairs-1080 4]
"""
import argparse
import json
import os
import sys
from concurrent.futures import ThreadPoolExecutor

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import oracle  # noqa: E402
import visnav_b200  # noqa: E402
from visnav_b200 import workloads as wl  # noqa: E402

try:
    import cv2
except Exception:  # pragma: no cover
    cv2 = None


def corners_cv2(img, nf, optimized):
    cv2.setUseOptimized(optimized)
    pts = cv2.goodFeaturesToTrack(img, nf, 0.01, 8, None, None, 3, False)
    cv2.setUseOptimized(True)
    pts = np.zeros((0, 2), np.float32) if pts is None else pts.reshape(-1, 2)
    h, w = img.shape
    keep = (pts[:, 0] >= 19) & (pts[:, 0] < w - 19) & (pts[:, 1] >= 19) & (pts[:, 1] < h - 19)
    return pts[keep].astype(np.float64)


def corner_diff(got, want):
    """(missing, extra, ordered_equal) between two ordered corner lists."""
    a = {(float(x), float(y)) for x, y in got}
    b = {(float(x), float(y)) for x, y in want}
    return len(b - a), len(a - b), bool(np.array_equal(got, want))


def run_size(ctx, o, ref, w, h, nf, n_pairs, first):
    pairs = wl.synth_stereo_batch(n_pairs, w, h, first=first)
    out = ctx.stereo_batch(pairs, nf, True, 70, 1.2)
    rep = dict(size="%dx%d" % (w, h), num_features=nf, pairs=n_pairs, images=2 * n_pairs)
    cpu = [None] * (2 * n_pairs)

    def one(k):
        cpu[k] = o.detect_describe(pairs[k // 2, k % 2], nf)
    with ThreadPoolExecutor(os.cpu_count() or 1) as ex:
        list(ex.map(one, range(2 * n_pairs)))
    tot_ref = missing = extra = unordered = 0
    ang_err = 0.0
    desc_bad = desc_bits = desc_total = 0
    for k in range(2 * n_pairs):
        xy, ang, desc = cpu[k]
        n = int(out["n_kp"][k])
        g = out["xy"][k, :n]
        m, e, same = corner_diff(g, xy)
        tot_ref += len(xy); missing += m; extra += e; unordered += (not same)
        if same:
            ang_err = max(ang_err, float(np.abs(out["angles"][k, :n] - ang).max()) if n else 0.0)
            x = out["desc"][k, :n] ^ desc
            bits = int(np.unpackbits(x.view(np.uint8)).sum())
            desc_bits += bits
            desc_bad += int((x != 0).any(axis=1).sum())
            desc_total += n
    rep["corners_vs_oracle"] = dict(reference_corners=tot_ref, missing=missing, extra=extra,
                                    images_with_any_difference=unordered,
                                    mismatch_rate=(missing + extra) / max(tot_ref, 1))
    rep["angles_vs_oracle"] = dict(max_abs_err_rad=ang_err, tolerance=1e-5)
    rep["descriptors_vs_oracle"] = dict(compared=desc_total, differing=desc_bad, differing_bits=desc_bits)
    # match lists: GPU list vs the oracle's matcher on the ORACLE's descriptors (identical if everything above is)
    bad_lists = total_matches = 0
    ref_bad = 0
    for p in range(n_pairs):
        want = o.match(cpu[2 * p][2], cpu[2 * p + 1][2], 70, 1.2)
        nm = int(out["n_matches"][p])
        bad_lists += not np.array_equal(out["matches"][p, :nm], want)
        total_matches += len(want)
        if ref is not None and p < 8:
            ref_bad += not np.array_equal(ref.match(cpu[2 * p][2], cpu[2 * p + 1][2], 70, 1.2), want)
    rep["matches_vs_oracle"] = dict(lists=n_pairs, lists_differing=int(bad_lists), reference_matches=total_matches)
    if ref is not None:
        rep["oracle_vs_reference_library_match_lists_differing"] = int(ref_bad)
    if cv2 is not None:
        for name, opt in (("scalar", False), ("default_dispatch", True)):
            tr = mi = exx = un = 0
            for k in range(min(2 * n_pairs, 32)):
                want = corners_cv2(pairs[k // 2, k % 2], nf, opt)
                n = int(out["n_kp"][k])
                m, e, same = corner_diff(out["xy"][k, :n], want)
                tr += len(want); mi += m; exx += e; un += (not same)
            rep["corners_vs_cv2_" + name] = dict(images=min(2 * n_pairs, 32), reference_corners=tr, missing=mi, extra=exx,
                                                 images_with_any_difference=un, mismatch_rate=(mi + exx) / max(tr, 1))
    return rep


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=64)
    ap.add_argument("--pairs-1080", type=int, default=4)
    a = ap.parse_args()
    ctx = visnav_b200.Context(0)
    o = oracle.load_oracle()
    ref = oracle.load_ref()
    rep = dict(cv2=(cv2.__version__ if cv2 is not None else None), reference_library=ref is not None, sizes=[])
    if cv2 is not None:
        # response MAP vs cv2.cornerMinEigenVal (scalar dispatch) pixel by pixel, in both kernel modes: OpenCV's box
        # filter keeps a RUNNING column sum in double (ColumnSum<double, float>); where an addition is inexact its
        # rounding residue travels down the column.  Full-height boxes (what launches that fill the GPU use) restate it
        # and must differ nowhere; row-banded launches (VSLAM_OPT_RESPONSE_BAND = 1: the opt-in latency mode) sum the three rows directly.
        from visnav_b200.keypoints import VSLAM_OPT_RESPONSE_BAND
        rm = []
        for (w, h, seed) in [(752, 480, 1234), (1920, 1080, 77), (1920, 1080, 302), (320, 240, 5)]:
            img = wl.synth_image(seed, w, h)
            cv2.setUseOptimized(False)
            want = cv2.cornerMinEigenVal(img, 3, ksize=3)
            cv2.setUseOptimized(True)
            thr = float(want.max()) * 0.01
            rec = dict(size="%dx%d" % (w, h), seed=seed, pixels=w * h,
                       oracle_differing=int((o.response(img).view(np.uint32) != want.view(np.uint32)).sum()))
            for mode, band in (("full_height_default", 0), ("banded_opt_in", 1)):
                ctx.set_option(VSLAM_OPT_RESPONSE_BAND, band)
                got = ctx.response(img)
                ctx.set_option(VSLAM_OPT_RESPONSE_BAND, 0)
                diff = got.view(np.int32).astype(np.int64) - want.view(np.int32).astype(np.int64)
                bad = np.nonzero(diff)
                rec[mode] = dict(differing=int(len(bad[0])), max_ulp=int(np.abs(diff).max()), rate=len(bad[0]) / (w * h),
                                 largest_differing_response_over_threshold=(float(want[bad].max()) / thr if len(bad[0]) else 0.0))
            rm.append(rec)
        rep["response_map_vs_cv2_scalar"] = rm
    rep["sizes"].append(run_size(ctx, o, ref, 752, 480, 1500, a.pairs, 7000))
    rep["sizes"].append(run_size(ctx, o, ref, 1920, 1080, 4000, a.pairs_1080, 8000))
    rep["sizes"].append(run_size(ctx, o, ref, 321, 243, 300, 8, 9000))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rep, open("gpurun_out/parity.json", "w"), indent=1)
    print(json.dumps(rep))
    bad = sum(s["corners_vs_oracle"]["missing"] + s["corners_vs_oracle"]["extra"] + s["descriptors_vs_oracle"]["differing"] +
              s["matches_vs_oracle"]["lists_differing"] for s in rep["sizes"])
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
