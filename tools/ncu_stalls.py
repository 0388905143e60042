#!/usr/bin/env python3
"""usage: ncu_stalls.py <report.ncu-rep> <kernel-regex> [top N]  -- stall mix and hottest SASS lines of one kernel"""
import csv, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--kernel-name", "regex:" + pat], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
H, U = rows[0], rows[1]
for r in rows[2:3]:
    items = []
    for i, h in enumerate(H):
        if h.startswith("smsp__pcsamp_warps_issue_stalled") and not h.endswith("_not_issued"):
            try: items.append((float(r[i].replace(",", "")), h))
            except ValueError: pass
    tot = sum(v for v, _ in items)
    print("stall mix:", "  ".join("%s %.1f%%" % (h.replace("smsp__pcsamp_warps_issue_stalled_", ""), 100 * v / tot) for v, h in sorted(items, reverse=True)[:9]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
H = rows[hi]
ia, isamp, iex = H.index("Source"), H.index("# Samples"), H.index("Instructions Executed")
st = [(i, h) for i, h in enumerate(H) if h.startswith("stall_") and "Not Issued" not in h]
data = []
for n, r in enumerate(rows[hi + 1:]):
    try: data.append((int(r[isamp]), n, r[ia].strip(), int(r[iex]), r))
    except (ValueError, IndexError): pass
tot = sum(d[0] for d in data)
print("total samples", tot, "instructions", len(data), "executed warp-instr", sum(d[3] for d in data))
for s, n, srcl, ex, r in sorted(data, reverse=True)[:top]:
    tops = sorted(((int(r[i] or 0), h) for i, h in st), reverse=True)[:2]
    print("%6d %4.1f%% #%4d ex=%9d %-52s %s" % (s, 100 * s / tot, n, ex, srcl[:52], " ".join("%s=%d" % (h[6:], v) for v, h in tops)))
if "--dump" in sys.argv:
    for s, n, srcl, ex, r in data:
        print("%5d %9d %s" % (s, ex, srcl))
