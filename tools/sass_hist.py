#!/usr/bin/env python3
"""usage: sass_hist.py <object/.so> <function-substring> [--full]  -- opcode histogram of one kernel's SASS"""
import collections, re, subprocess, sys
out = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
on, cnt = False, collections.Counter()
full = "--full" in sys.argv
for ln in out.splitlines():
    if "Function :" in ln:
        on = sys.argv[2] in ln
        continue
    if not on:
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
    if m:
        op = m.group(1)
        cnt[op if full else op.split(".")[0]] += 1
for k, v in cnt.most_common():
    print("%6d %s" % (v, k))
print("total", sum(cnt.values()))
