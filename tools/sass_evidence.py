#!/usr/bin/env python3
"""usage: sass_evidence.py <libvisnav_b200.so> > profiles/sass_opcodes.txt
Per kernel of the built library: instruction count and the count of the opcodes that prove the Blackwell-native paths
(B200_PROFILING.md: tcgen05.mma -> UTC*MMA, tcgen05.ld -> LDTM, TMA -> UTMALDG / UBLKCP, cp.async -> LDGSTS, packed fp32
-> FMUL2 / FADD2 / FFMA2, warp reductions -> REDUX/CREDUX, mbarriers -> SYNCS)."""
import collections, re, subprocess, sys
out = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
KEYS = ["UTCIMMA", "UTCHMMA", "UTCQMMA", "LDTM", "UTMALDG", "UBLKCP", "LDGSTS", "SYNCS", "FMUL2", "FADD2", "FFMA2", "CREDUX",
        "REDUX", "VIMNMX", "VIMNMX3", "POPC", "IMMA", "HMMA", "DADD", "F2F", "IDP", "LDL", "STL"]
kern, cnt, arch = None, collections.OrderedDict(), None
for ln in out.splitlines():
    m = re.search(r"arch = (sm_\w+)", ln)
    if m:
        arch = m.group(1)
    if "Function :" in ln:
        name = ln.split("Function :")[1].strip()
        dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
        dem = dem.replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
        m2 = re.search(r"([A-Za-z_]\w*(?:<[^()]*>)?)\(", dem)
        kern = (m2.group(1) if m2 else dem) + " [" + str(arch) + "]"
        cnt[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
    if m and kern:
        cnt[kern]["total"] += 1
        if m.group(1) in KEYS:
            cnt[kern][m.group(1)] += 1
print("# %s: evidence opcodes per kernel (cuobjdump -sass)" % sys.argv[1].split("/")[-1])
for k, c in cnt.items():
    ev = "  ".join("%s=%d" % (op, c[op]) for op in KEYS if c[op])
    print("%-70s total=%5d  %s" % (k[:70], c["total"], ev))
