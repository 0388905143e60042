#!/usr/bin/env python3
"""Summarise ncu outputs brought back in gpurun_out/ into small text files for profiles/.
  ncu_summary.py launches <launches.csv>            per-kernel launch count, total time, share
  ncu_summary.py raw <report.ncu-rep> [metric ...]  selected raw metrics of every captured launch
  ncu_summary.py traffic <report.ncu-rep> <kernel-regex> <units-per-launch> <source-file> <out.json> [also-kernel-regex]
        dram__bytes_read/write of the captured launch of that kernel, per unit, with the sha256 of the kernel's source
        file: bench.py reads the JSON and marks the roofline's traffic stale when the source has changed since
"""
import collections
import csv
import subprocess
import sys

DEFAULT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
           "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
           "smsp__inst_executed.sum", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
           "sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
           "l1tex__data_bank_conflicts_pipe_lsu.sum", "lts__t_bytes.sum"]


def launches(path):
    rows = list(csv.reader(open(path)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    H = rows[hdr]
    ki, vi = H.index("Kernel Name"), H.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[hdr + 1:]:
        if len(r) <= vi:
            continue
        name = r[ki].split("(")[0].split("::")[-1]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += float(r[vi].replace(",", ""))
    tot = sum(v[1] for v in agg.values())
    print("# per-launch gpu__time_duration.sum under ncu (cold-cache, serialised): compare SHARES")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-36s launches=%5d total=%10.3f ms mean=%9.3f us share=%5.1f%%" % (k, v[0], v[1] / 1e6, v[1] / v[0] / 1e3, 100 * v[1] / tot))


def raw(path, metrics):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    H, U = rows[0], rows[1]
    for r in rows[2:]:
        print("## kernel:", r[H.index("Kernel Name")].split("(")[0], " grid", r[H.index("Grid Size")], " block", r[H.index("Block Size")])
        for m in metrics:
            if m in H:
                print("%-70s %-14s %s" % (m, U[H.index(m)], r[H.index(m)]))


def traffic(path, pattern, units, source, out_json, also=None):
    import hashlib, json, os, re
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    H = rows[0]
    best = None
    for r in rows[2:]:
        if re.search(pattern, r[H.index("Kernel Name")]):
            best = r
    if best is None:
        sys.exit("no launch of %s in %s" % (pattern, path))
    def val(m):
        v, u = float(best[H.index(m)].replace(",", "")), rows[1][H.index(m)]
        return v * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
    rd, wr = val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
    extra = None
    if also:  # a second kernel that shares the launch's work (response_multi_kernel: the narrow strips): its bytes count too
        for r in rows[2:]:
            if re.search(also, r[H.index("Kernel Name")]):
                extra = r
        if extra is not None:
            def val2(m):
                return float(extra[H.index(m)].replace(",", "")) * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[rows[1][H.index(m)]]
            rd += val2("dram__bytes_read.sum")
            wr += val2("dram__bytes_write.sum")
    def opt(m):
        try:
            return float(best[H.index(m)].replace(",", ""))
        except (ValueError, IndexError):
            return None
    rec = {"kernel": best[H.index("Kernel Name")].split("(")[0], "report": os.path.basename(path), "units_per_launch": int(units),
           "warp_instructions": opt("smsp__inst_executed.sum"),
           "issue_active_pct": opt("smsp__issue_active.avg.pct_of_peak_sustained_active"),
           "pipe_pct": {"alu": opt("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
                        "fma": opt("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
                        "fp64": opt("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
                        "xu": opt("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
                        "tensor": opt("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")},
           "dram_bytes_read": rd, "dram_bytes_write": wr, "dram_bytes_per_unit": (rd + wr) / int(units),
           "gpu_time_us": float(best[H.index("gpu__time_duration.sum")].replace(",", "")) *
                          {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(rows[1][H.index("gpu__time_duration.sum")], 1.0),
           "also": None if extra is None else extra[H.index("Kernel Name")].split("(")[0],
           "source": source, "source_sha256": hashlib.sha256(open(source, "rb").read()).hexdigest()}
    json.dump(rec, open(out_json, "w"), indent=1)
    print(json.dumps(rec))


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    elif sys.argv[1] == "traffic":
        traffic(*sys.argv[2:8])
    else:
        raw(sys.argv[2], sys.argv[3:] or DEFAULT)
