"""Per-kernel table from an ncu CSV with several metrics per launch (time, instructions, issue-active, DRAM bytes):
    python tools/summarize_metrics.py gpurun_out/x.csv"""
import collections
import csv
import re
import sys

rows = list(csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")))
hdr = rows[0]
ki, kn, mv, mn, mu = hdr.index("ID"), hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name"), hdr.index("Metric Unit")
per = collections.defaultdict(dict)
names = {}
for r in rows[1:]:
    if len(r) <= mv:
        continue
    v = float(r[mv].replace(",", ""))
    u = r[mu]
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}.get(u, 1.0)
    per[r[ki]][r[mn]] = v
    name = r[kn]
    m = re.search(r"void (k_ijk?(?:_occ)?)<(?:\d+, )?(\w+)\(", name)
    names[r[ki]] = "%s<%s>" % (m.group(1), m.group(2)) if m else re.sub(r"\(.*", "", name).replace("void ", "")[-44:]
agg = collections.defaultdict(lambda: collections.defaultdict(float))
for i, d in per.items():
    a = agg[names[i]]
    t = d.get("gpu__time_duration.sum", 0.0)
    a["n"] += 1
    a["ms"] += t
    a["inst"] += d.get("smsp__inst_executed.sum", 0.0)
    a["bytes"] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
    a["issue_t"] += d.get("smsp__issue_active.avg.pct_of_peak_sustained_active", 0.0) * t
    a["warps_t"] += d.get("sm__warps_active.avg.pct_of_peak_sustained_active", 0.0) * t
    a["fp64_t"] += d.get("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", 0.0) * t
tot = sum(a["ms"] for a in agg.values())
print("total %.1f ms; columns: launches, ms, share, GB/s (DRAM), warp-instr (G), issue-active %%, warps-active %%, fp64 pipe %%" % tot)
for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["ms"]):
    ms = a["ms"] or 1e-9
    print("%-46s n=%4d %8.2f ms %5.1f%% %7.0f GB/s %7.2f Ginst  issue %4.1f  warps %4.1f  fp64 %4.1f" % (
        k, a["n"], a["ms"], 100 * a["ms"] / tot, a["bytes"] / ms / 1e6, a["inst"] / 1e9, a["issue_t"] / ms, a["warps_t"] / ms, a["fp64_t"] / ms))
