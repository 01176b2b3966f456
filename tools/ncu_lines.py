"""Per-source-line executed warp instructions and stall samples of an ncu report (dev helper):
    python tools/ncu_lines.py gpurun_out/x.ncu-rep [min_pct]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
minpct = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur = None
lines = []
hdr = None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        ii = hdr.index("Instructions Executed")
        si = hdr.index("# Samples")
        continue
    if hdr and r[0].isdigit():
        try:
            lines.append((cur, int(r[0]), r[1], int(r[ii]), int(r[si])))
        except ValueError:
            pass
tot = sum(l[3] for l in lines) or 1
tots = sum(l[4] for l in lines) or 1
print("total warp instructions %d, samples %d" % (tot, tots))
for f, n, src, ins, smp in lines:
    if 100.0 * ins / tot >= minpct or 100.0 * smp / tots >= minpct:
        print("%-22s %4d  inst %5.1f%%  stall %5.1f%%  %s" % (f, n, 100.0 * ins / tot, 100.0 * smp / tots, src.strip()[:110]))
