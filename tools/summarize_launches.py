"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (and, for the generic
k_ijk / k_ij launchers, per module function the stencil body belongs to)."""
import collections
import csv
import re
import sys


def main(path, header=""):
    rows = list(csv.reader(l for l in open(path) if not l.startswith("==")))
    hdr = rows[0]
    kn, mv, mn, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name"), hdr.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        if len(r) <= mv or r[mn] != "gpu__time_duration.sum":
            continue
        name = r[kn]
        m = re.search(r"void (k_ijk?)<(\w+)\(", name)
        key = "%s<%s>" % (m.group(1), m.group(2)) if m else re.sub(r"\(.*", "", name).replace("void ", "")[:60]
        scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0}[r[mu]]
        agg[key][0] += 1
        agg[key][1] += float(r[mv].replace(",", "")) * scale
    tot = sum(v[1] for v in agg.values())
    n = sum(v[0] for v in agg.values())
    out = [header, "total kernel time %.2f ms in %d launches" % (tot, n)]
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append("%-46s n=%4d %10.3f ms %5.1f%% %9.1f us/launch" % (k, c, t, 100 * t / tot, 1e3 * t / c))
    return "\n".join(out) + "\n"


if __name__ == "__main__":
    sys.stdout.write(main(sys.argv[1], " ".join(sys.argv[2:])))
