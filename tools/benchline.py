"""Print the interesting numbers of a bench.py JSON line read from stdin (dev helper)."""
import json
import sys

d = json.loads(sys.stdin.read().strip().splitlines()[-1])
lab = d["roofline"]["labelled_kernels_ms_per_step"] if d.get("roofline") else {}
print(" ".join(sys.argv[1:]), "ms/step %.1f" % d["ms_per_step"], "launches", d.get("gpu_launches"),
      {k: round(v, 1) for k, v in lab.items()})
