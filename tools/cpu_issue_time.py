"""Dev helper: host time to ENQUEUE one dycore step (no synchronisation) vs. the GPU time of the step."""
import sys
import time

import torch

sys.path.insert(0, ".")
from pace_b200.synthetic import make_engine

n = int(sys.argv[1]) if len(sys.argv) > 1 else 48
eng = make_engine(n, 1, n_split=6)
for _ in range(2):
    eng.step_dynamics()
torch.cuda.synchronize()
t0 = time.time()
eng.step_dynamics()
t1 = time.time()
torch.cuda.synchronize()
t2 = time.time()
print("C%d: enqueue %.1f ms, until done %.1f ms" % (n, (t1 - t0) * 1e3, (t2 - t0) * 1e3))
