"""TEST INFRASTRUCTURE ONLY — compile all reference stencils once (single process) so that multi-rank
runs read a warm GT4Py cache instead of racing to write it."""
import sys
import time

sys.path.insert(0, __file__.rsplit("/oracle/", 1)[0])
from oracle.refshim.bootstrap import bootstrap  # noqa: E402

bootstrap()
from ndsl.comm.null_comm import NullComm  # noqa: E402

from oracle.refshim.run_reference import build_dycore, default_config_kwargs  # noqa: E402

layout = int(sys.argv[1]) if len(sys.argv) > 1 else 1
npx = int(sys.argv[2]) if len(sys.argv) > 2 else 13
n_split = int(sys.argv[3]) if len(sys.argv) > 3 else 1
k_split = int(sys.argv[4]) if len(sys.argv) > 4 else 1
ranks = range(layout * layout) if layout > 1 else [0]
for r in ranks:
    t0 = time.time()
    build_dycore(NullComm(rank=r, total_ranks=6 * layout * layout, fill_value=0.0),
                 default_config_kwargs(npx=npx, layout=(layout, layout), n_split=n_split, k_split=k_split))
    print("rank", r, "built in %.1f s" % (time.time() - t0), flush=True)
