"""The driver-coupling parity test on 2 GPUs (library-driven peer-store halo exchange):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/probe/coupling_2gpu.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import torch.distributed as dist

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda:%d" % rank))
from pace_b200.distributed import Placement, TorchComm  # noqa: E402
from tests.test_coupling import run_coupling  # noqa: E402

run_coupling("cuda:%d" % rank, comm=TorchComm(), mine=Placement(6, world).ranks_of(rank))
dist.barrier()
print("rank %d: coupling on %d GPUs bit-identical to the 6-rank reference" % (rank, world), flush=True)
dist.destroy_process_group()
