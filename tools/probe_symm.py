"""Probe: does torch symmetric memory (peer buffers over NVLink + device-side barrier) work on this box?"""
import os

import torch
import torch.distributed as dist

rank = int(os.environ["RANK"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda:%d" % rank))
import torch.distributed._symmetric_memory as symm_mem

t = symm_mem.empty(1024, dtype=torch.float64, device="cuda")
t.fill_(-1.0)
hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
print(rank, "ptrs", [hex(p) for p in hdl.buffer_ptrs], "signal", [hex(p) for p in hdl.signal_pad_ptrs], flush=True)
peer = hdl.get_buffer(1 - rank, (1024,), torch.float64)
hdl.barrier()
peer.fill_(float(rank + 1))
hdl.barrier()
print(rank, "local buffer now", t[:3].tolist(), flush=True)
dist.barrier()
dist.destroy_process_group()
