#!/usr/bin/env python
"""Probe (2+ ranks under torchrun): can this box give NVLink multicast (NVLS) mappings through torch symmetric memory?"""
import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
try:
    t = symm_mem.empty(1 << 20, dtype=torch.float32, device=torch.device("cuda", local))
    h = symm_mem.rendezvous(t, group=dist.group.WORLD)
    print(rank, "buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "multicast_ptr", hex(h.multicast_ptr), flush=True)
except Exception as e:
    print(rank, "symm_mem failed:", repr(e)[:500], flush=True)
dist.barrier()
dist.destroy_process_group()
