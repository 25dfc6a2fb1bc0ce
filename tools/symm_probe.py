#!/usr/bin/env python
"""Does this box give us NVLink multicast (NVLS) memory?  Run under torchrun with >= 2 ranks.  torch's symmetric-memory
allocator does the cuMem / multicast-object / fd plumbing; we only need the pointers it ends up with."""
import os
import sys

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
try:
    t = symm_mem.empty(1 << 20, dtype=torch.uint8, device=torch.device("cuda", local))
    h = symm_mem.rendezvous(t, dist.group.WORLD)
    info = {"rank": rank, "world": h.world_size, "buffer_ptrs": [hex(p) for p in h.buffer_ptrs], "multicast_ptr": hex(h.multicast_ptr),
            "signal_pad_ptrs": [hex(p) for p in h.signal_pad_ptrs], "backend": symm_mem.get_backend(torch.device("cuda", local))}
    print(info, flush=True)
except Exception as e:
    print(f"rank {rank}: symmetric memory failed: {type(e).__name__}: {e}", flush=True)
import ctypes
try:
    cu = ctypes.CDLL("libcuda.so.1")
    v = ctypes.c_int()
    dev = ctypes.c_int()
    cu.cuDeviceGet(ctypes.byref(dev), local)
    cu.cuDeviceGetAttribute(ctypes.byref(v), 132, dev)  # CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED
    print(f"rank {rank}: CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED = {v.value}", flush=True)
except Exception as e:
    print("cuda driver query failed", e)
dist.barrier()
dist.destroy_process_group()
