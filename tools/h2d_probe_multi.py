#!/usr/bin/env python
"""Concurrent host->device copy rates with N ranks (one per GPU), to find what caps the e2e upload at N > 1.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 tools/h2d_probe_multi.py

Variants per rank, all ranks copying at the same time (barrier before, max over ranks after):
  one stream / 4 streams over interleaved 64 MiB chunks / write-combined pinned memory / each rank alone (the others idle).
Rank 0 prints one JSON line."""
import ctypes
import json
import os
import subprocess
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
GB = 1 << 30
n = 4 * GB
d = torch.empty(n, dtype=torch.uint8, device=dev)
rt = ctypes.CDLL("libcudart.so.12")


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def gmax(x):
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def copy_rate(h, streams=1, chunk=256 << 20, alone=None):
    """GB/s per rank = bytes / max-over-ranks time."""
    ss = [torch.cuda.Stream(device=dev) for _ in range(streams)]
    best = 1e9
    for _ in range(3):
        barrier()
        t0 = time.perf_counter()
        if alone is None or alone == rank:
            for i, off in enumerate(range(0, n, chunk)):
                with torch.cuda.stream(ss[i % streams]):
                    d[off:off + chunk].copy_(h[off:off + chunk], non_blocking=True)
            torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        best = min(best, gmax(dt) if alone is None else dt)
    return n / best / 1e9


out = {"world": world}
if rank == 0:
    out["topology"] = subprocess.run("nvidia-smi topo -m | head -14; lscpu | grep -E 'NUMA|Model name|^CPU\\(s\\)|Socket'; free -g | head -2", shell=True,
                                     capture_output=True, text=True).stdout
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h.fill_(1)
out["one_stream_256MiB_chunks"] = copy_rate(h)
out["four_streams_64MiB_chunks"] = copy_rate(h, streams=4, chunk=64 << 20)
out["two_streams_1GiB_chunks"] = copy_rate(h, streams=2, chunk=1 << 30)
if world > 1:
    alone = []
    for r in range(world):
        x = copy_rate(h, alone=r)
        alone.append(gmax(x if r == rank else 0.0))
    out["each_rank_alone"] = alone
    # pairs: rank 0 together with rank r (which GPUs share an uplink?)
    pairs = {}
    for r in range(1, world):
        best = 1e9
        for _ in range(2):
            barrier()
            t0 = time.perf_counter()
            if rank in (0, r):
                d.copy_(h, non_blocking=True)
                torch.cuda.synchronize()
            best = min(best, gmax(time.perf_counter() - t0))
        pairs[f"0+{r}"] = n / best / 1e9
    out["rank0_with_one_other_per_rank_GBps"] = pairs
del h
# write-combined pinned memory (cudaHostAllocWriteCombined = 4)
p = ctypes.c_void_p()
if rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(n), ctypes.c_uint(4)) == 0:
    ctypes.memset(p, 1, n)
    best = 1e9
    for _ in range(3):
        barrier()
        t0 = time.perf_counter()
        rt.cudaMemcpyAsync(ctypes.c_void_p(d.data_ptr()), p, ctypes.c_size_t(n), ctypes.c_int(1), ctypes.c_void_p(0))
        torch.cuda.synchronize()
        best = min(best, gmax(time.perf_counter() - t0))
    out["write_combined_one_copy"] = n / best / 1e9
    rt.cudaFreeHost(p)
if rank == 0:
    print(json.dumps(out))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
