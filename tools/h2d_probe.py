#!/usr/bin/env python
"""Host->device copy-rate probe for the e2e path: pinned memory, chunk sizes, NUMA placement."""
import os
import subprocess
import sys
import time

import torch

print(subprocess.run("nvidia-smi topo -m; lscpu | grep -E 'NUMA|Model name|^CPU\\(s\\)|Socket'; nproc; free -g | head -2", shell=True,
                     capture_output=True, text=True).stdout)
dev = torch.device("cuda", 0)
GB = 1 << 30
n = 4 * GB
d = torch.empty(n, dtype=torch.uint8, device=dev)


def rate(h, chunks=1, reps=3):
    best = 1e9
    step = n // chunks
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for c in range(chunks):
            d[c * step:(c + 1) * step].copy_(h[c * step:(c + 1) * step], non_blocking=True)
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return n / best / 1e9


ncpu = os.cpu_count()
all_cpus = sorted(os.sched_getaffinity(0))
print("affinity", len(all_cpus), "cpus")
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h.fill_(1)
for ch in (1, 4, 16, 64):
    print(f"default placement, {ch} chunks: {rate(h, ch):.1f} GB/s", flush=True)
del h
# try first-touch on each half of the cpu list (a crude NUMA sweep)
halves = [all_cpus[: len(all_cpus) // 2], all_cpus[len(all_cpus) // 2:]]
for i, cp in enumerate(halves):
    try:
        os.sched_setaffinity(0, cp)
        h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
        h.fill_(1)
        print(f"cpus {cp[0]}..{cp[-1]}: {rate(h, 1):.1f} GB/s", flush=True)
        del h
    except Exception as e:
        print("affinity sweep failed", e)
os.sched_setaffinity(0, all_cpus)
# D2H for completeness
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
torch.cuda.synchronize()
t0 = time.perf_counter()
h.copy_(d, non_blocking=True)
torch.cuda.synchronize()
print(f"D2H: {n / (time.perf_counter() - t0) / 1e9:.1f} GB/s")
