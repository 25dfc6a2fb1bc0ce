#!/usr/bin/env python
"""Multi-GPU parity of the sharded path with the fused peer-store gather (run under torchrun, one rank per GPU).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/check_gather.py

Every rank reduces its y-band of a global synthetic cube and stores the result band into every rank's copy of the full
image (aflak_b200/csrc/shard.cpp).  Each rank then checks its own copy of the FULL image, bit for bit in sequential
mode, against (a) the same global cube reduced unsharded on that rank's GPU and (b) the oracle on sampled rows.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
import torch
import torch.distributed as dist

import aflak_b200 as afk
import oracle

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
afk.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
afk.set_stream(stream.cuda_stream)


def exchange(h):
    lst = [None] * world
    dist.all_gather_object(lst, h)
    return lst


SEED = 0xAF1A0005
ok = True
for (nz, ny, nx) in [(256, 1024 + 3, 512), (64, 74, 74), (128, 4096, 1024)]:
    g = afk.Gather(world, rank, ny, nx, exchange)
    y0, rows = g.band()
    band = afk.Image.synthetic(SEED, nz, ny, nx, y0=y0, rows=rows)
    whole = afk.Image.synthetic(SEED, nz, ny, nx)
    on, off1, off2 = (nz // 2 - 8, nz // 2 + 8), (8, 24), (nz - 24, nz - 8)
    z, z1, z2 = [(a + b) / 2 for a, b in (on, off1, off2)]
    r1 = np.float32(np.float32(z - z1) / np.float32(z2 - z1))
    cases = [("integral", [(0, nz)], afk.EPI_SUM, None),
             ("average/average/ratio", [(0, nz // 2), (nz // 2, nz)], afk.EPI_MEAN_RATIO, None),
             ("equivalent width", [on, off1, off2], afk.EPI_EW, [float(np.float32(1.0) - r1), float(r1), 20.0, 1e10, -1.0])]
    for name, windows, epi, params in cases:
        for mode in (afk.SUM_SEQUENTIAL, afk.SUM_COMPENSATED):
            for rep in range(3):  # exercise both result slots and the epoch flags
                g.reduce(band.device_ptr, nz, windows, epi, params, sum_mode=mode)
                g.wait()
            got = g.result_image().download()
            ref = afk.Image.alloc((ny, nx))
            afk.k_reduce(whole.device_ptr, nz, ny * nx, windows, epi, ref.device_ptr, params, sum_mode=mode)
            want = ref.download()
            if mode == afk.SUM_SEQUENTIAL:
                same = np.array_equal(got.view(np.uint32), want.view(np.uint32))
            else:
                same = np.allclose(got, want, rtol=1e-5, atol=1e-5)
            if not same:
                ok = False
                print(f"rank {rank}: MISMATCH {name} mode {mode} shape {(nz, ny, nx)}: {(got != want).sum()} pixels", flush=True)
    g.check()
    # oracle on sampled rows (sequential integral)
    g.reduce(band.device_ptr, nz, [(0, nz)], afk.EPI_SUM, sum_mode=afk.SUM_SEQUENTIAL)
    g.wait()
    got = g.result_image().download()
    for yy in (0, ny // 2, ny - 2):
        hb = oracle.gen_cube(SEED, nz, ny, nx, y0=yy, rows=2, threads=2)
        if not np.array_equal(got[yy:yy + 2], oracle.integral(hb, 0, nz)):
            ok = False
            print(f"rank {rank}: oracle mismatch rows {yy} shape {(nz, ny, nx)}", flush=True)
    torch.cuda.synchronize()
    dist.barrier()
    g.close()
    del band, whole

t = torch.tensor([0 if ok else 1], device="cuda")
dist.all_reduce(t)
if rank == 0:
    print("GATHER CHECK", "OK" if int(t.item()) == 0 else "FAILED", f"world={world}", flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if int(t.item()) == 0 else 1)
