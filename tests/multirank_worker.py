#!/usr/bin/env python
"""Worker of tests/test_gpu_multirank.py: run under torchrun, one rank per GPU (also usable by hand:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 tests/multirank_worker.py)

Every rank reduces its y-band of a global synthetic cube with the fused peer-store gather (aflak_b200/csrc/shard.cpp,
kernels.cu reduce_kernel epilogue) and then checks ITS copy of the full image against (a) the same cube reduced unsharded on
its own GPU and (b) the oracle, for the three epilogues x two sum modes, with uneven band splits (ny % world != 0), an empty
band (ny < world) and enough repetitions to use both result slots and wrap the epoch parity.  Also exercises the sharded
display path: the min/max exchange and the RGBA band gather.
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
# only the 64-byte IPC handles and the verdict cross the process group (gloo); the NCCL half exists for torch's symmetric-memory
# allocator, which supplies the multicast mapping of the second pass
dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
oracle.build()


def exchange(h):
    lst = [None] * world
    dist.all_gather_object(lst, h)
    return lst


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def same_bits(a, b):
    nan = np.isnan(a)
    return np.array_equal(nan, np.isnan(b)) and np.array_equal(bits(a)[~nan], bits(b)[~nan])


SEED = 0xAF1A0005
failures = []
notes = []
# (nz, ny, nx): uneven split; small image (k-split variant of the kernel); an empty band on the last rank(s); vec=1 lanes.
# Every shape runs twice: library-allocated buffers + CUDA IPC peer stores, then host-supplied symmetric memory whose NVLink
# multicast mapping (where the fabric has one) turns the epilogue's stores into one multimem.st per tile.
for memory, (nz, ny, nx) in [(m, s) for m in ("ipc", "symmetric") for s in [(96, 257, 128), (64, 74 + 1, 74), (48, max(world - 1, 1), 256), (40, 33, 37)]]:
    if memory == "symmetric":
        if world < 2:
            continue
        try:
            sym = afk.symmetric_memory_torch(afk.gather_bytes(ny, nx))
        except Exception as e:  # no symmetric-memory support in this torch / driver: say so, the IPC pass above stands
            notes.append(f"rank {rank}: symmetric memory unavailable ({type(e).__name__}: {str(e)[:120]})")
            break
        if not sym[1]:
            notes.append(f"rank {rank}: no multicast mapping on this fabric; symmetric pass uses unicast peer stores")
        g = afk.Gather(world, rank, ny, nx, symmetric=sym)
    else:
        g = afk.Gather(world, rank, ny, nx, exchange if world > 1 else None)
    y0, rows = g.band()
    assert (y0, rows) == afk.shard_band(ny, world, rank)
    band = afk.Image.synthetic(SEED, nz, ny, nx, y0=y0, rows=rows) if rows > 0 else afk.Image.alloc((nz, 0, nx))
    whole = afk.Image.synthetic(SEED, nz, ny, nx)
    host = oracle.gen_cube(SEED, nz, ny, nx)
    on, off1, off2 = (nz // 2 - 8, nz // 2 + 8), (4, 16), (nz - 16, nz - 4)
    z, z1, z2 = [np.float32((a + b) / 2) for a, b in (on, off1, off2)]
    r1 = np.float32(np.float32(z - z1) / np.float32(z2 - z1))
    ewp = [float(np.float32(1.0) - r1), float(r1), 20.0, 1e10, -1.0]
    cases = [("integral", [(0, nz)], afk.EPI_SUM, None), ("average", [(3, nz - 5)], afk.EPI_MEAN, None),
             ("average/average/ratio", [(0, nz // 2), (nz // 2, nz)], afk.EPI_MEAN_RATIO, None),
             ("equivalent width", [on, off1, off2], afk.EPI_EW, ewp)]
    for name, windows, epi, params in cases:
        for mode in (afk.SUM_SEQUENTIAL, afk.SUM_COMPENSATED):
            for rep in range(3):  # both result slots, epoch parity wraps
                g.reduce(band.device_ptr, nz, windows, epi, params, sum_mode=mode)
                g.wait()
            got = g.result_image().download()
            ref = afk.Image.alloc((ny, nx))
            afk.k_reduce(whole.device_ptr, nz, ny * nx, windows, epi, ref.device_ptr, params, sum_mode=mode)
            unsharded = ref.download()
            if epi == afk.EPI_SUM:
                want = oracle.integral(host, *windows[0])
            elif epi == afk.EPI_MEAN:
                want = oracle.average(host, *windows[0])
            elif epi == afk.EPI_MEAN_RATIO:
                want = oracle.ratio(oracle.average(host, *windows[0]), oracle.average(host, *windows[1]))
            else:
                off = oracle.linear_composition(oracle.average(host, *off1), oracle.average(host, *off2), ewp[0], ewp[1])
                want = oracle.equivalent_width(off, oracle.average(host, *on), 20.0, 1e10, True)
            if mode == afk.SUM_SEQUENTIAL:
                # the sequential sum never splits a pixel: sharded == unsharded == oracle, bit for bit
                if not same_bits(got, unsharded):
                    failures.append(f"rank {rank} {memory} {(nz, ny, nx)} {name} sequential: sharded != unsharded")
                if not same_bits(got, want):
                    failures.append(f"rank {rank} {memory} {(nz, ny, nx)} {name} sequential: sharded != oracle")
            else:
                # compensated: the split of the spectral axis between CTAs depends on the band shape, so sharded and
                # unsharded agree to the tier-C tolerance, not bit for bit
                tol = dict(rtol=2e-4, atol=2e-4) if epi == afk.EPI_EW else dict(rtol=1e-5, atol=0)
                if not np.allclose(got, want, equal_nan=True, **tol) or not np.allclose(got, unsharded, equal_nan=True, **tol):
                    failures.append(f"rank {rank} {memory} {(nz, ny, nx)} {name} compensated: outside tolerance")
    g.check()
    # protocol enforcement: a second reduce without a wait is refused
    g.reduce(band.device_ptr, nz, [(0, nz)], afk.EPI_SUM)
    try:
        g.reduce(band.device_ptr, nz, [(0, nz)], afk.EPI_SUM)
        failures.append(f"rank {rank}: a reduce without a wait in between was accepted")
    except afk.AflakError as e:
        if "afk_gather_wait" not in str(e):
            failures.append(f"rank {rank}: unexpected message {e}")
    g.wait()

    # ---- sharded display path: min/max exchange + RGBA band gather ----
    if nx % 4 == 0:
        img = oracle.integral(host, 0, nz)
        ref = afk.Image.alloc((ny, nx))
        afk.k_reduce(whole.device_ptr, nz, ny * nx, [(0, nz)], afk.EPI_SUM, ref.device_ptr, sum_mode=afk.SUM_SEQUENTIAL)
        lut = afk.Lut("Flame")
        lut.set_lims(0.2, 0.3, 0.9)
        table = oracle.lut_build("Flame")
        tone = oracle.create_curve([(0.0, 0.0), (0.4, 0.7), (1.0, 1.0)])
        lmax = np.float32(np.log(np.float32(1000.0) + np.float32(1.0)) / np.log(np.float32(1000.0)))
        chain = {"logscale_a": 1000.0, "logscale_vmin": float("nan"), "logscale_vmax": float("nan"), "normalize_min": 0.0,
                 "normalize_max": float(lmax / np.float32(65535.0)), "tone_table": tone}
        for rep in range(3):  # both RGBA slots
            for flip in (True, False):
                g.reduce(band.device_ptr, nz, [(0, nz)], afk.EPI_SUM, sum_mode=afk.SUM_SEQUENTIAL)
                g.wait()
                res = g.result_image()
                l0 = afk.launch_count()
                mn, mx = afk.call("image_min_max", res)
                if afk.launch_count() != l0:
                    failures.append(f"rank {rank} {(nz, ny, nx)}: image_min_max of the gathered image made a pass (no range memo)")
                if (np.float32(mn), np.float32(mx)) != tuple(np.float32(v) for v in oracle.image_min_max(img)):
                    failures.append(f"rank {rank} {(nz, ny, nx)}: exchanged min/max {(mn, mx)} != {oracle.image_min_max(img)}")
                # plain colour map, limits = get_vmin / get_vmax of the WHOLE image (lims.rs:7-35)
                rgba = g.display(lut, flip_y=flip)
                # the oracle's make_raw_image includes the vertical flip of image.rs:123-129
                want = oracle.make_raw_image(img if flip else img[::-1], oracle.get_vmin(img), oracle.get_vmax(img), table,
                                             lims=(0.2, 0.3, 0.9), channels=4)
                if not np.array_equal(rgba.reshape(-1), np.asarray(want).reshape(-1)):
                    bad = np.argwhere((rgba != want).any(axis=2))
                    failures.append(f"rank {rank} {(nz, ny, nx)} flip={flip}: gathered RGBA differs from the oracle at {len(bad)} pixels, "
                                    f"rows {sorted(set(bad[:, 0].tolist()))[:8]}, first {bad[0].tolist()} got {rgba[tuple(bad[0])].tolist()} want {want[tuple(bad[0])].tolist()}")
                # fused chain (BASELINE config 4) on the band, logscale range from the exchange: same bytes as one GPU on the whole image
                rgba = g.display(lut, vmin=0.0, vmax=65535.0, flip_y=flip, chain=chain)
                single = afk.colormap(ref, lut, vmin=0.0, vmax=65535.0, flip_y=flip, channels=4, chain=chain)
                if not np.array_equal(rgba, single):
                    failures.append(f"rank {rank} {(nz, ny, nx)} flip={flip}: gathered chain RGBA differs from the single-GPU chain")
        g.check()
    torch.cuda.synchronize()
    dist.barrier()
    g.close()
    del band, whole

allf = [None] * world
dist.all_gather_object(allf, failures)
flat = [f for fs in allf for f in fs]
allnotes = [None] * world
dist.all_gather_object(allnotes, notes)
if rank == 0:
    for f in flat + sorted({n for ns in allnotes for n in ns}):
        print(f, flush=True)
    print("MULTICAST", "used" if world > 1 and not any(ns for ns in allnotes) else "not used", flush=True)
    print("MULTIRANK", "OK" if not flat else "FAILED", f"world={world}", flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if not flat else 1)
