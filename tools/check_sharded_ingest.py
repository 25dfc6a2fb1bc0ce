#!/usr/bin/env python
"""Multi-GPU end-to-end check of the sharded path from a FITS file (run under torchrun, one rank per GPU):

    file on disk -> every rank reads and decodes only its own y-band (afk_fits_to_image_band)
                 -> fused reduce + peer-store gather of the 2-D result (afk_gather_*)
                 -> display statistics and colour mapping on the gathered image

checked on every rank against the same file ingested whole on that rank's GPU, and on rank 0 against the oracle.
Also times the sharded ingest (aggregate file GB/s) and the per-band axis-0 operators added in SURVEY 8f.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/check_sharded_ingest.py
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))

import numpy as np
import torch
import torch.distributed as dist

import aflak_b200 as afk
import oracle
from util import fits_hdu_bytes

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


ok = True
report = {"world": world, "cases": []}
tmp = os.environ.get("AFK_BENCH_TMP", "/tmp")
for tag, (nz, ny, nx), dt, cards in [
    ("i16 BSCALE/BZERO/BLANK", (96, 1027, 512), ">i2", [("BSCALE", 0.25), ("BZERO", 8192), ("BLANK", -32768), ("BUNIT", "adu"), ("CTYPE3", "WAVE"),
                                                          ("CUNIT3", "Angstrom"), ("CRVAL3", 3621.6), ("CRPIX3", 1), ("CDELT3", 0.83)]),
    ("f32", (64, 2048, 1024), ">f4", [("BUNIT", "Jy/beam")]),
]:
    path = os.path.join(tmp, f"afk_sharded_{tag.split()[0]}.fits")
    if rank == 0:
        rng = np.random.default_rng(12)
        a = (rng.integers(-2000, 30000, (nz, ny, nx)) if dt == ">i2" else rng.normal(5.0, 2.0, (nz, ny, nx))).astype(dt)
        if dt == ">i2":
            a[::7, ::13, ::5] = -32768
        with open(path, "wb") as f:
            f.write(fits_hdu_bytes(a, cards))
    dist.barrier()
    fits = afk.Fits.open(path)
    whole = fits.to_image(0, "")
    t0 = time.perf_counter()
    band = fits.to_image(0, "", world=world, rank=rank)
    afk.synchronize()
    t_band = time.perf_counter() - t0
    g = afk.Gather(world, rank, ny, nx, exchange)
    y0, rows = g.band()
    assert band.shape == (nz, rows, nx), (band.shape, rows)
    if not np.array_equal(band.download().view(np.uint32), whole.download()[:, y0:y0 + rows].view(np.uint32)):
        ok = False
        print(f"rank {rank}: band decode differs from the whole-file decode ({tag})", flush=True)
    # fused reduce + gather, sequential mode: bit-identical to the unsharded reduction
    for windows, epi in [([(0, nz)], afk.EPI_SUM), ([(0, nz // 2), (nz // 2, nz)], afk.EPI_MEAN_RATIO)]:
        g.reduce(band.device_ptr, nz, windows, epi, sum_mode=afk.SUM_SEQUENTIAL)
        g.wait()
        got = g.result_image().download()
        ref = afk.Image.alloc((ny, nx))
        afk.k_reduce(whole.device_ptr, nz, ny * nx, windows, epi, ref.device_ptr, sum_mode=afk.SUM_SEQUENTIAL)
        want = ref.download()
        nan_ok = np.array_equal(np.isnan(got), np.isnan(want))
        if not (nan_ok and np.array_equal(got[~np.isnan(got)], want[~np.isnan(want)])):
            ok = False
            print(f"rank {rank}: gathered result differs ({tag}, epilogue {epi})", flush=True)
    g.check()
    # the gathered integral is an ordinary image on every rank: statistics + colour mapping must agree with the unsharded ones
    g.reduce(band.device_ptr, nz, [(0, nz)], afk.EPI_SUM, sum_mode=afk.SUM_SEQUENTIAL)
    g.wait()
    full = g.result_image()
    afk.set_sum_mode(afk.SUM_SEQUENTIAL)
    ref = afk.call("integral", whole, 0, nz)[0]
    afk.set_sum_mode(afk.SUM_COMPENSATED)
    # (the BLANK pixels of the i16 case are NaN in the integral; the statistics skip them the same way on both sides)
    vm_a, vm_b = full.vmin_vmax(), ref.vmin_vmax()
    st_a, st_b = full.vmed_vmad(), ref.vmed_vmad()
    h_a, h_b = full.histogram(float(vm_a[0]), float(vm_a[1]))[0], ref.histogram(float(vm_b[0]), float(vm_b[1]))[0]
    lut = afk.Lut("Flame")
    if not (vm_a == vm_b and st_a == st_b and np.array_equal(h_a, h_b) and np.array_equal(afk.colormap(full, lut, channels=4), afk.colormap(ref, lut, channels=4))):
        ok = False
        print(f"rank {rank}: display statistics / colour bytes differ on the gathered image ({tag})", flush=True)
    # per-band axis-0 operators (no collective): each rank's band result equals the rows of the unsharded result
    band.set_axis(2, "WAVE", "Angstrom")
    whole.set_axis(2, "WAVE", "Angstrom")
    for name, args in [("variance", (0, nz)), ("median", (3, nz - 2)), ("extract_argmin_max_wavelength", (0, nz, False))]:
        try:
            b = afk.call(name, band, *args)
            w = afk.call(name, whole, *args)
        except afk.AflakError as e:  # median panics on the NaN of BLANK pixels, on both sides
            if "NaN" not in str(e):
                raise
            continue
        b, w = (b[0], w[0]) if isinstance(b, tuple) else (b, w)
        bb, ww = b.download(), w.download()[y0:y0 + rows]
        if not (np.array_equal(np.isnan(bb), np.isnan(ww)) and np.allclose(bb[~np.isnan(bb)], ww[~np.isnan(ww)], rtol=1e-6, atol=0)):
            ok = False
            print(f"rank {rank}: {name} on the band differs from the unsharded rows ({tag})", flush=True)
    if rank == 0:  # oracle on the first rows of the whole decode
        raw = np.fromfile(path, dtype=np.uint8, offset=2880, count=nz * ny * nx * int(dt[2]))
        fmt = {">i2": 2, ">f4": 0}[dt]
        kw = dict(blank=-32768, bscale=0.25, bzero=8192.0) if dt == ">i2" else {}
        want = oracle.fits_decode(raw[: ny * nx * int(dt[2])], fmt, ny * nx, **kw).reshape(ny, nx)
        got = whole.download()[0]
        if not (np.array_equal(np.isnan(got), np.isnan(want)) and np.array_equal(got[~np.isnan(got)], want[~np.isnan(want)])):
            ok = False
            print("rank 0: whole-file decode differs from the oracle", flush=True)
    tt = torch.tensor([t_band], device="cuda", dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    report["cases"].append({"case": tag, "shape": [nz, ny, nx], "file_bytes": nz * ny * nx * int(dt[2]), "band_ingest_ms_max_over_ranks": float(tt.item()) * 1e3,
                            "aggregate_file_GBps": nz * ny * nx * int(dt[2]) / float(tt.item()) / 1e9})
    torch.cuda.synchronize()
    dist.barrier()
    g.close()
    del band, whole, full, ref
    if rank == 0:
        os.remove(path)

t = torch.tensor([0 if ok else 1], device="cuda")
dist.all_reduce(t)
if rank == 0:
    report["ok"] = int(t.item()) == 0
    print("SHARDED INGEST CHECK", "OK" if report["ok"] else "FAILED", json.dumps(report), flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if int(t.item()) == 0 else 1)
