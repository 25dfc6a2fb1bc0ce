#!/usr/bin/env python
"""FITS ingest measurements on one B200.

  kernel   the decode kernel alone on device-resident raw bytes (HBM roofline: element bytes + 4 per voxel)
  memory   afk_image_ingest from pinned host memory (PCIe roofline: element bytes per voxel), vs the plain f32 upload
  file     open_fits + fits_to_image of a FITS file on local disk (page cache warm), wall clock
  cpu      the reference's host path restated (oracle decode: byte swap + cast + *bscale + bzero, one thread like fitrs/ndarray)

    python tools/bench_ingest.py > gpurun_out/ingest.json
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))

import numpy as np
import torch

import aflak_b200 as afk
import oracle
from util import fits_hdu_bytes

afk.set_device(0)
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
afk.set_stream(stream.cuda_stream)
PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
FORMATS = [(afk.FITS_F32, ">f4", "BITPIX -32"), (afk.FITS_F64, ">f8", "BITPIX -64"), (afk.FITS_I16, ">i2", "BITPIX 16"), (afk.FITS_I32, ">i4", "BITPIX 32")]
out = {"peak_gbs": PEAK, "kernel": [], "memory": [], "file": [], "cpu": []}


def ev_time(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        fn()
        e1.record(stream)
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


# ---- decode kernel alone: 2 Gi voxels (8 GiB f32 out) ----
n = 1 << 31
dst = torch.empty(n, dtype=torch.float32, device="cuda")
for fmt, dt, name in FORMATS:
    es = int(dt[2])
    raw = torch.randint(0, 255, (n * es,), dtype=torch.uint8, device="cuda")
    ms = ev_time(lambda: afk.k_ingest(raw.data_ptr(), fmt, n, dst.data_ptr(), bscale=0.5, bzero=1.0))
    nb = n * (es + 4)
    e = {"format": name, "voxels": n, "algorithmic_bytes": nb, "ms": ms, "GBps": nb / ms / 1e6, "frac_of_measured_peak": nb / ms / 1e6 / PEAK}
    if es == 4:
        ms2 = ev_time(lambda: afk.k_ingest(raw.data_ptr(), fmt, n, raw.data_ptr(), bscale=0.5, bzero=1.0))
        e["in_place_ms"] = ms2
        e["in_place_GBps"] = nb / ms2 / 1e6
    out["kernel"].append(e)
    print(json.dumps(e), file=sys.stderr, flush=True)
    del raw
del dst
torch.cuda.empty_cache()

# ---- from pinned host memory: 1 Gi voxels ----
n = 1 << 30
shape = (1024, 1024, 1024)
for fmt, dt, name in FORMATS:
    es = int(dt[2])
    host = torch.empty(n * es, dtype=torch.uint8).pin_memory()
    host.fill_(7)
    def run():
        img = afk.ingest(host.numpy(), fmt, shape, bscale=0.5, bzero=1.0, sync=False)
        return img
    ms = ev_time(run, reps=5, warm=2)
    e = {"format": name, "voxels": n, "pcie_bytes": n * es, "ms": ms, "pcie_GBps": n * es / ms / 1e6, "Gvoxels_per_s": n / ms / 1e6}
    out["memory"].append(e)
    print(json.dumps(e), file=sys.stderr, flush=True)
    del host
hostf = torch.empty(n, dtype=torch.float32).pin_memory()
hostf.fill_(1.0)
ms = ev_time(lambda: afk.Image.upload(hostf.numpy(), sync=False), reps=5, warm=2)
e = {"format": "plain f32 upload (afk_image_upload, already decoded on the host)", "voxels": n, "pcie_bytes": n * 4, "ms": ms, "pcie_GBps": n * 4 / ms / 1e6, "Gvoxels_per_s": n / ms / 1e6}
out["memory"].append(e)
print(json.dumps(e), file=sys.stderr, flush=True)
del hostf

# ---- from a file: MaNGA-shape f32 cube (100 MB) and a 4 GiB i16 / f32 cube ----
tmp = os.environ.get("AFK_BENCH_TMP", "/tmp")
rng = np.random.default_rng(2)
cases = [("MaNGA-shape 74x74x4563 BITPIX -32", (4563, 74, 74), ">f4", 3), ("1024x1024x1024 BITPIX -32 (4 GiB)", (1024, 1024, 1024), ">f4", 3),
         ("1024x1024x1024 BITPIX 16 (2 GiB)", (1024, 1024, 1024), ">i2", 3)]
for name, shp, dt, reps in cases:
    path = os.path.join(tmp, "afk_bench_ingest.fits")
    nvox = int(np.prod(shp))
    block = rng.integers(0, 30000, 1 << 22).astype(dt)
    with open(path, "wb") as f:
        hdr = fits_hdu_bytes(np.zeros((1, 1, 1), dtype=dt), [("BSCALE", 0.5), ("BZERO", 1.0), ("BUNIT", "Jy/beam")])[:2880]
        txt = hdr.decode()
        for i, nax in enumerate(reversed(shp)):
            txt = txt.replace(f"NAXIS{i + 1}  = {1:>20}", f"NAXIS{i + 1}  = {nax:>20}")
        f.write(txt.encode())
        left = nvox
        while left > 0:
            m = min(left, block.size)
            f.write(block[:m].tobytes())
            left -= m
        f.write(b"\0" * (-(nvox * int(dt[2])) % 2880))
    ts, stats = [], None
    for r in range(reps + 1):
        t0 = time.perf_counter()
        fits = afk.call("open_fits", afk.Paths([path]), 0)
        img = afk.call("fits_to_image", fits, 0, "")
        afk.synchronize()
        dt_s = time.perf_counter() - t0
        if r > 0:
            ts.append(dt_s)
            stats = afk.ingest_stats()
        del img, fits
    nb = nvox * int(dt[2])
    e = {"case": name, "file_bytes": nb, "wall_ms_median": float(np.median(ts)) * 1e3, "file_GBps": nb / float(np.median(ts)) / 1e9,
         "Gvoxels_per_s": nvox / float(np.median(ts)) / 1e9, "last_stats": stats}
    out["file"].append(e)
    print(json.dumps(e), file=sys.stderr, flush=True)
    # CPU: what the reference does after fitrs has the bytes: decode + `as f32` + * bscale + bzero (single thread), on a bounded sample
    sample = min(nvox, 1 << 27)
    raw = np.fromfile(path, dtype=np.uint8, count=sample * int(dt[2]), offset=2880)
    fmtc = {">f4": 0, ">i2": 2}[dt]
    t0 = time.perf_counter()
    oracle.fits_decode(raw, fmtc, sample, bscale=0.5, bzero=1.0)
    cs = time.perf_counter() - t0
    c = {"case": name, "kind": "port (oracle.fits_decode, 1 thread)", "sample_voxels": sample, "ms": cs * 1e3, "Gvoxels_per_s": sample / cs / 1e9}
    out["cpu"].append(c)
    print(json.dumps(c), file=sys.stderr, flush=True)
    os.remove(path)
print(json.dumps(out, indent=1))
