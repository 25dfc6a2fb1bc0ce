#!/usr/bin/env python
"""Median over the ALMA-shape cube: device time of the kernel the environment selects (AFK_MEDIAN_SINGLE_PASS=0|1) for a few
window lengths, with parity on sampled row bands against the oracle.  One JSON line.

    AFK_MEDIAN_SINGLE_PASS=1 python tools/bench_median.py
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
import torch

import aflak_b200 as afk
import oracle

afk.set_device(0)
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
afk.set_stream(stream.cuda_stream)
PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
nz, ny, nx = 3840, 1024, 1024
SEED = 0xAF1A0003
cube = afk.Image.synthetic(SEED, nz, ny, nx)
out = {"env": {k: v for k, v in os.environ.items() if k.startswith("AFK_")}, "cube": [nz, ny, nx], "peak_GBps": PEAK, "windows": {}}
for s, e in [(0, 3840), (0, 3839), (100, 1125), (1000, 3000)]:
    ts = []
    for i in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        med = afk.call("median", cube, s, e)
        e1.record(stream)
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts[1:]))
    nb = (e - s) * ny * nx * 4 + ny * nx * 4
    got = med.download()
    ok = True
    for y0 in (0, 511, 1021):
        band = oracle.gen_cube(SEED, nz, ny, nx, y0=y0, rows=3, threads=oracle.max_threads())
        want = oracle.median(band, s, e)
        ok &= bool(np.array_equal(got[y0:y0 + 3].view(np.uint32), want.view(np.uint32)))
    out["windows"][f"[{s},{e})"] = {"ms": round(ms, 3), "GBps": round(nb / ms / 1e6, 1), "frac_of_peak": round(nb / ms / 1e6 / PEAK, 3), "parity_sampled_rows": ok}
print(json.dumps(out))
