#!/usr/bin/env python
"""Device-time measurements of the other axis-0 reductions on the ALMA-shape (C3) and MaNGA-shape (C1)
cubes.  CUDA events on the library's stream; algorithmic bytes = 4 B per voxel read + 4 B per output pixel.

    python tools/bench_axis0.py > gpurun_out/axis0.json
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
import torch

import aflak_b200 as afk

afk.set_device(0)
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
afk.set_stream(stream.cuda_stream)
PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, reps, warm=3, cold=False):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        if cold:
            flush_buf.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        fn()
        e1.record(stream)
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts = np.array(ts)
    return float(np.median(ts)), float(ts.min())


def entry(name, nbytes, fn, reps=10, cold=False):
    l0 = afk.launch_count()
    fn()
    nl = afk.launch_count() - l0
    med, best = timed(fn, reps, cold=cold)
    e = {"config": name, "algorithmic_bytes": nbytes, "kernel_launches": nl, "ms_median": med, "ms_best": best,
         "GBps": nbytes / (med * 1e-3) / 1e9, "frac_of_measured_peak": nbytes / (med * 1e-3) / 1e9 / PEAK, "l2": "cold" if cold else "n/a (input >> L2)"}
    print(json.dumps(e), file=sys.stderr, flush=True)
    return e


def suite(tag, nz, ny, nx, windows, reps, cold):
    out = []
    cube = afk.Image.synthetic(0xAF1A0003, nz, ny, nx)
    cube.set_axis(2, "WAVE", "Angstrom")
    cube.set_wcs_axis(2, 3621.6, 1.0, 0.83)
    n2 = ny * nx
    for (s, e) in windows:
        nb = (e - s) * n2 * 4 + n2 * 4
        w = f"[{s},{e})"
        out.append(entry(f"{tag} integral{w} (reference kernel)", nb, lambda: afk.call("integral", cube, s, e), reps, cold))
        for mode, mname in ((afk.SUM_COMPENSATED, "default"), (afk.SUM_SEQUENTIAL, "sequential")):
            afk.set_sum_mode(mode)
            out.append(entry(f"{tag} variance{w} {mname}", nb, lambda: afk.call("variance", cube, s, e), reps, cold))
            out.append(entry(f"{tag} extract_centrobaric_wavelength{w} {mname}", nb, lambda: afk.call("extract_centrobaric_wavelength", cube, s, e), reps, cold))
        afk.set_sum_mode(afk.SUM_COMPENSATED)
        out.append(entry(f"{tag} extract_argmin_max_wavelength{w} (2 outputs)", nb + n2 * 4, lambda: afk.call("extract_argmin_max_wavelength", cube, s, e, False), reps, cold))
        out.append(entry(f"{tag} median{w}", nb, lambda: afk.call("median", cube, s, e), max(3, reps // 3), cold))
    del cube
    return out


out = []
out += suite("C1 MaNGA 74x74x4563", 4563, 74, 74, [(0, 4563), (3134, 3154)], 20, True)
out += suite("C3 ALMA 1024x1024x3840", 3840, 1024, 1024, [(0, 3840), (1900, 1940)], 6, False)
print(json.dumps({"peak_gbs": PEAK, "entries": out}, indent=1))
