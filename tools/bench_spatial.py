#!/usr/bin/env python
"""Device time of the spatial reductions / gathers and remaining elementwise operators.

    python tools/bench_spatial.py > gpurun_out/spatial.json
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
out = {"peak_gbs": PEAK, "entries": []}


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
    return float(np.median(ts))


def entry(name, nbytes, fn, reps=10, cold=False):
    l0 = afk.launch_count()
    fn()
    nl = afk.launch_count() - l0
    ms = timed(fn, reps, cold=cold)
    e = {"config": name, "algorithmic_bytes": nbytes, "kernel_launches": nl, "ms_median": ms, "GBps": nbytes / ms / 1e6,
         "frac_of_measured_peak": nbytes / ms / 1e6 / PEAK, "l2": "cold" if cold else "input >> L2"}
    out["entries"].append(e)
    print(json.dumps(e), file=sys.stderr, flush=True)


# ---- extract_wave / extrude / slice on the ALMA-shape cube ----
nz, ny, nx = 3840, 1024, 1024
cube = afk.Image.synthetic(0xAF1A0003, nz, ny, nx)
n2 = ny * nx
entry("C3 extract_wave ROI::All (whole cube -> 3840-point spectrum)", nz * n2 * 4 + nz * 4, lambda: afk.call("extract_wave", cube, afk.Roi()), 6)
afk.set_sum_mode(afk.SUM_SEQUENTIAL)
entry("C3 extract_wave ROI::All, sequential (bit-exact) mode", nz * n2 * 4 + nz * 4, lambda: afk.call("extract_wave", cube, afk.Roi()), 3)
afk.set_sum_mode(afk.SUM_COMPENSATED)
box = afk.Roi([(x, y) for y in range(500, 532) for x in range(500, 532)])
entry("C3 extract_wave 32x32-pixel region (interactive waveform), new Roi value per call", nz * 1024 * 4 + nz * 4, lambda: afk.call("extract_wave", cube, box), 20)
entry("C3 extrude 32x32-pixel region, new Roi value per call", 2 * nz * 1024 * 4, lambda: afk.call("extrude", cube, box), 20)
boxv = afk.Value(box)  # a constant node of the graph: the value keeps the device copy of its offsets
entry("C3 extract_wave 32x32-pixel region, Roi value reused (re-evaluation after another input changed)", nz * 1024 * 4 + nz * 4,
      lambda: afk.call("extract_wave", cube, boxv), 20)
entry("C3 extrude 32x32-pixel region, Roi value reused", 2 * nz * 1024 * 4, lambda: afk.call("extrude", cube, boxv), 20)
pv = afk.Value(afk.call("make_plane3d", (0.0, 512.0, 0.0), (0.0, 0.0, 1.0), (1.0, 0.0, 0.0), nx, nz))
entry("C3 slice_3d_to_2d position-velocity cut (3840 x 1024), mesh value reused", 2 * nz * nx * 4 + nz * nx * 12, lambda: afk.call("slice_3d_to_2d", cube, pv), 10)
del cube
# ---- elementwise on the 4096 x 4096 image ----
N = 4096 * 4096
img = afk.Image.synthetic(0xAF1A0004, 1, 4096, 4096)
img2 = afk.call("slice_one_frame", img, 0)
entry("C4 create_velocity_field_map 4096x4096", 8 * N, lambda: afk.call("create_velocity_field_map", img2, 1.05), 20, cold=True)
entry("C4 histogram_transformation 4096x4096 (min/max pass + map pass)", 12 * N, lambda: afk.call("histogram_transformation", img2, (0.1, 0.25, 0.9), (0.0, 1.0)), 20, cold=True)
rgb = afk.Image.synthetic(0xAF1A0005, 3, 4096, 4096)
entry("C4 apply_arcsinh_stretch 3x4096x4096", 24 * N, lambda: afk.call("apply_arcsinh_stretch", rgb, 0.01), 20, cold=True)
entry("C4 color_image_to_hsv 3x4096x4096", 24 * N, lambda: afk.call("color_image_to_hsv", rgb), 20, cold=True)
entry("C4 histogram_transformation_rgb 3x4096x4096", 36 * N, lambda: afk.call("histogram_transformation_rgb", rgb, (0.0, 0.5, 1.0, 0.1, 0.3, 0.8, 0.2, 0.7, 0.95), (0.0, 1.0)), 20, cold=True)
print(json.dumps(out, indent=1))
