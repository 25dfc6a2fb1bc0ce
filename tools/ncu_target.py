#!/usr/bin/env python
"""Short program for ncu: the fused average/average/ratio reduce kernel on the ALMA-shape cube, a few launches.

    python tools/ncu_target.py [--workload alma|manga|ew|lut] [--launches N]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np

import aflak_b200 as afk

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="alma")
ap.add_argument("--launches", type=int, default=6)
args = ap.parse_args()

assert afk.device_count() > 0
afk.set_device(0)
if args.workload == "alma":
    nz, ny, nx = 3840, 1024, 1024
    cube = afk.Image.synthetic(0xAF1A0003, nz, ny, nx)
    out = afk.Image.alloc((ny, nx))
    for _ in range(args.launches):
        afk.k_reduce(cube.device_ptr, nz, ny * nx, [(0, nz // 2), (nz // 2, nz)], afk.EPI_MEAN_RATIO, out.device_ptr)
elif args.workload == "manga":
    nz, ny, nx = 4563, 74, 74
    cube = afk.Image.synthetic(0xAF1A0001, nz, ny, nx)
    out = afk.Image.alloc((ny, nx))
    for _ in range(args.launches):
        afk.k_reduce(cube.device_ptr, nz, ny * nx, [(0, nz)], afk.EPI_SUM, out.device_ptr)
elif args.workload == "ew":
    nz, ny, nx = 4563, 74, 74
    cube = afk.Image.synthetic(0xAF1A0001, nz, ny, nx)
    for _ in range(args.launches):
        afk.fused_equivalent_width(cube, (3134, 3154), (3099, 3119), (3174, 3194), 20.0, 1e10, True)
elif args.workload == "lut":
    rng = np.random.default_rng(4)
    img = afk.Image.upload(np.exp(rng.standard_normal((4096, 4096))).astype(np.float32))
    lut = afk.Lut("Flame")
    outb = afk.Image.alloc((4096, 4096))  # 64 MiB of device bytes for RGBA8
    vmin, vmax = img.vmin_vmax()
    for _ in range(args.launches):
        afk.colormap_device(img, lut, float(vmin), float(vmax), outb.device_ptr, channels=4)
afk.synchronize()
print("ncu target ok:", args.workload, afk.launch_count(), "launches")
