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
elif args.workload == "c4":  # BASELINE configs[3]: min/max + fused chain, non-default limits; then the plain map and the tone curve
    rng = np.random.default_rng(4)
    img = afk.Image.upload(np.exp(rng.standard_normal((4096, 4096))).astype(np.float32))
    lut = afk.Lut("Flame")
    lut.set_lims(0.2, 0.3, 0.9)
    outb = afk.Image.alloc((4096, 4096))
    lmax = float(np.float32(np.log(np.float32(1001.0)) / np.log(np.float32(1000.0))))
    chain = {"logscale_a": 1000.0, "logscale_vmin": float("nan"), "logscale_vmax": float("nan"), "normalize_min": 0.0,
             "normalize_max": lmax / 65535.0, "tone_table": np.arange(256, dtype=np.float32)}
    vmin, vmax = img.vmin_vmax()
    scaled = afk.call("normalize_image", img, 0.0, float(vmax) / 65535.0)
    for _ in range(args.launches):
        afk.colormap_device(img, lut, 0.0, 65535.0, outb.device_ptr, channels=4, chain=chain)
        afk.colormap_device(img, lut, float(vmin), float(vmax), outb.device_ptr, channels=4)
        afk.call("apply_tone_curve", scaled, afk.ToneCurve(np.arange(256, dtype=np.float32)))
    afk.synchronize()
elif args.workload == "variance":  # axis0_kernel<VarComp> and <VarSeq> on the ALMA-shape cube
    nz, ny, nx = 3840, 1024, 1024
    cube = afk.Image.synthetic(0xAF1A0003, nz, ny, nx)
    for _ in range(args.launches):
        afk.call("variance", cube, 0, nz)
    afk.set_sum_mode(afk.SUM_SEQUENTIAL)
    for _ in range(args.launches):
        afk.call("variance", cube, 0, nz)
    afk.set_sum_mode(afk.SUM_COMPENSATED)
elif args.workload == "argmax":
    nz, ny, nx = 3840, 1024, 1024
    cube = afk.Image.synthetic(0xAF1A0003, nz, ny, nx)
    cube.set_axis(2, "WAVE", "Angstrom")
    for _ in range(args.launches):
        afk.call("extract_argmin_max_wavelength", cube, 0, nz, False)
elif args.workload == "median":
    nz, ny, nx = 3840, 1024, 1024
    cube = afk.Image.synthetic(0xAF1A0003, nz, ny, nx)
    for _ in range(args.launches):
        afk.call("median", cube, 0, nz)
elif args.workload == "wave":
    nz, ny, nx = 3840, 1024, 1024
    cube = afk.Image.synthetic(0xAF1A0003, nz, ny, nx)
    for _ in range(args.launches):
        afk.call("extract_wave", cube, afk.Roi())
elif args.workload == "ingest":  # decode kernel on device-resident raw bytes: BITPIX -32 then BITPIX 16
    n = 1 << 31
    dst = afk.Image.alloc((n,))
    raw = afk.Image.synthetic(0xAF1A0006, 1, 1, n)  # any bytes will do
    for _ in range(args.launches):
        afk.k_ingest(raw.device_ptr, afk.FITS_F32, n, dst.device_ptr, bscale=0.5, bzero=1.0)
    for _ in range(args.launches):
        afk.k_ingest(raw.device_ptr, afk.FITS_I16, n, dst.device_ptr, bscale=0.5, bzero=1.0)
elif args.workload == "c4":  # the 4096^2 display chain (BASELINE configs[3]) + the per-pixel operators, kernel by kernel
    N = 4096
    rng = np.random.default_rng(4)
    img = afk.Image.upload(np.exp(rng.standard_normal((N, N))).astype(np.float32))
    v = afk.Image.upload(np.exp(rng.standard_normal((N, N))).astype(np.float32))
    rgba = afk.Image.alloc((N, N))
    lut = afk.Lut("Flame")
    lut.set_lims(0.2, 0.3, 0.9)
    vmin, vmax = afk.call("image_min_max", img)
    tone = np.arange(256, dtype=np.float32)
    lmax = float(np.float32(np.log(np.float32(1001.0)) / np.log(np.float32(1000.0))))
    chain = {"logscale_a": 1000.0, "logscale_vmin": float(vmin), "logscale_vmax": float(vmax), "normalize_min": 0.0,
             "normalize_max": lmax / 65535.0, "tone_table": tone}
    for _ in range(args.launches):
        afk.call("image_min_max", img)
        afk.colormap_device(img, lut, float(vmin), float(vmax), rgba.device_ptr, channels=4)
        afk.colormap_device(img, lut, float(vmin), float(vmax), rgba.device_ptr, channels=3)
        afk.colormap_device(img, lut, 0.0, 65535.0, rgba.device_ptr, channels=4, chain=chain)
        afk.call("linear_composition", img, v, 0.4, 0.6)
        afk.call("ratio", img, v)
        afk.call("convert_to_logscale", img, 1000.0, float(vmin), float(vmax))
        afk.call("apply_tone_curve", img, afk.ToneCurve(tone))
        afk.call("down_sampling", img, 2)
elif args.workload == "stats":
    rng = np.random.default_rng(4)
    img = afk.Image.upload(np.exp(rng.standard_normal((4096, 4096))).astype(np.float32))
    vmin, vmax = img.vmin_vmax()
    for _ in range(args.launches):
        img.histogram(float(vmin), float(vmax))
        img.vmed_vmad()
afk.synchronize()
print("ncu target ok:", args.workload, afk.launch_count(), "launches")
