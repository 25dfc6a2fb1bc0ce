#!/usr/bin/env python
"""Per-config measurements for BASELINE.json configs C1..C4 and the individual operators (bench.py covers the headline
C3 line).  Device time via CUDA events on the library's stream; small configs are reported both L2-warm (back-to-back)
and L2-cold (a 256 MiB buffer is overwritten between launches).  Prints one JSON object; run on a B200:

    python tools/bench_configs.py > gpurun_out/configs.json
"""
import json
import os
import sys
import time

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


def timed(fn, reps=50, warm=5, cold=False):
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


def batched(fn, batch=20, rounds=5):
    """ms per call with `batch` calls enqueued back to back between two events: the device-side rate of a chained DAG
    (the host's per-call enqueue latency, ~10-20 us through ctypes, is hidden unless the call itself synchronises)."""
    best = float("inf")
    for _ in range(rounds):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(batch):
            fn()
        e1.record(stream)
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1) / batch)
    return best


def entry(name, nbytes, fn, reps=50, both=True, launches=None):
    l0 = afk.launch_count()
    fn()
    nl = afk.launch_count() - l0
    med, best = timed(fn, reps)
    bt = batched(fn, 20 if med < 1.0 else 5)
    e = {"config": name, "algorithmic_bytes": nbytes, "kernel_launches": nl, "ms_median": med, "ms_best": best,
         "GBps": nbytes / (med * 1e-3) / 1e9, "frac_of_measured_peak": nbytes / (med * 1e-3) / 1e9 / PEAK,
         "ms_batched": bt, "GBps_batched": nbytes / (bt * 1e-3) / 1e9, "frac_of_measured_peak_batched": nbytes / (bt * 1e-3) / 1e9 / PEAK}
    if both:
        cmed, cbest = timed(fn, max(10, reps // 2), cold=True)
        e.update({"cold_l2_ms_median": cmed, "cold_l2_GBps": nbytes / (cmed * 1e-3) / 1e9,
                  "cold_l2_frac_of_measured_peak": nbytes / (cmed * 1e-3) / 1e9 / PEAK})
    print(json.dumps(e), file=sys.stderr, flush=True)
    return e


out = []
# ---------------- C1 / C2: MaNGA ----------------
nz, ny, nx = 4563, 74, 74
n2 = ny * nx
cube = afk.Image.synthetic(0xAF1A0001, nz, ny, nx)
res = afk.Image.alloc((ny, nx))
out.append(entry("C1 MaNGA integral[0,4563) (k_reduce)", nz * n2 * 4 + n2 * 4,
                 lambda: afk.k_reduce(cube.device_ptr, nz, n2, [(0, nz)], afk.EPI_SUM, res.device_ptr)))
out.append(entry("C1 MaNGA integral[0,4563) via registry call", nz * n2 * 4 + n2 * 4, lambda: afk.call("integral", cube, 0, nz)))
out.append(entry("C1 MaNGA integral[3134,3154) template-sized", 20 * n2 * 4 + n2 * 4,
                 lambda: afk.k_reduce(cube.device_ptr, nz, n2, [(3134, 3154)], afk.EPI_SUM, res.device_ptr)))
W = ((3134, 3154), (3099, 3119), (3174, 3194))
out.append(entry("C2 MaNGA EW template fused (3x20 frames)", 60 * n2 * 4 + n2 * 4, lambda: afk.fused_equivalent_width(cube, *W, 20.0, 1e10, True)))
WW = ((2000, 2500), (1400, 1900), (2600, 3100))
out.append(entry("C2 MaNGA EW template fused, wide (3x500 frames)", 1500 * n2 * 4 + n2 * 4, lambda: afk.fused_equivalent_width(cube, *WW, 20.0, 1e10, True)))


def ew_graph(fusion):
    g = afk.Graph()
    g.set_fusion(fusion)
    c = g.add_constant(cube)
    avg = []
    for s, e in W:
        t = g.add_transform("average")
        g.connect(c, 0, t, 0)
        g.connect(g.add_constant(s), 0, t, 1)
        g.connect(g.add_constant(e), 0, t, 2)
        avg.append(t)
    rb = g.add_transform("ratio_from_bands")
    for i, t in enumerate(avg):
        g.connect(t, 1, rb, i)
    lc = g.add_transform("linear_composition")
    g.connect(avg[1], 0, lc, 0); g.connect(avg[2], 0, lc, 1); g.connect(rb, 0, lc, 2); g.connect(rb, 1, lc, 3)
    ew = g.add_transform("create_equivalent_width")
    g.connect(lc, 0, ew, 0); g.connect(avg[0], 0, ew, 1)
    g.connect(g.add_constant(20.0), 0, ew, 2); g.connect(g.add_constant(1e10), 0, ew, 3); g.connect(g.add_constant(True), 0, ew, 4)
    mm = g.add_transform("image_min_max")
    g.connect(ew, 0, mm, 0)
    ls = g.add_transform("convert_to_logscale")
    g.connect(ew, 0, ls, 0); g.connect(g.add_constant(1000.0), 0, ls, 1); g.connect(mm, 0, ls, 2); g.connect(mm, 1, ls, 3)
    start = g.add_constant(3134)
    g.connect(start, 0, avg[0], 1)
    return g, ls, start


for fusion in (True, False):
    g, ls, start = ew_graph(fusion)
    state = {"i": 0}

    def step(g=g, ls=ls, start=start, state=state):
        state["i"] ^= 1
        g.set_constant(start, 3134 - state["i"])  # the user drags the range: everything downstream re-evaluates
        g.compute(ls)

    t0 = time.perf_counter()
    for _ in range(200):
        step()
    afk.synchronize()
    wall = (time.perf_counter() - t0) / 200
    e = entry(f"C2 MaNGA EW template graph -> min/max -> logscale ({'fused' if fusion else 'unfused'}), re-evaluation", 60 * n2 * 4 + 3 * n2 * 4, step, both=False)
    e["host_wall_ms_per_reevaluation"] = wall * 1e3
    out.append(e)
f = [afk.call("slice_one_frame", cube, k) for k in (3109, 3144, 3184)]
out.append(entry("C2 linear_composition of 3 slices (2 chained calls)", 4 * n2 * 4 + 2 * n2 * 4,
                 lambda: afk.call("linear_composition", afk.call("linear_composition", f[0], f[1], 0.5, 0.25), f[2], 1.0, 0.25), both=False))
del cube

# ---------------- C3 variants: ALMA ----------------
nz, ny, nx = 3840, 1024, 1024
n2 = ny * nx
cube = afk.Image.synthetic(0xAF1A0003, nz, ny, nx)
res = afk.Image.alloc((ny, nx))
cb = nz * n2 * 4 + n2 * 4
out.append(entry("C3 ALMA average+average+ratio fused", cb, lambda: afk.k_reduce(cube.device_ptr, nz, n2, [(0, nz // 2), (nz // 2, nz)], afk.EPI_MEAN_RATIO, res.device_ptr), reps=30, both=False))
out.append(entry("C3 ALMA integral[0,3840)", cb, lambda: afk.k_reduce(cube.device_ptr, nz, n2, [(0, nz)], afk.EPI_SUM, res.device_ptr), reps=30, both=False))
out.append(entry("C3 ALMA integral[0,3840) sequential mode (reference bits)", cb, lambda: afk.k_reduce(cube.device_ptr, nz, n2, [(0, nz)], afk.EPI_SUM, res.device_ptr, sum_mode=afk.SUM_SEQUENTIAL), reps=30, both=False))
EW3 = [(1280, 2560), (0, 1280), (2560, 3840)]
out.append(entry("C3 ALMA EW fused, whole cube in 3 windows", cb, lambda: afk.fused_equivalent_width(cube, *EW3, 20.0, 1e10, True), reps=30, both=False))
out.append(entry("C3 ALMA unfused: average, average, ratio (3 launches)", cb + 4 * n2 * 4,
                 lambda: afk.call("ratio", afk.call("average", cube, 0, nz // 2)[0], afk.call("average", cube, nz // 2, nz)[0]), reps=30, both=False))
del cube

# ---------------- C4: 4096^2 image chain ----------------
N = 4096
rng = np.random.default_rng(4)
img = afk.Image.upload(np.exp(rng.standard_normal((N, N))).astype(np.float32))
npx = N * N
rgba = afk.Image.alloc((N, N))
lut = afk.Lut("Flame")
lut.set_lims(0.2, 0.3, 0.9)
vmin, vmax = afk.call("image_min_max", img)
tone = np.arange(256, dtype=np.float32)
lmax = float(np.float32(np.log(np.float32(1001.0)) / np.log(np.float32(1000.0))))
chain = {"logscale_a": 1000.0, "logscale_vmin": float(vmin), "logscale_vmax": float(vmax), "normalize_min": 0.0, "normalize_max": lmax / 65535.0, "tone_table": tone}
out.append(entry("C4 image_min_max 4096^2", npx * 4, lambda: afk.call("image_min_max", img)))
out.append(entry("C4 colormap RGBA8, limits (0.2, 0.3, 0.9) (given vmin/vmax)", npx * 8, lambda: afk.colormap_device(img, lut, float(vmin), float(vmax), rgba.device_ptr, channels=4)))
lut_dflt = afk.Lut("Flame")  # the reference's default limits (0, 0.5, 1)
out.append(entry("C4 colormap RGBA8, default limits (0, 0.5, 1) (given vmin/vmax)", npx * 8, lambda: afk.colormap_device(img, lut_dflt, float(vmin), float(vmax), rgba.device_ptr, channels=4)))
out.append(entry("C4 colormap RGB8 (given vmin/vmax)", npx * 7, lambda: afk.colormap_device(img, lut, float(vmin), float(vmax), rgba.device_ptr, channels=3)))
out.append(entry("C4 fused chain logscale->tone->LUT->RGBA8 (given min/max)", npx * 8, lambda: afk.colormap_device(img, lut, 0.0, 65535.0, rgba.device_ptr, channels=4, chain=chain)))


def c4_full():
    mn, mx = afk.call("image_min_max", img)
    ch = dict(chain, logscale_vmin=float(mn), logscale_vmax=float(mx))
    afk.colormap_device(img, lut, 0.0, 65535.0, rgba.device_ptr, channels=4, chain=ch)


out.append(entry("C4 full: min/max pass + fused chain -> RGBA8 (12 B/px)", npx * 12, c4_full))
nan_chain = dict(chain, logscale_vmin=float("nan"), logscale_vmax=float("nan"))
out.append(entry("C4 full, min/max kept on the device (chain limits = NaN): 2 kernels, 1 read-back (12 B/px)", npx * 12,
                 lambda: afk.colormap_device(img, lut, 0.0, 65535.0, rgba.device_ptr, channels=4, chain=nan_chain)))


def c4_unfused():
    mn, mx = afk.call("image_min_max", img)
    ls = afk.call("convert_to_logscale", img, 1000.0, float(mn), float(mx))
    nm = afk.call("normalize_image", ls, 0.0, lmax / 65535.0)
    tc = afk.call("apply_tone_curve", nm, afk.ToneCurve(tone))
    afk.colormap_device(tc, lut, 0.0, 65535.0, rgba.device_ptr, channels=4)


out.append(entry("C4 unfused: min/max, logscale, normalize, tone, colormap (36 B/px moved)", npx * 12, c4_unfused, both=False))
# individual operators at 4096^2
v = afk.Image.upload(np.exp(rng.standard_normal((N, N))).astype(np.float32))
for name, nb, fn in [
    ("linear_composition 4096^2", 12, lambda: afk.call("linear_composition", img, v, 0.4, 0.6)),
    ("ratio 4096^2", 12, lambda: afk.call("ratio", img, v)),
    ("create_equivalent_width 4096^2", 12, lambda: afk.call("create_equivalent_width", img, v, 20.0, 1e10, True)),
    ("convert_to_logscale 4096^2", 8, lambda: afk.call("convert_to_logscale", img, 1000.0, float(vmin), float(vmax))),
    ("log10 4096^2", 8, lambda: afk.call("log10", img)),
    ("image_multiplier(1,-1) 4096^2", 12, lambda: afk.call("image_multiplier", img, v, 1.0, -1.0)),
    ("down_sampling n=2 4096^2", 2, lambda: afk.call("down_sampling", img, 2)),
]:
    out.append(entry(name, npx * nb, fn))

# ---------------- the reference's CPU path (oracle port) on this box's host cores, for C1 / C2 / C4 ----------------
# Threaded the way the reference threads it: reductions are one serial frame-accumulate each (ndarray sum_axis), the
# sibling nodes of a DAG level run concurrently (rayon::scope), par_iter_mut operators use every core, the LUT mapping
# runs on the UI thread.  bench.py carries the same comparison for the headline C3 line.
import threading

import oracle

NCPU = oracle.max_threads()


def cpu_best(fn, reps=3):
    best = float("inf")
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
    return best * 1e3


def concurrently(*fns):
    res = [None] * len(fns)

    def run(i):
        res[i] = fns[i]()

    ts = [threading.Thread(target=run, args=(i,)) for i in range(len(fns))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    return res


cpu = []
hm = oracle.gen_cube(0xAF1A0001, 4563, 74, 74, threads=NCPU)
cpu.append({"config": "C1 MaNGA integral[0,4563)", "cores": 1, "ms": cpu_best(lambda: oracle.integral(hm, 0, 4563))})


def c2_cpu():
    on, o1, o2 = concurrently(lambda: oracle.average(hm, 3134, 3154), lambda: oracle.average(hm, 3099, 3119), lambda: oracle.average(hm, 3174, 3194))
    r2, r1 = oracle.ratio_from_bands(3144.0, 3109.0, 3184.0)
    ew = oracle.equivalent_width(oracle.linear_composition(o1, o2, r2, r1), on, 20.0, 1e10, True)
    mn, mx = oracle.image_min_max(ew)
    return oracle.convert_to_logscale(ew, 1000.0, mn, mx)  # 5476 pixels: one core (an OpenMP team costs more than the work)


cpu.append({"config": "C2 MaNGA EW template -> min/max -> logscale", "cores": 3, "ms": cpu_best(c2_cpu)})
rng = np.random.default_rng(4)
himg = np.exp(rng.standard_normal((N, N))).astype(np.float32)
table = oracle.lut_build("Flame", lims=(0.2, 0.3, 0.9))
tone_h = np.arange(256, dtype=np.float32)


def c4_cpu():
    mn, mx = oracle.image_min_max(himg)
    ls = oracle.convert_to_logscale(himg, 1000.0, mn, mx, threads=NCPU)
    nm = ((ls - np.float32(0.0)) / np.float32(lmax / 65535.0)).astype(np.float32)
    tc = oracle.apply_tone_curve(nm, tone_h)
    return oracle.make_raw_image(tc, 0.0, 65535.0, table, lims=(0.2, 0.3, 0.9), channels=4)


cpu.append({"config": "C4 4096^2 chain: min/max, logscale (all cores), normalize, tone curve, colour LUT (serial, UI thread) -> RGBA8",
            "cores": NCPU, "ms": cpu_best(c4_cpu, reps=2)})
for e in cpu:
    print(json.dumps(e), file=sys.stderr, flush=True)
print(json.dumps({"measured_hbm_peak_GBps": PEAK, "entries": out, "cpu_reference_path": {"kind": "port (oracle/)", "host_threads": NCPU, "entries": cpu}}, indent=1))
