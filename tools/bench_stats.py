#!/usr/bin/env python
"""Display statistics on the 4096x4096 image of BASELINE config 4: device time of the histogram, the
normalised median + MAD and zscale, next to the reference's CPU path restated (oracle: two full sorts + a serial
histogram loop, one thread like the UI thread).

    python tools/bench_stats.py [--gpu-only] > gpurun_out/stats.json
"""
import json
import os
import sys
import time

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
flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, reps=20, warm=3, cold=False):
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


rng = np.random.default_rng(4)
a = rng.lognormal(0.0, 1.0, (4096, 4096)).astype(np.float32)
a[::1000, ::3] = np.nan
img = afk.Image.upload(a)
vmin, vmax = img.vmin_vmax()
N = a.size
out = {"image": "4096x4096 f32 lognormal, 0.03% NaN", "peak_gbs": PEAK, "gpu": [], "cpu": []}


def entry(name, passes, fn):
    l0 = afk.launch_count()
    fn()
    nl = afk.launch_count() - l0
    warm, cold = timed(fn), timed(fn, reps=10, cold=True)
    nb = passes * N * 4
    e = {"op": name, "kernel_launches": nl, "image_passes": passes, "algorithmic_bytes": nb, "ms_l2_warm": warm, "ms_l2_cold": cold,
         "GBps_cold": nb / cold / 1e6, "note": "includes the host read-back of the result (one sync)"}
    out["gpu"].append(e)
    print(json.dumps(e), file=sys.stderr, flush=True)


entry("vmin/vmax (lims::get_vmin, get_vmax)", 1, lambda: img.vmin_vmax())
entry("histogram 100 bins (hist::histogram)", 1, lambda: img.histogram(vmin, vmax))
entry("vmed + vmad (lims::get_vmed_normalized, get_vmad_normalized): 1 min/max pass + 2 x (sample kernel, ONE bracket pass over the image, finish over the candidates); one read-back", 3, lambda: img.vmed_vmad())
entry("zscale(1000, 0.25)", 0, lambda: img.zscale(1000, 0.25))


def cpu(name, fn):
    t0 = time.perf_counter()
    fn()
    dt = time.perf_counter() - t0
    e = {"op": name, "kind": "port (oracle, 1 thread like the UI thread)", "ms": dt * 1e3}
    out["cpu"].append(e)
    print(json.dumps(e), file=sys.stderr, flush=True)


if "--gpu-only" not in sys.argv:  # the two sorts take ~5 s
    cpu("vmin/vmax", lambda: (oracle.get_vmin(a), oracle.get_vmax(a)))
    cpu("histogram 100 bins", lambda: oracle.histogram(a, vmin, vmax))
    cpu("vmed + vmad (two full sorts)", lambda: oracle.vmed_vmad(a))
    cpu("zscale(1000, 0.25)", lambda: oracle.zscale(a, 1000, 0.25))
print(json.dumps(out, indent=1))
