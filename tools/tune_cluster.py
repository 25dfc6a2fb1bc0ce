#!/usr/bin/env python
"""Small-image reduce variants on the MaNGA shape (C1 integral, C2 EW template): device time per call for the plan the
environment selects (AFK_REDUCE_CLUSTER=0|2|4|8|16, AFK_REDUCE_CLUSTER_JMAX).  One JSON line per run.

    AFK_REDUCE_CLUSTER=8 python tools/tune_cluster.py
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
flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def single(fn, reps=40, cold=True):
    for _ in range(5):
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
    return round(float(np.median(ts)) * 1e3, 2)


def b2b(fn, batch=20, rounds=5):
    best = 1e9
    for _ in range(rounds):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(batch):
            fn()
        e1.record(stream)
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1) / batch)
    return round(best * 1e3, 2)


def tier_c(got, s64, a64):
    err = np.abs(got.astype(np.float64) - s64)
    return bool(np.all((err <= 1e-5 * np.abs(s64)) | (err <= 4 * 2.0 ** -24 * a64)))


out = {"env": {k: v for k, v in os.environ.items() if k.startswith("AFK_")}}
shapes = [(4563, 74, 74), (2000, 128, 128), (1000, 300, 300)] if "--shapes" in sys.argv else [(4563, 74, 74)]
for nz, ny, nx in shapes:
    n2 = ny * nx
    cube = afk.Image.synthetic(0xAF1A0001, nz, ny, nx)
    res = afk.Image.alloc((ny, nx))
    fn = lambda: afk.k_reduce(cube.device_ptr, nz, n2, [(0, nz)], afk.EPI_SUM, res.device_ptr)
    e = {"kernel": afk.k_reduce_describe(cube.device_ptr, n2, [(0, nz)], res.device_ptr), "MB": round(nz * n2 * 4 / 1e6, 1),
         "cold_us": single(fn), "warm_us": single(fn, cold=False), "b2b_us": b2b(fn)}
    e["GBps_cold"] = round(nz * n2 * 4 / e["cold_us"] / 1e3, 1)
    if nz * n2 <= 30_000_000:
        host = oracle.gen_cube(0xAF1A0001, nz, ny, nx, threads=oracle.max_threads())
        s64, a64 = oracle.sum_f64(host, 0, nz)
        e["parity"] = tier_c(res.download(), s64, a64)
    out[f"integral_{nz}x{ny}x{nx}"] = e
    if (nz, ny, nx) == (4563, 74, 74):
        W = ((3134, 3154), (3099, 3119), (3174, 3194))
        fn2 = lambda: afk.k_reduce(cube.device_ptr, nz, n2, list(W), afk.EPI_EW, res.device_ptr, params=[0.5, 0.5, 20.0, 1e10, -1.0])
        out["ew_3x20"] = {"kernel": afk.k_reduce_describe(cube.device_ptr, n2, list(W), res.device_ptr), "cold_us": single(fn2),
                          "warm_us": single(fn2, cold=False), "b2b_us": b2b(fn2)}
        fn3 = lambda: afk.k_reduce(cube.device_ptr, nz, n2, [(100, 700), (1000, 1500)], afk.EPI_MEAN_RATIO, res.device_ptr)
        out["ratio_600_500"] = {"kernel": afk.k_reduce_describe(cube.device_ptr, n2, [(100, 700), (1000, 1500)], res.device_ptr),
                                "cold_us": single(fn3), "b2b_us": b2b(fn3)}
print(json.dumps(out))
