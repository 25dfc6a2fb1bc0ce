#!/usr/bin/env python
"""BASELINE.json configs[4] (SURVEY 8d "C5"): the ALMA-scale cube [nz=2048, ny=4096, nx=4096] f32 (137.4 GB) sharded by
y-bands over N = 2/4/8 B200s of one box; integral, average/average/ratio and the equivalent-width template, each one
fused reduce launch per rank whose epilogue gathers the 2-D result into every rank's full image over NVLink.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/bench_c5.py

STRONG scaling: the global cube is fixed, rank r holds the band afk_shard_band(4096, N, r) generated on the device
(same counter-based generator as the oracle, so any row of any band can be checked on the host without moving
128 GiB).  Times are CUDA events on the launching stream, max over ranks.  Rank 0 prints one JSON document.
Parity inside the run: sampled rows of the gathered full image against the oracle (sequential mode: bit for bit;
compensated mode: SURVEY 8c tier C), on every rank's own copy.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
import torch
import torch.distributed as dist

import aflak_b200 as afk
import oracle

ap = argparse.ArgumentParser()
ap.add_argument("--nz", type=int, default=2048)
ap.add_argument("--ny", type=int, default=4096)
ap.add_argument("--nx", type=int, default=4096)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--warmup", type=int, default=3)
args = ap.parse_args()

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
afk.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)
afk.set_stream(stream.cuda_stream)


def exchange(h):
    lst = [None] * world
    dist.all_gather_object(lst, h)
    return lst


SEED = 0xAF1A0005
nz, ny, nx = args.nz, args.ny, args.nx
try:
    with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
        PEAK = float(json.load(f)["hbm_gbs"])
except Exception:
    PEAK = 6650.0

g = afk.Gather(world, rank, ny, nx, exchange)
y0, rows = g.band()
band = afk.Image.synthetic(SEED, nz, ny, nx, y0=y0, rows=rows)
torch.cuda.synchronize()

# the EW template with windows scaled to nz = 2048 (20-frame windows), and a wide variant
c = nz // 2
narrow = [(c - 10, c + 10), (c - 45, c - 25), (c + 30, c + 50)]
w3 = nz // 3
wide = [(w3, 2 * w3), (0, w3), (2 * w3, 3 * w3)]


def ew_params(win):
    on, off1, off2 = win
    z, z1, z2 = [(a + b) / 2 for a, b in (on, off1, off2)]
    r1 = np.float32(np.float32(z - z1) / np.float32(z2 - z1))
    return [float(np.float32(1.0) - r1), float(r1), 20.0, 1e10, -1.0]


cases = [
    ("integral[0,nz)", [(0, nz)], afk.EPI_SUM, None),
    ("average[0,nz/2) / average[nz/2,nz) -> ratio", [(0, nz // 2), (nz // 2, nz)], afk.EPI_MEAN_RATIO, None),
    ("equivalent width template, 3 x 20 frames", narrow, afk.EPI_EW, ew_params(narrow)),
    ("equivalent width template, wide (whole cube in 3 windows)", wide, afk.EPI_EW, ew_params(wide)),
]


def barrier():
    dist.barrier()
    torch.cuda.synchronize()


def timed(windows, epi, params, mode):
    for _ in range(args.warmup):
        g.reduce(band.device_ptr, nz, windows, epi, params, sum_mode=mode)
        g.wait()
    barrier()
    afk.profile_reduce(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        g.reduce(band.device_ptr, nz, windows, epi, params, sum_mode=mode)
        g.wait()
    e1.record(stream)
    barrier()
    k_ms, k_n = afk.profile_reduce_read()
    afk.profile_reduce(False)
    t = torch.tensor([e0.elapsed_time(e1) / args.steps, k_ms / max(k_n, 1)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0].item()), float(t[1].item())


def check_rows(windows, epi, params, name):
    """sampled rows of this rank's copy of the gathered image vs the oracle"""
    ok = True
    sample = sorted({0, ny // world - 1, ny // 2, ny // 2 + 1, ny - 1})
    for mode in (afk.SUM_SEQUENTIAL, afk.SUM_COMPENSATED):
        g.reduce(band.device_ptr, nz, windows, epi, params, sum_mode=mode)
        g.wait()
        torch.cuda.synchronize()
        got = g.result_image().download()
        for yy in sample:
            hb = oracle.gen_cube(SEED, nz, ny, nx, y0=yy, rows=1, threads=4)
            if epi == afk.EPI_SUM:
                a, b = windows[0]
                if mode == afk.SUM_SEQUENTIAL:
                    same = np.array_equal(got[yy:yy + 1], oracle.integral(hb, a, b))
                else:
                    s64, a64 = oracle.sum_f64(hb, a, b)
                    err = np.abs(got[yy:yy + 1].astype(np.float64) - s64)
                    same = bool(np.all((err <= 1e-5 * np.abs(s64)) | (err <= 4 * 2.0 ** -24 * a64)))
            elif epi == afk.EPI_MEAN_RATIO:
                want = oracle.ratio(oracle.average(hb, *windows[0]), oracle.average(hb, *windows[1]))
                same = np.array_equal(got[yy:yy + 1], want) if mode == afk.SUM_SEQUENTIAL else np.allclose(got[yy:yy + 1], want, rtol=1e-5)
            else:
                on, off1, off2 = windows
                off = oracle.linear_composition(oracle.average(hb, *off1), oracle.average(hb, *off2), params[0], params[1])
                want = oracle.equivalent_width(off, oracle.average(hb, *on), 20.0, 1e10, True)
                same = np.array_equal(got[yy:yy + 1], want) if mode == afk.SUM_SEQUENTIAL else np.allclose(got[yy:yy + 1], want, rtol=1e-4, atol=1e-4)
            if not same:
                ok = False
                print(f"rank {rank}: C5 parity MISMATCH {name} mode {mode} row {yy}", flush=True)
    return ok


entries, all_ok = [], True
for name, windows, epi, params in cases:
    ok = check_rows(windows, epi, params, name)
    t = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(t)
    ok = int(t.item()) == 0
    all_ok &= ok
    frames = sum(b - a for a, b in windows)
    for mode, mname in ((afk.SUM_COMPENSATED, "compensated"), (afk.SUM_SEQUENTIAL, "sequential (reference bits)")):
        ms, k_ms = timed(windows, epi, params, mode)
        bytes_global = frames * ny * nx * 4 + ny * nx * 4
        gbs = bytes_global / (ms * 1e-3) / 1e9
        k_gbs_per_gpu = (frames * rows * nx * 4 + rows * nx * 4) / (k_ms * 1e-3) / 1e9
        entries.append({
            "case": name, "sum": mname, "windows": windows, "ms_per_step_max_over_ranks": ms, "aggregate_GBps": gbs,
            "aggregate_voxels_per_s": frames * ny * nx / (ms * 1e-3), "algorithmic_bytes_global": bytes_global,
            "reduce_kernel_ms_max_over_ranks": k_ms, "per_gpu_kernel_GBps": k_gbs_per_gpu,
            "per_gpu_frac_of_measured_peak": k_gbs_per_gpu / PEAK, "per_gpu_frac_of_nominal_8TBs": k_gbs_per_gpu / 8000.0,
            "gathered_bytes_over_nvlink": ny * nx * 4 * (world - 1), "parity_sampled_rows": "ok" if ok else "MISMATCH",
        })
g.check()
if rank == 0:
    print(json.dumps({
        "config": "C5 ALMA-scale cube, strong scaling", "global_cube": [nz, ny, nx], "cube_bytes": nz * ny * nx * 4, "n_gpus": world,
        "band_rows_per_gpu": rows, "band_bytes_per_gpu": nz * rows * nx * 4, "steps": args.steps, "warmup": args.warmup,
        "measured_hbm_peak_GBps": PEAK, "gather": "peer stores fused into the reduce kernel epilogue (NVLink P2P), stream-ordered wait",
        "parity": "OK" if all_ok else "FAILED", "entries": entries}, indent=1), flush=True)
barrier()
g.close()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if all_ok else 1)
