#!/usr/bin/env python
"""bench.py -- aflak hot path on B200: spectral average+ratio over an ALMA-shape cube (BASELINE config 3).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload alma|manga|ew|lut]

One "step" = one pass of the hot path over one cube: `average[0,nz/2)`, `average[nz/2,nz)` and their `ratio`
(aflak_primitives/src/lib.rs:398-411, unit.rs:863-880), which the evaluator runs as ONE fused kernel that streams the
whole cube once.  Metric: cube GB/s = algorithmic bytes (cube read + 2-D result write) / device time.

N>1 (torchrun, one process per GPU): weak scaling.  The global cube is [nz, 1024*N, 1024]; rank r owns the y-band
[1024 r, 1024 (r+1)), reduces it with the same kernel and the 2-D result bands are gathered (the only
NVLink traffic).  value = bytes of all ranks / max-over-ranks device time.

`e2e`: the same step through the registry / graph evaluator with HOST (pinned) input: upload + operators + read-back
of the 2-D result inside the timed region.
`--impl reference`: the CPU restatement of the reference's own code path (oracle/, kind "port": the reference is
pure Rust and cannot be built in this image) timed on the host cores on a bounded band of the same cube.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np

SEED = 0xAF1A0003
WORKLOADS = {
    # name: (nz, ny per rank, nx, description)
    "alma": (3840, 1024, 1024, "ALMA-shape cube 1024x1024x3840 f32: average[0,1920) / average[1920,3840) -> ratio (fused, one pass)"),
    "manga": (4563, 74, 74, "MaNGA-shape cube 74x74x4563 f32: integral[0,4563)"),
}
METRIC = "cube GB/s, spectral average+ratio (fused) over ALMA-shape cube"
UNIT = "GB/s"


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def algorithmic_bytes(nz, ny, nx, k_frames=None):
    k = nz if k_frames is None else k_frames
    return k * ny * nx * 4 + ny * nx * 4


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md clocks line)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                mx.append(float(p[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle (restatement of the reference's ndarray path) on a bounded band of the same cube
# ---------------------------------------------------------------------------------------------------------------
def cpu_sample_shape(nz, ny, nx):
    rows = max(1, min(ny, 64))  # [3840, 64, 1024] = 1.0 GB: ~seconds per pass on a few cores
    return nz, rows, nx


def cpu_step(oracle, host, nz, threads):
    a = oracle.average(host, 0, nz // 2, threads=threads)
    b = oracle.average(host, nz // 2, nz, threads=threads)
    return oracle.ratio(a, b)


def cpu_baseline(nz, ny, nx, reps=3, budget_s=25.0):
    """Times the oracle with all host threads on a bounded band; returns (dict, per-rep seconds)."""
    import oracle

    oracle.build()
    threads = oracle.max_threads()
    snz, srows, snx = cpu_sample_shape(nz, ny, nx)
    host = oracle.gen_cube(SEED, nz, ny, nx, y0=0, rows=srows)
    nbytes = algorithmic_bytes(snz, srows, snx)
    times = []
    t_all = time.perf_counter()
    for _ in range(reps):
        t0 = time.perf_counter()
        cpu_step(oracle, host, nz, threads)
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_all > budget_s:
            break
    best = min(times)
    return {
        "value": nbytes / best / 1e9, "unit": UNIT, "cores": threads, "kind": "port",
        "sample": f"band [{snz},{srows},{snx}] of the cube ({nbytes / 1e9:.2f} GB), oracle average x2 + ratio, "
                  f"pixel-tiled over {threads} threads, best of {len(times)}",
        "voxels_per_s": snz * srows * snx / best,
    }, times


def run_reference(args):
    """--impl reference: rank 0 times the CPU path; other ranks exit 0 without work."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import oracle

    oracle.build()
    nz, ny, nx, desc = WORKLOADS[args.workload]
    threads = oracle.max_threads()
    snz, srows, snx = cpu_sample_shape(nz, ny, nx)
    host = oracle.gen_cube(SEED, nz, ny, nx, y0=0, rows=srows)
    nbytes = algorithmic_bytes(snz, srows, snx)
    for _ in range(args.warmup):
        cpu_step(oracle, host, nz, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_step(oracle, host, nz, threads)
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    value = nbytes / dt / 1e9
    sample = (f"band [{snz},{srows},{snx}] of the cube ({nbytes / 1e9:.2f} GB) per step, oracle average x2 + ratio, "
              f"{threads} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "cube": [nz, ny, nx], "timing": "host wall clock; CPU path, no GPU"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------
def build_c3_graph(afk, nz):
    """average[0,nz/2), average[nz/2,nz) -> ratio, wired like an aflak node graph (aflak_cake DST)."""
    g = afk.Graph()
    cube = g.add_constant(0)  # placeholder; replaced by the uploaded cube every step (update_constant_node)
    avg = []
    for s, e in ((0, nz // 2), (nz // 2, nz)):
        t = g.add_transform("average")
        cs, ce = g.add_constant(int(s)), g.add_constant(int(e))
        g.connect(cs, 0, t, 1)
        g.connect(ce, 0, t, 2)
        avg.append(t)
    ratio = g.add_transform("ratio")
    return g, cube, avg, ratio


def run_gpu(args):
    import torch

    import aflak_b200 as afk

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    if afk.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    afk.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    afk.set_stream(stream.cuda_stream)  # the library launches on this stream, so torch events time it

    nz, ny, nx, desc = WORKLOADS[args.workload]
    n2 = ny * nx
    windows = [(0, nz // 2), (nz // 2, nz)]
    cube = afk.Image.synthetic(SEED, nz, ny * world, nx, y0=rank * ny, rows=ny)
    out = torch.empty((ny, nx), dtype=torch.float32, device=dev)
    full = torch.empty((ny * world, nx), dtype=torch.float32, device=dev) if world > 1 else out
    step_bytes = algorithmic_bytes(nz, ny, nx)

    def step():
        afk.k_reduce(cube.device_ptr, nz, n2, windows, afk.EPI_MEAN_RATIO, out.data_ptr())
        if world > 1:
            dist.all_gather_into_tensor(full, out)  # the one exchange step: 2-D result bands over NVLink

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        t_wait = time.time()
        while not sampler.lines and time.time() - t_wait < 3.0:
            time.sleep(0.05)
    for _ in range(args.warmup):
        step()
    barrier()
    launches0 = afk.launch_count()
    afk.profile_reduce(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    ms_total = e0.elapsed_time(e1)
    k_ms, k_n = afk.profile_reduce_read()
    afk.profile_reduce(False)
    launches = afk.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    value = step_bytes * world / (ms_step * 1e-3) / 1e9

    # ---- e2e: host (pinned) cube -> upload -> registry operators via the graph evaluator -> host result ----
    e2e = run_e2e(args, afk, torch, dist, cube, nz, ny, nx, rank, world, dev)

    if rank == 0:
        peak, peak_src = measured_peak()
        k_avg_ms = k_ms / max(k_n, 1)
        achieved = step_bytes / (k_avg_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": desc, "cube_per_gpu": [nz, ny, nx], "global_cube": [nz, ny * world, nx],
                       "sharding": "y-bands, one per GPU; all_gather of the 2-D result" if world > 1 else "single GPU",
                       "sum": "compensated (Kahan) f32", "l2": "inputs (16.1 GB per step) larger than L2, no flush needed",
                       "voxels_per_s": nz * ny * nx * world / (ms_step * 1e-3)},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": None, "kernel": "reduce_kernel<4,true,2>", "kernel_ms": k_avg_ms, "launches_timed": k_n,
                         "algorithmic_bytes_per_launch": step_bytes, "peak_source": peak_src,
                         "frac_of_nominal_8TBs": achieved / 8000.0},
            "e2e": e2e, "gpu_launches": int(launches * world), "clocks": clocks,
        }
        if world == 1 and not args.no_cpu:
            cb, _ = cpu_baseline(nz, ny, nx)
            line["cpu_baseline"] = cb
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def run_e2e(args, afk, torch, dist, cube, nz, ny, nx, rank, world, dev):
    import psutil

    cube_bytes = nz * ny * nx * 4
    avail = psutil.virtual_memory().available
    rows = ny
    while rows > 16 and nz * rows * nx * 4 * world * 2 > avail:  # keep pinned host memory bounded on small hosts
        rows //= 2
    n_steps = max(1, min(args.steps, args.e2e_steps))
    host_t = torch.empty((nz, rows, nx), dtype=torch.float32, pin_memory=True)
    # fill the pinned host cube from the device-resident synthetic cube (rows [0, rows) of this rank's band)
    tmp = afk.Image.synthetic(SEED, nz, ny * world, nx, y0=rank * ny, rows=rows)
    afk.lib.afk_image_download(tmp.handle, host_t.data_ptr())
    del tmp
    host = host_t.numpy()
    res_t = torch.empty((rows, nx), dtype=torch.float32, pin_memory=True)
    g, c_cube, avg, ratio = build_c3_graph(afk, nz)

    def step():
        img = afk.Image.upload(host, sync=False)
        g.set_constant(c_cube, img)
        res = g.compute(ratio)
        res.download(out=res_t.numpy())  # blocks until the 2-D result is on the host

    # wire the graph once with a real image so the edges type-check (Image -> average.image)
    img0 = afk.Image.upload(host, sync=True)
    g.set_constant(c_cube, img0)
    for t in avg:
        g.connect(c_cube, 0, t, 0)
    g.connect(avg[0], 0, ratio, 0)
    g.connect(avg[1], 0, ratio, 1)
    del img0
    for _ in range(min(2, args.warmup)):
        step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n_steps):
        step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / n_steps
    tt = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dt = float(tt.item())
    sbytes = algorithmic_bytes(nz, rows, nx)
    return {"value": sbytes * world / dt / 1e9, "unit": UNIT, "h2d_bytes_per_step": nz * rows * nx * 4 * world,
            "d2h_bytes_per_step": rows * nx * 4 * world, "ms_per_step": dt * 1e3, "steps": n_steps,
            "cube_per_gpu": [nz, rows, nx],
            "path": "pinned host cube -> afk_image_upload -> graph(average, average, ratio; fused) -> afk_image_download"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="alma", choices=sorted(WORKLOADS))
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
