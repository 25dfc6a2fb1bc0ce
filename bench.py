#!/usr/bin/env python
"""bench.py -- aflak hot path on B200: spectral average+ratio over an ALMA-shape cube (BASELINE.json configs[2]), plus
every other BASELINE config and the multi-rank parity check in the same JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--skip c5,configs,e2e,cpu,parity]

Headline (`value`, unchanged round over round): one "step" = one pass of the hot path over one cube: `average[0,nz/2)`,
`average[nz/2,nz)` and their ratio (aflak_primitives/src/lib.rs:398-411, unit.rs:863-880), which the evaluator runs as
ONE fused kernel that streams the whole cube once.  Metric: cube GB/s = algorithmic bytes (cube read + 2-D result
write) / device time.  N>1 (torchrun, one process per GPU): weak scaling, global cube [nz, 1024*N, 1024], rank r owns
the y-band [1024 r, 1024 (r+1)); the kernel's epilogue stores the 2-D result band into every rank's copy of
the full image through NVLink peer mappings (fused gather, aflak_b200/csrc/shard.cpp).  `--gather nccl` uses
torch.distributed all_gather instead (the checked fallback).  value = bytes of all ranks / max-over-ranks device time.

Further blocks of the same line:
  gather_parity  after the timed region every rank compares sampled rows of ITS copy of the gathered image with the
                 oracle (sequential mode bit for bit, compensated mode SURVEY 8c tier C) and the whole image with the
                 NCCL-gathered result; any mismatch makes the run exit non-zero.
  e2e            the same step with HOST (pinned) cubes: upload + operators + read-back inside the timed region.
                 N=1: operator registry / graph evaluator; N>1: per-rank band upload -> fused reduce+gather -> rank 0
                 reads the ONE assembled image back.
  configs        (N=1) BASELINE configs[0], [1], [3]: C1 MaNGA integral, C2 EW template + linear_composition of 3
                 slices, C4 4096^2 display chain -- device time single-shot (cold and warm L2) and back to back,
                 algorithmic bytes, fraction of the measured HBM peak, and the CPU port timed beside each.
  c5_strong      (every N) BASELINE configs[4]: the fixed [2048,4096,4096] cube (128 GiB) sharded by y-bands, strong
                 scaling: integral and the EW template (whole-cube windows and 3 x 20 frames), in-run parity.
`--impl reference`: the CPU restatement of the reference's own code path (oracle/, kind "port": the reference is pure
Rust and there is no rustc in this image) timed on the host cores on the FULL cube, threaded the way the reference
threads it; the all-cores variant is reported beside it.
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
NZ, NY, NX = 3840, 1024, 1024
DESC = "ALMA-shape cube 1024x1024x3840 f32: average[0,1920) / average[1920,3840) -> ratio (fused, one pass)"
METRIC = "cube GB/s, spectral average+ratio (fused) over ALMA-shape cube"
UNIT = "GB/s"
C5_SEED = 0xAF1A0005
C5_SHAPE = (2048, 4096, 4096)


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs: device copy, read+write bytes)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel_label):
    """dram bytes per launch of the reduce kernel from the committed `ncu --set full` capture (profiles/), or None when
    the capture is of another instantiation than the one that ran (template arguments compared)."""
    import re

    ran = re.findall(r"=(\d+)", kernel_label.split(">")[0])
    for name in ("r2_reduce_kernel_alma_ncu_full.json", "r1_reduce_kernel_alma_ncu_full.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                d = json.load(f)
            captured = re.findall(r"\d+", d["kernel"].split("<")[1].split(">")[0])
            if captured != ran:
                continue
            return {"bytes": float(d["dram_bytes_read"]) + float(d["dram_bytes_write"]), "source": "profiles/" + name,
                    "kernel": d["kernel"]}
        except Exception:
            continue
    return None


def algorithmic_bytes(nz, ny, nx):
    return nz * ny * nx * 4 + ny * nx * 4


def config_dict(world, gather_mode="p2p"):
    """The workload description.  Identical in the GPU arm and the reference arm (a pure function of N)."""
    return {
        "workload": DESC,
        "cube_per_gpu": f"{NZ}x{NY}x{NX} f32 ({NZ * NY * NX * 4 / 1e9:.2f} GB)",
        "global_cube": f"{NZ}x{NY * world}x{NX} f32",
        "sharding": ("single GPU" if world == 1 else
                     "y-bands, one per GPU; 2-D result gathered by " +
                     ("peer stores fused into the reduce kernel (NVLink P2P)" if gather_mode == "p2p" else "NCCL all_gather")),
        "l2": "inputs (16.1 GB per step per GPU) are 128x larger than L2: no flush needed",
    }


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
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, sm_all = [], [], set(), []
        for ts, ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 9:
                continue
            try:
                v, m = float(p[1]), float(p[2])
            except ValueError:
                continue
            sm_all.append(v)
            mx.append(m)
            if t0 is not None and not (t0 - 0.06 <= ts <= t1 + 0.06):
                continue
            sm.append(v)
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            sm = sm_all
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle (restatement of the reference's ndarray path)
# ---------------------------------------------------------------------------------------------------------------
def cpu_step_reference_threading(oracle, host, nf):
    """What aflak does: ndarray sum_axis is a serial frame-accumulate loop; the two sibling `average`
    nodes are evaluated concurrently by rayon::scope (aflak_cake/src/dst/compute.rs:206-224) -> 2 threads; the
    elementwise division runs on the caller's thread."""
    res = [None, None]

    def work(i, s, e):
        res[i] = oracle.average(host, s, e)  # ctypes releases the GIL

    ts = [threading.Thread(target=work, args=(0, 0, nf // 2)), threading.Thread(target=work, args=(1, nf // 2, nf))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    return oracle.ratio(res[0], res[1])


def cpu_step_all_cores(oracle, host, nf, threads):
    a = oracle.average(host, 0, nf // 2, threads=threads)
    b = oracle.average(host, nf // 2, nf, threads=threads)
    return oracle.ratio(a, b)


def cpu_time(fn, reps, budget_s):
    times = []
    t_all = time.perf_counter()
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_all > budget_s:
            break
    return times


def cpu_baseline(frames=256, budget_s=12.0):
    """cpu_baseline leg of the GPU arm (rank 0, N=1): ~10-25 s of CPU work on a bounded sample (the rate is flat in the
    frame count; the reference arm runs the full cube)."""
    import oracle

    oracle.build()
    threads = oracle.max_threads()
    host = oracle.gen_cube(SEED, frames, NY, NX, threads=threads)
    nbytes = algorithmic_bytes(frames, NY, NX)
    t_ref = cpu_time(lambda: cpu_step_reference_threading(oracle, host, frames), 5, budget_s)
    t_all = cpu_time(lambda: cpu_step_all_cores(oracle, host, frames, threads), 5, budget_s / 2)
    ref, allc = min(t_ref), min(t_all)
    return {
        "value": nbytes / ref / 1e9, "unit": UNIT, "cores": 2, "kind": "port",
        "sample": f"[{frames},{NY},{NX}] frames of the cube ({nbytes / 1e9:.2f} GB), oracle average x2 (one thread each, serial "
                  f"frame-accumulate like ndarray sum_axis; siblings concurrent like rayon::scope) + ratio, best of {len(t_ref)}",
        "voxels_per_s": frames * NY * NX / ref,
        "all_cores": {"value": nbytes / allc / 1e9, "unit": UNIT, "cores": threads,
                      "sample": f"same sample, reductions pixel-tiled over {threads} OpenMP threads (not what the reference does), best of {len(t_all)}"},
    }


def run_reference(args):
    """--impl reference: rank 0 times the CPU path on the full cube; other ranks exit 0 without work."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import psutil

    import oracle

    oracle.build()
    threads = oracle.max_threads()
    # the full cube (16.1 GB of host memory, 0.67 s per step on the round-1 box); a smaller host falls back to as many
    # full 1024x1024 frames as fit in a third of the free memory, and says so
    frames = NZ
    avail = psutil.virtual_memory().available
    while frames > 32 and frames * NY * NX * 4 > avail / 3:
        frames //= 2
    host = oracle.gen_cube(SEED, frames, NY, NX, threads=threads)
    nbytes = algorithmic_bytes(frames, NY, NX)
    t0 = time.perf_counter()
    cpu_step_reference_threading(oracle, host, frames)
    first = time.perf_counter() - t0
    # keep the whole run within a few minutes whatever --steps says (the full-cube step is ~0.7 s)
    steps = max(1, min(args.steps, int(150.0 / max(first, 1e-3)) - args.warmup))
    for _ in range(args.warmup):
        cpu_step_reference_threading(oracle, host, frames)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_step_reference_threading(oracle, host, frames)
    dt = (time.perf_counter() - t0) / steps
    value = nbytes / dt / 1e9
    t_all = cpu_time(lambda: cpu_step_all_cores(oracle, host, frames, threads), 3, 10.0)
    whole = "the full cube" if frames == NZ else f"the first {frames} frames of the cube (host memory bound)"
    sample = (f"{whole}: [{frames},{NY},{NX}] ({nbytes / 1e9:.2f} GB) per step; oracle average x2 (serial frame-accumulate "
              f"like ndarray 0.12.1 sum_axis, the two sibling nodes concurrent like rayon::scope: 2 threads) + ratio")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": config_dict(args.gpus, args.gather),
        "detail": {"timing": "host wall clock; CPU path, no GPU; rank 0 processes one GPU's cube (the metric is a rate)",
                   "host_threads_available": threads, "frames_per_step": frames, "voxels_per_s": frames * NY * NX / dt},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 2, "kind": "port", "sample": sample,
                         "all_cores": {"value": nbytes / min(t_all) / 1e9, "unit": UNIT, "cores": threads,
                                       "sample": f"same cube, reductions pixel-tiled over {threads} OpenMP threads "
                                                 "(a parallelisation the reference does not have)"}},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ---------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------
class Ctx:
    """what every block of the GPU arm needs"""

    def __init__(self, args, afk, torch, dist, rank, world, local, dev, stream):
        self.args, self.afk, self.torch, self.dist = args, afk, torch, dist
        self.rank, self.world, self.local, self.dev, self.stream = rank, world, local, dev, stream
        self.peak, self.peak_src = measured_peak()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def exchange(self, h):
        lst = [None] * self.world
        self.dist.all_gather_object(lst, h)
        return lst

    def max_over_ranks(self, *vals):
        t = self.torch.tensor(list(vals), dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t.tolist()]

    def all_ok(self, ok):
        t = self.torch.tensor([0 if ok else 1], device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t)
        return int(t.item()) == 0


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


def bind_to_gpu_numa(torch, local):
    """Pin this rank's host threads (and therefore its first-touch / pinned allocations) to the NUMA node the GPU hangs
    off: with one rank per GPU the H2D copies of all ranks otherwise cross the socket interconnect.  Best effort."""
    try:
        p = torch.cuda.get_device_properties(local)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            spec = f.read().strip()
        cpus = set()
        for part in spec.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return {"numa_node": node, "cpus": len(cpus)}
    except Exception:
        return None


def tier_c_ok(got, s64, a64):
    """SURVEY 8c tier C: |gpu - o64| <= 1e-5 |o64| or <= 4 * 2^-24 * sum|x|; NaN/Inf positions exact."""
    got = got.astype(np.float64)
    fin = np.isfinite(s64)
    if not np.array_equal(np.isnan(got), np.isnan(s64)) or not np.array_equal(got[~fin & ~np.isnan(s64)], s64[~fin & ~np.isnan(s64)]):
        return False
    err = np.abs(got[fin] - s64[fin])
    return bool(np.all((err <= 1e-5 * np.abs(s64[fin])) | (err <= 4 * 2.0 ** -24 * a64[fin])))


def bits_equal(a, b):
    return np.array_equal(np.ascontiguousarray(a, dtype=np.float32).view(np.uint32), np.ascontiguousarray(b, dtype=np.float32).view(np.uint32))


def ew_params(win):
    """ratio_from_bands (aflak_primitives/src/lib.rs:226-246) of the window centres + the template's constants"""
    on, off1, off2 = win
    z, z1, z2 = [np.float32((a + b) / 2) for a, b in (on, off1, off2)]
    r1 = np.float32(np.float32(z - z1) / np.float32(z2 - z1))
    return [float(np.float32(1.0) - r1), float(r1), 20.0, 1e10, -1.0]


def oracle_rows(oracle, seed, shape, yy, windows, epi, params, afk):
    """the oracle's result for row yy of the global cube, sequential f32 (the reference's bits), and for the integral the
    f64 sum / sum of magnitudes of tier C"""
    nz, ny, nx = shape
    hb = oracle.gen_cube(seed, nz, ny, nx, y0=yy, rows=1, threads=4)
    if epi == afk.EPI_SUM:
        a, b = windows[0]
        return oracle.integral(hb, a, b), oracle.sum_f64(hb, a, b)
    if epi == afk.EPI_MEAN_RATIO:
        return oracle.ratio(oracle.average(hb, *windows[0]), oracle.average(hb, *windows[1])), None
    on, off1, off2 = windows
    off = oracle.linear_composition(oracle.average(hb, *off1), oracle.average(hb, *off2), params[0], params[1])
    return oracle.equivalent_width(off, oracle.average(hb, *on), 20.0, 1e10, True), None


def check_gathered_rows(cx, oracle, gather, band_ptr, seed, shape, windows, epi, params, rows_sample, what):
    """Both sum modes: reduce + gather, then sampled rows of THIS rank's copy of the full image against the oracle.
    Returns (ok, list of messages)."""
    afk = cx.afk
    nz = shape[0]
    ok, msgs = True, []
    for mode in (afk.SUM_SEQUENTIAL, afk.SUM_COMPENSATED):
        gather.reduce(band_ptr, nz, windows, epi, params, sum_mode=mode)
        gather.wait()
        cx.torch.cuda.synchronize()
        got = gather.result_image().download()
        for yy in rows_sample:
            want, f64 = oracle_rows(oracle, seed, shape, yy, windows, epi, params, afk)
            row = got[yy:yy + 1]
            if mode == afk.SUM_SEQUENTIAL:
                same = bits_equal(row, want)
            elif f64 is not None:
                same = tier_c_ok(row, f64[0], f64[1])
            else:  # quotients of compensated means: 1e-5 relative to the reference's own result scale
                same = bool(np.allclose(row, want, rtol=1e-4 if epi == afk.EPI_EW else 1e-5, atol=1e-4 if epi == afk.EPI_EW else 0.0, equal_nan=True))
            if not same:
                ok = False
                msgs.append(f"rank {cx.rank}: {what}: mode {'sequential' if mode == afk.SUM_SEQUENTIAL else 'compensated'} row {yy} differs from the oracle")
    return ok, msgs


def run_gather_parity(cx, gather, cube, nz, ny, nx, windows, out, full):
    """In-run multi-rank parity of the headline workload."""
    import oracle

    oracle.build()
    afk, torch, dist = cx.afk, cx.torch, cx.dist
    world, rank = cx.world, cx.rank
    shape = (nz, ny * world, nx)
    rows_sample = sorted({0, ny - 1, (ny * world) // 2, min(ny * world - 1, (ny * world) // 2 + 1), ny * world - 1, rank * ny + 7})
    ok, msgs = check_gathered_rows(cx, oracle, gather, cube.device_ptr, SEED, shape, windows, afk.EPI_MEAN_RATIO, None, rows_sample,
                                   "C3 average/average/ratio")
    # the whole gathered image against the NCCL all_gather of the unfused kernel result, both modes, bit for bit
    nccl_same = True
    for mode in (afk.SUM_SEQUENTIAL, afk.SUM_COMPENSATED):
        gather.reduce(cube.device_ptr, nz, windows, afk.EPI_MEAN_RATIO, sum_mode=mode)
        p = gather.wait()
        afk.k_reduce(cube.device_ptr, nz, ny * nx, windows, afk.EPI_MEAN_RATIO, out.data_ptr(), sum_mode=mode)
        if world > 1:
            dist.all_gather_into_tensor(full, out)
        torch.cuda.synchronize()
        mine = afk.Image.wrap_device(p, (ny * world, nx)).download()
        theirs = full.cpu().numpy() if world > 1 else out.cpu().numpy()
        if not bits_equal(mine, theirs):
            nccl_same = False
            msgs.append(f"rank {rank}: fused gather differs from NCCL all_gather (mode {mode}): {(mine.view(np.uint32) != theirs.view(np.uint32)).sum()} pixels")
    gather.check()
    ok = cx.all_ok(ok and nccl_same)
    for m in msgs:
        print(m, file=sys.stderr, flush=True)
    return {"status": "ok" if ok else "FAILED", "ranks_checked": world, "rows_sampled_per_rank": len(rows_sample),
            "oracle": "sequential mode bit for bit, compensated mode tier C (1e-5 relative to the f64 path)",
            "vs_nccl_all_gather": "whole image, both sum modes, bit for bit",
            "each_rank_checks": "its own copy of the gathered [ny*N, nx] image"}, ok


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
    numa = bind_to_gpu_numa(torch, local) if world > 1 else None  # N=1 keeps every host core for the cpu_baseline leg
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    afk.set_stream(stream.cuda_stream)  # the library launches on this stream, so torch events time it
    cx = Ctx(args, afk, torch, dist, rank, world, local, dev, stream)
    skip = set(x for x in args.skip.split(",") if x)

    nz, ny, nx = NZ, NY, NX
    n2 = ny * nx
    windows = [(0, nz // 2), (nz // 2, nz)]
    cube = afk.Image.synthetic(SEED, nz, ny * world, nx, y0=rank * ny, rows=ny)
    step_bytes = algorithmic_bytes(nz, ny, nx)

    # the fused-gather object exists at every N (world 1: the same kernel path with one destination); the timed step at
    # N=1 is the plain kernel, as in round 1
    gather = afk.Gather(world, rank, ny * world, nx, cx.exchange if world > 1 else None)
    assert gather.band() == (rank * ny, ny)
    use_gather = world > 1 and args.gather == "p2p"
    out = torch.empty((ny, nx), dtype=torch.float32, device=dev)
    full = torch.empty((ny * world, nx), dtype=torch.float32, device=dev) if world > 1 else out

    def step():
        if use_gather:
            gather.reduce(cube.device_ptr, nz, windows, afk.EPI_MEAN_RATIO)  # band -> every rank's full image, + flag
            gather.wait()                                                     # one-warp stream-ordered wait for all bands
        else:
            afk.k_reduce(cube.device_ptr, nz, n2, windows, afk.EPI_MEAN_RATIO, out.data_ptr())
            if world > 1:
                dist.all_gather_into_tensor(full, out)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        t_wait = time.time()
        while not sampler.lines and time.time() - t_wait < 3.0:
            time.sleep(0.05)
    for _ in range(args.warmup):
        step()
    cx.barrier()
    launches0 = afk.launch_count()
    afk.profile_reduce(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_mark0 = time.time()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    cx.barrier()
    t_mark1 = time.time()
    ms_total = e0.elapsed_time(e1)
    k_ms, k_n = afk.profile_reduce_read()
    afk.profile_reduce(False)
    launches = afk.launch_count() - launches0
    gather.check()
    if rank == 0 and (t_mark1 - t_mark0) < 1.0:
        # keep the GPU under the same load a little longer so nvidia-smi (>= 50 ms period) samples it under load
        t_end = time.time() + 1.0
        while time.time() < t_end:
            for _ in range(20):
                afk.k_reduce(cube.device_ptr, nz, n2, windows, afk.EPI_MEAN_RATIO, out.data_ptr())
            torch.cuda.synchronize()
        t_mark1 = time.time()
    clocks = sampler.stop(t_mark0, t_mark1) if rank == 0 else None
    (ms_max,) = cx.max_over_ranks(ms_total)
    ms_step = ms_max / args.steps
    value = step_bytes * world / (ms_step * 1e-3) / 1e9
    kernel_label = afk.k_reduce_describe(cube.device_ptr, n2, windows, out.data_ptr())

    failed = False
    parity = None
    if "parity" not in skip:
        parity, ok = run_gather_parity(cx, gather, cube, nz, ny, nx, windows, out, full)
        failed |= not ok

    # ---- the same step with the gather's stores going through an NVLink-SHARP multicast mapping (A/B, N > 1) ----
    mc = None
    if world > 1 and "mc" not in skip:
        mc, ok = run_gather_multicast(cx, gather, cube, nz, ny, nx, windows, step_bytes, ms_step)
        failed |= not ok

    # ---- e2e: host (pinned) cube -> upload -> operators -> host result ----
    e2e = None
    if "e2e" not in skip:
        e2e = run_e2e(cx, gather, nz, ny, nx)
    cx.barrier()
    gather.close()
    del cube, gather

    configs = None
    if world == 1 and "configs" not in skip:
        configs = run_configs(cx)
    c5 = None
    if "c5" not in skip:
        c5, ok = run_c5_strong(cx)
        failed |= not ok

    if rank == 0:
        k_avg_ms = k_ms / max(k_n, 1)
        achieved = step_bytes / (k_avg_ms * 1e-3) / 1e9
        tr = ncu_traffic(kernel_label)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": config_dict(world, args.gather),
            "detail": {"sum": "compensated (Kahan) f32", "voxels_per_s": nz * ny * nx * world / (ms_step * 1e-3),
                       "timing": "CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": cx.peak, "unit": "GB/s", "frac": achieved / cx.peak,
                         "traffic": tr["bytes"] if tr else None, "traffic_source": (tr["source"] + " (" + tr["kernel"] + ")") if tr else None,
                         "kernel": kernel_label, "kernel_ms": k_avg_ms, "launches_timed": k_n,
                         "algorithmic_bytes_per_launch": step_bytes, "peak_source": cx.peak_src,
                         "frac_of_nominal_8TBs": achieved / 8000.0},
            "e2e": e2e, "gpu_launches": int(launches * world), "clocks": clocks,
            "gather_parity": parity["status"] if parity else "skipped", "gather_parity_detail": parity,
            "c5_strong": c5,
        }
        if mc is not None:
            line["gather_multicast"] = mc
        if configs is not None:
            line["configs"] = configs
        if numa:
            line["detail"]["host_affinity"] = f"rank 0 bound to NUMA node {numa['numa_node']} ({numa['cpus']} cpus), every rank to its GPU's node"
        if world == 1 and "cpu" not in skip:
            line["cpu_baseline"] = cpu_baseline()
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 1 if failed else 0


def run_gather_multicast(cx, gather, cube, nz, ny, nx, windows, step_bytes, ms_p2p):
    """Headline step with host-supplied symmetric memory (torch's allocator: cuMem + multicast object) whose multicast mapping
    turns the epilogue's N stores per tile into one multimem.st.  Timed like the headline; the gathered image must equal the
    peer-store gather's bit for bit (same kernel plan, same arithmetic)."""
    afk, torch, dist = cx.afk, cx.torch, cx.dist
    rank, world, stream, args = cx.rank, cx.world, cx.stream, cx.args
    err, g2 = "", None
    try:
        sym = afk.symmetric_memory_torch(afk.gather_bytes(ny * world, nx))
        g2 = afk.Gather(world, rank, ny * world, nx, symmetric=sym)
    except Exception as e:
        err = f"{type(e).__name__}: {str(e)[:200]}"
    if not cx.all_ok(err == ""):
        return {"status": "unavailable", "error": err}, True
    info = {"status": "ok", "multicast_mapping": bool(g2.multicast),
            "memory": "torch symmetric memory (cuMem + NVLink multicast object), passed to afk_gather_create_symmetric"}
    for _ in range(args.warmup):
        g2.reduce(cube.device_ptr, nz, windows, afk.EPI_MEAN_RATIO)
        g2.wait()
    cx.barrier()
    afk.profile_reduce(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        g2.reduce(cube.device_ptr, nz, windows, afk.EPI_MEAN_RATIO)
        g2.wait()
    e1.record(stream)
    cx.barrier()
    k_ms, k_n = afk.profile_reduce_read()
    afk.profile_reduce(False)
    ms, kms = cx.max_over_ranks(e0.elapsed_time(e1) / args.steps, k_ms / max(k_n, 1))
    # parity: one more step on both gathers, whole image, bit for bit
    gather.reduce(cube.device_ptr, nz, windows, afk.EPI_MEAN_RATIO)
    gather.wait()
    a = gather.result_image().download()
    g2.reduce(cube.device_ptr, nz, windows, afk.EPI_MEAN_RATIO)
    g2.wait()
    b = g2.result_image().download()
    same = bits_equal(a, b)
    try:
        g2.check()
    except Exception as e:
        same = False
        info["error"] = str(e)[:200]
    ok = cx.all_ok(same)
    info.update({"ms_per_step": ms, "value_GBps": step_bytes * world / (ms * 1e-3) / 1e9, "reduce_kernel_ms_max_over_ranks": kms,
                 "peer_store_gather_ms_per_step": ms_p2p, "parity_vs_peer_store_gather": "ok (whole image, bit for bit, every rank)" if ok else "FAILED"})
    cx.barrier()
    g2.close()
    return info, ok


# ---------------------------------------------------------------------------------------------------------------
# e2e
# ---------------------------------------------------------------------------------------------------------------
def run_e2e(cx, gather, nz, ny, nx):
    import psutil

    args, afk, torch, dist = cx.args, cx.afk, cx.torch, cx.dist
    rank, world, dev = cx.rank, cx.world, cx.dev
    avail = psutil.virtual_memory().available
    rows = ny
    while rows > 16 and nz * rows * nx * 4 * world * 1.5 > avail:  # keep pinned host memory bounded on small hosts
        rows //= 2
    n_steps = max(1, min(args.steps, args.e2e_steps))
    host_t = torch.empty((nz, rows, nx), dtype=torch.float32, pin_memory=True)
    # fill the pinned host cube from a device-generated synthetic cube (rows [0, rows) of this rank's band)
    tmp = afk.Image.synthetic(SEED, nz, ny * world, nx, y0=rank * ny, rows=rows)
    afk.lib.afk_image_download(tmp.handle, host_t.data_ptr())
    del tmp
    host = host_t.numpy()
    sbytes = algorithmic_bytes(nz, rows, nx)
    windows = [(0, nz // 2), (nz // 2, nz)]
    h2d_ms = []

    if world == 1:
        res_t = torch.empty((rows, nx), dtype=torch.float32, pin_memory=True)
        g, c_cube, avg, ratio = build_c3_graph(afk, nz)
        phases = []

        def step():
            t0 = time.perf_counter()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record(cx.stream)
            img = afk.Image.upload(host, sync=False)  # pinned cudaMemcpyAsync, stream-ordered
            ev1.record(cx.stream)
            t1 = time.perf_counter()
            g.set_constant(c_cube, img)
            res = g.compute(ratio)                    # fused average/average/ratio kernel on the same stream
            t2 = time.perf_counter()
            res.download(out=res_t.numpy())           # blocks until the 2-D result is on the host
            t3 = time.perf_counter()
            phases.append((t1 - t0, t2 - t1, t3 - t2))
            h2d_ms.append(ev0.elapsed_time(ev1))

        # wire the graph once with a real image so the edges type-check (Image -> average.image)
        img0 = afk.Image.upload(host, sync=True)
        g.set_constant(c_cube, img0)
        for t in avg:
            g.connect(c_cube, 0, t, 0)
        g.connect(avg[0], 0, ratio, 0)
        g.connect(avg[1], 0, ratio, 1)
        del img0
        path = "pinned host cube -> afk_image_upload -> graph(average, average, ratio; fused) -> afk_image_download"
        d2h = rows * nx * 4
    else:
        # the sharded cube end to end: every rank uploads ITS band of the one global cube, reduces it with the fused
        # gather, and rank 0 reads the ONE assembled [rows*N, nx] image back to the host
        if rows != ny:
            return {"value": None, "unit": UNIT, "skipped": "not enough host memory for one pinned band per rank"}
        res_t = torch.empty((ny * world, nx), dtype=torch.float32, pin_memory=True) if rank == 0 else None

        def step():
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record(cx.stream)
            img = afk.Image.upload(host, sync=False)
            ev1.record(cx.stream)
            gather.reduce(img.device_ptr, nz, windows, afk.EPI_MEAN_RATIO)
            p = gather.wait()
            if rank == 0:
                afk.Image.wrap_device(p, (ny * world, nx)).download(out=res_t.numpy())
            else:
                torch.cuda.synchronize()
            h2d_ms.append(ev0.elapsed_time(ev1))

        path = ("per-rank pinned band of the global cube -> afk_image_upload -> afk_gather_reduce (fused reduce + NVLink gather) -> "
                "afk_gather_wait -> rank 0 downloads the one assembled image")
        d2h = ny * world * nx * 4  # rank 0 only; the other ranks read nothing back

    for _ in range(2):
        step()
    cx.barrier()
    del h2d_ms[:]
    t0 = time.perf_counter()
    for _ in range(n_steps):
        step()
    torch.cuda.synchronize()
    dt_local = (time.perf_counter() - t0) / n_steps
    h2d_local = float(np.median(h2d_ms)) if h2d_ms else 0.0
    dt, h2d_max = cx.max_over_ranks(dt_local, h2d_local)
    (neg_h2d_min,) = cx.max_over_ranks(-h2d_local)
    band_bytes = nz * rows * nx * 4
    out = {"value": sbytes * world / dt / 1e9, "unit": UNIT, "h2d_bytes_per_step": band_bytes * world,
           "d2h_bytes_per_step": d2h, "ms_per_step": dt * 1e3, "steps": n_steps, "cube_per_gpu": [nz, rows, nx],
           "h2d_GBps_per_gpu": {"slowest_rank": band_bytes / (h2d_max * 1e-3) / 1e9 if h2d_max > 0 else None,
                                "fastest_rank": band_bytes / (-neg_h2d_min * 1e-3) / 1e9 if neg_h2d_min < 0 else None,
                                "how": "CUDA events around the rank's own upload (pinned cudaMemcpyAsync, 256 MiB chunks), median over the timed steps"},
           "path": path}
    if world == 1:
        ph = np.array(phases[-n_steps:]).mean(0) * 1e3
        out["host_call_ms"] = {"upload_enqueue": float(ph[0]), "graph_compute_enqueue": float(ph[1]), "download_wait": float(ph[2])}
        # the interactive case aflak is built around: the cube stays cached on the device, the user drags
        # the range -> only the Integer constants change, the graph re-evaluates and the 2-D result is read back
        g2, c2, avg2, ratio2 = build_c3_graph(afk, nz)
        img = afk.Image.upload(host, sync=True)
        g2.set_constant(c2, img)
        for t in avg2:
            g2.connect(c2, 0, t, 0)
        g2.connect(avg2[0], 0, ratio2, 0)
        g2.connect(avg2[1], 0, ratio2, 1)
        bound = g2.add_constant(int(nz // 2))
        g2.connect(bound, 0, avg2[0], 2)
        g2.compute(ratio2).download(out=res_t.numpy())
        torch.cuda.synchronize()
        n_int = 20
        t0 = time.perf_counter()
        for i in range(n_int):
            g2.set_constant(bound, int(nz // 2 - (i % 2)))  # update_constant_node: invalidates average #0 and ratio
            g2.compute(ratio2).download(out=res_t.numpy())
        dt_int = (time.perf_counter() - t0) / n_int
        out["resident_cube"] = {"value": sbytes / dt_int / 1e9, "unit": UNIT, "ms_per_step": dt_int * 1e3, "steps": n_int,
                                "h2d_bytes_per_step": 8, "d2h_bytes_per_step": rows * nx * 4,
                                "path": "cube cached as a device handle in the graph; per step: set_constant(end) -> compute -> download 2-D result"}
        del g2, img
    return out


# ---------------------------------------------------------------------------------------------------------------
# BASELINE configs[0], [1], [3] (C1, C2, C4) at N=1
# ---------------------------------------------------------------------------------------------------------------
def run_configs(cx):
    import oracle

    oracle.build()
    afk, torch, stream = cx.afk, cx.torch, cx.stream
    PEAK = cx.peak
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=cx.dev)  # 2x the 126 MB L2

    def single(fn, reps=40, warm=5, cold=False):
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
        return float(np.median(ts)) * 1e3

    def back_to_back(fn, batch=20, rounds=5):
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
        return best * 1e3

    def host_wall(fn, n=200):
        fn()
        afk.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        afk.synchronize()
        return (time.perf_counter() - t0) / n * 1e6

    def entry(what, nbytes, fn, cold=True):
        l0 = afk.launch_count()
        fn()
        nl = afk.launch_count() - l0
        e = {"what": what, "algorithmic_bytes": nbytes, "kernel_launches": nl, "floor_us_at_measured_peak": nbytes / PEAK / 1e3}
        warm = single(fn)
        e["device_us_single_warm_l2"] = warm
        if cold:
            c = single(fn, reps=20, cold=True)
            e["device_us_single_cold_l2"] = c
            e["frac_single_cold_l2"] = nbytes / (c * 1e-6) / 1e9 / PEAK
        e["frac_single_warm_l2"] = nbytes / (warm * 1e-6) / 1e9 / PEAK
        b = back_to_back(fn)
        e["device_us_back_to_back"] = b
        e["frac_back_to_back"] = nbytes / (b * 1e-6) / 1e9 / PEAK
        return e

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

    ncpu = oracle.max_threads()
    out = {"timing": "CUDA events on the launching stream; single = one call between two events (median), cold = a 256 MiB "
                     "buffer is overwritten before each call, back_to_back = 20 calls between two events (best of 5)",
           "measured_hbm_peak_GBps": PEAK}

    # ---------------- C1: MaNGA-shape cube, integral over the whole range (BASELINE configs[0]) ----------------
    nz, ny, nx = 4563, 74, 74
    n2 = ny * nx
    cube = afk.Image.synthetic(0xAF1A0001, nz, ny, nx)
    res = afk.Image.alloc((ny, nx))
    c1b = nz * n2 * 4 + n2 * 4
    c1 = entry("MaNGA-shape cube 74x74x4563: integral[0,4563) (raw kernel layer)", c1b,
               lambda: afk.k_reduce(cube.device_ptr, nz, n2, [(0, nz)], afk.EPI_SUM, res.device_ptr))
    c1["kernel"] = afk.k_reduce_describe(cube.device_ptr, n2, [(0, nz)], res.device_ptr)
    c1["registry_call"] = entry("the same through the operator registry (afk_transform_call \"integral\")", c1b,
                                lambda: afk.call("integral", cube, 0, nz), cold=False)
    c1["registry_call"]["host_wall_us_per_call"] = host_wall(lambda: afk.call("integral", cube, 0, nz))
    hm = oracle.gen_cube(0xAF1A0001, nz, ny, nx, threads=ncpu)
    c1["cpu_port"] = {"ms": cpu_best(lambda: oracle.integral(hm, 0, nz)), "cores": 1, "kind": "port",
                      "threading": "ndarray sum_axis: one serial frame-accumulate loop"}
    # parity of what was timed
    s64, a64 = oracle.sum_f64(hm, 0, nz)
    c1["parity"] = "ok" if tier_c_ok(res.download(), s64, a64) else "FAILED"
    out["C1"] = c1

    # ---------------- C2: EW template + linear_composition of 3 slices (BASELINE configs[1]) ----------------
    W = ((3134, 3154), (3099, 3119), (3174, 3194))
    c2b = 60 * n2 * 4 + n2 * 4
    c2 = entry("MaNGA-shape cube: equivalent-width template, 3 x 20 frames, fused kernel (afk_fused_equivalent_width)", c2b,
               lambda: afk.fused_equivalent_width(cube, *W, 20.0, 1e10, True))

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
        g.connect(avg[1], 0, lc, 0)
        g.connect(avg[2], 0, lc, 1)
        g.connect(rb, 0, lc, 2)
        g.connect(rb, 1, lc, 3)
        ew = g.add_transform("create_equivalent_width")
        g.connect(lc, 0, ew, 0)
        g.connect(avg[0], 0, ew, 1)
        g.connect(g.add_constant(20.0), 0, ew, 2)
        g.connect(g.add_constant(1e10), 0, ew, 3)
        g.connect(g.add_constant(True), 0, ew, 4)
        mm = g.add_transform("image_min_max")
        g.connect(ew, 0, mm, 0)
        ls = g.add_transform("convert_to_logscale")
        g.connect(ew, 0, ls, 0)
        g.connect(g.add_constant(1000.0), 0, ls, 1)
        g.connect(mm, 0, ls, 2)
        g.connect(mm, 1, ls, 3)
        start = g.add_constant(3134)
        g.connect(start, 0, avg[0], 1)
        return g, ls, start

    for fusion in (True, False):
        g, ls, start = ew_graph(fusion)
        state = {"i": 0}

        def reeval(g=g, ls=ls, start=start, state=state):
            state["i"] ^= 1
            g.set_constant(start, 3134 - state["i"])  # the user drags the range: everything downstream re-evaluates
            return g.compute(ls)

        e = entry("EW template graph (3 average -> ratio_from_bands -> linear_composition -> create_equivalent_width -> image_min_max -> "
                  f"convert_to_logscale), re-evaluation after a constant change, {'fused' if fusion else 'unfused'} evaluator",
                  60 * n2 * 4 + 3 * n2 * 4, reeval, cold=False)
        e["host_wall_us_per_reevaluation"] = host_wall(reeval)
        c2["graph_reevaluation_fused" if fusion else "graph_reevaluation_unfused"] = e
    f = [afk.call("slice_one_frame", cube, k) for k in (3109, 3144, 3184)]
    c2["linear_composition_3_slices"] = entry("linear_composition of 3 slice_one_frame views (two chained calls)", 4 * n2 * 4 + 2 * n2 * 4,
                                              lambda: afk.call("linear_composition", afk.call("linear_composition", f[0], f[1], 0.5, 0.25), f[2], 1.0, 0.25), cold=False)

    def c2_cpu():
        on, o1, o2 = concurrently(lambda: oracle.average(hm, *W[0]), lambda: oracle.average(hm, *W[1]), lambda: oracle.average(hm, *W[2]))
        r2, r1 = oracle.ratio_from_bands(3144.0, 3109.0, 3184.0)
        ew = oracle.equivalent_width(oracle.linear_composition(o1, o2, r2, r1), on, 20.0, 1e10, True)
        mn, mx = oracle.image_min_max(ew)
        return ew, oracle.convert_to_logscale(ew, 1000.0, mn, mx)

    c2["cpu_port"] = {"ms": cpu_best(c2_cpu), "cores": 3, "kind": "port",
                      "threading": "the three sibling average nodes concurrent (rayon::scope), each a serial loop; the 5476-pixel tail on one core"}
    afk.set_sum_mode(afk.SUM_SEQUENTIAL)
    got = afk.fused_equivalent_width(cube, *W, 20.0, 1e10, True).download()
    afk.set_sum_mode(afk.SUM_INHERIT)
    c2["parity"] = "ok" if bits_equal(got, c2_cpu()[0]) else "FAILED"
    out["C2"] = c2
    del cube, res, f, g

    # ---------------- C4: 4096^2 display chain (BASELINE configs[3]) ----------------
    N = 4096
    rng = np.random.default_rng(4)
    himg = np.exp(rng.standard_normal((N, N))).astype(np.float32)
    img = afk.Image.upload(himg)
    npx = N * N
    rgba = afk.Image.alloc((N, N))
    tone = np.arange(256, dtype=np.float32)
    lmax = float(np.float32(np.log(np.float32(1001.0)) / np.log(np.float32(1000.0))))
    nan_chain = {"logscale_a": 1000.0, "logscale_vmin": float("nan"), "logscale_vmax": float("nan"), "normalize_min": 0.0,
                 "normalize_max": lmax / 65535.0, "tone_table": tone}
    c4 = {}
    for key, lims in (("lims_0.2_0.3_0.9", (0.2, 0.3, 0.9)), ("lims_default", (0.0, 0.5, 1.0))):
        lut = afk.Lut("Flame")
        lut.set_lims(*lims)
        e = entry(f"4096^2 image: image_min_max -> convert_to_logscale(1000) -> normalize -> apply_tone_curve(256) -> Flame LUT limits {lims} "
                  "-> RGBA8, flip_y; min/max pass + one fused chain kernel, the range stays on the device", npx * 12,
                  lambda lut=lut: afk.colormap_device(img, lut, 0.0, 65535.0, rgba.device_ptr, channels=4, chain=nan_chain))
        c4[key] = e
    lut = afk.Lut("Flame")
    lut.set_lims(0.2, 0.3, 0.9)
    vmin, vmax = afk.call("image_min_max", img)
    c4["colormap_only"] = entry("make_raw_image alone (given vmin/vmax) -> RGBA8", npx * 8,
                                lambda: afk.colormap_device(img, lut, float(vmin), float(vmax), rgba.device_ptr, channels=4))
    c4["image_min_max"] = entry("image_min_max alone (incl. the 8-byte read-back)", npx * 4, lambda: afk.call("image_min_max", img))
    table = oracle.lut_build("Flame")  # built with the default limits; set_lims only changes the look-up (lut.rs:232-236)

    def c4_cpu():
        mn, mx = oracle.image_min_max(himg)
        ls = oracle.convert_to_logscale(himg, 1000.0, mn, mx, threads=ncpu)
        nm = ((ls - np.float32(0.0)) / np.float32(lmax / 65535.0)).astype(np.float32)
        tc = oracle.apply_tone_curve(nm, tone)
        return oracle.make_raw_image(tc, 0.0, 65535.0, table, lims=(0.2, 0.3, 0.9), channels=4)

    t0 = time.perf_counter()
    want = c4_cpu()
    c4["cpu_port"] = {"ms": (time.perf_counter() - t0) * 1e3, "cores": ncpu, "kind": "port",
                      "threading": "logscale on all cores (par_iter_mut), min/max and the LUT mapping serial (UI thread)"}
    got = afk.colormap(img, lut, 0.0, 65535.0, channels=4, chain=nan_chain)
    # tier B: the chain calls logf; CUDA's and glibc's differ in the last bit now and then, which moves a pixel by one
    # tone-table entry (tests/test_gpu_display.py)
    diff = np.abs(got.astype(np.int32) - np.asarray(want).reshape(got.shape).astype(np.int32))
    frac = float((diff.max(axis=2) > 0).mean())
    gpu_exact = None
    try:  # the same chain with every pixel on the reference's own IEEE sequence (no guard band, no fast divisions)
        afk.set_display_exact(True)
        gpu_exact = afk.colormap(img, lut, 0.0, 65535.0, channels=4, chain=nan_chain)
    finally:
        afk.set_display_exact(False)
    byte_exact = bool(np.array_equal(got, gpu_exact))
    c4["parity"] = "ok" if byte_exact and frac < 1e-3 else "FAILED"
    c4["parity_detail"] = {"fast_paths_vs_all_ieee_device_sequence": "byte-exact" if byte_exact else "DIFFERENT",
                           "pixels_differing_from_cpu_port": frac,
                           "bar": "byte-exact against the device's all-IEEE sequence; tier B against the CPU port (logf: CUDA libm vs glibc "
                                  "differ in the last bit now and then, which moves a pixel by one tone-table entry): < 0.1 % of the pixels"}
    # the display statistics aflak_plot computes for the same image (hist.rs:10-66, lims.rs:37-106, zscale.rs:45-242);
    # guarded: a failure here is recorded, it does not take the headline down
    try:
        st = {}
        fvmin, fvmax = float(vmin), float(vmax)
        st["histogram_100_bins"] = entry("hist::histogram, 100 bins (incl. the read-back of the counts)", npx * 4, lambda: img.histogram(fvmin, fvmax))
        st["vmed_vmad"] = entry("lims::get_vmed_normalized + get_vmad_normalized: exact order statistics, one min/max pass + one bracket "
                                "pass per statistic (the reference sorts the image twice)", npx * 12, lambda: img.vmed_vmad())
        st["zscale_1000_0.25"] = entry("Image::zscale(1000, 0.25): strided sample on the device, line fit on the host", 0, lambda: img.zscale(1000, 0.25),
                                       cold=False)
        t0 = time.perf_counter()
        want_stats = oracle.vmed_vmad(himg)
        st["cpu_port"] = {"vmed_vmad_ms": (time.perf_counter() - t0) * 1e3, "cores": 1, "kind": "port", "threading": "two full sorts on the UI thread"}
        got_stats = img.vmed_vmad()
        exact = got_stats[0].tobytes() == want_stats[0].tobytes() and got_stats[1].tobytes() == want_stats[1].tobytes()
        st["parity"] = "ok" if exact else "FAILED"
        st["parity_how"] = "vmed and vmad bit for bit against the oracle's two sorts"
        c4["display_statistics"] = st
        if not exact:
            c4["parity"] = "FAILED"
    except Exception as e:  # noqa: BLE001
        c4["display_statistics"] = {"error": repr(e)[:300]}
    out["C4"] = c4
    del img, rgba, flush_buf
    return out


# ---------------------------------------------------------------------------------------------------------------
# BASELINE configs[4] (C5): fixed 128 GiB cube, strong scaling, every N
# ---------------------------------------------------------------------------------------------------------------
def run_c5_strong(cx):
    import oracle

    oracle.build()
    afk, torch, dist = cx.afk, cx.torch, cx.dist
    rank, world, dev, stream = cx.rank, cx.world, cx.dev, cx.stream
    nz, ny, nx = C5_SHAPE
    steps, warm = max(3, min(cx.args.steps, 10)), 3
    base = {"workload": "BASELINE configs[4]: ALMA-scale cube 4096x4096x2048 f32 (137.4 GB), y-bands over the GPUs, fused reduce + NVLink gather",
            "global_cube": [nz, ny, nx], "cube_bytes": nz * ny * nx * 4, "n_gpus": world, "scaling": "strong", "steps": steps, "warmup": warm}
    afk.trim_pools()  # the band needs most of the device at N=1 (128 GiB of ~179 GiB)
    try:
        g = afk.Gather(world, rank, ny, nx, cx.exchange if world > 1 else None)
        y0, rows = g.band()
        band = afk.Image.synthetic(C5_SEED, nz, ny, nx, y0=y0, rows=rows)
        torch.cuda.synchronize()
        alloc_ok = True
        err = ""
    except Exception as e:  # out of memory on a small device: record it, do not fail the headline
        alloc_ok, err = False, str(e)
    if not cx.all_ok(alloc_ok):
        base["status"] = "not run: the band does not fit on the device"
        base["error"] = err[:200]
        band = None
        if "g" in locals():
            g.close()
        return base, True
    base["band_rows_per_gpu"] = rows
    base["band_bytes_per_gpu"] = nz * rows * nx * 4
    free_b, total_b = torch.cuda.mem_get_info()
    base["device_memory"] = {"total_GiB": total_b / 2**30, "free_after_band_GiB": free_b / 2**30}

    c = nz // 2
    narrow = [(c - 10, c + 10), (c - 45, c - 25), (c + 30, c + 50)]
    w3 = nz // 3
    wide = [(w3, 2 * w3), (0, w3), (2 * w3, 3 * w3)]
    cases = [
        ("integral", "integral[0,2048)", [(0, nz)], afk.EPI_SUM, None),
        ("ew_whole_cube", "equivalent-width template, the whole cube in 3 windows", wide, afk.EPI_EW, ew_params(wide)),
        ("ew_3x20", "equivalent-width template, 3 x 20 frames", narrow, afk.EPI_EW, ew_params(narrow)),
    ]
    rows_sample = sorted({0, ny // world - 1, ny // 2, ny // 2 + 1, ny - 1})
    all_ok = True
    msgs = []
    entries = {}
    for key, name, windows, epi, params in cases:
        ok, m = check_gathered_rows(cx, oracle, g, band.device_ptr, C5_SEED, (nz, ny, nx), windows, epi, params, rows_sample, "C5 " + name)
        msgs += m
        ok = cx.all_ok(ok)
        all_ok &= ok
        for _ in range(warm):
            g.reduce(band.device_ptr, nz, windows, epi, params)
            g.wait()
        cx.barrier()
        afk.profile_reduce(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            g.reduce(band.device_ptr, nz, windows, epi, params)
            g.wait()
        e1.record(stream)
        cx.barrier()
        k_ms, k_n = afk.profile_reduce_read()
        afk.profile_reduce(False)
        ms, kms = cx.max_over_ranks(e0.elapsed_time(e1) / steps, k_ms / max(k_n, 1))
        frames = sum(b - a for a, b in windows)
        bytes_global = frames * ny * nx * 4 + ny * nx * 4
        per_gpu = (frames * rows * nx * 4 + rows * nx * 4) / (kms * 1e-3) / 1e9
        entries[key] = {"case": name, "windows": windows, "sum": "compensated", "ms_per_step_max_over_ranks": ms,
                        "aggregate_GBps": bytes_global / (ms * 1e-3) / 1e9, "aggregate_voxels_per_s": frames * ny * nx / (ms * 1e-3),
                        "algorithmic_bytes_global": bytes_global, "reduce_kernel_ms_max_over_ranks": kms,
                        "per_gpu_kernel_GBps": per_gpu, "per_gpu_frac_of_measured_peak": per_gpu / cx.peak,
                        "gathered_bytes_over_nvlink": ny * nx * 4 * (world - 1), "parity_sampled_rows": "ok" if ok else "FAILED"}
    # ---- the same cases with the result stored through the NVLink multicast mapping (A/B against the peer stores) ----
    if world > 1 and "mc" not in [x.strip() for x in cx.args.skip.split(",")]:
        err, g2 = "", None
        try:
            g2 = afk.Gather(world, rank, ny, nx, symmetric=afk.symmetric_memory_torch(afk.gather_bytes(ny, nx)))
        except Exception as e:
            err = f"{type(e).__name__}: {str(e)[:200]}"
        if cx.all_ok(err == ""):
            mc_ok = True
            for key, name, windows, epi, params in cases:
                g.reduce(band.device_ptr, nz, windows, epi, params)
                g.wait()
                ref_img = g.result_image().download()
                for _ in range(warm):
                    g2.reduce(band.device_ptr, nz, windows, epi, params)
                    g2.wait()
                same = bits_equal(ref_img, g2.result_image().download())
                del ref_img
                cx.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(steps):
                    g2.reduce(band.device_ptr, nz, windows, epi, params)
                    g2.wait()
                e1.record(stream)
                cx.barrier()
                (ms2,) = cx.max_over_ranks(e0.elapsed_time(e1) / steps)
                same = cx.all_ok(same)
                mc_ok &= same
                entries[key]["multicast_gather"] = {"ms_per_step_max_over_ranks": ms2, "multicast_mapping": bool(g2.multicast),
                                                    "parity_vs_peer_store_gather": "ok (whole image, bit for bit)" if same else "FAILED"}
            all_ok &= mc_ok
            g2.close()
        else:
            base["multicast_gather"] = {"status": "unavailable", "error": err}

    # ---- sharded display chain: min/max exchange + every rank colour-maps its band + RGBA band gather ----
    lut = afk.Lut("Flame")
    lut.set_lims(0.2, 0.3, 0.9)
    tone = oracle.create_curve([(0.0, 0.0), (0.4, 0.7), (1.0, 1.0)])
    lmax = np.float32(np.log(np.float32(1000.0) + np.float32(1.0)) / np.log(np.float32(1000.0)))
    chain = {"logscale_a": 1000.0, "logscale_vmin": float("nan"), "logscale_vmax": float("nan"), "normalize_min": 0.0,
             "normalize_max": float(lmax / np.float32(65535.0)), "tone_table": tone}
    d_err, memo, mn, mx, res = "", False, float("nan"), float("nan"), None
    try:  # local work only: the collectives below are reached by every rank whatever happens here
        g.reduce(band.device_ptr, nz, [(0, nz)], afk.EPI_SUM, None)
        g.wait()
        res = g.result_image()
        l0 = afk.launch_count()
        mn, mx = afk.call("image_min_max", res)
        memo = afk.launch_count() == l0
        got = g.display(lut, vmin=0.0, vmax=65535.0, chain=chain)
        single = afk.colormap(res, lut, vmin=0.0, vmax=65535.0, channels=4, chain=chain)  # one GPU, the whole gathered image
        if not np.array_equal(got, single):
            d_err = "gathered RGBA differs from the one-GPU chain"
        elif not memo:
            d_err = "image_min_max of the gathered image made a pass over it"
        del got, single
    except Exception as e:
        d_err = str(e)[:300]
    d_ok = cx.all_ok(d_err == "")
    all_ok &= d_ok
    disp = {"what": "gathered 4096x4096 integral -> logscale(1000) -> normalize -> tone curve(256) -> Flame LUT -> RGBA8, flip_y: every rank maps its "
                    "band (range from the min/max exchange, no pass over the image) and stores it into every rank's RGBA image (peer stores)",
            "parity": "ok" if d_ok else "FAILED", "error": d_err,
            "parity_how": "every rank's gathered RGBA image == the one-GPU chain on its copy of the gathered image, byte for byte"}
    if d_ok:
        for _ in range(warm):
            g.colormap(lut, 0.0, 65535.0, True, chain)
            g.wait_rgba()
        cx.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            g.colormap(lut, 0.0, 65535.0, True, chain)
            g.wait_rgba()
        e1.record(stream)
        cx.barrier()
        t_sharded = e0.elapsed_time(e1) / steps
        out1 = torch.empty(ny * nx * 4, dtype=torch.uint8, device=cx.dev)
        for _ in range(warm):
            afk.colormap_device(res, lut, 0.0, 65535.0, out1.data_ptr(), channels=4, chain=chain)
        e0.record(stream)
        for _ in range(steps):
            afk.colormap_device(res, lut, 0.0, 65535.0, out1.data_ptr(), channels=4, chain=chain)
        e1.record(stream)
        e1.synchronize()
        dms, sms = cx.max_over_ranks(t_sharded, e0.elapsed_time(e1) / steps)
        disp.update({"ms_per_step_max_over_ranks": dms, "same_chain_on_one_gpu_whole_image_ms": sms,
                     "rgba_bytes_received_per_rank": ny * nx * 4 * (world - 1) // world,
                     "image_min_max_from_exchange_without_a_pass": memo, "image_min_max": [float(mn), float(mx)]})
        del out1
    base["sharded_display"] = disp
    del res
    g.check()
    for m in msgs:
        print(m, file=sys.stderr, flush=True)
    base["cases"] = entries
    base["parity"] = "ok" if all_ok else "FAILED"
    base["parity_how"] = (f"rows {rows_sample} of every rank's copy of the gathered image vs the oracle on the same rows of the global cube "
                          "(sequential mode bit for bit, compensated mode tier C)")
    # strong-scaling efficiency against the smallest N that ran earlier on this box (the driver runs N = 1, 2, 4, 8 back to back)
    if rank == 0:
        hist_path = os.path.join(os.environ.get("TMPDIR", "/tmp"), "aflak_b200_c5_strong.json")
        try:
            with open(hist_path) as f:
                hist = json.load(f)
        except Exception:
            hist = {}
        hist[str(world)] = {k: v["ms_per_step_max_over_ranks"] for k, v in entries.items()}
        try:
            with open(hist_path, "w") as f:
                json.dump(hist, f)
        except Exception:
            pass
        n0 = min(int(k) for k in hist)
        if n0 < world:
            base["strong_scaling_efficiency"] = {"vs_n_gpus": n0, **{k: hist[str(n0)][k] * n0 / (entries[k]["ms_per_step_max_over_ranks"] * world)
                                                                     for k in entries if k in hist[str(n0)]}}
        else:
            base["strong_scaling_efficiency"] = {"vs_n_gpus": n0, "note": "this is the smallest N that ran on this box so far"}
    base["status"] = "ok"
    cx.barrier()
    g.close()
    del band
    return base, all_ok


_REAL_STDOUT = None


def _quiet_stdout():
    """The contract is ONE JSON line on stdout.  Libraries loaded below (NCCL prints its version banner to stdout when
    NCCL_DEBUG is set in the environment) write to file descriptor 1 directly: point fd 1 at stderr for the whole run and
    keep the real stdout for the final line."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT if _REAL_STDOUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    _quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--gather", default="p2p", choices=["p2p", "nccl"], help="N>1: fused peer-store gather (default) or NCCL all_gather")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--skip", default="", help="comma list of blocks to leave out: c5, configs, e2e, cpu, parity, mc")
    ap.add_argument("--no-cpu", action="store_true", help="same as --skip cpu")
    args = ap.parse_args()
    if args.no_cpu:
        args.skip = (args.skip + ",cpu").strip(",")
    if args.impl == "reference":
        return run_reference(args)
    if args.warmup < 3:
        args.warmup = 3
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
