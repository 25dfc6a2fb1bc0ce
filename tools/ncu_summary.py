#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box with `ncu -i`) into profiles/<name>.json + a trimmed raw CSV.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/r1_reduce_kernel_alma_ncu_full
"""
import csv
import io
import json
import subprocess
import sys

KEEP = ["Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__bytes_read.sum.per_second", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__occupancy_limit_registers", "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum",
        "l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
        "smsp__inst_executed.sum", "launch__shared_mem_per_block_static", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = [hdr.index(k) for k in KEEP if k in hdr]
    with open(out + "_raw.csv", "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([hdr[i] for i in idx])
        w.writerow([units[i] for i in idx])
        for r in data:
            w.writerow([r[i] for i in idx])

    def val(r, name):
        i = hdr.index(name)
        return float(r[i].replace(",", "")) * SCALE.get(units[i], 1.0)

    n = len(data)
    summ = {
        "report": rep, "kernel": data[0][hdr.index("Kernel Name")], "launches_captured": n,
        "grid": data[0][hdr.index("Grid Size")], "block": data[0][hdr.index("Block Size")],
        "gpu_time_ms": sum(val(r, "gpu__time_duration.sum") for r in data) / n,
        "dram_bytes_read": sum(val(r, "dram__bytes_read.sum") for r in data) / n,
        "dram_bytes_write": sum(val(r, "dram__bytes_write.sum") for r in data) / n,
        "dram_throughput_pct_of_peak": sum(val(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed") for r in data) / n,
        "sm_throughput_pct": sum(val(r, "sm__throughput.avg.pct_of_peak_sustained_elapsed") for r in data) / n,
        "warps_active_pct": sum(val(r, "sm__warps_active.avg.pct_of_peak_sustained_active") for r in data) / n,
        "registers_per_thread": val(data[0], "launch__registers_per_thread"),
        "l2_hit_rate_pct": sum(val(r, "lts__t_sector_hit_rate.pct") for r in data) / n,
        "note": "ncu --set full --clock-control none; per-launch averages; times are under the profiler (not bench values)",
    }
    with open(out + ".json", "w") as f:
        json.dump(summ, f, indent=1)
    print(json.dumps(summ, indent=1))


def main_grouped(rep, out):
    """One summary entry per kernel name (reports that captured several different kernels)."""
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]

    def val(r, name):
        i = hdr.index(name)
        try:
            return float(r[i].replace(",", "")) * SCALE.get(units[i], 1.0)
        except ValueError:
            return float("nan")

    groups = {}
    for r in data:
        groups.setdefault(r[hdr.index("Kernel Name")], []).append(r)
    summ = {"report": rep, "note": "ncu --set full --clock-control none; per-launch averages per kernel; times are under the profiler (not bench values)", "kernels": []}
    for name, rs in groups.items():
        n = len(rs)
        avg = lambda k: sum(val(r, k) for r in rs) / n
        t_ns = avg("gpu__time_duration.sum") * (1e6 if units[hdr.index("gpu__time_duration.sum")] == "ms" else 1e3 if units[hdr.index("gpu__time_duration.sum")] == "us" else 1.0)
        rd, wr = avg("dram__bytes_read.sum"), avg("dram__bytes_write.sum")
        summ["kernels"].append({
            "kernel": name, "launches_captured": n, "grid": rs[0][hdr.index("Grid Size")], "block": rs[0][hdr.index("Block Size")],
            "gpu_time_ms": t_ns / 1e6, "dram_bytes_read": rd, "dram_bytes_write": wr, "dram_GBps": (rd + wr) / t_ns if t_ns else None,
            "dram_throughput_pct_of_peak": avg("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
            "sm_throughput_pct": avg("sm__throughput.avg.pct_of_peak_sustained_elapsed"),
            "warps_active_pct": avg("sm__warps_active.avg.pct_of_peak_sustained_active"),
            "registers_per_thread": val(rs[0], "launch__registers_per_thread"),
            "l2_hit_rate_pct": avg("lts__t_sector_hit_rate.pct"),
            "issue_active_pct": avg("smsp__issue_active.avg.pct_of_peak_sustained_active"),
            "warp_instructions": avg("smsp__inst_executed.sum"),
        })
    with open(out + ".json", "w") as f:
        json.dump(summ, f, indent=1)
    for k in summ["kernels"]:
        print(f"{k['kernel'][:70]:70s} x{k['launches_captured']} {k['gpu_time_ms']:9.3f} ms  rd {k['dram_bytes_read']/1e9:8.3f} GB wr {k['dram_bytes_write']/1e9:7.3f} GB "
              f"dram {k['dram_throughput_pct_of_peak']:5.1f}% sm {k['sm_throughput_pct']:5.1f}% regs {k['registers_per_thread']:.0f}")


if __name__ == "__main__":
    if len(sys.argv) > 3 and sys.argv[3] == "--grouped":
        main_grouped(sys.argv[1], sys.argv[2])
    else:
        main(sys.argv[1], sys.argv[2])
