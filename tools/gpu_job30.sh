#!/bin/bash
# round-end style check: the whole GPU suite, smoke, the default bench line and its reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit=$?"
tail -3 gpurun_out/pytest_gpu.log | cut -c1-250
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit=$?"
timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/bench.json")); r=json.load(open("gpurun_out/bench_ref.json"))
print("value", round(d["value"],1), d["unit"], "frac", round(d["roofline"]["frac"],3), "e2e", round(d["e2e"]["value"],1), "launches", d["gpu_launches"], "clocks", d["clocks"])
print("reference", round(r["value"],2), r["unit"], "cores", r["cpu_baseline"]["cores"], "all-cores", round(r["cpu_baseline"]["all_cores"]["value"],1))
PY
