#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log | cut -c1-250
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit=$?"
cut -c1-600 gpurun_out/bench.json
timeout 600 python tools/bench_configs.py > gpurun_out/configs.json 2> gpurun_out/configs.err; echo "configs exit=$?"
for w in c4 median; do
  timeout 300 python tools/ncu_target.py --workload $w --launches 2 > gpurun_out/ncu_plain_$w.log 2>&1 || { echo "plain $w failed"; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -c 24 -o gpurun_out/prof_$w -f python tools/ncu_target.py --workload $w --launches 2 > gpurun_out/ncu_full_$w.log 2>&1
  echo "ncu $w exit=$?"; tail -1 gpurun_out/ncu_full_$w.log | cut -c1-200
done
ls -la gpurun_out/*.ncu-rep
