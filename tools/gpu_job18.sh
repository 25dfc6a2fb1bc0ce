#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log | cut -c1-250
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit=$?"
cut -c1-1500 gpurun_out/bench.json
for w in variance argmax wave ingest; do
  timeout 300 python tools/ncu_target.py --workload $w --launches 3 > gpurun_out/ncu_plain_$w.log 2>&1 || { echo "plain $w failed"; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -s 1 -c 8 -o gpurun_out/prof_$w -f python tools/ncu_target.py --workload $w --launches 3 > gpurun_out/ncu_full_$w.log 2>&1
  echo "ncu $w exit=$?"; tail -1 gpurun_out/ncu_full_$w.log | cut -c1-200
done
ls -la gpurun_out/*.ncu-rep
