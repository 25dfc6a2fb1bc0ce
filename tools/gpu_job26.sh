#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/ncu_target.py --workload median --launches 2 > gpurun_out/ncu_plain_median.log 2>&1 || { echo "plain failed"; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:median -c 2 -o gpurun_out/prof_median2 -f python tools/ncu_target.py --workload median --launches 2 > gpurun_out/ncu_full_median2.log 2>&1
echo "ncu exit=$?"; tail -1 gpurun_out/ncu_full_median2.log | cut -c1-200
