#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/ncu_target.py --workload c4 --launches 2 > gpurun_out/ncu_plain_c4.log 2>&1 || { echo "plain c4 failed"; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -c 24 -o gpurun_out/prof_c4 -f python tools/ncu_target.py --workload c4 --launches 2 > gpurun_out/ncu_full_c4.log 2>&1
echo "ncu c4 exit=$?"
