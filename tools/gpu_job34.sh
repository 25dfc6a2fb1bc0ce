#!/bin/bash
# ncu --set full with the source page of the streamed median kernel (after the plain run has exited 0)
mkdir -p gpurun_out
timeout 300 python tools/ncu_target.py --workload median --launches 2 > gpurun_out/ncu_plain_median.log 2>&1 || { echo "plain failed"; tail -3 gpurun_out/ncu_plain_median.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -c 4 -o gpurun_out/prof_median3 -f python tools/ncu_target.py --workload median --launches 2 > gpurun_out/ncu_full_median.log 2>&1
echo "ncu exit=$?"; tail -1 gpurun_out/ncu_full_median.log | cut -c1-200
ls -la gpurun_out/prof_median3.ncu-rep
