#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stat_pass -c 6 -o gpurun_out/prof_stats -f python tools/ncu_stats_target.py > gpurun_out/ncu_stats_full.log 2>&1
echo "ncu exit=$?"; ls -la gpurun_out/prof_stats.ncu-rep
