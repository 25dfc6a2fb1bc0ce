#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_stats.py -m gpu -q -x --timeout 300 > gpurun_out/pytest_stats.log 2>&1; echo "pytest stats exit=$?" >> gpurun_out/pytest_stats.log
tail -25 gpurun_out/pytest_stats.log | cut -c1-300
timeout 900 python tools/bench_stats.py > gpurun_out/stats.json 2> gpurun_out/stats.err; echo "stats bench exit=$?"
cut -c1-330 gpurun_out/stats.err | tail -20
