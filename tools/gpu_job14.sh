#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_axis0.py -m gpu -q -x --timeout 300 > gpurun_out/pytest_axis0.log 2>&1; echo "pytest axis0 exit=$?" >> gpurun_out/pytest_axis0.log
tail -25 gpurun_out/pytest_axis0.log | cut -c1-300
timeout 900 python tools/bench_axis0.py > gpurun_out/axis0.json 2> gpurun_out/axis0.err; echo "axis0 bench exit=$?"
grep median gpurun_out/axis0.err | cut -c1-200
