#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_axis0.py tests/test_gpu_fullsize.py -m gpu -q -x --timeout 600 -k "median or fullsize or c3" > gpurun_out/pytest_med.log 2>&1; echo "pytest exit=$?"
tail -15 gpurun_out/pytest_med.log | cut -c1-300
timeout 900 python tools/bench_axis0.py > gpurun_out/axis0.json 2> gpurun_out/axis0.err; echo "axis0 bench exit=$?"
grep median gpurun_out/axis0.err | cut -c1-260
