#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_spatial.py tests/test_gpu_fullsize.py -m gpu -q -x --timeout 900 > gpurun_out/pytest_sp.log 2>&1; echo "pytest exit=$?"
tail -8 gpurun_out/pytest_sp.log | cut -c1-300
timeout 900 python tools/bench_spatial.py > gpurun_out/spatial.json 2> gpurun_out/spatial.err; echo "spatial bench exit=$?"
grep -o '"config": "[^"]*"\|"ms_median": [0-9.]*' gpurun_out/spatial.err | paste - - | head -30
