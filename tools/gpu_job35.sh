#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_stats.py -m gpu -q -x --timeout 600 > gpurun_out/pytest_st.log 2>&1; echo "pytest exit=$?"
tail -12 gpurun_out/pytest_st.log | cut -c1-300
timeout 900 python tools/bench_stats.py > gpurun_out/stats.json 2> gpurun_out/stats.err; echo "stats bench exit=$?"
grep -o '"config": "[^"]*"\|"ms_median": [0-9.]*\|"kernel_launches": [0-9]*' gpurun_out/stats.err | paste - - - | head -20
bash tools/gpu_job36.sh | tail -10
