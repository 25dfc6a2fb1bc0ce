#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_elementwise.py -m gpu -q -x --timeout 600 > gpurun_out/pytest_el.log 2>&1; echo "pytest exit=$?"
tail -12 gpurun_out/pytest_el.log | cut -c1-300
timeout 900 python tools/bench_configs.py > gpurun_out/configs.json 2> gpurun_out/configs.err; echo "configs exit=$?"
grep -o '"config": "[^"]*"\|"ms_median": [0-9.]*' gpurun_out/configs.err | paste - - | grep -i "multiplier\|ratio 4096\|min_max\|chain\|colormap" | head
