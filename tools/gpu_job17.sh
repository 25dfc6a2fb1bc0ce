#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_spatial.py -m gpu -q -x --timeout 300 > gpurun_out/pytest_spatial.log 2>&1; echo "pytest spatial exit=$?" >> gpurun_out/pytest_spatial.log
tail -40 gpurun_out/pytest_spatial.log | cut -c1-300
timeout 900 python tools/bench_spatial.py > gpurun_out/spatial.json 2> gpurun_out/spatial.err; echo "spatial bench exit=$?"
cut -c1-260 gpurun_out/spatial.err | tail -20
