#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_axis0.py tests/test_gpu_ingest.py -m gpu -q -x --timeout 300 > gpurun_out/pytest_a.log 2>&1; echo "pytest exit=$?"
tail -4 gpurun_out/pytest_a.log | cut -c1-250
timeout 900 python tools/bench_axis0.py > gpurun_out/axis0.json 2> gpurun_out/axis0.err; echo "axis0 bench exit=$?"
grep median gpurun_out/axis0.err | cut -c1-210
timeout 1200 python tools/bench_ingest.py > gpurun_out/ingest.json 2> gpurun_out/ingest.err; echo "ingest bench exit=$?"
grep "case" gpurun_out/ingest.err | cut -c1-330
