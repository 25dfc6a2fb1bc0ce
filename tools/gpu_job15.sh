#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ingest.py -m gpu -q -x --timeout 300 > gpurun_out/pytest_ingest.log 2>&1; echo "pytest ingest exit=$?" >> gpurun_out/pytest_ingest.log
tail -25 gpurun_out/pytest_ingest.log | cut -c1-300
df -h /tmp | tail -1; free -g | head -2; nproc
timeout 1200 python tools/bench_ingest.py > gpurun_out/ingest.json 2> gpurun_out/ingest.err; echo "ingest bench exit=$?"
cut -c1-330 gpurun_out/ingest.err | tail -40
