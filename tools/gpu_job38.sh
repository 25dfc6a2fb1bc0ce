#!/bin/bash
mkdir -p gpurun_out
for u in 8 16; do
  AFK_REDUCE_KG_UNROLL=$u timeout 600 python tools/bench_configs.py > gpurun_out/configs_u$u.json 2> gpurun_out/configs_u$u.err; echo "u=$u exit=$?"
  grep -o '"config": "C[12] [^"]*"\|"ms_median": [0-9.]*' gpurun_out/configs_u$u.err | paste - - | head -8
done
AFK_REDUCE_KG_UNROLL=16 timeout 900 python -m pytest tests/test_gpu_reductions.py tests/test_gpu_graph.py -m gpu -q -x --timeout 600 2>&1 | tail -2
