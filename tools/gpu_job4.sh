#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
for u in 4 8; do for c in 4 6; do
  echo "== unroll $u ctas $c"
  AFK_REDUCE_UNROLL=$u AFK_REDUCE_CTAS=$c python bench.py --steps 100 --warmup 5 --no-cpu --e2e-steps 1 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(d['value'], d['roofline']['achieved'], d['roofline']['kernel_ms'], d['clocks'])
    else: print(l.rstrip())
"
done; done
python tools/e2e_probe.py 2>&1 | tail -12
