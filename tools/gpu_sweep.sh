#!/bin/bash
# sweep of the reduce-kernel tuning knobs on the ALMA-shape workload (bench.py, no CPU leg, 1 e2e step)
mkdir -p gpurun_out
for cfg in "8 4" "8 3" "8 2" "16 2" "16 3" "4 4"; do
  set -- $cfg
  echo "== unroll $1 ctas $2"
  AFK_REDUCE_UNROLL=$1 AFK_REDUCE_CTAS=$2 python bench.py --steps 100 --warmup 5 --no-cpu --e2e-steps 1 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(d['value'], d['roofline']['achieved'], d['roofline']['kernel_ms'], d['clocks'])
    else: print(l.rstrip())
"
done
