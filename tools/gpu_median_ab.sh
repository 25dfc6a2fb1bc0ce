#!/bin/bash
# A/B of the median kernels on the ALMA cube (under gpurun).  MEDIAN_SWEEP: space-separated VAR=value settings, each timed
# with tools/bench_median.py and traced (AFK_MEDIAN_TRACE=1).  Output: gpurun_out/median_*.json / .txt
mkdir -p gpurun_out
for kv in ${MEDIAN_SWEEP:-AFK_X=0}; do
  name=$(echo "$kv" | tr -c 'A-Za-z0-9.\n' '_')
  env $(echo "$kv" | tr ',' ' ') python tools/bench_median.py > gpurun_out/median_$name.json 2> gpurun_out/median_$name.err; echo "$kv exit=$?"
  python - "$kv" gpurun_out/median_$name.json <<'PY'
import json, sys
d = json.load(open(sys.argv[2]))
print("   ", {k: (v["ms"], v["parity_sampled_rows"]) for k, v in d["windows"].items()})
PY
  if [ -z "$MEDIAN_NOTRACE" ]; then
    env $(echo "$kv" | tr ',' ' ') AFK_MEDIAN_TRACE=1 python tools/bench_median.py > /dev/null 2> gpurun_out/median_${name}_trace.txt; grep "afk median trace" gpurun_out/median_${name}_trace.txt | head -2
  fi
done
