#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log | cut -c1-250
python tools/bench_configs.py > gpurun_out/configs.json 2> gpurun_out/configs.err; echo "configs exit=$?"
AFK_REDUCE_KGROUPS=0 python tools/bench_configs.py > gpurun_out/configs_kg0.json 2> gpurun_out/configs_kg0.err
python - <<'PY'
import json
for f in ('gpurun_out/configs.json','gpurun_out/configs_kg0.json'):
    d=json.load(open(f)); print(f)
    for e in d['entries'][:9]:
        print(f"  {e['config'][:74]:74s} L{e['kernel_launches']:2d} {e['ms_median']*1e3:9.1f} us {e['GBps']:8.1f} GB/s" + (f" | cold {e['cold_l2_ms_median']*1e3:8.1f} us" if 'cold_l2_ms_median' in e else ""))
PY
