#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_display.py tests/test_gpu_elementwise.py tests/test_gpu_graph.py -m gpu -q -x --timeout 600 > gpurun_out/pytest_disp.log 2>&1; echo "pytest exit=$?"
tail -12 gpurun_out/pytest_disp.log | cut -c1-300
timeout 900 python tools/bench_configs.py > gpurun_out/configs.json 2> gpurun_out/configs.err; echo "configs exit=$?"
python - <<'PY'
import json
for e in json.load(open("gpurun_out/configs.json"))["entries"]:
    print(f'{e["config"][:66]:66s} {e["ms_median"]*1e3:8.1f} us single  {e["ms_batched"]*1e3:8.1f} us batched {e["GBps_batched"]:7.0f} GB/s')
PY
