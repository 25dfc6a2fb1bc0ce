#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_axis0.py tests/test_gpu_fullsize.py -m gpu -q -x --timeout 900 > gpurun_out/pytest_med.log 2>&1; echo "pytest exit=$?"
tail -8 gpurun_out/pytest_med.log | cut -c1-300
timeout 900 python tools/bench_axis0.py > gpurun_out/axis0.json 2> gpurun_out/axis0.err; echo "axis0 bench exit=$?"
python - <<'PY'
import json
for e in json.load(open("gpurun_out/axis0.json"))["entries"]:
    if "[0," in e["config"]: print(e["config"][:78], round(e["ms_median"],3), "ms", round(e["GBps"]), "GB/s")
PY
