#!/bin/bash
# usage: gpu_job21.sh N  -- kernel-change validation on GPU 0, then the C5 strong-scaling run on N GPUs
mkdir -p gpurun_out
N=${1:-2}
timeout 900 python -m pytest tests/test_gpu_elementwise.py tests/test_gpu_display.py tests/test_gpu_graph.py tests/test_gpu_fullsize.py -m gpu -q -x --timeout 600 > gpurun_out/pytest_b.log 2>&1; echo "pytest exit=$?"
tail -3 gpurun_out/pytest_b.log | cut -c1-250
timeout 600 python tools/bench_configs.py > gpurun_out/configs.json 2> gpurun_out/configs.err; echo "configs exit=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 tools/bench_c5.py > gpurun_out/c5_n$N.json 2> gpurun_out/c5_n$N.err; echo "c5 exit=$?"
grep -E "parity|MISMATCH|Error|error" gpurun_out/c5_n$N.json gpurun_out/c5_n$N.err | head -20
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/c5_n$N.json"))
    for e in d["entries"]:
        print(e["case"][:40], e["sum"][:5], round(e["ms_per_step_max_over_ranks"],3), "ms", round(e["aggregate_GBps"]), "GB/s agg", round(e["per_gpu_frac_of_measured_peak"],3))
except Exception as ex:
    print("no c5 json", ex)
PY
