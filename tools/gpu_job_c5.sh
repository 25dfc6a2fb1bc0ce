#!/bin/bash
# usage: gpu_job_c5.sh N -- the C5 strong-scaling run (tools/bench_c5.py) on N GPUs of one box
mkdir -p gpurun_out
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 tools/bench_c5.py 2> gpurun_out/c5_n$N.err | sed -n '/^{/,$p' > gpurun_out/c5_n$N.json; echo "c5 exit=${PIPESTATUS[0]}"
grep -E "MISMATCH|Error|error" gpurun_out/c5_n$N.err | head -10
python - <<PY
import json
d=json.load(open("gpurun_out/c5_n$N.json"))
print("parity", d["parity"], "n_gpus", d["n_gpus"])
for e in d["entries"]:
    print(e["case"][:40], e["sum"][:5], round(e["ms_per_step_max_over_ranks"],3), "ms", round(e["aggregate_GBps"]), "GB/s agg", round(e["per_gpu_frac_of_measured_peak"],3))
PY
