#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
nvidia-smi topo -m 2>/dev/null | head -14 > gpurun_out/topo_n$N.txt
lscpu | grep -i "numa\|socket\|^CPU(s)" >> gpurun_out/topo_n$N.txt
python -c "import os; print('affinity', len(os.sched_getaffinity(0)), sorted(os.sched_getaffinity(0))[:40])" >> gpurun_out/topo_n$N.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench exit=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_n$N.json")); e=d["e2e"]
print(d["n_gpus"], round(d["value"]), "GB/s dev;", "e2e", round(e["value"],1), "GB/s", round(e["ms_per_step"],1), "ms", d["config"].get("host_affinity"))
PY
cat gpurun_out/topo_n$N.txt | cut -c1-160
