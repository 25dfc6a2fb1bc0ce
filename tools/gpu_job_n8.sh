#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 tools/check_sharded_ingest.py > gpurun_out/sharded_ingest_n$N.log 2>&1; echo "sharded ingest exit=$?"
grep "SHARDED INGEST" gpurun_out/sharded_ingest_n$N.log | cut -c1-900
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 tools/check_gather.py > gpurun_out/check_gather_n$N.log 2>&1; echo "gather exit=$?"
grep "GATHER CHECK" gpurun_out/check_gather_n$N.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench exit=$?"
cut -c1-700 gpurun_out/bench_n$N.json
