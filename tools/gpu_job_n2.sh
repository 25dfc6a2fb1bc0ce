#!/bin/bash
# N-GPU job: parity of the fused gather, then bench.py at N with both gather implementations
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/check_gather.py > gpurun_out/check_gather_n$N.log 2>&1; echo "check exit=$?" >> gpurun_out/check_gather_n$N.log
tail -5 gpurun_out/check_gather_n$N.log
for ga in p2p nccl; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 100 --warmup 5 --gather $ga > gpurun_out/bench_n${N}_$ga.json 2> gpurun_out/bench_n${N}_$ga.err; echo "bench $ga exit=$?"
tail -2 gpurun_out/bench_n${N}_$ga.err; cat gpurun_out/bench_n${N}_$ga.json
done
