#!/bin/bash
mkdir -p gpurun_out
python tools/bench_configs.py > gpurun_out/configs.json 2> gpurun_out/configs.err; echo "configs exit=$?"
tail -3 gpurun_out/configs.err | cut -c1-300
python bench.py --steps 100 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit=$?"
cat gpurun_out/bench.json | cut -c1-600
python bench.py --steps 20 --warmup 3 --no-cpu --e2e-steps 1 > gpurun_out/ncu_plain1.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 20 --warmup 3 --no-cpu --e2e-steps 1 > gpurun_out/ncu_launch.log 2>&1
python tools/ncu_target.py > gpurun_out/ncu_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:reduce_kernel -s 2 -c 2 -o gpurun_out/prof_reduce_alma_tuned -f python tools/ncu_target.py > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
