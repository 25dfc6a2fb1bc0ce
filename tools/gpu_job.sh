#!/bin/bash
# One parameterised GPU job (replaces the per-call scripts of round 1).  Usage under gpurun:
#   gpurun --timeout 1500 -- 'bash tools/gpu_job.sh tests bench ncu_reduce'
# Steps: tests | bench | bench_ref | ncu_reduce | ncu_launches | ncu_c4 | configs | n2 (multi-rank checks, needs --gpus 2)
mkdir -p gpurun_out
for step in "$@"; do
  case "$step" in
    tests)
      python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest_gpu.log
      tail -5 gpurun_out/pytest_gpu.log ;;
    tests_all)
      python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest_gpu.log
      grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | tail -30 ;;
    bench)
      python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit=$?"
      tail -3 gpurun_out/bench_n1.err; head -c 1500 gpurun_out/bench_n1.json; echo ;;
    bench_ref)
      python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref_n1.json 2> gpurun_out/bench_ref_n1.err; echo "bench_ref exit=$?"
      head -c 600 gpurun_out/bench_ref_n1.json; echo ;;
    ncu_reduce)
      ncu --set full --clock-control none --import-source on -k regex:reduce_kernel -c 3 -o gpurun_out/reduce_alma -f \
          python tools/ncu_target.py --workload alma --launches 4 > gpurun_out/ncu_reduce.log 2>&1; echo "ncu_reduce exit=$?" ;;
    ncu_launches)
      ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv \
          python bench.py --steps 5 --warmup 3 --skip cpu > gpurun_out/ncu_launches.log 2>&1; echo "ncu_launches exit=$?" ;;
    ncu_c4)
      ncu --set full --clock-control none --import-source on -k regex:"colormap_kernel|minmax_kernel|tone_kernel" -c 9 -o gpurun_out/c4_kernels -f \
          python tools/ncu_target.py --workload c4 --launches 2 > gpurun_out/ncu_c4.log 2>&1; echo "ncu_c4 exit=$?" ;;
    configs)
      python tools/bench_configs.py > gpurun_out/configs.json 2> gpurun_out/configs.err; echo "configs exit=$?"; tail -3 gpurun_out/configs.err ;;
    n2|n4|n8)
      n=${step#n}
      python -m pytest tests/test_gpu_multirank.py -m gpu -x -q > gpurun_out/pytest_multirank.log 2>&1; echo "multirank exit=$?"; tail -15 gpurun_out/pytest_multirank.log
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 20 --warmup 5 \
          > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "bench n=$n exit=$?"
      tail -5 gpurun_out/bench_n$n.err; head -c 1200 gpurun_out/bench_n$n.json; echo ;;
    *) echo "unknown step $step" ;;
  esac
done
