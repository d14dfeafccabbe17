#!/bin/bash
mkdir -p gpurun_out/bench
timeout 600 python -m pytest tests/test_act_norm.py tests/test_gpu_umma.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_r13.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_r13.log
timeout 300 python tools/bench_act_norm.py > gpurun_out/bench_act_norm.txt 2>&1
for cfg in "qwen3 int8 128" "qwen3 bf16 128" "qwen3 int8 64" "qwen3 bf16 256" "qwen3 int8 4096"; do
  set -- $cfg
  timeout 300 python bench.py --shape $1 --fmt $2 --tokens $3 --steps 50 --warmup 5 --no-sweep --no-cpu-baseline \
     > gpurun_out/bench/e_${1}_${2}_${3}.json 2> gpurun_out/bench/e_${1}_${2}_${3}.err
done
