#!/bin/bash
# item-descriptor prefetch in the TMA producer / gather warps: parity + decode bench lines
mkdir -p gpurun_out/bench
timeout 600 python -m pytest tests/test_gpu_umma.py tests/test_gpu_parity.py tests/test_linear.py -m gpu -x -q > gpurun_out/pytest_pref.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_pref.log
for cfg in "qwen3 int8 128" "qwen3 bf16 128" "qwen3 int8 512" "qwen3 int8 4096" "qwen3 bf16 8192" "dsv3 fp8 128"; do
  set -- $cfg
  timeout 300 python bench.py --shape $1 --fmt $2 --tokens $3 --steps 50 --warmup 5 --no-sweep --no-cpu-baseline \
     > gpurun_out/bench/p_${1}_${2}_${3}.json 2> gpurun_out/bench/p_${1}_${2}_${3}.err
done
