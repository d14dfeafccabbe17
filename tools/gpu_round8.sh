#!/bin/bash
# runtime weight-block rows (96 / 64) for umma_moe_kernel: parity at forced sizes + decode bench lines
mkdir -p gpurun_out/bench
for r in 96 64 0; do
  SGLB_WROWS=$r timeout 600 python -m pytest tests/test_gpu_umma.py tests/test_gpu_parity.py tests/test_linear.py -m gpu -x -q -k "bf16 or int8 or shared or linear or graphed" > gpurun_out/pytest_wrows_$r.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_wrows_$r.log
done
for r in 128 96 64 0; do
for cfg in "qwen3 int8 128" "qwen3 bf16 128" "qwen3 int8 16" "qwen3 int8 512"; do
  set -- $cfg
  SGLB_WROWS=$r timeout 300 python bench.py --shape $1 --fmt $2 --tokens $3 --steps 50 --warmup 5 --no-sweep --no-cpu-baseline \
     > gpurun_out/bench/w${r}_${1}_${2}_${3}.json 2> gpurun_out/bench/w${r}_${1}_${2}_${3}.err
done; done
