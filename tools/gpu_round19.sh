#!/bin/bash
mkdir -p gpurun_out/final2
timeout 900 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/final2/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final2/pytest_gpu.log
timeout 300 python tools/dbg_perm.py > gpurun_out/final2/dbg_perm.txt 2>&1
for cfg in "qwen3 bf16 8192" "qwen3 bf16 4096"; do
  set -- $cfg
  timeout 300 python bench.py --shape $1 --fmt $2 --tokens $3 --steps 30 --warmup 5 --no-sweep --no-cpu-baseline \
     > gpurun_out/final2/${1}_${2}_${3}.json 2> gpurun_out/final2/${1}_${2}_${3}.err
done
