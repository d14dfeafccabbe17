#!/bin/bash
# end-of-round acceptance pass: what the driver runs (GPU tests, smoke, both bench arms) + per-config lines
mkdir -p gpurun_out/final
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/final/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/final/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/final/smoke.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final/bench_reference.json 2> gpurun_out/final/bench_reference.err
timeout 600 python bench.py > gpurun_out/final/bench_default.json 2> gpurun_out/final/bench_default.err
for cfg in "qwen3 bf16 128" "qwen3 bf16 8192" "qwen3 int8 4096" "dsv3 fp8 16" "dsv3 fp8 256" "dsv3 fp8 8192" "gptoss mxfp4 128" "gptoss mxfp4 4096"; do
  set -- $cfg
  timeout 300 python bench.py --shape $1 --fmt $2 --tokens $3 --steps 30 --warmup 5 --no-sweep --no-cpu-baseline \
     > gpurun_out/final/${1}_${2}_${3}.json 2> gpurun_out/final/${1}_${2}_${3}.err
done
