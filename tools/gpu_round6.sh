#!/bin/bash
# CUDA-graph launch mode: parity test + default bench (graph) + eager for comparison + decode lines
mkdir -p gpurun_out/bench
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "graphed" > gpurun_out/pytest_graph.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_graph.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_default_graph.json 2> gpurun_out/bench_default_graph.err
timeout 600 python bench.py --no-cpu-baseline --no-sweep --launch eager > gpurun_out/bench_default_eager.json 2> gpurun_out/bench_default_eager.err
for cfg in "dsv3 fp8 1" "dsv3 fp8 16" "dsv3 fp8 256" "gptoss mxfp4 1" "gptoss mxfp4 16" "gptoss mxfp4 128" "qwen3 bf16 128" "qwen3 bf16 16"; do
  set -- $cfg
  timeout 300 python bench.py --shape $1 --fmt $2 --tokens $3 --steps 30 --warmup 5 --no-sweep --no-cpu-baseline \
     > gpurun_out/bench/g_${1}_${2}_${3}.json 2> gpurun_out/bench/g_${1}_${2}_${3}.err
done
