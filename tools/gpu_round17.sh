#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q --durations=8 > gpurun_out/pytest_fullsize.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fullsize.log
timeout 300 python bench.py --shape qwen3 --fmt bf16 --tokens 8192 --steps 20 --warmup 5 --no-sweep --no-cpu-baseline > gpurun_out/bench_tensorbound.json 2> gpurun_out/bench_tensorbound.err
