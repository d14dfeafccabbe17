#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python tools/e2e_profile.py > gpurun_out/e2e_profile.txt 2>&1
timeout 600 python bench.py --no-sweep > gpurun_out/bench_r14.json 2> gpurun_out/bench_r14.err
timeout 300 python bench.py --no-sweep --no-cpu-baseline --shape dsv3 --fmt fp8 --tokens 128 > gpurun_out/bench_r14_dsv3.json 2> gpurun_out/bench_r14_dsv3.err
timeout 300 python bench.py --no-sweep --no-cpu-baseline --shape gptoss --fmt mxfp4 --tokens 128 > gpurun_out/bench_r14_mx4.json 2> gpurun_out/bench_r14_mx4.err
