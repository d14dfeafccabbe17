#!/bin/bash
mkdir -p gpurun_out/final3
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/final3/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final3/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/final3/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/final3/smoke.log
timeout 600 python bench.py > gpurun_out/final3/bench_default.json 2> gpurun_out/final3/bench_default.err
