#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/e2e_profile.py > gpurun_out/e2e_profile.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
