#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_act_norm.py -m gpu -x -q > gpurun_out/pytest_act.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_act.log
timeout 300 python tools/bench_act_norm.py > gpurun_out/bench_act_norm.txt 2>&1
