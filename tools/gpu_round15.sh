#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/e2e_profile_mx4.py > gpurun_out/e2e_profile_mx4.txt 2>&1
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "host_graph or graphed" > gpurun_out/pytest_hostgraph.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_hostgraph.log
