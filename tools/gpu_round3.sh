#!/bin/bash
# dense linear parity + ncu capture of the fp8 tensor-native kernel at prefill (DSV3 M=8192)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_linear.py -m gpu -x -q > gpurun_out/pytest_linear.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_linear.log
SKIP=6 timeout 900 bash tools/ncu_umma.sh fp8 8192 f8_m8192 dsv3 umma_f8 > gpurun_out/ncu_f8_sh.log 2>&1
