#!/bin/bash
# fp8 kernel change: parity (fp8 tests of both GPU suites + linears) and the DSV3 bench lines
mkdir -p gpurun_out/bench
timeout 900 python -m pytest tests/test_gpu_umma.py tests/test_gpu_parity.py tests/test_linear.py -m gpu -x -q -k "fp8 or f8 or linear" > gpurun_out/pytest_f8.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_f8.log
for cfg in "dsv3 fp8 8192" "dsv3 fp8 128" "dsv3 fp8 1024"; do
  set -- $cfg
  timeout 300 python bench.py --shape $1 --fmt $2 --tokens $3 --steps 30 --warmup 5 --no-sweep --no-cpu-baseline \
     > gpurun_out/bench/${1}_${2}_${3}.json 2> gpurun_out/bench/${1}_${2}_${3}.err
done
python tools/sumbench.py gpurun_out/bench/dsv3*.json > gpurun_out/bench_summary_f8.txt 2>&1
