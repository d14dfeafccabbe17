#!/bin/bash
# pair fp8 kernel with the pipelined promotion: parity + DSV3 prefill bench on both tcgen05 paths
mkdir -p gpurun_out/bench
timeout 900 python -m pytest tests/test_gpu_umma.py -m gpu -x -q -k "umma2" > gpurun_out/pytest_pair.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_pair.log
for path in umma2 umma; do
for tok in 8192 2048; do
  timeout 300 python bench.py --shape dsv3 --fmt fp8 --tokens $tok --path $path --steps 20 --warmup 5 --no-sweep --no-cpu-baseline \
     > gpurun_out/bench/dsv3_fp8_${tok}_$path.json 2> gpurun_out/bench/dsv3_fp8_${tok}_$path.err
done; done
python tools/sumbench.py gpurun_out/bench/dsv3_fp8_*_umma*.json > gpurun_out/bench_summary_pair.txt 2>&1
