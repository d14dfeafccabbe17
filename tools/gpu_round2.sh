#!/bin/bash
# Re-validation pass on a fresh box: TMEM read microbenchmark, parity tests, smoke, default bench,
# the reference arm, and the key per-config bench lines.
mkdir -p gpurun_out/bench
timeout 60 tools/tmem_ld_bw.bin > gpurun_out/tmem_ld_bw.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
for cfg in "qwen3 bf16 128" "qwen3 bf16 8192" "qwen3 int8 4096" "dsv3 fp8 128" "dsv3 fp8 8192" "gptoss mxfp4 128" "gptoss mxfp4 4096"; do
  set -- $cfg
  timeout 300 python bench.py --shape $1 --fmt $2 --tokens $3 --steps 30 --warmup 5 --no-sweep --no-cpu-baseline \
     > gpurun_out/bench/${1}_${2}_${3}.json 2> gpurun_out/bench/${1}_${2}_${3}.err
done
python tools/sumbench.py gpurun_out/bench/*.json > gpurun_out/bench_summary.txt 2>&1
