#!/bin/bash
# experiment: persistent grid size vs time for the decode GEMMs (is an SM's rate capped, i.e. does the last wave matter?)
mkdir -p gpurun_out/bench
for g in 148 128 111 96 74; do
for cfg in "qwen3 int8 128" "qwen3 bf16 128"; do
  set -- $cfg
  SGLB_GRID=$g timeout 300 python bench.py --shape $1 --fmt $2 --tokens $3 --steps 50 --warmup 5 --no-sweep --no-cpu-baseline --launch eager \
     > gpurun_out/bench/g${g}_${1}_${2}_${3}.json 2> gpurun_out/bench/g${g}_${1}_${2}_${3}.err
done; done
