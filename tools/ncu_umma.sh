#!/bin/bash
# ncu capture of the tcgen05 kernels (one stage-1 and one stage-2 launch).
# usage: tools/ncu_umma.sh <fmt> <tokens> <tag> [shape] [kernel regex]
FMT=${1:-bf16}; TOK=${2:-4096}; TAG=${3:-umma}; SHAPE=${4:-qwen3}; KRE=${5:-umma_}
CMD="python bench.py --steps 2 --warmup 3 --no-sweep --no-cpu-baseline --fmt $FMT --tokens $TOK --shape $SHAPE ${EXTRA:-}"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$KRE -s ${SKIP:-6} -c 2 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_$TAG.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_$TAG.log
