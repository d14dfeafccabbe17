#!/bin/bash
# Experiments on the block-scaled mxfp4 kernel at decode: shape / tile cap / stage count, then ncu.
mkdir -p gpurun_out/mx4e
B="python bench.py --fmt mxfp4 --steps 30 --warmup 5 --no-sweep --no-cpu-baseline"
run() { tag=$1; shift; "$@" > gpurun_out/mx4e/$tag.json 2> gpurun_out/mx4e/$tag.err; }
run dsv3_mxfp4_128 $B --shape dsv3 --tokens 128
run dsv3_mxfp4_16 $B --shape dsv3 --tokens 16
run qwen3_mxfp4_128 $B --shape qwen3 --tokens 128
for cap in 16 32 64 112; do
  SGLB_MX4_CAP=$cap run gptoss_mxfp4_128_cap$cap $B --shape gptoss --tokens 128
  mv gpurun_out/mx4e/gptoss_mxfp4_128_cap$cap.json gpurun_out/mx4e/cap${cap}_gptoss_mxfp4_128.json
done
for st in 2 3 4; do
  SGLB_MX4_STAGES=$st run st${st}_gptoss_mxfp4_128 $B --shape gptoss --tokens 128
done
python tools/sumbench.py gpurun_out/mx4e/*.json > gpurun_out/mx4e/summary.txt 2>&1
SKIP=6 tools/ncu_umma.sh mxfp4 128 mx4_m128 gptoss umma_mx4 > gpurun_out/mx4e/ncu.log 2>&1
