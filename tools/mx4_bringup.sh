#!/bin/bash
# Bring-up of the block-scaled mxfp4 kernel: layout test, parity, then bench lines.
mkdir -p gpurun_out/mx4
timeout 300 python -m pytest tests/test_gpu_umma.py -q -x -k "mx4_scale_tiles" > gpurun_out/mx4/tiles.log 2>&1
timeout 600 python -m pytest tests/test_gpu_umma.py -q -k "umma_mxfp4" > gpurun_out/mx4/parity.log 2>&1
if ! grep -q " passed" gpurun_out/mx4/parity.log || grep -q "failed" gpurun_out/mx4/parity.log; then
  SGLB_MX4_TX=16384 timeout 600 python -m pytest tests/test_gpu_umma.py -q -k "umma_mxfp4" > gpurun_out/mx4/parity_tx16k.log 2>&1
fi
for M in 1 16 128 1024 4096; do
  timeout 300 python bench.py --shape gptoss --fmt mxfp4 --tokens $M --steps 30 --warmup 5 --no-sweep --no-cpu-baseline \
     > gpurun_out/mx4/gptoss_mxfp4_$M.json 2> gpurun_out/mx4/gptoss_mxfp4_$M.err
done
python tools/sumbench.py gpurun_out/mx4/gptoss_mxfp4_*.json > gpurun_out/mx4/summary.txt 2>&1
