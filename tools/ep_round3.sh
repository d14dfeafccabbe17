#!/bin/bash
# 8-GPU sanity: EP parity check + default bench + the DSV3 decode line
N=${1:-8}
mkdir -p gpurun_out/ep$N
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 300 $TR tools/ep_check.py > gpurun_out/ep$N/ep_check.log 2>&1; echo "rc=$?" >> gpurun_out/ep$N/ep_check.log
timeout 200 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-sweep --no-cpu-baseline > gpurun_out/ep$N/bench_default.json 2> gpurun_out/ep$N/bench_default.err
timeout 200 $TR bench.py --gpus $N --shape dsv3 --fmt fp8 --tokens 256 --steps 20 --warmup 5 --no-sweep --no-cpu-baseline > gpurun_out/ep$N/dsv3_fp8_256.json 2> gpurun_out/ep$N/dsv3_fp8_256.err
