#!/bin/bash
# short N-GPU validation: EP parity check, default bench, one DSV3 line, reference arm
N=${1:-2}
mkdir -p gpurun_out/ep$N
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 400 $TR tools/ep_check.py > gpurun_out/ep$N/ep_check.log 2>&1; echo "rc=$?" >> gpurun_out/ep$N/ep_check.log
timeout 300 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/ep$N/bench_default.json 2> gpurun_out/ep$N/bench_default.err
timeout 300 $TR bench.py --gpus $N --shape dsv3 --fmt fp8 --tokens 256 --steps 20 --warmup 5 --no-sweep --no-cpu-baseline > gpurun_out/ep$N/dsv3_fp8_256.json 2> gpurun_out/ep$N/dsv3_fp8_256.err
timeout 300 $TR bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/ep$N/ref.json 2> gpurun_out/ep$N/ref.err
timeout 300 python -m pytest tests/test_gpu_ep.py -m gpu -x -q > gpurun_out/ep$N/pytest_ep.log 2>&1; echo "rc=$?" >> gpurun_out/ep$N/pytest_ep.log
