#!/bin/bash
# N-GPU pass: expert-parallel parity check + bench lines (run with gpurun --gpus N).
N=${1:-2}
mkdir -p gpurun_out/ep$N
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR tools/ep_check.py > gpurun_out/ep$N/ep_check.log 2>&1
timeout 300 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/ep$N/bench_default.json 2> gpurun_out/ep$N/bench_default.err
for cfg in "dsv3 fp8 256" "dsv3 fp8 8192" "qwen3 bf16 4096"; do
  set -- $cfg
  timeout 300 $TR bench.py --gpus $N --shape $1 --fmt $2 --tokens $3 --steps 20 --warmup 5 --no-sweep --no-cpu-baseline \
    > gpurun_out/ep$N/${1}_${2}_${3}.json 2> gpurun_out/ep$N/${1}_${2}_${3}.err
done
timeout 300 $TR bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/ep$N/ref.json 2> gpurun_out/ep$N/ref.err
