mkdir -p gpurun_out/r2d; O=gpurun_out/r2d
export SGLB_EP_TIMEOUT_S=20
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531"
timeout 900 $TR bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench_n2 rc=$?" >> $O/rc.txt
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 --ep-mode exact --no-configs --no-sweep > $O/bench_n2_exact.json 2> $O/bench_n2_exact.err; echo "bench_n2_exact rc=$?" >> $O/rc.txt
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 --shape qwen3 --fmt int8 --tokens 128 --no-configs --no-sweep > $O/bench_n2_qwen.json 2> $O/bench_n2_qwen.err; echo "bench_n2_qwen rc=$?" >> $O/rc.txt
timeout 900 python -m pytest tests/test_gpu_ep.py -x -q -m gpu > $O/pytest_ep.log 2>&1; echo "pytest_ep rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -c 600 $O/bench_n2.err; tail -c 1500 $O/bench_n2.json
