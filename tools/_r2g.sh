mkdir -p gpurun_out/r2g; O=gpurun_out/r2g
export SGLB_EP_TIMEOUT_S=20
timeout 1800 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/rc.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531"
timeout 900 $TR bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench_n2 rc=$?" >> $O/rc.txt
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 --ep-mode exact --no-configs --no-sweep > $O/bench_n2_exact.json 2> $O/bench_n2_exact.err; echo "bench_n2_exact rc=$?" >> $O/rc.txt
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 --shape qwen3 --fmt int8 --tokens 128 --no-configs --no-sweep > $O/bench_n2_qwen.json 2> $O/bench_n2_qwen.err; echo "bench_n2_qwen rc=$?" >> $O/rc.txt
timeout 300 $TR bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > $O/bench_n2_ref.json 2> $O/bench_n2_ref.err; echo "bench_n2_ref rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -5 $O/pytest.log; tail -c 400 $O/bench_n2.err; tail -c 2500 $O/bench_n2.json
