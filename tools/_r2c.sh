mkdir -p gpurun_out/r2c; O=gpurun_out/r2c
nvidia-smi -L > $O/gpus.txt; lscpu | head -20 > $O/cpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/rc.txt
timeout 1500 python tools/run_reference_scripts.py --bench --out $O/reference_scripts.txt > $O/refscripts.log 2>&1; echo "refscripts rc=$?" >> $O/rc.txt
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/rc.txt
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench_ref rc=$?" >> $O/rc.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --windows 1 --no-configs --no-sweep --no-cpu-baseline --no-e2e > $O/ncu_bench.log 2>&1; echo "ncu rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -3 $O/pytest.log; tail -c 1500 $O/bench.json
