mkdir -p gpurun_out/r2h; O=gpurun_out/r2h
nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/gather4_probe tools/gather4_probe.cu && timeout 60 /tmp/gather4_probe > $O/gather4.txt 2>&1; echo "gather4 rc=$?" >> $O/rc.txt
timeout 600 python -m pytest tests/test_gpu_layer.py -q -m gpu > $O/pytest_layer.log 2>&1; echo "pytest_layer rc=$?" >> $O/rc.txt
timeout 1500 python tools/run_reference_scripts.py --bench --out $O/reference_scripts.txt > $O/refscripts.log 2>&1; echo "refscripts rc=$?" >> $O/rc.txt
timeout 1200 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/rc.txt
B="python bench.py --steps 2 --warmup 3 --windows 1 --no-configs --no-sweep --no-cpu-baseline --no-e2e --launch eager"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv $B > $O/ncu_launches.log 2>&1; echo "ncu_launches rc=$?" >> $O/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:umma_f8_kernel -s 6 -c 2 -o $O/f8_m256 $B > $O/ncu_full.log 2>&1; echo "ncu_full rc=$?" >> $O/rc.txt
cat $O/rc.txt; cat $O/gather4.txt; tail -3 $O/pytest_layer.log
