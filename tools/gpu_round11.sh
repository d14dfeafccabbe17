#!/bin/bash
# final evidence of the round: default bench (plain), its ncu launch list, one --set full capture of the top kernel
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1g_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-sweep --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:umma_moe -s 6 -c 2 -f -o gpurun_out/prof_r1g \
  python bench.py --steps 2 --warmup 3 --no-sweep --no-cpu-baseline --launch eager > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/prof_r1g.ncu-rep --page raw --csv > gpurun_out/r1g_umma_int8_m128_raw.csv 2>/dev/null
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/smoke.log
