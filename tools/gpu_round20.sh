#!/bin/bash
timeout 300 python tools/dbg_pair_determinism.py > gpurun_out/dbg_pair_release.txt 2>&1
timeout 200 python bench.py --shape qwen3 --fmt bf16 --tokens 8192 --path umma2 --steps 30 --warmup 5 --no-sweep --no-cpu-baseline > gpurun_out/bench_pair_release.json 2> gpurun_out/bench_pair_release.err
