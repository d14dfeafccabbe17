#!/bin/bash
# A/B/C of the decode chain fusion levels 0 / 1 / 2: tools/ab_fusion.sh <tag>  (outputs gpurun_out/<tag>/ab.txt)
O=gpurun_out/$1; mkdir -p $O
run() {  # shape fmt tokens
  for f in "--fusion 0" "--fusion 1" "--fusion 2"; do
    python bench.py --shape $1 --fmt $2 --tokens $3 --steps 30 --warmup 5 --windows 3 --no-sweep --no-configs --no-cpu-baseline --no-e2e $f 2>>$O/ab.err |
      python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
st=d['step_roofline']['stage_us'].get('routed', d['step_roofline']['stage_us'])
print('$1 $2 M=$3 $f', 'ms %.4f' % d['ms_per_step'], 'step %.3f' % d['step_roofline']['frac'], 'kernel %.3f' % d['roofline']['frac'], {k: round(v,1) for k,v in st.items() if isinstance(v,(int,float)) and v>0}, 'launches', d['gpu_launches'])" >> $O/ab.txt
  done
}
run qwen3 int8 1; run qwen3 int8 16; run qwen3 int8 128; run qwen3 int8 1024
run qwen3 bf16 128; run dsv3 fp8 256; run dsv3 fp8 16; run gptoss mxfp4 128; run refbench int8 4
cat $O/ab.txt
