mkdir -p gpurun_out/r2e; O=gpurun_out/r2e
export SGLB_EP_TIMEOUT_S=20
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531"
timeout 300 $TR tools/dbg_ep_parity.py > $O/par.log 2>&1; echo "par rc=$?" >> $O/rc.txt
for f in bf16 int8 fp8; do timeout 300 $TR tools/dbg_graphed_ep.py $f > $O/dbg_$f.log 2>&1; echo "dbg_$f rc=$?" >> $O/rc.txt; done
CUDA_LAUNCH_BLOCKING=1 timeout 300 $TR tools/dbg_ep_parity.py > $O/par_block.log 2>&1; echo "par_block rc=$?" >> $O/rc.txt
cat $O/rc.txt; grep "ep_parity\|Error" $O/par.log | tail -30
