"""Multi-GPU check of the expert-parallel layer over real NCCL / NVLink peer memory: every rank's EPMoE output
(fixed, exact, p2p, captured p2p graphs; ragged token counts incl. a rank without tokens; with and without the shared
expert) must be torch.equal to the single-GPU fused_experts (+ shared_expert) over the full expert set.  The same
check runs inside `bench.py --gpus N` before anything is timed (bench.ep_parity); this is the stand-alone form.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/ep_check.py"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench  # noqa: E402


def main():
    rank, world, local = bench.dist_env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    res = bench.ep_parity(None, world, rank, dev)
    ok = res["p2p_eq_nccl"] and res["eq_1gpu"]
    if rank == 0:
        print(json.dumps(res))
        print("EP CHECK", "PASSED" if ok else "FAILED", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
