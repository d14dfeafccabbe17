"""Debug: bench.ep_parity with progress lines."""
import os, sys, types
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
torch.cuda.set_device(dev)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
print(rank, bench.ep_parity(types.SimpleNamespace(), world, rank, dev, verbose=True), flush=True)
dist.barrier(); dist.destroy_process_group()
