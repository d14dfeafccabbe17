"""Debug: GraphedEP's warm-up calls on static buffers, step by step with synchronisation after each kernel."""
import os
import sys
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import sgl_cpu_tests_b200 as K
from sgl_cpu_tests_b200 import ep

rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29544")
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
torch.cuda.set_device(dev)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
fmt = sys.argv[1] if len(sys.argv) > 1 else "fp8"
E, topk, H, I, Is = 8 * world, 4, 1024, 512, 512
g = torch.Generator(device=dev).manual_seed(7)
w = bench.make_weights(fmt, E, H, I, dev, g)
sh = bench.make_shared(fmt, H, Is, dev, g)
fid = bench.FMT_ID[fmt]
el = E // world
local_w = {k: (v[rank * el:(rank + 1) * el].contiguous() if torch.is_tensor(v) else v) for k, v in w.items()}
M = 300
layer = ep.EPMoE(local_w, fid, E, topk, shared=sh, routed_scaling_factor=2.5, path=K.PATH_UMMA)
gt = torch.Generator(device=dev).manual_seed(100 + rank)
x = (torch.randn(M, H, device=dev, generator=gt) / 8).bfloat16()
tw, ids = torch.topk(torch.softmax(torch.randn(M, E, device=dev, generator=gt), -1), topk)
ids = ids.to(torch.int32)
ref = K.fused_experts_impl(x, w["w1"], w["w2"], tw, ids, False, fid, w["s1"], w["s2"], w["block"], False, path=K.PATH_UMMA)
ref = K.shared_expert_impl(x, sh["w1"], sh["w2"], ref, 2.5, False, fid, sh["s1"], sh["s2"], sh["block"], False)
for i in range(3):
    out = layer.forward(x, tw, ids, mode="p2p")
    torch.cuda.synchronize()
    print(rank, "eager p2p", i, torch.equal(out, ref), flush=True)
# static buffers as GraphedEP leaves them: zeros / -1
z = torch.zeros_like(x); mi = torch.full_like(ids, -1); zw = torch.zeros_like(tw)
out = layer.forward(z, zw, mi, mode="p2p")
torch.cuda.synchronize()
print(rank, "eager p2p on empty routing ok", float(out.float().abs().max()), flush=True)
s2 = torch.cuda.Stream(device=dev)
s2.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s2):
    out = layer.forward(x, tw, ids, mode="p2p")
s2.synchronize()
print(rank, "eager p2p on a second stream", torch.equal(out, ref), flush=True)
gph = layer.graphed(M, H, dev)
torch.cuda.synchronize()
print(rank, "graphs captured", flush=True)
for i in range(4):
    out = gph(x, tw, ids).clone()
    torch.cuda.synchronize()
    print(rank, "replay", i, torch.equal(out, ref), flush=True)
layer.close()
dist.barrier()
dist.destroy_process_group()
