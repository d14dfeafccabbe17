"""Multi-GPU check of the expert-parallel layer over real NCCL: every rank compares EPMoE.forward on its
own tokens with the single-GPU fused_experts over the full expert set (same seeded weights everywhere).
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/ep_check.py"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import sgl_cpu_tests_b200 as K  # noqa: E402
from sgl_cpu_tests_b200 import ep  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ok = True
    for fmt, (E, topk, H, I), Ms in (("bf16", (32, 4, 512, 256), (1, 37, 700)), ("fp8", (16, 4, 1024, 512), (5, 300)),
                                     ("int8", (32, 8, 512, 256), (64,))):
        g = torch.Generator(device=dev).manual_seed(7)       # same weights on every rank
        rnd = lambda *s: torch.randn(*s, device=dev, generator=g)
        if fmt == "bf16":
            w = dict(w1=(rnd(E, 2 * I, H) / 10).bfloat16(), w2=(rnd(E, H, I) / 10).bfloat16(), s1=None, s2=None, block=None)
        elif fmt == "fp8":
            w = dict(w1=(rnd(E, 2 * I, H) * 100).clamp(-400, 400).to(torch.float8_e4m3fn),
                     w2=(rnd(E, H, I) * 100).clamp(-400, 400).to(torch.float8_e4m3fn),
                     s1=rnd(E, 2 * I // 128, H // 128) * 1e-3, s2=rnd(E, H // 128, I // 128) * 1e-3, block=[128, 128])
        else:
            w = dict(w1=torch.randint(-127, 128, (E, 2 * I, H), device=dev, generator=g, dtype=torch.int8),
                     w2=torch.randint(-127, 128, (E, H, I), device=dev, generator=g, dtype=torch.int8),
                     s1=torch.rand(E, 2 * I, device=dev, generator=g) * 1e-2,
                     s2=torch.rand(E, H, device=dev, generator=g) * 1e-2, block=None)
        fid = {"bf16": K.FMT_BF16, "fp8": K.FMT_FP8_W8A16, "int8": K.FMT_INT8_W8A8}[fmt]
        el = E // world
        sl = slice(rank * el, (rank + 1) * el)
        local_w = {k: (v[sl].contiguous() if torch.is_tensor(v) else v) for k, v in w.items()}
        layer = ep.EPMoE(local_w, fid, E, topk)
        for M in Ms:
            gt = torch.Generator(device=dev).manual_seed(100 + rank)   # different tokens per rank
            x = (torch.randn(M, H, device=dev, generator=gt) / 8).bfloat16()
            score = torch.randn(M, E, device=dev, generator=gt)
            tw, ids = torch.topk(torch.softmax(score, -1), topk)
            ids = ids.to(torch.int32)
            ref = K.fused_experts_impl(x, w["w1"], w["w2"], tw, ids, False, fid, w["s1"], w["s2"], w["block"], False)
            outs = {}
            for mode in ("fixed", "exact", "p2p", "p2p", "p2p"):   # p2p thrice: both buffer parities and their reuse
                out = layer.forward(x, tw, ids, mode=mode)
                if mode == "p2p" and layer.last_mode != "p2p":
                    ok = False
                    print(f"[rank {rank}] peer-memory path unavailable", flush=True)
                if mode in outs and mode == "p2p":
                    ok &= torch.equal(out, outs[mode])
                outs[mode] = out
                if mode == "p2p":   # same per-rank partial rows, same rank-order f32 sum as the NCCL path
                    same = torch.equal(out, outs["fixed"])
                    ok &= same
                    if not same and rank == 0:
                        print(f"[rank {rank}] {fmt} M={M}: p2p differs bitwise from the NCCL fixed-capacity path", flush=True)
                d = (out.float() - ref.float()).abs().max().item()
                if fmt == "int8":   # the reference's int8 criterion (test_moe_int8.py:134-137): partial sums cancel
                    good = float((out.float() - ref.float()).abs().mean() / ref.float().abs().mean()) < 0.01
                else:
                    good = torch.allclose(out.float(), ref.float(), rtol=1e-2, atol=1e-2)
                ok &= good
                if rank == 0 or not good:
                    print(f"[rank {rank}/{world}] {fmt} M={M} mode={mode}: max diff {d:.5f} {'ok' if good else 'MISMATCH'}",
                          flush=True)
    t = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("EP CHECK", "PASSED" if int(t.item()) else "FAILED", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(t.item()) else 1)


if __name__ == "__main__":
    main()
