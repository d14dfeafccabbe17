#!/usr/bin/env python
"""One expert-parallel rank's expert phase, emulated on ONE GPU with per-stage CUDA events (sglb_profile_*): the rank
receives world x cap token slots whose ids are -1 except for its own E / world experts (ep.py: PeerExchange.enqueue ->
experts_into) and writes one bf16 row per routed slot.  Shows where the time of the `experts` phase of bench.py
--gpus N goes without paying for N GPUs.

  python tools/ep_rank_profile.py [--world 8] [--cap 256] [--shape dsv3] [--fmt fp8] [--steps 20]"""
import argparse
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import sgl_cpu_tests_b200 as K  # noqa: E402
from sgl_cpu_tests_b200 import _lib as L  # noqa: E402

STAGES = ["align", "quant_x", "w13_gemm", "quant_act", "w2_gemm", "topk_reduce"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--cap", type=int, default=256)
    ap.add_argument("--shape", default="dsv3")
    ap.add_argument("--fmt", default="fp8")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--hint", type=float, default=-1.0)
    a = ap.parse_args()
    E, topk, H, I = bench.SHAPES[a.shape][:4]
    El = E // a.world
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(1)
    w = bench.make_weights(a.fmt, El, H, I, dev, g)
    fid = bench.FMT_ID[a.fmt]
    M = a.world * a.cap
    x = (torch.randn(M, H, device=dev, generator=g) / H ** 0.5).bfloat16()
    ids_g = torch.stack([torch.randperm(E, device=dev, generator=g)[:topk] for _ in range(M)]).int()
    ids = torch.where(ids_g < El, ids_g, torch.full_like(ids_g, -1)).contiguous()   # rank 0 owns experts [0, El)
    tw = torch.rand(M, topk, device=dev, generator=g)
    y = torch.empty(M * topk, H, dtype=torch.bfloat16, device=dev)
    hint = a.hint if a.hint >= 0 else a.cap * topk * a.world / E   # ep.py _hint: expected rows per local expert
    lib = L.load()

    def call():
        K.fused_experts_impl(x, w["w1"], w["w2"], tw, ids, False, fid, w["s1"], w["s2"], w["block"], False, slot_out=y,
                             rows_per_expert_hint=hint)

    for _ in range(3):
        call()
    torch.cuda.synchronize()
    lib.sglb_profile_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms, cnt = (C.c_float * 6)(), (C.c_int * 6)()
    L.check(lib.sglb_profile_read(ms, cnt, 6))
    lib.sglb_profile_enable(0)
    st = {STAGES[i]: round(1e3 * ms[i] / max(cnt[i], 1), 1) for i in range(6) if cnt[i]}
    wbytes = El * 3 * H * I * {"bf16": 2, "int8": 1, "fp8": 1, "mxfp4": 0.5}[a.fmt]
    print(f"{a.shape} {a.fmt} world={a.world} cap={a.cap}: {M} slots x top{topk}, {int((ids >= 0).sum())} routed rows on {El} experts, "
          f"hint {hint:.1f}; eager call {1e3 * e0.elapsed_time(e1) / a.steps:.1f} us; stages {st}; "
          f"weights {wbytes / 1e9:.2f} GB = {wbytes / 6.45e12 * 1e6:.0f} us at 6.45 TB/s")


if __name__ == "__main__":
    main()
