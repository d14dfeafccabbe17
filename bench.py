#!/usr/bin/env python
"""bench.py -- MoE-layer throughput of the fused-experts hot path on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference ...                     # CPU arm: the oracle port on the host cores

Workload (configs[1] of BASELINE.json): Qwen3-30B-A3B MoE layer, int8 w8a8 per-channel weights,
E=128, topk=8, H=2048, I=768, synthetic data generated as bench_moe.py:14-47 does.  One step = one
fused_experts call over one batch of M tokens.  `value` is timed with inputs resident in HBM, `e2e`
goes through torch.ops.sgl_kernel.fused_experts_cpu with HOST buffers (H2D/D2H inside the timed
region; weights stay resident in HBM after their first upload).  With N > 1 the experts are sharded
expert-parallel (sgl_cpu_tests_b200/ep.py) and every rank contributes M tokens (weak scaling).
"""
import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

CFG = dict(model="Qwen3-30B-A3B MoE layer", E=128, topk=8, H=2048, I=768, fmt="int8_w8a8")
STAGES = ["align", "quant_x", "w13_gemm", "quant_act", "w2_gemm", "topk_reduce"]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--tokens", type=int, default=128, help="M, tokens per step and per GPU")
    ap.add_argument("--fmt", default="int8", choices=["bf16", "int8", "fp8", "mxfp4"])
    ap.add_argument("--shape", default="qwen3", choices=["qwen3", "dsv3", "gptoss", "refbench"])
    ap.add_argument("--sweep", default="1,16,128,1024,4096", help="extra M values reported under 'sweep'")
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--path", default="auto", choices=["auto", "stream", "umma", "umma2"], help="GEMM path (auto = by rows/expert)")
    ap.add_argument("--no-mx4-native", action="store_true", help="mxfp4: dequantising kernels instead of the block-scaled tensor cores")
    ap.add_argument("--block-k", type=int, default=128, help="fp8 scale block along K (experiments; the layer's is 128)")
    ap.add_argument("--launch", default="auto", choices=["auto", "eager", "graph"],
                    help="how the value leg launches a step: eager C-ABI calls, or one CUDA-graph replay per step "
                         "(auto = graph on one GPU up to 1024 tokens, where the step is launch bound)")
    ap.add_argument("--copies", type=int, default=2, help="weight copies rotated between steps (bench_moe.py:50-58)")
    return ap.parse_args()


SHAPES = {  # E, topk, H, I
    "qwen3": (128, 8, 2048, 768),
    "dsv3": (256, 8, 7168, 2048),
    "gptoss": (128, 4, 2880, 2880),
    "refbench": (256, 8, 7168, 384),   # bench_moe.py:144-145
}
BLOCK_K = 128
WBYTES = {"bf16": 2.0, "int8": 1.0, "fp8": 1.0, "mxfp4": 0.5}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d.get("bf16_tflops_sustained", d.get("bf16_tflops")), "measured"
    return 6650.0, 1590.0, "fallback"


def make_weights(fmt, E, H, I, device, gen):
    """Synthetic weights in the value ranges of the reference tests (finite outputs)."""
    def rnd(*shape):
        return torch.randn(*shape, device=device, generator=gen)
    if fmt == "bf16":
        return dict(w1=(rnd(E, 2 * I, H) / 10).bfloat16(), w2=(rnd(E, H, I) / 10).bfloat16(), s1=None, s2=None, block=None)
    if fmt == "int8":
        w1 = torch.randint(-127, 128, (E, 2 * I, H), device=device, generator=gen, dtype=torch.int8)
        w2 = torch.randint(-127, 128, (E, H, I), device=device, generator=gen, dtype=torch.int8)
        return dict(w1=w1, w2=w2, s1=torch.rand(E, 2 * I, device=device, generator=gen) * 1e-2,
                    s2=torch.rand(E, H, device=device, generator=gen) * 1e-2, block=None)
    if fmt == "fp8":
        w1 = (rnd(E, 2 * I, H) * 100).clamp(-400, 400).to(torch.float8_e4m3fn)
        w2 = (rnd(E, H, I) * 100).clamp(-400, 400).to(torch.float8_e4m3fn)
        bk = BLOCK_K
        return dict(w1=w1, w2=w2, s1=rnd(E, 2 * I // 128, H // bk) * 1e-3, s2=rnd(E, H // 128, I // bk) * 1e-3,
                    block=[128, bk])
    def nibbles(*shape):  # 64 spare bytes behind the tensor: the 4-bit TMA boxes of a K % 128 != 0 weight read past it
        n = math.prod(shape)
        buf = torch.zeros(n + 64, device=device, dtype=torch.uint8)
        buf[:n] = torch.randint(0, 256, (n,), device=device, generator=gen, dtype=torch.uint8)
        return buf[:n].view(*shape)
    w1 = nibbles(E, 2 * I, H // 2)
    w2 = nibbles(E, H, I // 2)
    s1 = torch.randint(118, 122, (E, 2 * I, H // 32), device=device, generator=gen, dtype=torch.uint8)
    s2 = torch.randint(118, 122, (E, H, I // 32), device=device, generator=gen, dtype=torch.uint8)
    return dict(w1=w1, w2=w2, s1=s1, s2=s2, block=None)


def make_batch(M, E, topk, H, device, gen):
    x = (torch.randn(M, H, device=device, generator=gen) / math.sqrt(H)).bfloat16()
    score = torch.randn(M, E, device=device, generator=gen).bfloat16()
    tw, ids = torch.topk(torch.softmax(score, dim=-1, dtype=torch.float32), topk)   # bench_moe.py:14-17
    return x, tw.contiguous(), ids.to(torch.int32).contiguous()


def algorithmic_bytes(fmt, E, H, I, M, topk, ids):
    """E_act*(3*H*I*b_w + scale bytes) + 2*M*H*2 + M*k*8, with the measured E_act."""
    e_act = int(torch.unique(ids[ids >= 0]).numel())
    bw = WBYTES[fmt]
    scale = {"bf16": 0, "int8": (2 * I + H) * 4, "fp8": 3 * (H // 128) * max(1, I // 128) * 4,
             "mxfp4": 3 * H * I // 32}[fmt]
    per_expert = 3 * H * I * bw + scale
    total = e_act * per_expert + 2 * M * H * 2 + M * topk * 8
    w13 = e_act * (2 * H * I * bw + scale * 2 / 3) + M * H * (1 if fmt == "int8" else 2) + M * topk * I * 2
    w2 = e_act * (H * I * bw + scale / 3) + M * topk * I * (1 if fmt == "int8" else 2) + M * topk * H * 2
    return dict(e_act=e_act, total=total, w13=w13, w2=w2)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: [self.lines.append(l) for l in self.proc.stdout], daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx = float(f[2])
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


FMT_ID = {"bf16": 0, "int8": 1, "fp8": 2, "mxfp4": 4}
PATHS = {"auto": 0, "stream": 1, "umma": 2, "umma2": 3}


def path_used(args, fmt, M, topk, E, H, I):
    """Mirrors the SGLB_PATH_AUTO rule of capi.cu (rows/expert >= 4 and a supported shape)."""
    if args.path != "auto":
        return args.path
    if fmt == "fp8" and H % 128 == 0 and I % 128 == 0:
        return "umma"          # tensor-native fp8 kernel at every M (capi.cu AUTO rule)
    if fmt == "mxfp4" and H % 64 == 0 and I % 64 == 0 and not args.no_mx4_native:
        return "umma"          # block-scaled tensor-native mxfp4 kernel at every M
    if fmt in ("fp8", "mxfp4"):
        ok = H % 64 == 0 and I % 64 == 0
    else:
        eb = {"bf16": 2, "int8": 1}[fmt]
        ok = (H * eb) % 128 == 0 and (I * eb) % 128 == 0
    # (the CTA-pair kernel is explicit-only: capi.cu no longer picks it under AUTO)
    return "umma" if ok and M * topk / E >= 4.0 else "stream"


PATH_DESC = {"stream": "stream (mma.sync weight streaming)", "umma": "umma (tcgen05 + TMEM + TMA)",
             "umma2": "umma2 (tcgen05 cta_group::2, CTA pair per tile)"}
KERNEL_DESC = {("umma2", "fp8"): "umma2_f8_kernel", ("umma2", "bf16"): "umma2_moe_kernel", ("umma2", "int8"): "umma2_moe_kernel", ("umma", "bf16"): "umma_moe_kernel", ("umma", "int8"): "umma_moe_kernel", ("umma", "fp8"): "umma_f8_kernel",
               ("umma", "mxfp4"): "umma_dq_kernel"}


def time_steps(fn, steps, warmup, world):
    """W warm-up steps, then exactly K steps between barrier+synchronize pairs, CUDA events on the
    launching stream; returns the max over ranks of the elapsed ms."""
    import torch.distributed as dist
    for i in range(warmup):
        fn(i)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        fn(warmup + i)
    b.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = torch.tensor([a.elapsed_time(b)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def run_b200(args):
    import sgl_cpu_tests_b200 as K
    from sgl_cpu_tests_b200 import _lib as L
    rank, world, local = dist_env()
    if args.gpus != world and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if args.gpus > 1 and world == 1:
        raise SystemExit("launch multi-GPU runs with torch.distributed.run (one rank per GPU)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    E, topk, H, I = SHAPES[args.shape]
    fmt = args.fmt
    gen = torch.Generator(device=dev)
    gen.manual_seed(1111 + rank)
    M = args.tokens
    hbm_peak, tc_peak, peak_src = peaks()
    lib = L.load()
    lib.sglb_set_mx4_native(0 if args.no_mx4_native else 1)

    if world > 1:
        from sgl_cpu_tests_b200 import ep
        layer_sets = [ep.EPMoE(make_weights(fmt, E // world, H, I, dev, gen), FMT_ID[fmt], E, topk) for _ in range(args.copies)]
    else:
        layer_sets = [make_weights(fmt, E, H, I, dev, gen) for _ in range(args.copies)]

    def use_graph(m):
        return world == 1 and (args.launch == "graph" or (args.launch == "auto" and m <= 1024))

    def make_graph_step(x, tw, ids):
        """One captured call per weight copy; a step is a single graph replay (K.GraphedExperts)."""
        graphs = []
        for w in layer_sets:
            g = K.GraphedExperts(x.shape[0], w["w1"], w["w2"], tw.shape[1], FMT_ID[fmt], w["s1"], w["s2"], w["block"],
                                 False, path=PATHS[args.path])
            g.hidden.copy_(x), g.topk_weights.copy_(tw), g.topk_ids.copy_(ids)
            graphs.append(g)
        torch.cuda.synchronize()

        def step(i):
            g = graphs[i % len(graphs)]
            g.graph.replay()
            return g.out
        step.kernels_per_step = graphs[0].kernels_per_replay
        return step

    def make_step(x, tw, ids):
        outs = [torch.empty_like(x) for _ in range(2)]

        def step(i):
            w = layer_sets[i % len(layer_sets)]
            if world > 1:
                return w.forward(x, tw, ids)
            return K.fused_experts_impl(x, w["w1"], w["w2"], tw, ids, False, FMT_ID[fmt], w["s1"], w["s2"], w["block"], False,
                                        path=PATHS[args.path])
        return step

    x, tw, ids = make_batch(M, E, topk, H, dev, gen)
    eager_step = make_step(x, tw, ids)
    graphed = use_graph(M)
    step = eager_step
    if graphed:
        try:
            step = make_graph_step(x, tw, ids)
        except Exception as e:  # capture refused (driver / allocator state): measure the eager launches instead
            sys.stderr.write(f"[bench] CUDA graph capture failed ({e}); falling back to eager launches\n")
            torch.cuda.synchronize()
            graphed = False

    # ---- headline: device-resident inputs ----
    clocks = ClockSampler(local)
    l0 = K.launch_count()
    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    l_w = K.launch_count()
    clocks.start()
    ms = time_steps(step, args.steps, 0, world)
    launches = K.launch_count() - l_w
    if graphed:  # kernels inside the replayed graphs (counted once, at capture, by the C ABI)
        launches = step.kernels_per_step * args.steps
    # The timed region lasts a few ms, less than nvidia-smi needs to report once: keep the same steps running
    # (untimed, same count on every rank) until the sampler has seen about half a second of this load.
    extra = int(min(20000, max(0.0, 0.6 - ms * 1e-3) / max(ms * 1e-3 / args.steps, 1e-6)))
    for i in range(extra):
        step(i)
    torch.cuda.synchronize()
    clk = clocks.stop()
    clk["sampled_over"] = f"the {args.steps} timed steps + {extra} identical untimed steps (nvidia-smi -lms 100)"
    ms_per_step = ms / args.steps
    tokens_per_s = M * world / (ms_per_step * 1e-3)

    out = {
        "metric": "MoE layer tokens/s (fused_experts, int8 w8a8)" if fmt == "int8" else f"MoE layer tokens/s (fused_experts, {fmt})",
        "value": tokens_per_s, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": {"bf16": "bf16", "int8": "s8", "fp8": "e4m3->bf16", "mxfp4": "e2m1->bf16"}[fmt], "data": "synthetic",
        "config": {"workload": f"{args.shape} MoE layer {fmt}: E={E} topk={topk} H={H} I={I}, M={M} tokens/step/GPU",
                   "tokens_per_step_per_gpu": M, "parallelism": "single GPU" if world == 1 else f"ep{world}",
                   "l2_policy": f"inputs larger than L2: {args.copies} weight copies rotated between steps",
                   "path": PATH_DESC[path_used(args, fmt, M, topk, E, H, I)],
                   "launch": "one CUDA-graph replay per step (K.GraphedExperts)" if graphed else "eager C-ABI calls"},
        "clocks": clk, "gpu_launches": int(launches),
    }
    if world > 1:
        out["config"]["exchange"] = {"p2p": "NVLink peer memory (ep_p2p.cu: stores into the owners' buffers, peer loads back; no NCCL on the data path)",
                                     "fixed": "NCCL all-to-all, fixed capacity", "exact": "NCCL all-to-all, exact splits"}.get(
                                         layer_sets[0].last_mode, str(layer_sets[0].last_mode))
    out["dtype"] = {"fp8": "e4m3 x e4m3(hi+lo) -> f32", "mxfp4": "e2m1 x e4m3(hi+lo) -> f32"}.get(fmt, out["dtype"]) if \
        path_used(args, fmt, M, topk, E, H, I) == "umma" and not args.no_mx4_native else out["dtype"]

    if world == 1:
        # ---- per-stage kernel timing for the roofline (second pass, same steps, CUDA events per stage) ----
        ab = algorithmic_bytes(fmt, E, H, I, M, topk, ids)
        lib.sglb_profile_enable(1)
        for i in range(args.steps):
            eager_step(i)
        torch.cuda.synchronize()
        import ctypes
        msv = (ctypes.c_float * 6)()
        cnt = (ctypes.c_int * 6)()
        L.check(lib.sglb_profile_read(msv, cnt, 6))
        lib.sglb_profile_enable(0)
        stage_us = {STAGES[i]: (msv[i] / cnt[i] * 1e3 if cnt[i] else 0.0) for i in range(6)}
        k_us = stage_us["w13_gemm"]
        achieved = ab["w13"] / (k_us * 1e-6) / 1e9 if k_us > 0 else 0.0
        used = path_used(args, fmt, M, topk, E, H, I)
        kdesc = "umma_mx4_kernel" if (fmt == "mxfp4" and used == "umma" and not args.no_mx4_native) else KERNEL_DESC.get((used, fmt), "stream_gemm_kernel")
        kname = (f"{kdesc}<stage 1> (tcgen05 w13 GEMM + silu_and_mul)" if used.startswith("umma")
                 else "stream_gemm_kernel<stage 1> (mma.sync w13 GEMM + silu_and_mul)")
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            ent = json.load(open(tpath)).get(f"{args.shape}/{fmt}/M={M}/{used}_stage1")
            traffic = ent["bytes"] if ent else None
        out["roofline"] = {"bound": "hbm", "kernel": kname,
                           "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                           "peak_source": peak_src, "traffic": traffic, "kernel_us": k_us,
                           "algorithmic_bytes": ab["w13"], "active_experts": ab["e_act"],
                           "kernel_timing": "second pass over the same steps with per-stage CUDA events on the launching stream"}
        step_gbs = ab["total"] / (ms_per_step * 1e-3) / 1e9
        out["step_roofline"] = {"bound": "hbm", "achieved": step_gbs, "peak": hbm_peak, "unit": "GB/s",
                                "frac": step_gbs / hbm_peak, "algorithmic_bytes": ab["total"], "stage_us": stage_us}
        # Which roof bounds this workload: time of the algorithmic FLOPs at the measured tensor peak against time
        # of the algorithmic bytes at the measured HBM peak.  Prefill batches are tensor bound; the
        # denominator is the sustained cuBLAS bf16 figure for every format (fp8 / mxfp4 issue two e4m3 passes).
        w13_flops = 4.0 * M * topk * H * I
        step_flops = 6.0 * M * topk * H * I
        if w13_flops / (tc_peak * 1e12) > ab["w13"] / (hbm_peak * 1e9):
            tf = w13_flops / (k_us * 1e-6) / 1e12 if k_us > 0 else 0.0
            out["roofline"].update({"bound": "tensor", "achieved": tf, "peak": tc_peak, "unit": "TFLOP/s",
                                    "frac": tf / tc_peak, "algorithmic_flops": w13_flops,
                                    "hbm_gbs": achieved, "peak_kind": "sustained cuBLAS bf16 (MEASURED_PEAKS.json)"})
            stf = step_flops / (ms_per_step * 1e-3) / 1e12
            out["step_roofline"].update({"bound": "tensor", "achieved": stf, "peak": tc_peak, "unit": "TFLOP/s",
                                         "frac": stf / tc_peak, "algorithmic_flops": step_flops, "hbm_gbs": step_gbs})

        # ---- sweep over M ----
        if not args.no_sweep:
            sweep = []
            for m in [int(v) for v in args.sweep.split(",") if v]:
                xs, tws, idss = make_batch(m, E, topk, H, dev, gen)
                st = make_graph_step(xs, tws, idss) if (graphed and use_graph(m)) else make_step(xs, tws, idss)
                t = time_steps(st, 20, 3, 1) / 20
                t_eager = time_steps(make_step(xs, tws, idss), 20, 3, 1) / 20 if (graphed and use_graph(m)) else t
                abm = algorithmic_bytes(fmt, E, H, I, m, topk, idss)
                flops = 6.0 * m * topk * H * I
                sweep.append({"M": m, "ms": t, "tokens_per_s": m / (t * 1e-3), "hbm_frac": abm["total"] / (t * 1e-3) / 1e9 / hbm_peak,
                              "tflops": flops / (t * 1e-3) / 1e12, "tc_frac_bf16_peak": flops / (t * 1e-3) / 1e12 / tc_peak,
                              "active_experts": abm["e_act"], "launch": "graph" if (graphed and use_graph(m)) else "eager",
                              "ms_eager": t_eager})
            out["sweep"] = sweep

    # ---- e2e: the drop-in op with HOST buffers ----
    out["e2e"] = run_e2e(args, K, world, rank, dev, layer_sets, x, tw, ids, M, E, topk, H, I, fmt)

    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(fmt, E, topk, H, I, M, budget_s=20.0)
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def run_e2e(args, K, world, rank, dev, layer_sets, x, tw, ids, M, E, topk, H, I, fmt):
    """Same metric through torch.ops.sgl_kernel.fused_experts_cpu with host tensors: per step the
    activations and routing go H2D from pinned memory and the result comes back D2H into the caller's
    hidden_states (inplace=True).  Host weights are uploaded once (first call) and stay in HBM."""
    sys.path.insert(0, K.DROPIN_PATH)
    import sgl_kernel  # noqa: F401  (registers torch.ops.sgl_kernel.*)
    op = torch.ops.sgl_kernel.fused_experts_cpu
    xh = x.cpu().pin_memory()
    twh, idh = tw.cpu().pin_memory(), ids.cpu().pin_memory()
    steps = max(5, min(args.steps, 50))
    if world > 1:
        layer = layer_sets[0]

        def call():
            return layer.forward_host(xh, twh, idh)
    else:
        w = layer_sets[0]
        w1h, w2h = w["w1"].cpu(), w["w2"].cpu()
        s1h = None if w["s1"] is None else w["s1"].cpu()
        s2h = None if w["s2"] is None else w["s2"].cpu()
        if fmt == "mxfp4":
            def call():
                return op(xh, w1h, w2h, twh, idh, True, 4, s1h, s2h, None, None, None, False)
        else:
            def call():
                return op(xh, w1h, w2h, twh, idh, True, fmt == "int8", fmt == "fp8", s1h, s2h, w["block"], None, None, False)
    for _ in range(3):
        xh.copy_(x)
        call()
    torch.cuda.synchronize()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        call()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], device=dev)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt = float(t.item())
    h2d = xh.numel() * 2 + twh.numel() * 4 + idh.numel() * 4
    d2h = xh.numel() * 2
    return {"value": M * world * steps / dt, "unit": "tokens/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
            "ms_per_step": dt / steps * 1e3, "steps": steps,
            "api": "torch.ops.sgl_kernel.fused_experts_cpu(host tensors, inplace=True); weights HBM-resident after first call"}


def cpu_workload(fmt, E, topk, H, I, m, seed=1111):
    g = torch.Generator().manual_seed(seed)
    w = make_weights(fmt, E, H, I, "cpu", g)
    x, tw, ids = make_batch(m, E, topk, H, "cpu", g)
    return w, x, tw, ids


def cpu_call(fmt, w, x, tw, ids):
    from oracle import moe_oracle as O
    if fmt == "mxfp4":
        E = w["w1"].shape[0]
        return O.fused_experts(x, w["w1"], w["w2"], tw, ids, "mxfp4", w["s1"], w["s2"])
    return O.fused_experts(x, w["w1"], w["w2"], tw, ids, fmt, w["s1"], w["s2"], w["block"])


def cpu_baseline(fmt, E, topk, H, I, M, budget_s):
    """The oracle port (torch-native restatement of the reference's own CPU oracle code) on the host
    cores.  Reported only."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    m = min(M, 16)
    w, x, tw, ids = cpu_workload(fmt, E, topk, H, I, m)
    t0 = time.perf_counter()
    cpu_call(fmt, w, x, tw, ids)
    one = time.perf_counter() - t0
    reps = max(1, min(10, int(budget_s / max(one, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(reps):
        cpu_call(fmt, w, x, tw, ids)
    dt = (time.perf_counter() - t0) / reps
    return {"value": m / dt, "unit": "tokens/s", "cores": cores, "kind": "port",
            "sample": f"{reps} calls of the torch oracle on the first {m} tokens of the workload ({dt * 1e3:.0f} ms/call)"}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  fused_experts_cpu itself is not
    in /root/reference (third-party sgl-kernel, absent), so this times the oracle port: the
    reference's own torch-native oracle code for this path, all host threads."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    E, topk, H, I = SHAPES[args.shape]
    fmt = args.fmt
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    m = min(args.tokens, 16)
    w, x, tw, ids = cpu_workload(fmt, E, topk, H, I, m)
    t0 = time.perf_counter()
    cpu_call(fmt, w, x, tw, ids)
    one = time.perf_counter() - t0
    total = args.steps + args.warmup
    budget = 150.0
    if one * total > budget:   # shrink the per-step sample so the run ends within a few minutes
        m = max(1, int(m * budget / (one * total)))
        w, x, tw, ids = cpu_workload(fmt, E, topk, H, I, m)
    for _ in range(args.warmup):
        cpu_call(fmt, w, x, tw, ids)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_call(fmt, w, x, tw, ids)
    dt = (time.perf_counter() - t0) / args.steps
    val = m / dt
    sample = f"each step = the torch oracle on {m} tokens of the workload (bounded sample)"
    print(json.dumps({
        "impl": "reference", "metric": "MoE layer tokens/s (fused_experts, int8 w8a8)" if fmt == "int8" else f"MoE layer tokens/s (fused_experts, {fmt})",
        "value": val, "unit": "tokens/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (torch CPU)", "data": "synthetic",
        "config": {"workload": f"{args.shape} MoE layer {fmt}: E={E} topk={topk} H={H} I={I}, M={args.tokens} tokens/step/GPU",
                   "sample_tokens_per_step": m},
        "cpu_baseline": {"value": val, "unit": "tokens/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


if __name__ == "__main__":
    a = parse()
    BLOCK_K = a.block_k
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
