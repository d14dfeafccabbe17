#!/usr/bin/env python
"""bench.py -- MoE-layer throughput of the fused-MoE hot path on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference ...                     # CPU arm: the oracle port on the host cores

Default workload = BASELINE.json configs[2] (the largest single-GPU configuration; with --gpus N it is configs[4]):
DeepSeek-V3 MoE layer, fp8 e4m3 w8a16 block-128 weights, E=256 topk=8 H=7168 I=2048, routed by
biased_grouped_topk (G=8, topk_group=2: test_biased_grouped_topk.py:94-95) with the shared expert
(I=2048, routed_scaling_factor 2.5: test_moe_fp8_ext.py:60-61) inside the step, M = 256 tokens per GPU.
One step = router -> fused_experts -> shared_expert over one batch of M tokens of synthetic data generated as
bench_moe.py:14-47 / the reference tests do.  `value` is timed with inputs resident in HBM (one CUDA-graph replay per
step up to 1024 tokens), `e2e` goes through torch.ops.sgl_kernel.{biased_grouped_topk,fused_experts,shared_expert}_cpu
with HOST buffers (H2D/D2H inside the timed region; weights stay resident in HBM after their first upload).  With
N > 1 the experts are sharded expert-parallel (sgl_cpu_tests_b200/ep.py), every rank brings M tokens (weak
scaling), the shared expert is replicated and overlaps the exchange, and the line carries `ep_parity`: a seeded
layer run through every exchange mode and compared bit for bit with the single-GPU result before anything is timed.
`configs` (one GPU) holds the other BASELINE configurations measured the same way.
"""
import argparse
import json
import math
import os
import platform
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

STAGES = ["align", "quant_x", "w13_gemm", "quant_act", "w2_gemm", "topk_reduce"]
SHAPES = {  # E, topk, H, I
    "qwen3": (128, 8, 2048, 768),
    "dsv3": (256, 8, 7168, 2048),
    "gptoss": (128, 4, 2880, 2880),
    "refbench": (256, 8, 7168, 384),   # bench_moe.py:144-145
}
# router of each layer: (biased, num_expert_group, topk_group, renormalize); shared expert intermediate size; rsf
LAYER = {
    "qwen3": dict(router=(False, 1, 1, True), shared=0, rsf=1.0),       # norm_topk_prob (config.json:23), test_moe.py:61-70
    # test_biased_grouped_topk.py:94-95, test_moe_fp8_ext.py:60-61.  bias_scale: the script draws the correction bias as
    # randn(E) (:55); added to sigmoid scores in (0, 1) that bias decides the routing almost alone and ~40 of the 256
    # experts receive every token (measured).  A trained DeepSeek-V3 bias is a small load-BALANCING term, and SURVEY 8d /
    # BASELINE.md size this configuration with all experts streaming (11.3 GB), so the headline draws randn(E) * 0.02;
    # the script's own generator is measured beside it (`--bias-scale 1`, configs entry "C3, randn bias").
    "dsv3": dict(router=(True, 8, 2, True), shared=2048, rsf=2.5, bias_scale=0.02),
    "gptoss": dict(router=(False, 1, 1, True), shared=0, rsf=1.0),
    "refbench": dict(router=(False, 1, 1, False), shared=0, rsf=1.0),   # bench_moe.py:14-17: softmax + topk
}
BLOCK_K = 128
WBYTES = {"bf16": 2.0, "int8": 1.0, "fp8": 1.0, "mxfp4": 0.5}
FMT_ID = {"bf16": 0, "int8": 1, "fp8": 2, "mxfp4": 4}
PATHS = {"auto": 0, "stream": 1, "umma": 2, "umma2": 3}
DTYPE = {"bf16": "bf16", "int8": "s8 x s8 -> s32", "fp8": "e4m3 x e4m3(hi+lo) -> f32", "mxfp4": "e2m1 x e4m3(hi+lo) -> f32"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--tokens", type=int, default=256, help="M, tokens per step and per GPU")
    ap.add_argument("--fmt", default="fp8", choices=["bf16", "int8", "fp8", "mxfp4"])
    ap.add_argument("--shape", default="dsv3", choices=list(SHAPES))
    ap.add_argument("--sweep", default=None, help="extra M values reported under 'sweep' (default: by shape)")
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the other BASELINE configurations ('configs')")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle check of the timed step ('parity_check')")
    ap.add_argument("--parity-tokens", type=int, default=32, help="tokens of the headline step compared with the oracle")
    ap.add_argument("--no-router", action="store_true", help="routing precomputed outside the step (experiments)")
    ap.add_argument("--no-shared", action="store_true", help="leave the shared expert out of the step (experiments)")
    ap.add_argument("--bias-scale", type=float, default=None, help="std of the router's correction bias (default: by shape)")
    ap.add_argument("--path", default="auto", choices=list(PATHS), help="GEMM path (auto = by rows/expert)")
    ap.add_argument("--no-mx4-native", action="store_true", help="mxfp4: dequantising kernels instead of the block-scaled tensor cores")
    ap.add_argument("--fusion", type=int, default=1, choices=[0, 1, 2],
                    help="decode chain fusion level (include/sglb200.h sglb_set_chain_fusion; 1 = the library default)")
    ap.add_argument("--block-k", type=int, default=128, help="fp8 scale block along K (experiments; the layer's is 128)")
    ap.add_argument("--launch", default="auto", choices=["auto", "eager", "graph"],
                    help="how the value leg launches a step: eager C-ABI calls, or one CUDA-graph replay per step "
                         "(auto = graph up to 1024 tokens, where the step is launch bound)")
    ap.add_argument("--copies", type=int, default=0, help="weight copies rotated between steps (0 = enough to exceed L2)")
    ap.add_argument("--windows", type=int, default=5, help="timed windows of --steps steps; the median is reported")
    ap.add_argument("--ep-mode", default=None, choices=[None, "p2p", "fixed", "exact"], help="N > 1: force an exchange mode")
    return ap.parse_args()


def peaks():
    """(hbm GB/s, {dtype: TFLOP/s}, source): MEASURED_PEAKS.json (driver-written) for HBM and bf16; int8 / fp8 / mxf8f6f4
    dense tensor peaks from profiles/peaks_r2.json (tools/mma_peak.cu, measured on this pool), else null."""
    hbm, bf16, src = 6650.0, 1590.0, "fallback (B200_PROFILING.md)"
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        hbm, bf16, src = d["hbm_gbs"], d.get("bf16_tflops_sustained", d.get("bf16_tflops")), "measured (MEASURED_PEAKS.json)"
    tc = {"bf16": bf16}
    p2 = os.path.join(ROOT, "profiles", "peaks_r2.json")
    if os.path.exists(p2):
        d = json.load(open(p2))
        for k in ("int8", "fp8", "mxf8f6f4"):
            if d.get(k + "_tops_sustained"):
                tc[k] = d[k + "_tops_sustained"]
    return hbm, tc, src


def _chunked(E, fn):
    """Builds a big [E, ...] tensor expert-chunk by expert-chunk so float32 temporaries stay small."""
    step = max(1, min(E, 16))
    return torch.cat([fn(min(step, E - e0)) for e0 in range(0, E, step)], dim=0)


def make_weights(fmt, E, H, I, device, gen):
    """Synthetic weights in the value ranges of the reference tests (finite outputs)."""
    def rnd(*shape):
        return torch.randn(*shape, device=device, generator=gen)
    if fmt == "bf16":
        return dict(w1=_chunked(E, lambda n: (rnd(n, 2 * I, H) / 10).bfloat16()), w2=_chunked(E, lambda n: (rnd(n, H, I) / 10).bfloat16()),
                    s1=None, s2=None, block=None)
    if fmt == "int8":
        w1 = torch.randint(-127, 128, (E, 2 * I, H), device=device, generator=gen, dtype=torch.int8)
        w2 = torch.randint(-127, 128, (E, H, I), device=device, generator=gen, dtype=torch.int8)
        return dict(w1=w1, w2=w2, s1=torch.rand(E, 2 * I, device=device, generator=gen) * 1e-2,
                    s2=torch.rand(E, H, device=device, generator=gen) * 1e-2, block=None)
    if fmt == "fp8":
        f8 = lambda *s: (rnd(*s) * 100).clamp(-400, 400).to(torch.float8_e4m3fn)  # noqa: E731
        bk = BLOCK_K
        return dict(w1=_chunked(E, lambda n: f8(n, 2 * I, H)), w2=_chunked(E, lambda n: f8(n, H, I)),
                    s1=rnd(E, -(-2 * I // 128), H // bk) * 1e-3, s2=rnd(E, H // 128, -(-I // bk)) * 1e-3, block=[128, bk])

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


def make_shared(fmt, H, Is, device, gen):
    w = make_weights(fmt, 1, H, Is, device, gen)
    return {k: (v[0].contiguous() if torch.is_tensor(v) else v) for k, v in w.items()}


def make_batch(M, E, H, device, gen):
    x = (torch.randn(M, H, device=device, generator=gen) / math.sqrt(H)).bfloat16()
    gating = torch.randn(M, E, device=device, generator=gen).bfloat16()     # bench_moe.py:14-15
    return x, gating


def scale_bytes(fmt, H, I):
    return {"bf16": 0, "int8": (2 * I + H) * 4, "fp8": 3 * (H // 128) * max(1, -(-I // 128)) * 4, "mxfp4": 3 * H * I // 32}[fmt]


def algorithmic_bytes(fmt, E, H, I, M, topk, ids, shared_I=0):
    """E_act*(3*H*I*b_w + scale bytes) + 2*M*H*2 + M*k*8 (+ the shared expert's weights), measured E_act."""
    e_act = int(torch.unique(ids[(ids >= 0) & (ids < E)]).numel())
    bw = WBYTES[fmt]
    scale = scale_bytes(fmt, H, I)
    per_expert = 3 * H * I * bw + scale
    shared = (3 * H * shared_I * bw + scale_bytes(fmt, H, shared_I)) if shared_I else 0
    total = e_act * per_expert + shared + 2 * M * H * 2 + M * topk * 8
    w13 = e_act * (2 * H * I * bw + scale * 2 / 3) + M * H * (1 if fmt == "int8" else 2) + M * topk * I * 2
    w2 = e_act * (H * I * bw + scale / 3) + M * topk * I * (1 if fmt == "int8" else 2) + M * topk * H * 2
    return dict(e_act=e_act, total=total, w13=w13, w2=w2, shared=shared)


def flops(M, topk, H, I, shared_I=0):
    return 6.0 * M * topk * H * I + 6.0 * M * H * shared_I


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
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def path_used(args, fmt, M, topk, E, H, I):
    """Mirrors the SGLB_PATH_AUTO rule of capi.cu (rows/expert >= 2 or at most 32 slots, and a supported shape)."""
    if args.path != "auto":
        return args.path
    rpe = M * topk / E
    if rpe >= 128 and (fmt == "bf16" and (H * 2) % 128 == 0 and (I * 2) % 128 == 0
                       or fmt == "fp8" and H % 128 == 0 and I % 128 == 0 and H * I >= 8.0e6):
        return "umma2"         # prefill: CTA-pair kernels over pre-gathered rows (capi.cu auto_path)
    if fmt == "fp8" and H % 128 == 0 and I % 128 == 0:
        return "umma"          # tensor-native fp8 kernel at every M (capi.cu AUTO rule)
    if fmt == "mxfp4" and H % 64 == 0 and I % 64 == 0 and not args.no_mx4_native:
        return "umma"          # block-scaled tensor-native mxfp4 kernel at every M
    if fmt in ("fp8", "mxfp4"):
        ok = H % 64 == 0 and I % 64 == 0
    else:
        eb = {"bf16": 2, "int8": 1}[fmt]
        ok = (H * eb) % 128 == 0 and (I * eb) % 128 == 0
    return "umma" if ok and (M * topk / E >= 2.0 or M * topk <= 32) else "stream"


PATH_DESC = {"stream": "stream (mma.sync weight streaming)", "umma": "umma (tcgen05 + TMEM + TMA)",
             "umma2": "umma2 (tcgen05 cta_group::2, CTA pair per tile)"}
KERNEL_DESC = {("umma2", "fp8"): "umma2_f8_kernel", ("umma2", "bf16"): "umma2_moe_kernel", ("umma2", "int8"): "umma2_moe_kernel",
               ("umma", "bf16"): "umma_moe_kernel", ("umma", "int8"): "umma_moe_kernel", ("umma", "fp8"): "umma_f8_kernel",
               ("umma", "mxfp4"): "umma_mx4_kernel"}


def time_window(fn, steps, world, start=0):
    """Exactly `steps` steps between barrier + synchronize pairs, CUDA events on the launching stream; returns the max
    over ranks of the elapsed ms."""
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        fn(start + i)
    b.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = torch.tensor([a.elapsed_time(b)], device="cuda")
    if world > 1:
        both = [torch.zeros_like(ms) for _ in range(world)]
        dist.all_gather(both, ms)
        per_rank = [float(t.item()) for t in both]
        return max(per_rank), per_rank
    return float(ms.item()), [float(ms.item())]


def time_steps(fn, steps, warmup, world=1, windows=None):
    """W warm-up steps, then `windows` windows of exactly K steps each; returns (median window ms, all windows,
    per-rank ms of the median window), or just the one window's ms when `windows` is not given."""
    for i in range(warmup):
        fn(i)
    if windows is None:
        return time_window(fn, steps, world, warmup)[0]
    res = [time_window(fn, steps, world, warmup + w * steps) for w in range(windows)]
    order = sorted(range(windows), key=lambda i: res[i][0])
    mid = order[len(order) // 2]
    return res[mid][0], [r[0] for r in res], res[mid][1]


# ---------------------------------------------------------------------------------------------------------------
class Workload:
    """One MoE layer configuration on this rank: weights (rotated copies), router, optional shared expert."""

    def __init__(self, args, shape, fmt, M, dev, gen, world=1, rank=0, copies=0, router=True, shared=True):
        import sgl_cpu_tests_b200 as K
        self.K, self.args, self.shape, self.fmt, self.M, self.dev, self.world, self.rank = K, args, shape, fmt, M, dev, world, rank
        self.E, self.topk, self.H, self.I = SHAPES[shape]
        cfg = LAYER[shape]
        self.biased, self.G, self.tg, self.renorm = cfg["router"]
        self.shared_I = cfg["shared"] if shared else 0
        self.rsf = cfg["rsf"] if self.shared_I else 1.0
        self.with_router = router
        e_local = self.E // world
        per_copy = e_local * (3 * self.H * self.I * WBYTES[fmt] + scale_bytes(fmt, self.H, self.I))
        l2 = torch.cuda.get_device_properties(dev).L2_cache_size
        self.copies = copies if copies > 0 else max(2, int(math.ceil(2.0 * l2 / max(per_copy, 1))) + 1)
        self.copies = min(self.copies, 8)
        self.bytes_per_copy = per_copy
        self.l2_bytes = l2
        self.weights = [make_weights(fmt, e_local, self.H, self.I, dev, gen) for _ in range(self.copies)]
        self.shared = [make_shared(fmt, self.H, self.shared_I, dev, gen) for _ in range(self.copies)] if self.shared_I else None
        self.bias_scale = args.bias_scale if getattr(args, "bias_scale", None) is not None else cfg.get("bias_scale", 1.0)
        self.bias = (torch.randn(self.E, device=dev, generator=gen) * self.bias_scale).bfloat16() if self.biased else None
        self.router = K.Router(self.topk, self.renorm, self.G, self.tg, self.bias)
        self.x, self.gating = make_batch(M, self.E, self.H, dev, gen)
        self.tw, self.ids = self.router(self.gating)
        self.path = PATHS[args.path]
        if world == 1:
            self.layers = [K.MoELayer(self.weights[c], FMT_ID[fmt], self.router, self.shared[c] if self.shared else None,
                                      self.rsf, self.path) for c in range(self.copies)]
        else:
            from sgl_cpu_tests_b200 import ep
            cap = max(16, 1 << max(0, M - 1).bit_length())
            self.layers = [ep.EPMoE(self.weights[c], FMT_ID[fmt], self.E, self.topk, capacity=cap,
                                    shared=self.shared[c] if self.shared else None, routed_scaling_factor=self.rsf,
                                    path=self.path) for c in range(self.copies)]

    def l2_policy(self):
        ws = self.copies * self.bytes_per_copy
        return (f"{self.copies} weight copies rotated between steps: {ws / 1e6:.0f} MB of weights per rank against "
                f"{self.l2_bytes / 1e6:.0f} MB of L2" + ("" if ws > 2 * self.l2_bytes else " (working set NOT above 2 x L2)"))

    def eager_step(self):
        K = self.K
        if self.world == 1:
            def step(i):
                layer = self.layers[i % self.copies]
                if self.with_router:
                    return layer.forward(self.x, self.gating)
                return layer.add_shared(self.x, layer.experts(self.x, self.tw, self.ids))
        else:
            def step(i):
                layer = self.layers[i % self.copies]
                if self.with_router:
                    tw, ids = self.router(self.gating)
                else:
                    tw, ids = self.tw, self.ids
                return layer.forward(self.x, tw, ids, mode=self.args.ep_mode)
        step.kernels_per_step = None
        return step

    def graph_step(self):
        graphs = []
        if self.world == 1:
            for layer in self.layers:
                g = layer.graphed(self.M)
                g.hidden.copy_(self.x), g.gating.copy_(self.gating)
                graphs.append(g)
        else:
            for layer in self.layers:
                g = layer.graphed(self.M, self.H, self.dev, router=self.router if self.with_router else None)
                if g is None:
                    raise RuntimeError("peer-memory exchange unavailable")
                g.hidden.copy_(self.x)
                if self.with_router:
                    g.gating.copy_(self.gating)
                else:
                    g.topk_weights.copy_(self.tw), g.topk_ids.copy_(self.ids)
                graphs.append(g)
        torch.cuda.synchronize()

        def step(i):
            g = graphs[i % len(graphs)]
            g()
            return g.out
        step.kernels_per_step = graphs[0].kernels_per_replay
        step.graphs = graphs
        return step

    def use_graph(self):
        a = self.args
        if a.launch == "eager" or (self.world > 1 and a.ep_mode not in (None, "p2p")):
            return False
        return a.launch == "graph" or self.M <= 1024

    def make_step(self):
        if self.use_graph():
            try:
                return self.graph_step(), True
            except Exception as e:  # capture refused (driver / allocator state, no peer access): eager launches instead
                sys.stderr.write(f"[bench] CUDA graph capture failed ({type(e).__name__}: {e}); falling back to eager launches\n")
                torch.cuda.synchronize()
        return self.eager_step(), False

    def stage_profile(self, steps):
        """Per-stage CUDA events recorded inside the C ABI around the routed experts' kernels (sglb_profile_enable)."""
        import ctypes
        from sgl_cpu_tests_b200 import _lib as L
        lib = L.load()
        layer = self.layers[0]
        rides = self.world == 1 and bool(self.shared_I) and layer.rides(self.M)   # shared expert inside the routed launches
        out = {"shared_rides": rides}
        for what in ("routed", "shared"):
            if what == "shared" and (not self.shared_I or rides):
                continue
            lib.sglb_profile_enable(1)
            for i in range(steps):
                lay = self.layers[i % self.copies]
                if what == "routed":
                    routed = lay.experts(self.x, self.tw, self.ids, with_shared=rides)
                else:
                    lay.add_shared(self.x, routed)
            torch.cuda.synchronize()
            msv, cnt = (ctypes.c_float * 6)(), (ctypes.c_int * 6)()
            L.check(lib.sglb_profile_read(msv, cnt, 6))
            lib.sglb_profile_enable(0)
            out[what] = {STAGES[i]: (msv[i] / cnt[i] * 1e3 if cnt[i] else 0.0) for i in range(6)}
        del layer
        return out


def roofline_of(wl, ms_per_step, stage_us, hbm_peak, tc, peak_src, traffic_key=None):
    """roofline (dominant kernel = routed w13 GEMM) and step_roofline objects of one workload."""
    fmt, E, H, I, M, topk = wl.fmt, wl.E, wl.H, wl.I, wl.M, wl.topk
    ab = algorithmic_bytes(fmt, E, H, I, M, topk, wl.ids, wl.shared_I)
    k_us = stage_us["routed"]["w13_gemm"]
    rides = bool(stage_us.get("shared_rides"))
    if rides:   # the shared expert's w13 tiles stream inside the same launch: its 2*H*I weight bytes and M more act rows
        ab["w13"] += ab["shared"] * 2 / 3 + M * wl.shared_I * 2
    gbs = ab["w13"] / (k_us * 1e-6) / 1e9 if k_us > 0 else 0.0
    used = path_used(wl.args, fmt, M, topk, E, H, I)
    kdesc = KERNEL_DESC.get((used, fmt), "stream_gemm_kernel")
    if fmt == "mxfp4" and wl.args.no_mx4_native:
        kdesc = "umma_dq_kernel"
    kname = (f"{kdesc}<stage 1> (tcgen05 w13 GEMM + silu_and_mul" + (", shared expert's tiles riding along)" if rides else ")")
             if used.startswith("umma") else "stream_gemm_kernel<stage 1> (mma.sync w13 GEMM + silu_and_mul)")
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        ent = json.load(open(tpath)).get(traffic_key or f"{wl.shape}/{fmt}/M={M}/{used}_stage1")
        traffic = ent["bytes"] if ent else None
    roof = {"bound": "hbm", "kernel": kname, "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak,
            "peak_source": peak_src, "traffic": traffic, "kernel_us": k_us, "algorithmic_bytes": ab["w13"],
            "active_experts": ab["e_act"],
            "kernel_timing": "pass over the same steps with per-stage CUDA events on the launching stream (C ABI sglb_profile_*)"}
    if traffic is None:
        roof["traffic_note"] = "no ncu --set full capture of this kernel at this configuration under profiles/"
    step_gbs = ab["total"] / (ms_per_step * 1e-3) / 1e9
    step = {"bound": "hbm", "achieved": step_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": step_gbs / hbm_peak,
            "algorithmic_bytes": ab["total"], "stage_us": stage_us}
    # Which roof bounds the workload: time of the algorithmic FLOPs at the tensor peak against time of the algorithmic
    # bytes at the HBM peak.  The tensor roof is the one of the arithmetic the reference performs: s8 x s8
    # for int8 w8a8 (test_moe_int8.py:37-56), bf16 for bf16 and for the weight-only formats (fp8 w8a16 / mxfp4 are
    # dequantised to bf16 there); the fp8 / mxfp4 kernels issue two e4m3 passes per bf16-equivalent FLOP, reported as
    # `pipe` against the measured f8f6f4 peak.
    roof_dt = "int8" if fmt == "int8" else "bf16"
    tc_peak = tc.get(roof_dt)
    w13_flops = 4.0 * M * topk * H * I
    step_flops = flops(M, topk, H, I, wl.shared_I)
    if tc_peak and w13_flops / (tc_peak * 1e12) > ab["w13"] / (hbm_peak * 1e9):
        tf = w13_flops / (k_us * 1e-6) / 1e12 if k_us > 0 else 0.0
        kind = ("measured dense s8 tcgen05 peak (profiles/peaks_r2.json)" if roof_dt == "int8"
                else "sustained cuBLAS bf16 (MEASURED_PEAKS.json)")
        roof.update({"bound": "tensor", "achieved": tf, "peak": tc_peak, "unit": "TFLOP/s" if roof_dt == "bf16" else "TOP/s",
                     "frac": tf / tc_peak, "algorithmic_flops": w13_flops, "hbm_gbs": gbs, "peak_kind": kind})
        if fmt in ("fp8", "mxfp4"):
            pk = tc.get("fp8" if fmt == "fp8" else "mxf8f6f4")
            roof["pipe"] = {"issued_tflops": 2 * tf, "peak": pk, "frac": (2 * tf / pk) if pk else None,
                            "note": "hi + lo e4m3 activation pieces: two f8f6f4 MMAs per bf16-equivalent FLOP"}
        stf = step_flops / (ms_per_step * 1e-3) / 1e12
        step.update({"bound": "tensor", "achieved": stf, "peak": tc_peak, "unit": roof["unit"], "frac": stf / tc_peak,
                     "algorithmic_flops": step_flops, "hbm_gbs": step_gbs})
    elif fmt == "int8" and not tc_peak:
        roof["tensor_peak_note"] = "int8 tensor peak not measured on this pool (profiles/peaks_r2.json absent)"
    return roof, step


def step_parity(wl, step, tokens):
    """Checks the step that is about to be timed -- same layer objects, same launch mode (graph replay or eager) --
    against the CPU oracle: router ids on EVERY token (torch.equal; weights bit-equal for the biased router, 1e-5 for
    the softmax one), then the layer output of `tokens` evenly spaced tokens against the oracle restricted to the
    experts those tokens use (the reference's tolerance, rtol = atol = 1e-2; int8: mean relative error < 1e-2 as in
    tests/test_gpu_parity.py).  The oracle is used here as the checker only; nothing of it is timed or shipped.
    Sequence matched: test_biased_grouped_topk.py:67-80 -> test_moe_fp8_ext.py:118 -> :60-61."""
    from oracle import moe_oracle as O
    t0 = time.time()
    out = step(0).clone()           # step 0 runs weight copy 0
    torch.cuda.synchronize()
    res = {"tokens": 0, "launch_checked": "graph" if hasattr(step, "graphs") else "eager"}
    gating = wl.gating.cpu()
    if wl.with_router:
        if wl.biased:
            tw_o, ids_o = O.biased_grouped_topk(gating, wl.bias.cpu(), wl.topk, wl.renorm, wl.G, wl.tg)
        else:
            tw_o, ids_o = O.grouped_topk(gating, wl.topk, wl.renorm, wl.G, wl.tg)
        tw_g, ids_g = wl.router(wl.gating)
        res["ids_equal"] = bool(torch.equal(ids_g.cpu(), ids_o))
        res["weights_equal"] = bool(torch.equal(tw_g.cpu(), tw_o) if wl.biased else
                                    torch.allclose(tw_g.cpu(), tw_o, rtol=1e-5, atol=1e-7))
        res["router_tokens"] = int(gating.shape[0])
    else:
        tw_o, ids_o = wl.tw.cpu(), wl.ids.cpu()
        res["ids_equal"] = res["weights_equal"] = None
    n = min(tokens, wl.M)
    idx = torch.linspace(0, wl.M - 1, n).round().long().unique()
    ids_s, tw_s = ids_o[idx], tw_o[idx]
    used = torch.unique(ids_s[ids_s >= 0]).long()
    remap = torch.full((wl.E,), -1, dtype=torch.int32)
    remap[used] = torch.arange(used.numel(), dtype=torch.int32)
    ids_r = torch.where(ids_s >= 0, remap[ids_s.clamp_min(0).long()], ids_s)
    w = wl.weights[0]
    du = used.to(wl.dev)
    sub = {k: (w[k][du].cpu() if torch.is_tensor(w[k]) else w[k]) for k in ("w1", "w2", "s1", "s2", "block")}
    x = wl.x[idx.to(wl.dev)].cpu()
    ref = O.fused_experts_lean(x, sub["w1"], sub["w2"], tw_s, ids_r, wl.fmt, sub["s1"], sub["s2"], sub["block"])
    if wl.shared_I:
        sh = {k: (v.cpu() if torch.is_tensor(v) else v) for k, v in wl.shared[0].items()}
        ref = O.shared_expert(x, sh["w1"], sh["w2"], ref, wl.rsf, wl.fmt, sh["s1"], sh["s2"], sh["block"])
    got, ref = out[idx.to(wl.dev)].float().cpu(), ref.float()
    res["tokens"] = int(idx.numel())
    res["experts_checked"] = int(used.numel())
    res["max_diff"] = float((got - ref).abs().max())
    res["mean_rel"] = float((got - ref).abs().mean() / ref.abs().mean().clamp_min(1e-30))
    if wl.fmt == "int8":
        res["criterion"] = "mean relative error < 1e-2 (tests/test_gpu_parity.py, int8)"
        res["outputs_ok"] = bool(torch.isfinite(got).all()) and res["mean_rel"] < 1e-2
    else:
        res["criterion"] = "allclose rtol = atol = 1e-2 (utils.py:3-25 of the reference)"
        res["outputs_ok"] = bool(torch.isfinite(got).all() and torch.allclose(got, ref, rtol=1e-2, atol=1e-2))
    if wl.fmt == "mxfp4":
        res["note"] = "parity unpinned: the reference has no fused-mxfp4 call site, the oracle is composed from test_mxfp4.py:65-128"
    res["ok"] = bool(res["outputs_ok"] and res["ids_equal"] is not False and res["weights_equal"] is not False)
    res["seconds"] = round(time.time() - t0, 2)
    return res


def measure_config(args, name, shape, fmt, M, dev, gen, hbm_peak, tc, peak_src, steps=10, e2e=True):
    """One entry of `configs`: the layer of `shape` in `fmt` at M tokens, measured like the headline."""
    wl = Workload(args, shape, fmt, M, dev, gen, router=not args.no_router, shared=not args.no_shared)
    step, graphed = wl.make_step()
    par = None
    if not args.no_parity:
        try:
            par = step_parity(wl, step, 4)
        except Exception as e:  # noqa: BLE001
            par = {"ok": False, "error": f"{type(e).__name__}: {e}"}
    ms = time_steps(step, steps, 3) / steps
    prof = wl.stage_profile(max(3, steps // 2))
    roof, srf = roofline_of(wl, ms, prof, hbm_peak, tc, peak_src)
    ent = {"name": name, "workload": describe(wl), "M": M, "ms": ms, "tokens_per_s": M / (ms * 1e-3),
           "launch": "graph" if graphed else "eager",
           "roofline": {k: roof[k] for k in ("bound", "frac", "kernel_us", "achieved", "peak", "unit", "kernel")},
           "step_roofline": {k: srf[k] for k in ("bound", "frac", "achieved", "peak", "unit")},
           "active_experts": roof["active_experts"]}
    if fmt == "mxfp4":
        ent["parity"] = "unpinned (the reference has no fused-mxfp4 call site; composed oracle)"
    if par is not None:
        ent["parity_check"] = par
    if e2e and not args.no_e2e:
        ent["e2e"] = run_e2e(args, wl, 1, 0, min(steps, 10))
    del step, wl
    torch.cuda.empty_cache()
    return ent


def describe(wl):
    r = ("biased_grouped_topk" if wl.biased else "grouped_topk") + f"(G={wl.G}, topk_group={wl.tg}, renormalize={wl.renorm})"
    if wl.biased:
        r += f" with correction bias ~ {getattr(wl, 'bias_scale', 1.0):g} * randn"
    s = f" + shared_expert(I={wl.shared_I}, rsf={wl.rsf})" if wl.shared_I else ""
    return (f"{wl.shape} MoE layer {wl.fmt}: E={wl.E} topk={wl.topk} H={wl.H} I={wl.I}, "
            f"{r if wl.with_router else 'routing precomputed'}{s}, M={wl.M} tokens/step/GPU")


# ---------------------------------------------------------------------------------------------------------------
def ep_parity(args, world, rank, dev, verbose=False):
    """Before anything is timed at N > 1: a small seeded layer through every exchange mode (fixed, exact, p2p, and the
    captured p2p graphs) on ragged token counts incl. a rank without tokens, compared BIT FOR BIT with the single-GPU
    fused_experts (+ shared_expert) over the full expert set on the same tokens.  Any mismatch fails the run."""
    import torch.distributed as dist
    import sgl_cpu_tests_b200 as K
    from sgl_cpu_tests_b200 import ep
    res = {"p2p_eq_nccl": True, "eq_1gpu": True, "max_diff": 0.0, "cases": 0, "modes": []}
    cases = [("fp8", (8 * world, 4, 1024, 512), 512), ("bf16", (4 * world, 4, 512, 256), 0), ("int8", (4 * world, 8, 512, 256), 256)]
    for fmt, (E, topk, H, I), Is in cases:
        g = torch.Generator(device=dev).manual_seed(7)       # same weights on every rank
        w = make_weights(fmt, E, H, I, dev, g)
        sh = make_shared(fmt, H, Is, dev, g) if Is else None
        fid = FMT_ID[fmt]
        el = E // world
        sl = slice(rank * el, (rank + 1) * el)
        local_w = {k: (v[sl].contiguous() if torch.is_tensor(v) else v) for k, v in w.items()}
        for Ms in ([37 + 3 * r for r in range(world)], [0 if r == world - 1 else 5 + r for r in range(world)], [300] * world):
            M = Ms[rank]
            layer = ep.EPMoE(local_w, fid, E, topk, shared=sh, routed_scaling_factor=2.5, path=K.PATH_UMMA)
            gt = torch.Generator(device=dev).manual_seed(100 + rank)   # different tokens per rank
            x = (torch.randn(M, H, device=dev, generator=gt) / 8).bfloat16()
            tw, ids = torch.topk(torch.softmax(torch.randn(M, E, device=dev, generator=gt), -1), topk)
            ids = ids.to(torch.int32)
            if M:
                ids[0, 0] = -1
            ref = K.fused_experts_impl(x, w["w1"], w["w2"], tw, ids, False, fid, w["s1"], w["s2"], w["block"], False, path=K.PATH_UMMA)
            if sh is not None and M:
                ref = K.shared_expert_impl(x, sh["w1"], sh["w2"], ref, 2.5, False, fid, sh["s1"], sh["s2"], sh["block"], False)
            outs = {}
            for mode in ("fixed", "exact", "p2p", "p2p", "p2p"):   # p2p thrice: both buffer parities and their reuse
                out = layer.forward(x, tw, ids, mode=mode)
                if layer.last_mode != mode:
                    res["eq_1gpu"] = False
                    res.setdefault("errors", []).append(f"rank {rank}: mode {mode} fell back to {layer.last_mode}")
                if mode == "p2p" and "fixed" in outs:
                    res["p2p_eq_nccl"] &= bool(torch.equal(out, outs["fixed"]))
                outs[mode] = out
                same = bool(torch.equal(out, ref))
                res["eq_1gpu"] &= same
                if M:
                    res["max_diff"] = max(res["max_diff"], float((out.float() - ref.float()).abs().max()))
            if verbose:
                torch.cuda.synchronize()
                print(f"[ep_parity] rank {rank} {fmt} Ms={Ms}: eager modes done, eq_1gpu={res['eq_1gpu']}", file=sys.stderr, flush=True)
            if max(Ms) == min(Ms):      # equal token counts: the captured graphs as well (fixed shape by construction)
                gph = layer.graphed(M, H, dev)
                if gph is None:
                    res["eq_1gpu"] = False
                else:
                    for _ in range(3):
                        out = gph(x, tw, ids).clone()
                        res["eq_1gpu"] &= bool(torch.equal(out, ref))
            res["cases"] += 1
            layer.close()
            torch.cuda.synchronize()
            dist.barrier()
    res["modes"] = ["fixed", "exact", "p2p", "p2p graph"]
    flags = torch.tensor([int(res["p2p_eq_nccl"]), int(res["eq_1gpu"])], device=dev)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    md = torch.tensor([res["max_diff"]], device=dev)
    dist.all_reduce(md, op=dist.ReduceOp.MAX)
    res["p2p_eq_nccl"], res["eq_1gpu"], res["max_diff"] = bool(flags[0].item()), bool(flags[1].item()), float(md.item())
    res["what"] = ("seeded fp8(+shared) / bf16 / int8(+shared) layers, ragged token counts incl. a rank with 0 tokens, every rank "
                   "torch.equal to the single-GPU fused_experts (+ shared_expert) with the same GEMM path")
    return res


def ep_phases(wl, steps):
    """Per-phase microseconds of the eager peer-memory call (CUDA events between the phases), median over steps."""
    from sgl_cpu_tests_b200 import _lib as L, ops
    layer = wl.layers[0]
    cap = layer._agreed_capacity(wl.M, wl.dev, None)
    px = layer._peer_exchange(cap, wl.H, wl.dev)
    if px is None:
        return None
    lib = L.load()
    B = layer.backend
    names = ["router", "dispatch", "wait", "experts", "combine"]
    samples = {n: [] for n in names}
    main = torch.cuda.current_stream()
    for it in range(steps + 2):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
        px.calls += 1
        par = px.calls & 1
        ev[0].record()
        tw, ids = wl.router(wl.gating)
        ev[1].record()
        s = main.cuda_stream
        L.check(lib.sglb_ep_p2p_dispatch(ops._ptr(wl.x), ops._ptr(ids), ops._ptr(tw), wl.M, wl.topk, wl.E, wl.world, wl.rank, wl.H,
                                         cap, px.peers, par, s))
        shared_y = None
        if B.shared is not None:
            px.side.wait_stream(main)
            with torch.cuda.stream(px.side):
                shared_y = B.shared_mlp(wl.x)
        ev[2].record()
        L.check(lib.sglb_ep_p2p_wait_dispatch(px.base, wl.world, cap, wl.H, wl.topk, s))
        ev[3].record()
        B.experts_into(px.recv_x[par], px.recv_tw[par], px.recv_ids[par], px.y[par], layer._hint(cap))
        ev[4].record()
        out = torch.empty_like(wl.x)
        if shared_y is not None:
            main.wait_stream(px.side)
        L.check(lib.sglb_ep_p2p_combine(ops._ptr(ids), wl.M, wl.topk, wl.E, wl.world, wl.rank, wl.H, cap, px.peers, par,
                                        ops._ptr(shared_y), float(layer.rsf), ops._ptr(out), s))
        ev[5].record()
        torch.cuda.synchronize()
        if it >= 2:
            for i, n in enumerate(names):
                samples[n].append(ev[i].elapsed_time(ev[i + 1]) * 1e3)
    return {n: statistics.median(v) for n, v in samples.items()}


# ---------------------------------------------------------------------------------------------------------------
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
    shape, fmt, M = args.shape, args.fmt, args.tokens
    gen = torch.Generator(device=dev)
    gen.manual_seed(1111 + rank)
    hbm_peak, tc, peak_src = peaks()
    lib = L.load()
    lib.sglb_set_mx4_native(0 if args.no_mx4_native else 1)
    lib.sglb_set_chain_fusion(args.fusion)

    parity = None
    if world > 1:
        parity = ep_parity(args, world, rank, dev)
        if not (parity["p2p_eq_nccl"] and parity["eq_1gpu"]):
            if rank == 0:
                print(json.dumps({"error": "expert-parallel parity check failed", "ep_parity": parity}))
            import torch.distributed as dist
            dist.destroy_process_group()
            raise SystemExit(3)

    wl = Workload(args, shape, fmt, M, dev, gen, world, rank, args.copies, router=not args.no_router, shared=not args.no_shared)
    E, topk, H, I = wl.E, wl.topk, wl.H, wl.I
    step, graphed = wl.make_step()

    # ---- the step that gets timed, checked against the oracle first (N = 1; N > 1 has ep_parity above) ----
    step_par = None
    if world == 1 and not args.no_parity:
        step_par = step_parity(wl, step, args.parity_tokens)
        if not step_par["ok"]:
            print(json.dumps({"error": "the timed step differs from the oracle", "parity_check": step_par}))
            raise SystemExit(3)

    # ---- headline: device-resident inputs ----
    clocks = ClockSampler(local)
    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    l0 = K.launch_count()
    clocks.start()
    ms, windows, per_rank = time_steps(step, args.steps, 0, world, windows=max(1, args.windows))
    launches = (K.launch_count() - l0) // max(1, args.windows)
    if graphed:  # kernels inside the replayed graphs (counted once, at capture, by the C ABI)
        launches = step.kernels_per_step * args.steps
    # The timed region lasts a few ms, less than nvidia-smi needs to report once: keep the same steps running
    # (untimed, same count on every rank) until the sampler has seen about half a second of this load.
    extra = int(min(20000, max(0.0, 0.6 - ms * 1e-3 * len(windows)) / max(ms * 1e-3 / args.steps, 1e-6)))
    for i in range(extra):
        step(i)
    torch.cuda.synchronize()
    clk = clocks.stop()
    clk["sampled_over"] = f"{len(windows)} timed windows of {args.steps} steps + {extra} identical untimed steps (nvidia-smi -lms 100)"
    ms_per_step = ms / args.steps
    tokens_per_s = M * world / (ms_per_step * 1e-3)
    used = path_used(args, fmt, M, topk, E, H, I)
    routed = "fused_experts" + (" + shared_expert" if wl.shared_I else "")
    out = {
        "metric": f"MoE layer tokens/s ({'biased_' if wl.biased else ''}grouped_topk -> {routed}, {fmt})",
        "value": tokens_per_s, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": DTYPE[fmt] if used != "stream" or fmt in ("bf16", "int8") else {"fp8": "e4m3 -> bf16", "mxfp4": "e2m1 -> bf16"}[fmt],
        "data": "synthetic",
        "config": {"workload": describe(wl), "tokens_per_step_per_gpu": M,
                   "parallelism": "single GPU" if world == 1 else f"ep{world} (E/{world} = {E // world} experts per GPU, shared expert replicated)",
                   "l2_policy": wl.l2_policy(), "path": PATH_DESC[used],
                   "launch": ("one CUDA-graph replay per step (router + experts + shared expert captured)" if graphed
                              else "eager C-ABI calls"),
                   "timing": f"median of {len(windows)} windows of {args.steps} steps, each the max over ranks of a CUDA-event interval"},
        "windows_ms_per_step": [w / args.steps for w in windows],
        "clocks": clk, "gpu_launches": int(launches),
    }
    if step_par is not None:
        out["parity_check"] = step_par
    if world > 1:
        out["per_rank_ms_per_step"] = [p / args.steps for p in per_rank]
        out["config"]["exchange"] = {"p2p": "NVLink peer memory (ep_p2p.cu: stores into the owners' buffers, per-slot rows read back with peer loads; no NCCL on the data path)",
                                     "fixed": "NCCL all-to-all, fixed capacity", "exact": "NCCL all-to-all, exact splits"}.get(
                                         wl.layers[0].last_mode, str(wl.layers[0].last_mode))
        out["ep_parity"] = parity
        # per-rank roofline: this rank streams its E/world experts + the replicated shared expert
        ids_local = wl.ids  # routing of this rank's tokens; the experts hit on THIS rank come from every rank's tokens
        e_local = E // world
        bw = WBYTES[fmt]
        per_rank_bytes = e_local * (3 * H * I * bw + scale_bytes(fmt, H, I)) + (3 * H * wl.shared_I * bw + scale_bytes(fmt, H, wl.shared_I) if wl.shared_I else 0) \
            + 2 * M * H * 2 + M * topk * 8
        gbs = per_rank_bytes / (ms_per_step * 1e-3) / 1e9
        xbytes = {"dispatch_rows": int(M * H * 2), "return_slot_rows": int(M * topk * H * 2),
                  "note": "upper bounds: a row travels once per owner rank, a slot row once"}
        # a rank computes the routed rows it RECEIVES (M * topk on average: every rank sends as many tokens) + its own
        # tokens through the replicated shared expert; the bound is whichever roof asks for more time
        fl = flops(M, topk, H, I, wl.shared_I)
        roof_dt = "int8" if fmt == "int8" else "bf16"
        t_hbm, t_tc = per_rank_bytes / (hbm_peak * 1e9), (fl / (tc[roof_dt] * 1e12) if tc.get(roof_dt) else 0.0)
        if t_tc > t_hbm:
            tf = fl / (ms_per_step * 1e-3) / 1e12
            out["roofline"] = {"bound": "tensor", "kernel": "whole step of one rank (received routed rows + local shared expert)",
                               "achieved": tf, "peak": tc[roof_dt], "unit": "TFLOP/s" if roof_dt == "bf16" else "TOP/s",
                               "frac": tf / tc[roof_dt], "peak_source": peak_src, "traffic": None, "algorithmic_flops": fl,
                               "hbm_gbs": gbs, "exchange_bytes_per_rank": xbytes}
        else:
            out["roofline"] = {"bound": "hbm", "kernel": "whole step of one rank (all local experts assumed active)", "achieved": gbs,
                               "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak, "peak_source": peak_src, "traffic": None,
                               "algorithmic_bytes": per_rank_bytes, "exchange_bytes_per_rank": xbytes}
        del ids_local
        try:
            ph = ep_phases(wl, 10)
        except Exception as e:  # noqa: BLE001
            ph = {"error": str(e)}
        out["phase_us"] = ph

    if world == 1:
        prof = wl.stage_profile(max(3, min(args.steps, 20)))
        out["roofline"], out["step_roofline"] = roofline_of(wl, ms_per_step, prof, hbm_peak, tc, peak_src)
        # ---- sweep over M on the headline layer ----
        if not args.no_sweep:
            default_sweep = {"dsv3": "1,16,64,256,8192", "qwen3": "1,16,128,1024,4096", "gptoss": "1,128,4096", "refbench": "4,3929"}[shape]
            sweep = []
            for m in [int(v) for v in (args.sweep or default_sweep).split(",") if v]:
                w2 = Workload.__new__(Workload)
                w2.__dict__.update(wl.__dict__)
                w2.M = m
                w2.x, w2.gating = make_batch(m, E, H, dev, gen)
                w2.tw, w2.ids = w2.router(w2.gating)
                st, gr = w2.make_step()
                t = time_steps(st, 10, 3) / 10
                abm = algorithmic_bytes(fmt, E, H, I, m, topk, w2.ids, wl.shared_I)
                fl = flops(m, topk, H, I, wl.shared_I)
                roof_dt = "int8" if fmt == "int8" else "bf16"
                ent = {"M": m, "ms": t, "tokens_per_s": m / (t * 1e-3), "hbm_frac": abm["total"] / (t * 1e-3) / 1e9 / hbm_peak,
                       "tflops": fl / (t * 1e-3) / 1e12, "active_experts": abm["e_act"], "launch": "graph" if gr else "eager"}
                if tc.get(roof_dt):
                    ent[f"tc_frac_{roof_dt}_peak"] = ent["tflops"] / tc[roof_dt]
                sweep.append(ent)
                del st
                torch.cuda.empty_cache()
            out["sweep"] = sweep

    # ---- e2e: the drop-in ops with HOST buffers ----
    if not args.no_e2e:
        out["e2e"] = run_e2e(args, wl, world, rank, max(5, min(args.steps, 30)))

    if world == 1 and not args.no_configs:
        del step
        wl.weights = wl.shared = wl.layers = None
        del wl
        torch.cuda.empty_cache()
        cfgs = []
        if (shape, fmt) == ("dsv3", "fp8") and args.bias_scale is None:
            args.bias_scale = 1.0
            try:
                cfgs.append(measure_config(args, "C3, randn bias as test_biased_grouped_topk.py:55 (routing collapses onto few experts)",
                                           "dsv3", "fp8", M, dev, gen, hbm_peak, tc, peak_src, e2e=False))
            except Exception as e:  # noqa: BLE001
                cfgs.append({"name": "C3, randn bias", "M": M, "error": f"{type(e).__name__}: {e}"})
                torch.cuda.empty_cache()
            args.bias_scale = None
        plan = [("C1 Qwen3-30B-A3B bf16", "qwen3", "bf16", [128, 8192]),
                ("C2 Qwen3-30B-A3B int8 w8a8", "qwen3", "int8", [1, 16, 128, 1024, 4096]),
                ("C4 gpt-oss mxfp4", "gptoss", "mxfp4", [128, 4096]),
                ("bench_moe.py:144-145 bf16", "refbench", "bf16", [4, 3929]),
                ("bench_moe.py:144-145 int8", "refbench", "int8", [4, 3929]),
                ("bench_moe.py:144-145 fp8", "refbench", "fp8", [4, 3929])]
        for name, shp, f, Ms in plan:
            if (shp, f) == (shape, fmt):
                continue
            for m in Ms:
                try:
                    cfgs.append(measure_config(args, name, shp, f, m, dev, gen, hbm_peak, tc, peak_src, e2e=(m == Ms[0])))
                except Exception as e:  # noqa: BLE001
                    cfgs.append({"name": name, "M": m, "error": f"{type(e).__name__}: {e}"})
                    torch.cuda.empty_cache()
        out["configs"] = cfgs
        bad = [c["name"] + f" M={c['M']}" for c in cfgs if c.get("parity_check") and not c["parity_check"].get("ok")]
        if bad:
            print(json.dumps({"error": "timed configs differ from the oracle", "configs": bad,
                              "parity_checks": [c.get("parity_check") for c in cfgs]}))
            raise SystemExit(3)

    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(shape, fmt, M, budget_s=25.0)
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def run_e2e(args, wl, world, rank, steps):
    """Same metric through the reference-facing ops with host tensors, the calls a user of the reference makes:
    [biased_]grouped_topk_cpu -> fused_experts_cpu(inplace=True) -> shared_expert_cpu(inplace=True), every activation /
    routing tensor a pinned HOST tensor: per step they go H2D and every op's result comes back D2H.  Host weights are
    uploaded once (first call) and stay in HBM.  N > 1: EPMoE.forward_host (host buffers in / out around the exchange)."""
    K = wl.K
    sys.path.insert(0, K.DROPIN_PATH)
    import sgl_kernel  # noqa: F401  (registers torch.ops.sgl_kernel.*)
    from sgl_kernel import common_ops as co
    ops = torch.ops.sgl_kernel
    M, E, topk, H = wl.M, wl.E, wl.topk, wl.H
    pin = lambda t: t.cpu().pin_memory()  # noqa: E731
    xh, gh = pin(wl.x), pin(wl.gating)
    # one pinned input buffer per call (warm-ups included): the inplace ops overwrite their input, and a memcpy that
    # restores it inside the timed loop would be bench overhead, not a cost of the API (0.3 ms of 3.0 at M = 256)
    # (prefill-sized inputs, > 1 GB of pinned ring, keep the restoring memcpy)
    use_ring = world == 1 and xh.numel() * 2 * (steps + 3) <= (1 << 30)
    ring = [xh.clone().pin_memory() for _ in range(steps + 3 if use_ring else 1)] if world == 1 else []
    per_op = {"router": 0.0, "fused_experts": 0.0, "shared_expert": 0.0}
    calls = [0]
    h2d = d2h = shared_uploads = 0
    if world > 1:
        layer = wl.layers[0]
        twh, idh = pin(wl.tw), pin(wl.ids)

        def call():
            return layer.forward_host(xh, twh, idh)
        h2d = xh.numel() * 2 + twh.numel() * 4 + idh.numel() * 4
        d2h = xh.numel() * 2
        api = "ep.EPMoE.forward_host(pinned host hidden / topk_weights / topk_ids) -> host tensor; routing precomputed"
    else:
        w = wl.weights[0]
        fmt = wl.fmt
        w1h, w2h = w["w1"].cpu(), w["w2"].cpu()
        s1h = None if w["s1"] is None else w["s1"].cpu()
        s2h = None if w["s2"] is None else w["s2"].cpu()
        bias_h = None if wl.bias is None else wl.bias.cpu()
        sh = None
        if wl.shared:
            s = wl.shared[0]
            sh = {k: (v.cpu() if torch.is_tensor(v) else v) for k, v in s.items()}
        twh = torch.empty(M, topk, dtype=torch.float32).pin_memory()
        idh = torch.empty(M, topk, dtype=torch.int32).pin_memory()

        def call():
            hs_work = ring[calls[0] % len(ring)]
            calls[0] += 1
            if not use_ring:
                hs_work.copy_(xh)
            t_a = time.perf_counter()
            if wl.with_router:
                if wl.biased:   # out-variant, test_biased_grouped_topk.py:67-80
                    co.biased_grouped_topk_cpu(twh, idh, hs_work, gh, bias_h, topk, wl.renorm, wl.G, wl.tg)
                    tw_, id_ = twh, idh
                else:           # functional, test_moe.py:61-70
                    tw_, id_ = ops.grouped_topk_cpu(hs_work, gh, topk, wl.renorm, wl.G, wl.tg, 0, None, None)
            else:
                tw_, id_ = wl_tw_h, wl_id_h
            t_b = time.perf_counter()
            if fmt == "mxfp4":
                routed = ops.fused_experts_cpu(hs_work, w1h, w2h, tw_, id_, not wl.shared_I, 4, s1h, s2h, None, None, None, False)
            else:
                routed = ops.fused_experts_cpu(hs_work, w1h, w2h, tw_, id_, not wl.shared_I, fmt == "int8", fmt == "fp8", s1h, s2h,
                                               w["block"], None, None, False)
            t_c = time.perf_counter()
            if sh is not None:
                routed = ops.shared_expert_cpu(hs_work, sh["w1"], sh["w2"], routed, wl.rsf, True, fmt == "int8", fmt == "fp8",
                                               sh["s1"], sh["s2"], sh["block"], None, None, False)
            t_d = time.perf_counter()
            per_op["router"] += t_b - t_a
            per_op["fused_experts"] += t_c - t_b
            per_op["shared_expert"] += t_d - t_c
            return routed
        wl_tw_h, wl_id_h = pin(wl.tw), pin(wl.ids)
        if wl.with_router:
            h2d += gh.numel() * 2 + (bias_h.numel() * 2 if bias_h is not None else 0)
            d2h += M * topk * 8
        h2d += xh.numel() * 2 + M * topk * 8
        d2h += xh.numel() * 2
        if sh is not None:
            shared_uploads = 2 * xh.numel() * 2      # hidden_states and fused_experts_out, unless still in HBM (below)
            d2h += xh.numel() * 2
        api = ("torch.ops.sgl_kernel." + ("biased_grouped_topk_cpu -> " if wl.biased else "grouped_topk_cpu -> ") * int(wl.with_router)
               + "fused_experts_cpu" + (" -> shared_expert_cpu" if sh is not None else "")
               + " on pinned host tensors (inplace); weights HBM-resident after the first call")
    for _ in range(3):
        call()
    torch.cuda.synchronize()
    for k in per_op:
        per_op[k] = 0.0
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    from sgl_cpu_tests_b200 import ops as _ops
    hits0 = _ops.shadow_hits
    t0 = time.perf_counter()
    for _ in range(steps):
        call()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    reused = (_ops.shadow_hits - hits0) / steps      # tensors per step shared_expert_cpu found still in HBM (ops._shadow_get)
    if shared_uploads:
        h2d += int(shared_uploads * max(0.0, 2.0 - reused) / 2.0)
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([dt], device=wl.dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    K.clear_weight_cache()
    return {"value": M * world * steps / dt, "unit": "tokens/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
            "ms_per_step": dt / steps * 1e3, "steps": steps, "api": api,
            "per_op_ms": {k: v / steps * 1e3 for k, v in per_op.items()} if world == 1 else None,
            "tensors_reused_in_hbm_per_step": reused,
            "note": "fixed token count per call: host-tensor calls replay a cached CUDA graph per (M, weights)"}


# ---------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port (torch-native restatement of the reference's own CPU oracle code) on the host cores.
def cpu_info():
    model, amx = platform.processor() or "unknown", False
    try:
        for l in open("/proc/cpuinfo"):
            if l.startswith("model name") and model in ("", "unknown", "x86_64"):
                model = l.split(":", 1)[1].strip()
            if l.startswith("flags"):
                amx = "amx_bf16" in l
                break
    except OSError:
        pass
    return {"model": model, "amx_bf16": amx, "cores": os.cpu_count() or 1}


def cpu_workload(shape, fmt, m, seed=1111):
    """Host copy of the workload.  Large weights are generated as raw bytes (torch.randn of 10^10 elements on the host
    would take minutes): e4m3 codes with the top exponent bit cleared (|w| < 2), int8 / e2m1 codes uniform."""
    E, topk, H, I = SHAPES[shape]
    cfg = LAYER[shape]
    g = torch.Generator().manual_seed(seed)

    def tiled(shape_, block_fn):
        """[shape_] filled with repeats of one random block of <= 2^26 elements (distinct memory, cheap to draw)."""
        n = math.prod(shape_)
        blk = block_fn(min(n, 1 << 26))
        reps = -(-n // blk.numel())
        return (blk.repeat(reps)[:n] if reps > 1 else blk[:n]).view(*shape_)

    def raw(shape_, mask=0xFF):
        return tiled(shape_, lambda k: torch.randint(0, 256, (k,), generator=g, dtype=torch.uint8) & mask)

    def weights(E_, I_):
        if fmt == "bf16":
            bf = lambda k: (torch.randn(k, generator=g) / 10).bfloat16()  # noqa: E731
            return dict(w1=tiled((E_, 2 * I_, H), bf), w2=tiled((E_, H, I_), bf), s1=None, s2=None, block=None)
        if fmt == "int8":
            return dict(w1=raw((E_, 2 * I_, H)).view(torch.int8), w2=raw((E_, H, I_)).view(torch.int8),
                        s1=torch.rand(E_, 2 * I_, generator=g) * 1e-2, s2=torch.rand(E_, H, generator=g) * 1e-2, block=None)
        if fmt == "fp8":
            return dict(w1=raw((E_, 2 * I_, H), 0xBF).view(torch.float8_e4m3fn), w2=raw((E_, H, I_), 0xBF).view(torch.float8_e4m3fn),
                        s1=torch.randn(E_, -(-2 * I_ // 128), H // 128, generator=g) * 1e-2,
                        s2=torch.randn(E_, H // 128, -(-I_ // 128), generator=g) * 1e-2, block=[128, 128])
        return dict(w1=raw((E_, 2 * I_, H // 2)), w2=raw((E_, H, I_ // 2)),
                    s1=torch.randint(118, 122, (E_, 2 * I_, H // 32), generator=g, dtype=torch.uint8),
                    s2=torch.randint(118, 122, (E_, H, I_ // 32), generator=g, dtype=torch.uint8), block=None)
    w = weights(E, I)
    sh = None
    if cfg["shared"]:
        s = weights(1, cfg["shared"])
        sh = {k: (v[0] if torch.is_tensor(v) else v) for k, v in s.items()}
    x, gating = make_batch(m, E, H, "cpu", g)
    bias = (torch.randn(E, generator=g) * cfg.get("bias_scale", 1.0)).bfloat16() if cfg["router"][0] else None
    return dict(w=w, shared=sh, x=x, gating=gating, bias=bias, cfg=cfg, E=E, topk=topk, fmt=fmt)


def cpu_call(c):
    """One step of the layer on the host cores with the oracle port: router -> fused experts -> shared expert."""
    from oracle import moe_oracle as O
    biased, G, tg, renorm = c["cfg"]["router"]
    if biased:
        tw, ids = O.biased_grouped_topk(c["gating"], c["bias"], c["topk"], renorm, G, tg)
    else:
        tw, ids = O.grouped_topk(c["gating"], c["topk"], renorm, G, tg)
    w = c["w"]
    out = O.fused_experts_lean(c["x"], w["w1"], w["w2"], tw, ids, c["fmt"], w["s1"], w["s2"], w["block"])
    if c["shared"] is not None:
        s = c["shared"]
        out = O.shared_expert(c["x"], s["w1"], s["w2"], out, c["cfg"]["rsf"], c["fmt"], s["s1"], s["s2"], s["block"])
    return out


def cpu_make_resident(c):
    """One-off, outside every timed loop: the weight-only formats (fp8, mxfp4) dequantised to bf16 on the host
    (DeepSeek-V3: 22.6 GB), so that a timed CPU step is GEMMs over resident bf16 weights -- the fair CPU figure.  The
    literal oracle re-dequantises every active expert to f32 on every call."""
    from oracle import moe_oracle as O
    if c["fmt"] not in ("fp8", "mxfp4"):
        return None
    t0 = time.perf_counter()
    w = c["w"]
    r = dict(w1=O.dequantise_to_bf16(w["w1"], w["s1"], c["fmt"], w["block"]),
             w2=O.dequantise_to_bf16(w["w2"], w["s2"], c["fmt"], w["block"]), shared=None)
    if c["shared"] is not None:
        s = c["shared"]
        r["shared"] = (O.dequantise_to_bf16(s["w1"], s["s1"], c["fmt"], s["block"]),
                       O.dequantise_to_bf16(s["w2"], s["s2"], c["fmt"], s["block"]))
    r["seconds"] = time.perf_counter() - t0
    return r


def cpu_call_resident(c, r):
    """One step on bf16-resident weights: router -> the reference's bf16 oracle (test_moe.py:22-54: bf16 GEMMs, oneDNN /
    AMX where the CPU has it) -> shared expert."""
    from oracle import moe_oracle as O
    biased, G, tg, renorm = c["cfg"]["router"]
    if biased:
        tw, ids = O.biased_grouped_topk(c["gating"], c["bias"], c["topk"], renorm, G, tg)
    else:
        tw, ids = O.grouped_topk(c["gating"], c["topk"], renorm, G, tg)
    out = O.torch_naive_moe(c["x"], r["w1"], r["w2"], tw, ids)
    if r["shared"] is not None:
        out = O.shared_expert_resident_bf16(c["x"], r["shared"][0], r["shared"][1], out, c["cfg"]["rsf"])
    return out


def cpu_time_resident(c, budget_s, max_steps=10):
    """-> (seconds per step, steps, dequantisation seconds) or None for formats with nothing to hoist."""
    r = cpu_make_resident(c)
    if r is None:
        return None
    t0 = time.perf_counter()
    cpu_call_resident(c, r)     # warm-up (page faults of the fresh 22.6 GB, oneDNN primitive creation)
    one = time.perf_counter() - t0
    reps = max(1, min(max_steps, int(budget_s / max(one, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(reps):
        cpu_call_resident(c, r)
    return (time.perf_counter() - t0) / reps, reps, r["seconds"]


def cpu_baseline(shape, fmt, M, budget_s):
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    c = cpu_workload(shape, fmt, M)
    t0 = time.perf_counter()
    cpu_call(c)
    one = time.perf_counter() - t0
    reps = max(1, min(10, int(budget_s / max(one, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(reps):
        cpu_call(c)
    dt = (time.perf_counter() - t0) / reps
    out = {"value": M / dt, "unit": "tokens/s", "cores": cores, "kind": "port", "cpu": cpu_info(),
           "sample": f"{reps} full steps of the torch oracle port (router + experts + shared expert) at the workload's own M = {M} "
                     f"({dt * 1e3:.0f} ms/step); weights dequantised per active expert per call, as the reference's oracle does; "
                     "tensors: same shapes, formats and generators' value ranges as the GPU arm, NOT the same bytes "
                     "(cpu_workload draws raw codes on the host with its own seed)"}
    try:
        res = cpu_time_resident(c, budget_s / 2)
        if res is not None:
            out["predequantised_bf16"] = {
                "value": M / res[0], "unit": "tokens/s", "cores": cores, "kind": "port, pre-dequantised bf16", "ms_per_step": res[0] * 1e3,
                "sample": f"{res[1]} full steps at M = {M} with the weights dequantised ONCE to bf16 outside the timed loop "
                          f"({res[2]:.1f} s, not timed) and the reference's bf16 oracle (test_moe.py:22-54) on them: the fair CPU figure"}
    except Exception as e:  # noqa: BLE001  (e.g. a host without 23 GB to spare)
        out["predequantised_bf16"] = {"error": f"{type(e).__name__}: {e}"}
    return out


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  fused_experts_cpu itself is not
    in /root/reference (third-party sgl-kernel, absent), so this times the oracle port: the
    reference's own torch-native oracle code for this path, all host threads, on the SAME configuration and the SAME
    token count as the GPU arm (shrunk only if one step exceeds the time budget, and then said so)."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    shape, fmt = args.shape, args.fmt
    # the GPU arm's step is `tokens` per GPU on every one of the N GPUs: the CPU arm does the whole job's tokens
    M = args.tokens * max(1, world)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    c = cpu_workload(shape, fmt, M)
    t0 = time.perf_counter()
    cpu_call(c)
    one = time.perf_counter() - t0
    budget = 180.0   # seconds for the whole arm: the driver runs it at N = 1 and again for every N of the scaling run
    resident = None
    if fmt in ("fp8", "mxfp4"):
        # The fair figure and this line's `value`: weights dequantised once, outside the timed steps.  The literal
        # oracle (the probing call above, repeated while it is cheap) is reported beside it.
        lit_reps = max(0, min(2, int(30.0 / max(one, 1e-3))))
        t0 = time.perf_counter()
        for _ in range(lit_reps):
            cpu_call(c)
        lit_dt = (time.perf_counter() - t0) / lit_reps if lit_reps else one
        try:
            resident = cpu_time_resident(c, budget - 60.0 - one * (1 + lit_reps), max_steps=args.steps)
        except Exception as e:  # noqa: BLE001
            sys.stderr.write(f"[bench] pre-dequantised CPU figure unavailable ({type(e).__name__}: {e}); literal oracle only\n")
    if resident is not None:
        dt, steps, deq_s = resident
        warmup, m = 1, M
        val = m / dt
        kind = "port, pre-dequantised bf16"
        sample = (f"each step = router + the reference's bf16 oracle (test_moe.py:22-54) + shared expert on {m} tokens "
                  f"({args.tokens} per GPU x {max(1, world)} GPUs), weights dequantised ONCE to bf16 outside the timed steps ({deq_s:.1f} s); "
                  f"the literal per-call-dequantising oracle takes {lit_dt * 1e3:.0f} ms/step on the same input; tensors: same shapes / "
                  "formats / value ranges as the GPU arm, not the same bytes")
        literal = {"value": m / lit_dt, "unit": "tokens/s", "ms_per_step": lit_dt * 1e3, "kind": "port", "steps": max(1, lit_reps)}
        _print_reference(args, shape, fmt, M, m, val, dt, steps, warmup, cores, kind, sample, literal)
        return
    total = args.steps + args.warmup
    m = M
    steps, warmup = args.steps, args.warmup
    if one * total > budget:   # the step cost is per active expert, not per token: fewer repetitions before fewer tokens
        warmup = 1
        steps = max(3, min(args.steps, int(budget / one) - 1))
        if one * (steps + 1) > 1.5 * budget:
            m = max(1, int(M * budget / (one * (steps + 1))))
            c = cpu_workload(shape, fmt, m)
    for _ in range(max(0, warmup - 1)):   # the probing call above was the first warm-up
        cpu_call(c)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_call(c)
    dt = (time.perf_counter() - t0) / steps
    val = m / dt
    sample = (f"each step = the torch oracle port (router + experts + shared expert) on {m} tokens of the workload"
              + ("" if m == M else f" (shrunk from {M}: one full step took {one:.1f} s)")
              + (f"; {steps} timed steps instead of {args.steps} to stay within the time budget" if steps != args.steps else "")
              + "; tensors: same shapes / formats / value ranges as the GPU arm, not the same bytes")
    _print_reference(args, shape, fmt, M, m, val, dt, steps, warmup, cores, "port", sample, None)


def _print_reference(args, shape, fmt, M, m, val, dt, steps, warmup, cores, kind, sample, literal):
    E, topk, H, I = SHAPES[shape]
    wl = argparse.Namespace(shape=shape, fmt=fmt, E=E, topk=topk, H=H, I=I, biased=LAYER[shape]["router"][0], G=LAYER[shape]["router"][1],
                            tg=LAYER[shape]["router"][2], renorm=LAYER[shape]["router"][3], shared_I=LAYER[shape]["shared"],
                            rsf=LAYER[shape]["rsf"], with_router=True, M=args.tokens, bias_scale=LAYER[shape].get("bias_scale", 1.0))
    routed = "fused_experts" + (" + shared_expert" if wl.shared_I else "")
    line = {
        "impl": "reference", "metric": f"MoE layer tokens/s ({'biased_' if wl.biased else ''}grouped_topk -> {routed}, {fmt})",
        "value": val, "unit": "tokens/s", "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (torch CPU)", "data": "synthetic",
        "config": {"workload": describe(wl), "tokens_per_step_per_gpu": args.tokens, "tokens_per_step": M,
                   "sample_tokens_per_step": m, "same_config": m == M},
        "cpu_baseline": {"value": val, "unit": "tokens/s", "cores": cores, "kind": kind, "sample": sample, "cpu": cpu_info()},
        "e2e": {"value": val, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    if literal is not None:
        line["literal_oracle"] = literal
        line["dtype"] = "bf16 GEMMs on pre-dequantised weights (torch CPU)"
    print(json.dumps(line))


if __name__ == "__main__":
    a = parse()
    BLOCK_K = a.block_k
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
