"""Generates tests/golden/*.pt by executing the REFERENCE's own oracle code.

Run in the build container only (it needs /root/reference; the GPU box never runs this):

    python tests/golden/make_golden.py

How: each reference script imports `sgl_kernel` (absent here) at the top and runs its cases at import.
We pre-seed sys.modules with recording stubs for `sgl_kernel`, `sgl_kernel.common_ops`,
`sglang.srt.layers.amx_utils` and `utils`, then exec the unmodified script with runpy.  The stubs
record the tensors handed to each op; the `compare` stub records the reference-side result the script
computed with its own torch oracle.  After the run the script's globals also give us its oracle
FUNCTIONS (torch_naive_moe, torch_w8a8_per_column_moe, native_fused_moe, grouped_topk_native, ...),
which we call on additional small seeded inputs.  Nothing from /root/reference is copied into the
repo: only input/output tensors are stored.
"""
import io
import os
import runpy
import sys
import types
from contextlib import redirect_stdout

import torch

REF = os.environ.get("SGLB_REFERENCE", "/root/reference")
OUT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(OUT, "..", "..")))
from oracle import moe_oracle as O  # noqa: E402  (only to cross-check the prepack formulas)

CALLS = []      # (op name, [args])
COMPARES = []   # (ref, out) in script order
MAX_CASE_BYTES = 3 << 19


def _clone(x):
    if isinstance(x, torch.Tensor):
        return x.detach().clone()
    if isinstance(x, (list, tuple)):
        return [_clone(v) for v in x]
    return x


def _nbytes(x):
    if isinstance(x, torch.Tensor):
        return x.numel() * x.element_size()
    if isinstance(x, (list, tuple)):
        return sum(_nbytes(v) for v in x)
    if isinstance(x, dict):
        return sum(_nbytes(v) for v in x.values())
    return 0


def _record(name, ret):
    def op(*args):
        CALLS.append((name, _clone(list(args))))
        return ret(*args)
    return op


def _install_stubs(pack_weight=None, pack_scale=None):
    ident = lambda w: w.clone()  # noqa: E731
    ops = {
        "convert_weight_packed": _record("convert_weight_packed", pack_weight or ident),
        "convert_scale_packed": _record("convert_scale_packed", pack_scale or ident),
        "fused_experts_cpu": _record("fused_experts_cpu", lambda hs, *a: torch.zeros_like(hs)),
        "shared_expert_cpu": _record("shared_expert_cpu", lambda hs, *a: torch.zeros_like(hs)),
        "mxfp4_scaled_mm_cpu": _record(
            "mxfp4_scaled_mm_cpu", lambda a, b, *r: torch.zeros(a.shape[0], b.shape[0], dtype=a.dtype)),
        "fp8_scaled_mm_cpu": _record("fp8_scaled_mm_cpu", lambda a, b, *r: torch.zeros(a.shape[0], b.shape[0], dtype=a.dtype)),
    }

    def grouped(*args):
        CALLS.append(("grouped_topk_cpu", _clone(list(args))))
        if len(args) == 8:  # out variant
            args[0].zero_(), args[1].zero_()
            return None
        m, topk = args[1].shape[0], args[2]
        return torch.zeros(m, topk), torch.zeros(m, topk, dtype=torch.int32)

    def biased(*args):
        CALLS.append(("biased_grouped_topk_cpu", _clone(list(args))))
        args[0].zero_(), args[1].zero_()
        return None

    ops["grouped_topk_cpu"] = grouped
    ops["biased_grouped_topk_cpu"] = biased

    sgl = types.ModuleType("sgl_kernel")
    common = types.ModuleType("sgl_kernel.common_ops")
    for k, v in ops.items():
        setattr(common, k, v)
    sgl.common_ops = common
    sys.modules["sgl_kernel"] = sgl
    sys.modules["sgl_kernel.common_ops"] = common

    class _NS:
        def __getattr__(self, name):
            return ops[name]

    torch.ops.__dict__["sgl_kernel"] = _NS()  # torch.ops.sgl_kernel.<op> -> stub

    for name in ("sglang", "sglang.srt", "sglang.srt.layers", "sglang.srt.layers.amx_utils"):
        sys.modules[name] = types.ModuleType(name)

    class CPUQuantMethod:
        UNQUANT = 0
    sys.modules["sglang.srt.layers.amx_utils"].CPUQuantMethod = CPUQuantMethod

    utils = types.ModuleType("utils")

    def compare(a, b, debug=False):
        COMPARES.append((_clone(a), _clone(b)))
    utils.compare = compare
    sys.modules["utils"] = utils


def run_script(name, **stub_kw):
    del CALLS[:], COMPARES[:]
    _install_stubs(**stub_kw)
    buf = io.StringIO()
    with redirect_stdout(buf):
        g = runpy.run_path(os.path.join(REF, name), run_name="__ref__")
    return g, list(CALLS), list(COMPARES), buf.getvalue()


def save(name, obj):
    path = os.path.join(OUT, name + ".pt")
    torch.save(obj, path)
    print(f"{name}.pt: {os.path.getsize(path) / 1024:.0f} KiB, {len(obj) if hasattr(obj, '__len__') else ''} cases")


def routing(m, e, topk, dtype=torch.bfloat16):
    score = torch.randn((m, e), dtype=dtype)
    score = torch.softmax(score, dim=-1, dtype=torch.float32)
    w, ids = torch.topk(score, topk)
    return w, ids.to(torch.int32)


def main():
    torch.manual_seed(20261104)

    # ---------------- routers ----------------
    g, calls, cmps, _ = run_script("test_grouped_topk.py")
    ref_fn = g["grouped_topk_native"]
    cases = []
    shapes = [(17, 8, 2, 2, 1), (17, 16, 4, 3, 2), (17, 32, 4, 3, 2), (33, 64, 1, 6, 1), (17, 256, 8, 4, 8),
              (17, 160, 8, 6, 2), (9, 128, 1, 8, 1), (5, 384, 1, 8, 1)]
    for (M, E, G, topk, tg) in shapes:
        for renorm in (True, False):
            for style in ("f32", "bf16x2M"):
                if style == "f32":
                    gating = torch.randn(M, E)
                else:  # the script's own conditioning (test_grouped_topk.py:44-46)
                    gating = torch.randn(M, E, dtype=torch.bfloat16) * 2 * M
                w, ids = ref_fn(torch.zeros(M, 4), gating.float(), topk, renorm, G, tg)
                cases.append(dict(M=M, E=E, G=G, topk=topk, topk_group=tg, renormalize=renorm, gating=gating,
                                  ref_weights=w, ref_ids=ids))
    save("router_softmax", cases)

    g, calls, cmps, _ = run_script("test_biased_grouped_topk.py")
    ref_fn = g["biased_grouped_topk"]
    cases = []
    for (M, E, G, topk, tg) in [(19, 256, 8, 8, 2), (7, 64, 4, 4, 2), (11, 32, 8, 3, 3), (6, 256, 8, 8, 4)]:
        for renorm in (True, False):
            for style in ("f32", "bf16x2M"):
                if style == "f32":
                    gating, bias = torch.randn(M, E), torch.randn(E)
                else:
                    gating = torch.randn(M, E, dtype=torch.bfloat16) * 2 * M
                    bias = torch.randn(E, dtype=torch.bfloat16)
                w, ids = ref_fn(torch.zeros(M, 4), gating.float(), bias.float(), topk, renorm, G, tg)
                cases.append(dict(M=M, E=E, G=G, topk=topk, topk_group=tg, renormalize=renorm, gating=gating,
                                  bias=bias, ref_weights=w, ref_ids=ids))
    save("router_biased", cases)

    # ---------------- bf16 MoE (test_moe.py) ----------------
    g, calls, cmps, _ = run_script("test_moe.py")
    naive = g["torch_naive_moe"]
    cases = []
    # the script's own case: inputs from the recorded fused_experts call, reference result from compare()
    fe = [c for c in calls if c[0] == "fused_experts_cpu"][0][1]
    gt = [c for c in calls if c[0] == "grouped_topk_cpu"][0][1]
    cases.append(dict(kind="script", a=fe[0], w1=fe[1], w2=fe[2], score=gt[1], topk=gt[2], renormalize=gt[3],
                      ref_out=cmps[0][0]))
    for (m, n, k, e, topk, renorm) in [(4, 32, 32, 4, 2, False), (2, 128, 32, 4, 2, True), (33, 64, 128, 8, 3, True),
                                       (7, 96, 64, 16, 6, False)]:
        a = torch.randn((m, k), dtype=torch.bfloat16) / 10
        w1 = torch.randn((e, 2 * n, k), dtype=torch.bfloat16) / 10
        w2 = torch.randn((e, k, n), dtype=torch.bfloat16) / 10
        score = torch.randn((m, e), dtype=torch.bfloat16)
        cases.append(dict(kind="fn", a=a, w1=w1, w2=w2, score=score, topk=topk, renormalize=renorm,
                          ref_out=naive(a, w1, w2, score, topk, renorm)))
    save("moe_bf16", cases)

    # ---------------- int8 w8a8 MoE (test_moe_int8.py) ----------------
    g, calls, cmps, _ = run_script("test_moe_int8.py")
    ref_fn, quant = g["torch_w8a8_per_column_moe"], g["per_token_quant_int8"]
    cases = []
    for (M, N, K, E, topk) in [(1, 128, 256, 8, 2), (9, 64, 128, 8, 3), (40, 96, 192, 4, 2)]:
        a = torch.randn((M, K), dtype=torch.bfloat16) / (K ** 0.5)
        w1 = ((torch.rand((E, 2 * N, K)) - 0.5) * 2 * 127).clamp(-128, 127).to(torch.int8)
        w2 = ((torch.rand((E, K, N)) - 0.5) * 2 * 127).clamp(-128, 127).to(torch.int8)
        w1_s, w2_s = torch.rand(E, 2 * N) * 1e-2, torch.rand(E, K) * 1e-2
        tw, ids = routing(M, E, topk)
        cases.append(dict(a=a, w1=w1, w2=w2, w1_s=w1_s, w2_s=w2_s, topk_weight=tw, topk_ids=ids,
                          ref_out=ref_fn(a, w1, w2, w1_s, w2_s, tw, ids.long(), topk)))
    x = torch.randn(13, 96, dtype=torch.bfloat16)
    q, s = quant(x)
    save("moe_int8", dict(cases=cases, quant=dict(x=x, q=q, scale=s)))

    # ---------------- fp8 block MoE + shared expert (test_moe_fp8_ext.py, BLOCK 64x128) ----------------
    g, calls, cmps, _ = run_script("test_moe_fp8_ext.py")
    native, scaled = g["native_fused_moe"], g["scaled_weight"]
    bn, bk = g["BLOCK_N"], g["BLOCK_K"]
    cases = []
    for (M, N, K, E, topk) in [(2, 128, 128, 8, 4), (21, 128, 256, 8, 2), (5, 256, 128, 4, 3)]:
        a = torch.randn(M, K, dtype=torch.bfloat16) / (K ** 0.5)
        w1 = (torch.randn(E, 2 * N, K) * 400).clamp(-400, 400).to(torch.float8_e4m3fn)
        w2 = (torch.randn(E, K, N) * 400).clamp(-400, 400).to(torch.float8_e4m3fn)
        w1s = torch.randn(E, 2 * N // bn, K // bk) * 1e-3
        w2s = torch.randn(E, K // bn, N // bk) * 1e-3
        tw, ids = routing(M, E, topk)
        ref = native(a, scaled(w1, w1s), scaled(w2, w2s), tw, ids.long(), topk)
        cases.append(dict(a=a, w1=w1, w2=w2, w1s=w1s, w2s=w2s, block=[bn, bk], topk_weight=tw, topk_ids=ids,
                          ref_out=ref.bfloat16()))
    # the script's own shared-expert runs (inline oracle): op inputs + compare() reference side
    shared = []
    se_calls = [c[1] for c in calls if c[0] == "shared_expert_cpu"]
    for args, (ref, _) in zip(se_calls, cmps[:len(se_calls)]):
        case = dict(a=args[0], w1=args[1], w2=args[2], fused_out=args[3], rsf=args[4], w1s=args[8], w2s=args[9],
                    block=list(args[10]), ref_out=ref)
        if _nbytes(case) <= MAX_CASE_BYTES:
            shared.append(case)
    save("moe_fp8", dict(cases=cases, shared=shared))

    # ---------------- masked / duplicate ids (test_moe_offloading_cpu.py, BLOCK 128x128) ----------------
    g, calls, cmps, _ = run_script("test_moe_offloading_cpu.py")
    native, scaled = g["native_fused_moe"], g["scaled_weight"]
    bn, bk = g["BLOCK_N"], g["BLOCK_K"]
    cases = []
    for (M, N, K, E, topk, ratio) in [(14, 128, 128, 8, 8, 8 / 128), (30, 128, 256, 8, 4, 0.3), (6, 256, 128, 4, 8, 0.9)]:
        ids = torch.randint(0, E, (M, topk), dtype=torch.int32)
        ids[~(torch.rand_like(ids.float()) < ratio)] = -1
        ids = ids[~(ids == -1).all(dim=1)]
        if ids.shape[0] == 0:
            ids = torch.full((1, topk), -1, dtype=torch.int32)
            ids[0, 0] = 0
        m = ids.shape[0]
        tw = torch.randn(m, topk)
        a = torch.randn(m, K, dtype=torch.bfloat16) / (K ** 0.5)
        w1 = (torch.randn(E, 2 * N, K) * 400).clamp(-400, 400).to(torch.float8_e4m3fn)
        w2 = (torch.randn(E, K, N) * 400).clamp(-400, 400).to(torch.float8_e4m3fn)
        w1s = torch.randn(E, 2 * N // bn, K // bk) * 1e-3
        w2s = torch.randn(E, K // bn, N // bk) * 1e-3
        ref = native(a, scaled(w1, w1s), scaled(w2, w2s), tw, ids, topk)
        cases.append(dict(a=a, w1=w1, w2=w2, w1s=w1s, w2s=w2s, block=[bn, bk], topk_weight=tw, topk_ids=ids,
                          ref_out=ref.bfloat16()))
    save("moe_offload", cases)

    # ---------------- shared expert bf16 / int8 (test_shared_experts.py) ----------------
    g, calls, cmps, _ = run_script("test_shared_experts.py")
    naive, w8a8, quant = g["torch_naive_moe"], g["torch_w8a8_per_column_moe"], g["per_token_quant_int8"]
    cases = []
    for (m, n, k, rsf) in [(2, 32, 32, 16), (21, 128, 64, 16), (40, 64, 96, 2.5)]:
        hs = torch.randn(m, k, dtype=torch.bfloat16) / k
        w1 = torch.randn(2 * n, k, dtype=torch.bfloat16)
        w2 = torch.randn(k, n, dtype=torch.bfloat16)
        fo = torch.randn(m, k, dtype=torch.bfloat16) / k
        ref = naive(hs.float(), w1.float(), w2.float(), fo.float(), rsf).to(torch.bfloat16)
        w1_q, w1_s = quant(w1)
        w2_q, w2_s = quant(w2)
        ref8 = w8a8(hs.float(), w1_q, w2_q, w1_s, w2_s, fo.float(), rsf).to(torch.bfloat16)
        cases.append(dict(hs=hs, w1=w1, w2=w2, fused_out=fo, rsf=rsf, ref_bf16=ref, w1_q=w1_q, w1_s=w1_s,
                          w2_q=w2_q, w2_s=w2_s, ref_int8=ref8))
    save("shared_expert", cases)

    # ---------------- fp8 shared expert with the real block quantiser (test_moe_fp8.py) ----------------
    g, calls, cmps, _ = run_script("test_moe_fp8.py")
    se_calls = [c[1] for c in calls if c[0] == "shared_expert_cpu"]
    cases = []
    for args, (ref, _) in zip(se_calls, cmps):
        case = dict(a=args[0], w1=args[1], w2=args[2], fused_out=args[3], rsf=args[4], w1s=args[8], w2s=args[9],
                    block=list(args[10]), ref_out=ref)
        if _nbytes(case) <= MAX_CASE_BYTES:
            cases.append(case)
    conv = g["convert_weight"]
    w = torch.randn(96, 200)  # ragged: pads to [128, 256]
    q, s, dq = conv(w, [64, 128], torch.bfloat16)
    save("moe_fp8_quantizer", dict(shared=cases, ragged=dict(w=w, q=q, scales=s, dq=dq, block=[64, 128])))

    # ---------------- mxfp4 (test_mxfp4.py): pack stubs = this repo's oracle formulas, so the script's own
    # "check weight scale packing format" line validates pack_scale against the reference expression ----
    g, calls, cmps, out = run_script("test_mxfp4.py", pack_scale=O.pack_scale)
    assert out.count("True") >= 3 and "False" not in out, out  # test_mxfp4.py:184-185 printed True each time
    util = g["MXFP4QuantizeUtil"]
    mm = [c[1] for c in calls if c[0] == "mxfp4_scaled_mm_cpu"]
    cws = [c[1] for c in calls if c[0] == "convert_scale_packed"]
    gemms = []
    for args, sc, (ref, _) in zip(mm, cws, cmps):
        gemms.append(dict(A=args[0], Bq=args[1], Bs=sc[0], Bs_packed=args[2], bias=args[3], ref_out=ref))
    x = torch.randn(6, 128) * torch.logspace(-3, 3, 6).unsqueeze(1)
    q, s = util.quantize(x)
    allcodes = torch.arange(256, dtype=torch.uint8).view(8, 32)  # every nibble pair
    sc = torch.tensor([[0], [1], [100], [126], [127], [128], [130], [200]] * 2, dtype=torch.uint8)
    save("mxfp4", dict(gemms=gemms, quant=dict(x=x, q=q, scale=s, dq=util.dequantize(q, torch.float32, s)),
                       codes=dict(q=allcodes, scale=sc, dq=util.dequantize(allcodes, torch.float32, sc))))

    # ---------------- prepack layouts: run the reference's own layout checks against oracle.pack_* ----------
    g, _, _, _ = run_script("test_moe.py", pack_weight=O.pack_weight_vnni)
    buf = io.StringIO()
    with redirect_stdout(buf):
        g["test_weight_prepack"](3, 64, 96)      # test_moe.py:115-124 (calls are commented out upstream)
    assert "True" in buf.getvalue(), buf.getvalue()
    # test_gemm_int8.py:80-93 needs ops that are out of scope at import time; evaluate its expression on the
    # single-block case it supports (oc = 32) directly:
    w = torch.randint(-128, 128, (32, 64), dtype=torch.int8)
    packed = O.pack_weight_vnni(w)
    ref = w.view(1, 32, 16, 4).permute(0, 2, 1, 3).contiguous().view(32, 64)
    assert torch.equal(ref, packed.view(-1).narrow(0, 0, 32 * 64).view(32, 64))
    comp = torch.matmul(torch.full((1, 64), 128, dtype=torch.int32), w.t().to(torch.int32))
    assert torch.equal(comp.view(-1), packed.view(-1)[32 * 64:].view(torch.int32))
    wb = torch.randn(3, 64, 96).to(torch.bfloat16)
    w8 = torch.randint(-128, 128, (2, 64, 128), dtype=torch.int8)
    wf = torch.randn(2, 32, 64).to(torch.float8_e4m3fn)
    sc = torch.randint(0, 256, (2, 64, 5), dtype=torch.uint8)
    save("prepack", dict(bf16=dict(w=wb, packed=O.pack_weight_vnni(wb)), int8=dict(w=w8, packed=O.pack_weight_vnni(w8)),
                         fp8=dict(w=wf, packed=O.pack_weight_vnni(wf)), scale=dict(s=sc, packed=O.pack_scale(sc))))


if __name__ == "__main__":
    main()
