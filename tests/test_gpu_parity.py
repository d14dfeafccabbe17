"""GPU parity tests: the CUDA path (through the C ABI, via the host mirror) against the CPU oracle on
the same seeded inputs, against the reference-generated golden fixtures, and through
size-independent properties at the BASELINE.json shapes.  Run with `-m gpu` on the B200 box."""
import math

import pytest
import torch

from conftest import load_golden
from oracle import moe_oracle as O

pytestmark = pytest.mark.gpu

BF16_TOL = dict(rtol=1e-2, atol=1e-2)  # utils.py:3-12 of the reference


@pytest.fixture(scope="module")
def K():
    import sgl_cpu_tests_b200 as K
    assert torch.cuda.is_available()
    return K


def dense(ids, w, E):
    out = torch.zeros(ids.shape[0], E)
    out.scatter_(1, ids.long().cpu(), w.float().cpu())
    return out


def routing(m, e, topk, dtype=torch.bfloat16):
    score = torch.softmax(torch.randn((m, e), dtype=dtype), dim=-1, dtype=torch.float32)
    w, ids = torch.topk(score, topk)
    return w, ids.to(torch.int32)


# ------------------------------------------------------------------------------------------ routers
ROUTER_SHAPES = [(123, 8, 2, 2, 1), (123, 16, 4, 3, 2), (123, 32, 4, 3, 2), (1123, 32, 4, 3, 2), (123, 64, 1, 6, 1),
                 (123, 256, 8, 4, 8), (123, 160, 8, 6, 2), (64, 128, 1, 8, 1), (3, 384, 1, 8, 1), (5, 1024, 8, 8, 4)]


@pytest.mark.parametrize("renorm", [True, False])
@pytest.mark.parametrize("shape", ROUTER_SHAPES)
def test_grouped_topk_bit_exact_ids(K, shape, renorm):
    M, E, G, topk, tg = shape
    torch.manual_seed(111)
    for gating in (torch.randn(M, E), torch.randn(M, E, dtype=torch.bfloat16),
                   torch.randn(M, E, dtype=torch.bfloat16) * 2 * M):  # last: test_grouped_topk.py:44-46
        w_ref, id_ref = O.grouped_topk(gating, topk, renorm, G, tg)
        w, ids = K.grouped_topk(torch.zeros(M, 4), gating.cuda(), topk, renorm, G, tg, 0, None, None)
        ids, w = ids.cpu(), w.cpu()
        # softmax is monotone in the logit, so picks (ids AND order) are bit-exact for every input
        # style, ties included: both sides break ties towards the lowest expert index.  Only picks
        # whose probability underflowed to (sub)zero may differ between expf implementations.
        live = (w_ref > 1e-30) & (w > 1e-30)
        assert torch.equal(ids[live], id_ref[live])
        if gating.dtype == torch.float32:
            assert torch.equal(ids, id_ref)
        assert torch.allclose(dense(ids, w, E), dense(id_ref, w_ref, E), rtol=1e-4, atol=1e-6)


def test_grouped_topk_out_variant_and_golden(K):
    for c in load_golden("router_softmax"):
        M, topk = c["M"], c["topk"]
        w = torch.empty(M, topk, dtype=torch.float32)
        ids = torch.empty(M, topk, dtype=torch.int32)
        K.grouped_topk_out(w, ids, torch.zeros(M, 4), c["gating"], topk, c["renormalize"], c["G"], c["topk_group"])
        w_o, id_o = O.grouped_topk(c["gating"], topk, c["renormalize"], c["G"], c["topk_group"])
        if c["gating"].dtype == torch.float32:
            assert torch.equal(ids, id_o)
            assert torch.allclose(w, w_o, rtol=1e-5, atol=1e-7)
            assert torch.allclose(dense(ids, w, c["E"]), dense(c["ref_ids"], c["ref_weights"], c["E"]), rtol=1e-5, atol=1e-6)
        else:
            # bf16 logits x 2M (test_grouped_topk.py:44-46) saturate the softmax: picks with a live probability are
            # bit-exact; picks whose probability underflowed are ties at zero, which both sides break identically
            # unless expf underflows one element earlier on one side.  The weights must agree either way.
            live = (w_o > 1e-30) & (w > 1e-30)
            assert torch.equal(ids[live], id_o[live])
            assert torch.allclose(dense(ids, w, c["E"]), dense(id_o, w_o, c["E"]), rtol=1e-4, atol=1e-6)


BIASED_SHAPES = [(122, 256, 8, 8, 2), (33, 64, 4, 4, 2), (7, 32, 8, 3, 3), (2048, 256, 8, 8, 4), (8192, 256, 8, 8, 4),
                 (100, 384, 1, 8, 1)]


@pytest.mark.parametrize("renorm", [True, False])
def test_biased_grouped_topk_bit_exact(K, renorm):
    """ids, their order AND the weights are bit-exact: both sides evaluate the correctly rounded sigmoid
    (csrc/router.cu header, O.sigmoid_f64), one f32 add for the bias, the same tie-break and the same f32 sum order."""
    torch.manual_seed(1111)
    for (M, E, G, topk, tg) in BIASED_SHAPES:
        for dt, scale in ((torch.float32, 1.0), (torch.bfloat16, 1.0), (torch.bfloat16, 2.0 * M)):
            gating = torch.randn(M, E, dtype=dt) * scale       # last style: test_biased_grouped_topk.py:53-55
            bias = torch.randn(E, dtype=dt)
            w_ref, id_ref = O.biased_grouped_topk(gating, bias, topk, renorm, G, tg)
            w = torch.empty(M, topk, dtype=torch.float32)
            ids = torch.empty(M, topk, dtype=torch.int32)
            K.biased_grouped_topk_out(w, ids, torch.zeros(M, 4), gating, bias, topk, renorm, G, tg)
            assert torch.equal(ids, id_ref), (M, E, dt, scale, int((ids != id_ref).any(dim=1).sum()))
            assert torch.equal(w, w_ref), (M, E, dt, scale, float((w - w_ref).abs().max()))


def test_biased_grouped_topk_golden(K):
    """The reference's own case (bf16 logits x 2M, bf16 bias, E 256 G 8 k 8 tg 2, M 122) and criterion
    (dense scatter allclose, test_biased_grouped_topk.py:87-91), plus bit-equality with the oracle."""
    for c in load_golden("router_biased"):
        w, ids = K.biased_grouped_topk(torch.zeros(c["M"], 4), c["gating"].cuda(), c["bias"].cuda(), c["topk"],
                                       c["renormalize"], c["G"], c["topk_group"])
        w, ids = w.cpu(), ids.cpu()
        w_o, id_o = O.biased_grouped_topk(c["gating"], c["bias"], c["topk"], c["renormalize"], c["G"], c["topk_group"])
        assert torch.equal(ids, id_o) and torch.equal(w, w_o)
        if c["gating"].dtype == torch.float32:
            assert torch.allclose(dense(ids, w, c["E"]), dense(c["ref_ids"], c["ref_weights"], c["E"]), rtol=1e-5, atol=1e-6)


def test_sigmoid_every_bf16_input(K):
    """All 2^16 bf16 bit patterns (finite ones) through the biased router with zero bias, topk = E = 32, no
    renormalisation: the returned weights are the kernel's sigmoid of each logit; bit-equal to O.sigmoid_f64."""
    bits = torch.arange(0, 65536, dtype=torch.int32).to(torch.int16).view(torch.bfloat16)
    gating = bits[torch.isfinite(bits.float())]
    gating = gating[: gating.numel() // 32 * 32].view(-1, 32)
    w, ids = K.biased_grouped_topk(torch.zeros(gating.shape[0], 4), gating.cuda(), torch.zeros(32).cuda(), 32, False, 1, 1)
    got = torch.zeros(gating.shape, dtype=torch.float32).scatter_(1, ids.long().cpu(), w.cpu())
    want = O.sigmoid_f64(gating.float())
    bad = (got.view(torch.int32) != want.view(torch.int32)).nonzero()
    assert bad.numel() == 0, (bad.shape[0], [(float(gating[i, j]), float(got[i, j]), float(want[i, j])) for i, j in bad[:8].tolist()])


def test_router_serving_args(K):
    """num_fused_shared_experts / routed_scaling_factor / num_token_non_padded of the 9-argument call
    (test_moe.py:61-70); parity unpinned, oracle = O._serving_args."""
    torch.manual_seed(5)
    M, E = 37, 64
    g, b = torch.randn(M, E, dtype=torch.bfloat16), torch.randn(E, dtype=torch.bfloat16)
    hs = torch.zeros(M, 4)
    for renorm in (True, False):
        for nfs, rsf, npad in ((0, 2.5, None), (1, 2.5, None), (2, None, torch.tensor([30])), (0, None, torch.tensor(5, dtype=torch.int64)),
                               (1, 16.0, 11)):
            topk = 4 + (nfs > 0)
            w, ids = K.biased_grouped_topk(hs, g.cuda(), b.cuda(), topk, renorm, 4, 2, nfs, rsf, npad if npad is None or isinstance(npad, int) else npad.cuda())
            w_o, id_o = O.biased_grouped_topk(g, b, topk, renorm, 4, 2, nfs, rsf, npad)
            assert torch.equal(ids.cpu(), id_o) and torch.equal(w.cpu(), w_o), (renorm, nfs, rsf)
            w, ids = K.grouped_topk(hs, g, topk, renorm, 4, 2, nfs, rsf, npad)
            w_o, id_o = O.grouped_topk(g, topk, renorm, 4, 2, nfs, rsf, npad)
            assert torch.equal(ids, id_o) and torch.allclose(w, w_o, rtol=1e-5, atol=1e-7), (renorm, nfs, rsf)


# ------------------------------------------------------------------------------------------ moe_align
@pytest.mark.parametrize("shape", [(1, 1, 1), (1, 8, 128), (14, 8, 8), (128, 8, 128), (129, 8, 256), (1024, 8, 256),
                                   (8192, 8, 256), (4096, 8, 128), (1000, 4, 1024), (5000, 1, 1), (0, 2, 8)])
def test_moe_align_bit_exact(K, shape):
    M, topk, E = shape
    torch.manual_seed(M + E)
    ids = torch.randint(-1, E, (M, topk), dtype=torch.int32)
    for chunk_rows in (8, 32):
        ref = O.moe_align(ids, E, chunk_rows, 128)
        got = K.moe_align(ids.cuda(), E, chunk_rows, 128)
        for k in ("expert_count", "expert_offset", "sorted_slot", "slot_pos", "counters"):
            assert torch.equal(got[k].cpu(), ref[k]), k
        for k, n in (("chunk_list", int(ref["counters"][1])), ("tile_list", int(ref["counters"][2]))):
            assert torch.equal(got[k].cpu()[:3 * n].view(-1, 3), ref[k]), k


# ------------------------------------------------------------------------------------------ fused experts
def check_close(out, ref, what=""):
    out, ref = out.float().cpu(), ref.float().cpu()
    assert torch.isfinite(out).all(), what
    assert torch.allclose(out, ref, **BF16_TOL), f"{what}: max diff {(out - ref).abs().max():.5f}"


def mean_rel(out, ref):
    out, ref = out.float().cpu(), ref.float().cpu()
    return float((out - ref).abs().mean() / ref.abs().mean())


@pytest.mark.parametrize("prepack", [False, True])
@pytest.mark.parametrize("shape", [(4, 32, 32, 4, 2), (2, 128, 32, 4, 2), (33, 64, 128, 8, 3), (224, 512, 1056, 8, 2),
                                   (128, 768, 2048, 128, 8), (1, 768, 2048, 128, 8), (17, 96, 72, 8, 8)])
def test_fused_experts_bf16(K, shape, prepack):
    m, n, k, e, topk = shape
    if prepack and ((2 * n) % 32 != 0 or k % 32 != 0):
        pytest.skip("prepack needs N % 32 == 0")
    torch.manual_seed(1111)
    a = torch.randn((m, k), dtype=torch.bfloat16) / 10           # test_moe.py:96-99
    w1 = torch.randn((e, 2 * n, k), dtype=torch.bfloat16) / 10
    w2 = torch.randn((e, k, n), dtype=torch.bfloat16) / 10
    score = torch.randn((m, e), dtype=torch.bfloat16)
    tw, ids = K.grouped_topk(a, score, topk, True, 1, 1, 0, None, None)
    ref = O.native_fused_moe(a, w1, w2, tw, ids)
    p1 = K.convert_weight_packed(w1) if prepack else w1
    p2 = K.convert_weight_packed(w2) if prepack else w2
    hs = a.clone()
    out = K.fused_experts(hs, p1, p2, tw, ids, True, K.CPUQuantMethod.UNQUANT, None, None, None, None, None, prepack)
    assert out.data_ptr() == hs.data_ptr()  # inplace: mutates and returns hidden_states
    check_close(out, ref, f"bf16 {shape}")


def test_fused_experts_bf16_golden(K):
    for c in load_golden("moe_bf16"):
        score = torch.softmax(c["score"], dim=-1, dtype=torch.float32)
        w, ids = torch.topk(score, c["topk"])
        if c["renormalize"]:
            w = w / w.sum(dim=-1, keepdim=True)
        out = K.fused_experts(c["a"].clone(), c["w1"], c["w2"], w, ids.to(torch.int32), False, False, False, None, None,
                              None, None, None, False)
        check_close(out, c["ref_out"], "bf16 golden")


@pytest.mark.parametrize("case", [(1, 128, 256, 8, 2, False), (39, 1280, 1024, 8, 3, True), (1024, 1280, 1024, 8, 3, True),
                                  (5, 64, 32, 4, 2, False), (64, 768, 2048, 128, 8, True)])
def test_fused_experts_int8(K, case):
    M, N, Kd, E, topk, prepack = case
    torch.manual_seed(1111)
    a = torch.randn((M, Kd), dtype=torch.bfloat16) / math.sqrt(Kd)   # test_moe_int8.py:97-122
    w1 = ((torch.rand((E, 2 * N, Kd)) - 0.5) * 2 * 127).clamp(-128, 127).to(torch.int8)
    w2 = ((torch.rand((E, Kd, N)) - 0.5) * 2 * 127).clamp(-128, 127).to(torch.int8)
    w1_s, w2_s = torch.rand(E, 2 * N) * 1e-2, torch.rand(E, Kd) * 1e-2
    tw, ids = routing(M, E, topk)
    ref = O.fused_experts(a, w1, w2, tw, ids, "int8", w1_s, w2_s)
    p1 = K.convert_weight_packed(w1) if prepack else w1
    p2 = K.convert_weight_packed(w2) if prepack else w2
    out = K.fused_experts(a.clone(), p1, p2, tw, ids, True, True, False, w1_s, w2_s, None, None, None, prepack)
    assert mean_rel(out, ref) < 0.01                 # test_moe_int8.py:134-137
    if M <= 64:
        check_close(out, ref, f"int8 {case}")        # the reference itself warns this may fail at large shapes (:131)


def test_fused_experts_int8_golden_and_quant(K):
    g = load_golden("moe_int8")
    for c in g["cases"]:
        out = K.fused_experts(c["a"].clone(), c["w1"], c["w2"], c["topk_weight"], c["topk_ids"], False, True, False,
                              c["w1_s"], c["w2_s"], None, None, None, False)
        check_close(out, c["ref_out"], "int8 golden")
        assert mean_rel(out, c["ref_out"]) < 0.01


@pytest.mark.parametrize("block", [[64, 128], [128, 128]])
@pytest.mark.parametrize("shape", [(2, 128, 128, 8, 4), (121, 512, 1024, 8, 2), (1212, 512, 1024, 8, 2), (16, 2048, 7168, 4, 2)])
def test_fused_experts_fp8(K, shape, block):
    M, N, Kd, E, topk = shape
    bn, bk = block
    torch.manual_seed(1111)
    a = torch.randn(M, Kd, dtype=torch.bfloat16) / math.sqrt(Kd)     # test_moe_fp8_ext.py:96-105
    w1 = (torch.randn(E, 2 * N, Kd) * 400).clamp(-400, 400).to(torch.float8_e4m3fn)
    w2 = (torch.randn(E, Kd, N) * 400).clamp(-400, 400).to(torch.float8_e4m3fn)
    w1s = torch.randn(E, 2 * N // bn, Kd // bk) * 1e-3
    w2s = torch.randn(E, Kd // bn, N // bk) * 1e-3
    tw, ids = routing(M, E, topk)
    ref = O.fused_experts(a, w1, w2, tw, ids, "fp8", w1s, w2s, block)
    out = K.fused_experts(a, K.convert_weight_packed(w1), K.convert_weight_packed(w2), tw, ids, False, False, True, w1s,
                          w2s, block, None, None, True)
    check_close(out, ref.bfloat16(), f"fp8 {shape} {block}")


def test_fused_experts_fp8_golden_and_offload(K):
    for c in load_golden("moe_fp8")["cases"] + load_golden("moe_offload"):
        out = K.fused_experts(c["a"], c["w1"], c["w2"], c["topk_weight"], c["topk_ids"], False, False, True, c["w1s"],
                              c["w2s"], c["block"], None, None, False)
        ref = c["ref_out"].float()
        if ref.abs().max() <= 1.5:
            check_close(out, ref, "fp8 golden")
        else:
            # The extra offload case (topk 8 over 4 experts, randn topk weights) sums terms of magnitude ~5
            # with both signs; one bf16 ulp of a term (0.03) already exceeds atol = 1e-2 on a cancelled
            # output.  The reference's atol is stated for O(1) outputs (utils.py:3-12), so scale it with
            # the output range here; rtol stays 1e-2.
            o = out.float().cpu()
            assert torch.allclose(o, ref, rtol=1e-2, atol=1e-2 * float(ref.abs().max())), float((o - ref).abs().max())
            assert mean_rel(out, ref) < 0.01
    # bf16 weights with -1 ids and duplicate experts in a row (test_moe_offloading_cpu.py:123-137)
    c = load_golden("moe_offload")[0]
    w1d = O.scaled_weight(c["w1"], c["w1s"], *c["block"]).bfloat16()
    w2d = O.scaled_weight(c["w2"], c["w2s"], *c["block"]).bfloat16()
    out = K.fused_experts(c["a"], w1d, w2d, c["topk_weight"], c["topk_ids"], False, False, False, None, None, None, None,
                          None, False)
    check_close(out, O.native_fused_moe(c["a"], w1d, w2d, c["topk_weight"], c["topk_ids"]), "bf16 offload")


def test_fused_experts_all_masked_and_empty(K):
    a = torch.randn(3, 64, dtype=torch.bfloat16)
    w1 = torch.randn(4, 64, 64, dtype=torch.bfloat16)
    w2 = torch.randn(4, 64, 32, dtype=torch.bfloat16)
    ids = torch.full((3, 2), -1, dtype=torch.int32)
    out = K.fused_experts(a, w1, w2, torch.ones(3, 2), ids, False, False, False, None, None, None, None, None, False)
    assert torch.equal(out, torch.zeros_like(out))
    out = K.fused_experts(a[:0], w1, w2, torch.ones(0, 2), ids[:0], False, False, False, None, None, None, None, None, False)
    assert out.shape == (0, 64)


@pytest.mark.parametrize("fmt", ["bf16", "int8", "fp8"])
@pytest.mark.parametrize("path", ["stream", "umma"])
def test_all_masked_on_a_poisoned_workspace(K, fmt, path):
    """Every id masked => the work lists are EMPTY and moe_align writes no entry at all: no kernel may look at entry 0.
    The workspace is filled with 0x7f first (torch's caching allocator hands out used memory; a fresh cudaMalloc is
    zeroed, which hid a lookahead of the gather warps that chased row indices out of an unwritten tile list)."""
    from sgl_cpu_tests_b200 import ops
    dev = torch.device("cuda", 0)
    M, E, topk, H, I = 300, 8, 4, 512, 256
    g = torch.Generator(device=dev).manual_seed(3)
    import bench
    w = bench.make_weights(fmt, E, H, I, dev, g)
    x = torch.randn(M, H, device=dev, generator=g).bfloat16()
    ids = torch.full((M, topk), -1, dtype=torch.int32, device=dev)
    tw = torch.ones(M, topk, device=dev)
    nbytes = ops.L.load().sglb_fused_experts_workspace_bytes(M, topk, E, H, I, bench.FMT_ID[fmt])
    for _ in range(2):
        ops._workspace(nbytes, dev).fill_(0x7f)
        out = K.fused_experts_impl(x, w["w1"], w["w2"], tw, ids, False, bench.FMT_ID[fmt], w["s1"], w["s2"], w["block"], False,
                                   path=K.PATH_STREAM if path == "stream" else K.PATH_UMMA)
        torch.cuda.synchronize()
        assert torch.equal(out, torch.zeros_like(out))


@pytest.mark.parametrize("shape", [(3, 64, 64, 4, 2), (37, 96, 352, 8, 4), (16, 2880, 2880, 8, 4)])
def test_fused_experts_mxfp4(K, shape):
    M, N, Kd, E, topk = shape
    torch.manual_seed(1111)
    a = torch.randn(M, Kd, dtype=torch.bfloat16) / 10
    w1q, w1s = O.mxfp4_quantize((torch.rand(E, 2 * N, Kd) - 0.5) * 2e-1)
    w2q, w2s = O.mxfp4_quantize((torch.rand(E, Kd, N) - 0.5) * 2e-1)
    w1s, w2s = w1s.view(E, 2 * N, Kd // 32), w2s.view(E, Kd, N // 32)
    tw, ids = routing(M, E, topk)
    ref = O.fused_experts(a, w1q, w2q, tw, ids, "mxfp4", w1s, w2s)
    for vnni in (False, True):
        if vnni and N % 32:
            continue
        args = (w1q, w2q, w1s, w2s) if not vnni else (K.convert_weight_packed(w1q), K.convert_weight_packed(w2q),
                                                      K.convert_scale_packed(w1s), K.convert_scale_packed(w2s))
        out = K.fused_experts(a, args[0], args[1], tw, ids, False, K.CPUQuantMethod.MXFP4, args[2], args[3], None, None,
                              None, vnni)
        check_close(out, ref, f"mxfp4 {shape} vnni={vnni}")


# gpt-oss: the 17-argument vintage (on-box vllm cpu_moe.py:308-326) -- per-expert biases and the clamped swiglu.
# PARITY UNPINNED: the reference holds no call site with these arguments; the oracle is O.native_fused_moe's
# bias / swiglu_oai form (gate = rows [0, I), up = rows [I, 2I)).
@pytest.mark.parametrize("shape", [(5, 64, 64, 4, 2), (37, 96, 352, 8, 4), (128, 2880, 2880, 8, 4), (600, 256, 512, 4, 2)])
@pytest.mark.parametrize("act", [(None, None), (1.702, 7.0), (1.702, None)])
def test_fused_experts_mxfp4_bias_swiglu_oai(K, shape, act):
    M, N, Kd, E, topk = shape
    alpha, limit = act
    torch.manual_seed(77)
    # inputs / 10 at the gpt-oss width, as the reference's own tests scale theirs (test_moe_fp8_ext.py:101): the
    # f32 oracle is held to rtol = atol = 1e-2 and bf16 intermediates of O(16) outputs sit two ulps from that
    a = torch.randn(M, Kd, dtype=torch.bfloat16) / (10 if Kd > 512 else 1)
    w1q, w1s = O.mxfp4_quantize((torch.rand(E, 2 * N, Kd) - 0.5) * 2e-1)
    w2q, w2s = O.mxfp4_quantize((torch.rand(E, Kd, N) - 0.5) * 2e-1)
    w1s, w2s = w1s.view(E, 2 * N, Kd // 32), w2s.view(E, Kd, N // 32)
    b1, b2 = torch.randn(E, 2 * N) * 0.5, torch.randn(E, Kd) * 0.5
    tw, ids = routing(M, E, topk)
    ids[0, 0] = -1
    ref = O.fused_experts(a, w1q, w2q, tw, ids, "mxfp4", w1s, w2s, w1_bias=b1, w2_bias=b2, alpha=alpha, limit=limit)
    out = K.fused_experts(a, w1q, w2q, tw, ids, False, K.CPUQuantMethod.MXFP4, w1s, w2s, None, None, None, b1, b2, alpha,
                          limit, False)
    check_close(out, ref, f"mxfp4 bias/act {shape} {act}")
    if Kd % 64 == 0 and N % 64 == 0:      # the dequantising tcgen05 kernel and the streaming kernel take the same epilogue
        import sgl_cpu_tests_b200._lib as L
        L.load().sglb_set_mx4_native(0)
        try:
            for path in (K.PATH_STREAM, K.PATH_UMMA):
                o2 = K.fused_experts_impl(a.cuda(), w1q.cuda(), w2q.cuda(), tw.cuda(), ids.cuda(), False, K.FMT_MXFP4, w1s.cuda(),
                                          w2s.cuda(), None, False, path=path, w1_bias=b1.cuda(), w2_bias=b2.cuda(), alpha=alpha,
                                          limit=limit)
                check_close(o2, ref, f"mxfp4 bias/act {shape} {act} path={path}")
        finally:
            L.load().sglb_set_mx4_native(1)


@pytest.mark.parametrize("fmt", ["bf16", "fp8"])
def test_fused_experts_bias_swiglu_oai_other_formats(K, fmt):
    """The same epilogue arguments through the bf16 (stream + tcgen05) and fp8 tensor-native kernels."""
    torch.manual_seed(78)
    M, I, H, E, topk = 90, 256, 512, 8, 3
    a = torch.randn(M, H, dtype=torch.bfloat16) / 4
    b1, b2 = torch.randn(E, 2 * I) * 0.3, torch.randn(E, H) * 0.3
    tw, ids = routing(M, E, topk)
    if fmt == "bf16":
        w1, w2 = (torch.randn(E, 2 * I, H) / 10).bfloat16(), (torch.randn(E, H, I) / 10).bfloat16()
        s1 = s2 = block = None
        fid = K.FMT_BF16
    else:
        w1 = (torch.randn(E, 2 * I, H) * 30).to(torch.float8_e4m3fn)
        w2 = (torch.randn(E, H, I) * 30).to(torch.float8_e4m3fn)
        s1, s2, block = torch.rand(E, 2 * I // 128, H // 128) * 4e-3, torch.rand(E, H // 128, I // 128) * 4e-3, [128, 128]
        fid = K.FMT_FP8_W8A16
    ref = O.fused_experts(a, w1, w2, tw, ids, fmt, s1, s2, block, w1_bias=b1, w2_bias=b2, alpha=1.702, limit=7.0)
    dev = lambda t: None if t is None else t.cuda()  # noqa: E731
    for path in (K.PATH_STREAM, K.PATH_UMMA):
        out = K.fused_experts_impl(dev(a), dev(w1), dev(w2), dev(tw), dev(ids), False, fid, dev(s1), dev(s2), block, False, path=path,
                                   w1_bias=dev(b1), w2_bias=dev(b2), alpha=1.702, limit=7.0)
        check_close(out, ref, f"{fmt} bias/act path={path}")


# ------------------------------------------------------------------------------------------ shared expert
@pytest.mark.parametrize("shape", [(2, 32, 32, 16), (121, 128, 64, 16), (1211, 128, 64, 16)])
def test_shared_expert_bf16_int8(K, shape):
    m, n, k, rsf = shape
    torch.manual_seed(1111)
    hs = torch.randn(m, k, dtype=torch.bfloat16) / k                  # test_shared_experts.py:57-60
    w1 = torch.randn(2 * n, k, dtype=torch.bfloat16)
    w2 = torch.randn(k, n, dtype=torch.bfloat16)
    fo = torch.randn(m, k, dtype=torch.bfloat16) / k
    hs2 = hs.clone()
    ref = O.shared_expert(hs, w1, w2, fo, rsf, "bf16")
    res = K.shared_expert(hs, w1, w2, fo, rsf, True, False, False, None, None, None, False)
    assert res.data_ptr() == hs.data_ptr()
    check_close(res, ref, "shared bf16")
    w1_q, w1_s = O.per_token_quant_int8(w1)
    w2_q, w2_s = O.per_token_quant_int8(w2)
    ref2 = O.shared_expert(hs2, w1_q, w2_q, fo, rsf, "int8", w1_s, w2_s)
    res2 = K.shared_expert(hs2, w1_q, w2_q, fo, rsf, True, True, False, w1_s, w2_s, None, False)
    check_close(res2, ref2, "shared int8")


def test_shared_expert_golden(K):
    for c in load_golden("shared_expert"):
        out = K.shared_expert(c["hs"].clone(), c["w1"], c["w2"], c["fused_out"], c["rsf"], False, False, False, None, None,
                              None, False)
        check_close(out, c["ref_bf16"], "shared golden bf16")
        out = K.shared_expert(c["hs"].clone(), c["w1_q"], c["w2_q"], c["fused_out"], c["rsf"], False, True, False,
                              c["w1_s"], c["w2_s"], None, None, None, False)
        check_close(out, c["ref_int8"], "shared golden int8")
    for c in load_golden("moe_fp8")["shared"] + load_golden("moe_fp8_quantizer")["shared"]:
        for vnni in (False, True):
            w1 = K.convert_weight_packed(c["w1"]) if vnni else c["w1"]
            w2 = K.convert_weight_packed(c["w2"]) if vnni else c["w2"]
            out = K.shared_expert(c["a"].clone(), w1, w2, c["fused_out"], c["rsf"], True, False, True, c["w1s"], c["w2s"],
                                  c["block"], None, None, vnni)
            check_close(out, c["ref_out"], "shared golden fp8")


# ------------------------------------------------------------------------------------------ mxfp4 linear + prepack
def test_mxfp4_scaled_mm_golden(K):
    for m in load_golden("mxfp4")["gemms"]:
        Bp = K.convert_weight_packed(m["Bq"])
        Sp = K.convert_scale_packed(m["Bs"])
        assert torch.equal(Sp, m["Bs_packed"])                       # test_mxfp4.py:185
        out = K.mxfp4_scaled_mm(m["A"], Bp, Sp, m["bias"], True)
        check_close(out, m["ref_out"], "mxfp4 mm")


def test_mxfp4_all_codes(K):
    c = load_golden("mxfp4")["codes"]          # every nibble pair x extreme e8m0 scales, as a GEMM against identity
    q, sc = c["q"], c["scale"]                 # [8, 32] bytes -> 64 values per row; scales [16,1] -> [8,2]
    dq = c["dq"]                               # [8, 64] f32
    rows = [i for i in range(8) if torch.isfinite(dq[i]).all() and dq[i].abs().max() < 1e30 and
            (dq[i] == 0).logical_or(dq[i].abs() > 1e-30).all()]
    eye = torch.eye(64, dtype=torch.bfloat16)
    out = K.mxfp4_scaled_mm(eye, q, sc.view(8, 2), None, False)     # out[k, n] = dq[n, k]
    got = out.float().t()
    assert torch.equal(got[rows], dq[rows].bfloat16().float())


def test_prepack_layouts_bit_exact(K):
    g = load_golden("prepack")
    for k in ("bf16", "int8", "fp8"):
        for dev in ("cpu", "cuda"):
            w = g[k]["w"].to(dev)
            p = K.convert_weight_packed(w)
            assert p.device == w.device
            assert torch.equal(p.cpu().view(torch.uint8), g[k]["packed"].view(torch.uint8)), k
    assert torch.equal(K.convert_scale_packed(g["scale"]["s"]), g["scale"]["packed"])
    w = torch.randn(256, 16 * 8, 32 * 24).to(torch.bfloat16)  # test_moe.py:126 (commented call), test_gemm.py:46 shape
    assert torch.equal(K.convert_weight_packed(w), O.pack_weight_vnni(w))


def test_building_blocks(K):
    import ctypes
    from sgl_cpu_tests_b200 import _lib as L
    lib = L.load()
    g = load_golden("moe_int8")["quant"]
    x = g["x"].cuda()
    q = torch.empty(x.shape, dtype=torch.int8, device="cuda")
    s = torch.empty(x.shape[0], dtype=torch.float32, device="cuda")
    L.check(lib.sglb_quant_rows_int8(x.data_ptr(), x.shape[0], x.shape[1], q.data_ptr(), s.data_ptr(), None))
    torch.cuda.synchronize()
    assert torch.equal(q.cpu(), g["q"]) and torch.equal(s.cpu(), g["scale"].view(-1))   # bit-exact


def test_errors_surface_as_runtime_error(K):
    a = torch.randn(4, 64, dtype=torch.bfloat16)
    w1 = torch.randn(4, 64, 64, dtype=torch.bfloat16)
    w2 = torch.randn(4, 64, 32, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError):
        K.fused_experts(a, w1, w2, torch.ones(4, 2), torch.zeros(4, 2, dtype=torch.int32), False, True, False, None, None,
                        None, None, None, False)     # int8 flag with bf16 weights / no scales
    with pytest.raises(RuntimeError):
        K.fused_experts(a, w1, w2[:, :32], torch.ones(4, 2), torch.zeros(4, 2, dtype=torch.int32), False, False, False,
                        None, None, None, None, None, False)
    with pytest.raises(RuntimeError):
        K.grouped_topk(a, torch.randn(5, 8), 2, False, 1, 1, 0, None, None)  # token count mismatch
    with pytest.raises(RuntimeError):
        K.grouped_topk(a, torch.randn(4, 8), 2, False, 3, 1, 0, None, None)  # G does not divide E


# ------------------------------------------------------------------------------------------ CUDA graph replay
@pytest.mark.parametrize("fmt", ["bf16", "int8"])
def test_graphed_experts_matches_eager(K, fmt):
    """K.GraphedExperts (one captured call, replayed) is bit-identical to the eager call, also after the
    static input buffers are refilled."""
    torch.manual_seed(7)
    M, E, topk, H, I = 33, 16, 4, 256, 128
    dev = torch.device("cuda")
    if fmt == "bf16":
        w1 = (torch.randn(E, 2 * I, H) / 10).bfloat16().to(dev)
        w2 = (torch.randn(E, H, I) / 10).bfloat16().to(dev)
        s1 = s2 = None
        f = K.FMT_BF16
    else:
        w1 = torch.randint(-127, 128, (E, 2 * I, H), dtype=torch.int8).to(dev)
        w2 = torch.randint(-127, 128, (E, H, I), dtype=torch.int8).to(dev)
        s1, s2 = (torch.rand(E, 2 * I) * 1e-2).to(dev), (torch.rand(E, H) * 1e-2).to(dev)
        f = K.FMT_INT8_W8A8
    g = K.GraphedExperts(M, w1, w2, topk, f, s1, s2)
    assert g.kernels_per_replay >= 3  # fused decode chain: prologue, w13, w2
    for seed in (1, 2):
        torch.manual_seed(seed)
        x = (torch.randn(M, H) / 4).bfloat16().to(dev)
        tw, ids = routing(M, E, topk)
        out = g(x, tw.to(dev), ids.to(dev)).clone()
        ref = K.fused_experts_impl(x, w1, w2, tw.to(dev), ids.to(dev), False, f, s1, s2, None, False)
        torch.cuda.synchronize()
        assert torch.equal(out, ref)


def test_host_graph_path_matches_eager(K):
    """Host-tensor calls replay a cached CUDA graph (ops._host_graph); results are bit-identical to the eager
    launches, for repeated calls with new inputs, for inplace and out-of-place forms and with masked ids."""
    torch.manual_seed(11)
    M, E, topk, H, I = 19, 8, 3, 128, 64
    w1 = (torch.randn(E, 2 * I, H) / 10).bfloat16()
    w2 = (torch.randn(E, H, I) / 10).bfloat16()
    assert K.ops._HOST_GRAPH
    for seed in (1, 2, 3):
        torch.manual_seed(seed)
        x = (torch.randn(M, H) / 4).bfloat16()
        tw, ids = routing(M, E, topk)
        if seed == 3:
            ids[::2, 0] = -1
        out_g = K.fused_experts_impl(x, w1, w2, tw, ids, False, K.FMT_BF16, None, None, None, False)
        xin = x.clone()
        ret = K.fused_experts_impl(xin, w1, w2, tw, ids, True, K.FMT_BF16, None, None, None, False)
        assert ret is xin and torch.equal(xin, out_g)
        K.ops._HOST_GRAPH = False
        try:
            out_e = K.fused_experts_impl(x, w1, w2, tw, ids, False, K.FMT_BF16, None, None, None, False)
        finally:
            K.ops._HOST_GRAPH = True
        assert torch.equal(out_g, out_e)
    assert any(g is not None for g in K.ops._host_graphs.values())
