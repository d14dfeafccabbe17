"""GPU parity of the tcgen05 (UMMA) grouped-GEMM path: forced with path=PATH_UMMA and compared with the
CPU oracle and with the weight-streaming path on the same inputs (the two CUDA paths share the align
pass, the int8 quantisation kernels and the top-k reduce, so they must agree far inside the bf16
tolerance).  Run with `-m gpu` on the B200 box."""
import math

import pytest
import torch

from oracle import moe_oracle as O

pytestmark = pytest.mark.gpu

BF16_TOL = dict(rtol=1e-2, atol=1e-2)  # utils.py:3-12 of the reference


@pytest.fixture(scope="module")
def K():
    import sgl_cpu_tests_b200 as K
    assert torch.cuda.is_available()
    return K


def describe(out, ref):
    d = (out.float() - ref.float()).abs()
    bad = d > (1e-2 + 1e-2 * ref.float().abs())
    rows = bad.any(dim=1).nonzero().flatten().tolist()
    cols = bad.any(dim=0).nonzero().flatten().tolist()
    return (f"max diff {d.max():.5f}, {int(bad.sum())}/{bad.numel()} bad, bad rows {rows[:12]}... ({len(rows)}), "
            f"bad cols {cols[:12]}... ({len(cols)}), out[0,:4]={out[0, :4].tolist()} ref[0,:4]={ref[0, :4].tolist()}")


def make_ids(M, E, topk, style, gen):
    if style == "uniform":
        score = torch.rand(M, E, generator=gen)
        return torch.topk(score, topk).indices.to(torch.int32)
    if style == "skewed":      # a few hot experts (many tiles each) + masked slots + duplicates in a row
        ids = torch.randint(0, max(1, E // 4), (M, topk), generator=gen, dtype=torch.int32)
        ids[torch.rand(M, topk, generator=gen) < 0.1] = -1
        return ids
    if style == "one":         # everything on one expert: rows = M*topk, several full 128-row tiles
        return torch.full((M, topk), E - 1, dtype=torch.int32)
    raise ValueError(style)


CASES = [  # M, I, H, E, topk, style
    (64, 128, 128, 4, 2, "uniform"),
    (300, 128, 256, 8, 2, "uniform"),
    (257, 192, 320, 8, 4, "skewed"),      # partial weight blocks in both stages, ragged tiles, -1 ids
    (130, 256, 128, 4, 2, "one"),
    (1, 128, 128, 2, 1, "uniform"),       # a single routed row
    (148, 128, 256, 4, 4, "skewed"),      # the expert-parallel local shape: one hot expert, masked ids
    (1024, 768, 2048, 128, 8, "uniform"),  # Qwen3-30B-A3B layer shape
]


@pytest.mark.parametrize("case", CASES)
def test_umma_bf16(K, case):
    M, I, H, E, topk, style = case
    g = torch.Generator().manual_seed(1111)
    a = torch.randn(M, H, generator=g).bfloat16() / 10
    w1 = (torch.randn(E, 2 * I, H, generator=g) / 10).bfloat16()
    w2 = (torch.randn(E, H, I, generator=g) / 10).bfloat16()
    ids = make_ids(M, E, topk, style, g)
    tw = torch.rand(M, topk, generator=g)
    ref = O.native_fused_moe(a, w1, w2, tw, ids)
    ad, w1d, w2d, twd, idd = a.cuda(), w1.cuda(), w2.cuda(), tw.cuda(), ids.cuda()
    out_s = K.fused_experts_impl(ad, w1d, w2d, twd, idd, False, K.FMT_BF16, None, None, None, False, path=K.PATH_STREAM)
    out_u = K.fused_experts_impl(ad, w1d, w2d, twd, idd, False, K.FMT_BF16, None, None, None, False, path=K.PATH_UMMA)
    torch.cuda.synchronize()
    out_s, out_u = out_s.cpu(), out_u.cpu()
    assert torch.isfinite(out_u).all(), describe(out_u, ref)
    assert torch.allclose(out_u.float(), ref.float(), **BF16_TOL), "vs oracle: " + describe(out_u, ref)
    assert torch.allclose(out_u.float(), out_s.float(), **BF16_TOL), "vs stream: " + describe(out_u, out_s)


INT8_CASES = [c for c in CASES if c[1] % 128 == 0 and c[2] % 128 == 0] + [
    (257, 256, 384, 8, 4, "skewed"),      # partial weight block in stage 2 (H = 384 = 256 + 128)
]


@pytest.mark.parametrize("case", INT8_CASES)
def test_umma_int8(K, case):
    M, I, H, E, topk, style = case
    g = torch.Generator().manual_seed(1111)
    a = (torch.randn(M, H, generator=g) / math.sqrt(H)).bfloat16()
    w1 = torch.randint(-127, 128, (E, 2 * I, H), generator=g, dtype=torch.int8)
    w2 = torch.randint(-127, 128, (E, H, I), generator=g, dtype=torch.int8)
    s1, s2 = torch.rand(E, 2 * I, generator=g) * 1e-2, torch.rand(E, H, generator=g) * 1e-2
    ids = make_ids(M, E, topk, style, g)
    tw = torch.rand(M, topk, generator=g)
    ad, w1d, w2d, twd, idd, s1d, s2d = (t.cuda() for t in (a, w1, w2, tw, ids, s1, s2))
    out_s = K.fused_experts_impl(ad, w1d, w2d, twd, idd, False, K.FMT_INT8_W8A8, s1d, s2d, None, False, path=K.PATH_STREAM)
    out_u = K.fused_experts_impl(ad, w1d, w2d, twd, idd, False, K.FMT_INT8_W8A8, s1d, s2d, None, False, path=K.PATH_UMMA)
    torch.cuda.synchronize()
    out_s, out_u = out_s.cpu(), out_u.cpu()
    assert torch.isfinite(out_u).all(), describe(out_u, out_s)
    # integer accumulation is exact on both paths and the epilogues use the same f32 expressions
    assert torch.equal(out_u, out_s), "vs stream: " + describe(out_u, out_s)
    if M * topk * H * I <= 300 * 2 * 256 * 128 * 8:
        ref = O.fused_experts(a, w1, w2, tw, ids, "int8", s1, s2)
        rel = float((out_u.float() - ref.float()).abs().mean() / ref.float().abs().mean().clamp_min(1e-12))
        assert rel < 0.01, f"mean rel {rel}: " + describe(out_u, ref)      # test_moe_int8.py:134-137


def test_umma_unsupported_shape_raises(K):
    a = torch.randn(8, 96, dtype=torch.bfloat16, device="cuda")    # 96 * 2 B is not a whole 128-B K slab
    w1 = torch.randn(2, 64, 96, dtype=torch.bfloat16, device="cuda")
    w2 = torch.randn(2, 96, 32, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(RuntimeError):
        K.fused_experts_impl(a, w1, w2, torch.ones(8, 1, device="cuda"), torch.zeros(8, 1, dtype=torch.int32, device="cuda"),
                             False, K.FMT_BF16, None, None, None, False, path=K.PATH_UMMA)


DQ_CASES = [  # M, I, H, E, topk, style
    (64, 128, 128, 4, 2, "uniform"),
    (300, 128, 256, 8, 2, "uniform"),
    (257, 192, 320, 8, 4, "skewed"),       # K = 320 / 192: 5 and 3 stages of 64; partial weight blocks
    (130, 256, 128, 4, 2, "one"),
    (148, 128, 256, 4, 4, "skewed"),
    (512, 512, 1024, 16, 4, "uniform"),
]


F8_CASES = [  # shapes the tensor-native fp8 kernel takes (H, I multiples of 128); M, I, H, E, topk, style
    (1, 128, 128, 2, 1, "uniform"),
    (33, 256, 384, 4, 2, "uniform"),       # K = 384 / 256: 3 and 2 scale blocks; partial weight block in stage 2
    (300, 128, 256, 8, 2, "uniform"),
    (130, 256, 128, 4, 2, "one"),          # 260 rows on one expert: two full tiles + a 4-row tile
    (148, 128, 256, 4, 4, "skewed"),
    (64, 2048, 7168, 4, 2, "uniform"),     # DeepSeek-V3 expert shape, 56 scale blocks along K
]


@pytest.mark.parametrize("native", [0, 1])
@pytest.mark.parametrize("block", [[128, 128], [64, 128], [64, 64], [128, 256]])
@pytest.mark.parametrize("case", DQ_CASES + F8_CASES)
def test_umma_fp8(K, case, block, native):
    """native = 1: e4m3 weights on the fp8 tensor cores with hi/lo-split activations (umma_f8.cu) wherever
    the shape allows (H, I % 128 == 0, block_k % 128 == 0, block_n % 32 == 0), else the dequantising kernel;
    native = 0: always the dequantising kernel (umma_dq.cu)."""
    M, I, H, E, topk, style = case
    bn, bk = block
    if native and (case in DQ_CASES) and not (H % 128 == 0 and I % 128 == 0 and bk % 128 == 0):
        pytest.skip("falls back to the dequantising kernel, covered by native = 0")
    if not native and case in F8_CASES and case[2] >= 4096:
        pytest.skip("large shape covered once")
    g = torch.Generator().manual_seed(1111)
    a = (torch.randn(M, H, generator=g) / math.sqrt(H)).bfloat16()
    a[0, :8] *= 2.0 ** -12                                                                 # far below the row maximum
    w1 = (torch.randn(E, 2 * I, H, generator=g) * 400).clamp(-448, 448).to(torch.float8_e4m3fn)
    w2 = (torch.randn(E, H, I, generator=g) * 400).clamp(-448, 448).to(torch.float8_e4m3fn)
    w1.view(torch.uint8)[0, :4, :8] = torch.arange(32, dtype=torch.uint8).view(4, 8)   # e4m3 subnormals / tiny codes
    cd = lambda x, y: (x + y - 1) // y
    w1s = torch.randn(E, cd(2 * I, bn), cd(H, bk), generator=g) * 1e-3
    w2s = torch.randn(E, cd(H, bn), cd(I, bk), generator=g) * 1e-3
    ids = make_ids(M, E, topk, style, g)
    tw = torch.rand(M, topk, generator=g) / topk     # keeps |out| < 1 so one bf16 ulp stays inside atol
    ref = O.fused_experts(a, w1, w2, tw, ids, "fp8", w1s, w2s, block)
    dev = [t.cuda() for t in (a, w1, w2, tw, ids, w1s, w2s)]
    outs = {}
    lib = K._lib.load()
    lib.sglb_set_fp8_native(native)
    try:
        for name, path in (("stream", K.PATH_STREAM), ("umma", K.PATH_UMMA)):
            outs[name] = K.fused_experts_impl(dev[0], dev[1], dev[2], dev[3], dev[4], False, K.FMT_FP8_W8A16, dev[5], dev[6],
                                              block, False, path=path).cpu()
    finally:
        lib.sglb_set_fp8_native(1)
    assert torch.isfinite(outs["umma"]).all(), describe(outs["umma"], ref)
    assert torch.allclose(outs["umma"].float(), ref.float(), **BF16_TOL), "vs oracle: " + describe(outs["umma"], ref)
    assert torch.allclose(outs["umma"].float(), outs["stream"].float(), **BF16_TOL), "vs stream: " + describe(outs["umma"], outs["stream"])


def test_split_rows_e4m3_is_exact_near_the_row_maximum(K):
    """The hi/lo split behind the tensor-native fp8 path: (hi + lo) * 2^e reproduces every bf16 element
    within 2^-9 of its row maximum exactly, and the rest to 2^-17 of the row maximum."""
    g = torch.Generator().manual_seed(7)
    x = torch.randn(64, 512, generator=g).bfloat16()
    x[1] *= 1e-6
    x[2] *= 1e6
    x[3] = 0
    x[4, ::2] *= 2.0 ** -11
    hi, lo, rexp = K.split_rows_e4m3(x.cuda())
    rec = (hi.cpu().view(torch.float8_e4m3fn).float() + lo.cpu().view(torch.float8_e4m3fn).float()) * rexp.cpu()[:, None]
    xf = x.float()
    rowmax = xf.abs().amax(dim=1, keepdim=True)
    near = xf.abs() >= rowmax * 2.0 ** -9
    assert torch.equal(rec[near], xf[near])
    assert ((rec - xf).abs() <= rowmax * 2.0 ** -17).all()


MX4_CASES = DQ_CASES + [
    (64, 2880, 2880, 4, 2, "uniform"),    # gpt-oss expert shape: K = 22.5 stages of 128, 22.5 row blocks
    (1, 128, 128, 2, 1, "uniform"),
    (5, 64, 192, 2, 2, "uniform"),        # I = 64: half a weight block per sub-tile
    (700, 256, 384, 2, 2, "one"),         # 1400 rows on one expert: 12 full 112-row tiles + a ragged one
]


@pytest.mark.parametrize("native", [0, 1])
@pytest.mark.parametrize("case", MX4_CASES)
def test_umma_mxfp4(K, case, native):
    """native = 1: e2m1 weights and e8m0 scales straight into tcgen05.mma kind::mxf8f6f4.block_scale
    (umma_mx4.cu); native = 0: the kernel that dequantises to bf16 in shared memory (umma_dq.cu)."""
    M, I, H, E, topk, style = case
    if not native and case not in DQ_CASES and case[2] != 2880:
        pytest.skip("dequantising kernel covered by the shared cases")
    g = torch.Generator().manual_seed(1111)
    a = torch.randn(M, H, generator=g).bfloat16() / 10
    a[0, :8] *= 2.0 ** -12
    w1q, w1s = O.mxfp4_quantize((torch.rand(E, 2 * I, H, generator=g) - 0.5) * 2e-1)
    w2q, w2s = O.mxfp4_quantize((torch.rand(E, H, I, generator=g) - 0.5) * 2e-1)
    w1s, w2s = w1s.view(E, 2 * I, H // 32).contiguous(), w2s.view(E, H, I // 32).contiguous()
    ids = make_ids(M, E, topk, style, g)
    tw = torch.rand(M, topk, generator=g)
    ref = O.fused_experts(a, w1q, w2q, tw, ids, "mxfp4", w1s, w2s)
    dev = [t.cuda() for t in (a, w1q, w2q, tw, ids, w1s, w2s)]
    outs = {}
    lib = K._lib.load()
    lib.sglb_set_mx4_native(native)
    try:
        for name, path in (("stream", K.PATH_STREAM), ("umma", K.PATH_UMMA)):
            outs[name] = K.fused_experts_impl(dev[0], dev[1], dev[2], dev[3], dev[4], False, K.FMT_MXFP4, dev[5], dev[6],
                                              None, False, path=path).cpu()
    finally:
        lib.sglb_set_mx4_native(1)
    assert torch.isfinite(outs["umma"]).all(), describe(outs["umma"], ref)
    assert torch.allclose(outs["umma"].float(), ref.float(), **BF16_TOL), "vs oracle: " + describe(outs["umma"], ref)
    assert torch.allclose(outs["umma"].float(), outs["stream"].float(), **BF16_TOL), "vs stream: " + describe(outs["umma"], outs["stream"])


@pytest.mark.parametrize("shape", [(3, 256, 0, 256), (1, 5760, 2880, 2880), (1, 200, 72, 96)])
def test_mx4_scale_tiles_layout(K, shape):
    """sglb_mx4_tile_scales against the layout written out in numpy: tile (e, block, k stage) holds the
    128 rows x 4 k-groups at byte (r % 32) * 16 + (r // 32) * 4 + g, rows tiled per [0, split) / [split, rows)
    segment, padding = 127 (2^0)."""
    E, rows, split, Kdim = shape
    G = Kdim // 32
    g = torch.Generator().manual_seed(5)
    sc = torch.randint(100, 140, (E, rows, G), generator=g, dtype=torch.uint8)
    lib = K._lib.load()
    n = lib.sglb_mx4_scale_tiles_bytes(E, rows, split, Kdim)
    tiles = torch.empty(n, dtype=torch.uint8, device="cuda")
    K._lib.check(lib.sglb_mx4_tile_scales(sc.cuda().data_ptr(), tiles.data_ptr(), E, rows, split, Kdim, None))
    torch.cuda.synchronize()
    cd = lambda x, y: (x + y - 1) // y
    segs = [(0, split), (split, rows)] if split else [(0, rows)]
    nkb = cd(Kdim, 128)
    ref = []
    for e in range(E):
        for lo, hi in segs:
            for b in range(cd(hi - lo, 128)):
                for kb in range(nkb):
                    t = torch.full((512,), 127, dtype=torch.uint8)
                    for r in range(128):
                        row = lo + b * 128 + r
                        for q in range(4):
                            if row < hi and kb * 4 + q < G:
                                t[(r % 32) * 16 + (r // 32) * 4 + q] = sc[e, row, kb * 4 + q]
                    ref.append(t)
    ref = torch.cat(ref)
    assert ref.numel() == n
    assert torch.equal(tiles.cpu(), ref)


@pytest.mark.parametrize("fmt", ["bf16", "int8", "fp8"])
@pytest.mark.parametrize("shape", [(300, 256, 512), (1024, 512, 1024)])
def test_shared_expert_umma(K, shape, fmt):
    """shared_expert with enough rows takes the tcgen05 kernels (E = 1 tile list): dense 2-GEMM MLP +
    fused_out * routed_scaling_factor epilogue vs the oracle (test_shared_experts.py:34-53)."""
    M, I, H = shape
    g = torch.Generator().manual_seed(1111)
    hs = (torch.randn(M, H, generator=g) / math.sqrt(H)).bfloat16()
    fo = (torch.randn(M, H, generator=g) / 16).bfloat16()
    rsf = 2.5
    if fmt == "bf16":
        w1 = (torch.randn(2 * I, H, generator=g) / 10).bfloat16()
        w2 = (torch.randn(H, I, generator=g) / 10).bfloat16()
        ref = O.shared_expert(hs, w1, w2, fo, rsf, "bf16")
        out = K.shared_expert(hs.clone(), w1, w2, fo, rsf, False, False, False, None, None, None, False)
    elif fmt == "int8":
        w1, s1 = O.per_token_quant_int8((torch.randn(2 * I, H, generator=g) / 10).bfloat16())
        w2, s2 = O.per_token_quant_int8((torch.randn(H, I, generator=g) / 10).bfloat16())
        ref = O.shared_expert(hs, w1, w2, fo, rsf, "int8", s1, s2)
        out = K.shared_expert(hs.clone(), w1, w2, fo, rsf, False, True, False, s1, s2, None, False)
    else:
        w1 = (torch.randn(2 * I, H, generator=g) * 100).clamp(-448, 448).to(torch.float8_e4m3fn)
        w2 = (torch.randn(H, I, generator=g) * 100).clamp(-448, 448).to(torch.float8_e4m3fn)
        s1 = torch.randn(2 * I // 128, H // 128, generator=g) * 1e-3
        s2 = torch.randn(H // 128, I // 128, generator=g) * 1e-3
        ref = O.shared_expert(hs, w1, w2, fo, rsf, "fp8", s1, s2, [128, 128])
        out = K.shared_expert(hs.clone(), w1, w2, fo, rsf, False, False, True, s1, s2, [128, 128], None, None, False)
    assert torch.allclose(out.float(), ref.float(), **BF16_TOL), describe(out, ref)


PAIR_CASES = [  # M, I, H, E, topk, style
    (64, 128, 128, 4, 2, "uniform"),
    (300, 128, 256, 8, 2, "uniform"),
    (257, 192, 320, 8, 4, "skewed"),      # partial weight blocks, ragged pair tiles, -1 ids
    (130, 256, 128, 4, 2, "one"),         # 260 rows on one expert: a full 256-row pair tile + a 4-row one
    (1, 128, 128, 2, 1, "uniform"),
    (700, 256, 384, 2, 2, "one"),         # 1400 rows: five full pair tiles + 120 rows
    (1024, 768, 2048, 128, 8, "uniform"),
]


@pytest.mark.parametrize("case", PAIR_CASES)
def test_umma2_bf16(K, case):
    """CTA-pair kernel (tcgen05.mma.cta_group::2, umma2_gemm.cu) vs the oracle and the single-CTA kernel."""
    M, I, H, E, topk, style = case
    g = torch.Generator().manual_seed(1111)
    a = torch.randn(M, H, generator=g).bfloat16() / 10
    w1 = (torch.randn(E, 2 * I, H, generator=g) / 10).bfloat16()
    w2 = (torch.randn(E, H, I, generator=g) / 10).bfloat16()
    ids = make_ids(M, E, topk, style, g)
    tw = torch.rand(M, topk, generator=g) / (topk if style == "one" else 1)
    ref = O.native_fused_moe(a, w1, w2, tw, ids)
    ad, w1d, w2d, twd, idd = a.cuda(), w1.cuda(), w2.cuda(), tw.cuda(), ids.cuda()
    out_u = K.fused_experts_impl(ad, w1d, w2d, twd, idd, False, K.FMT_BF16, None, None, None, False, path=K.PATH_UMMA)
    out_p = K.fused_experts_impl(ad, w1d, w2d, twd, idd, False, K.FMT_BF16, None, None, None, False, path=K.PATH_UMMA2)
    torch.cuda.synchronize()
    out_u, out_p = out_u.cpu(), out_p.cpu()
    assert torch.isfinite(out_p).all(), describe(out_p, ref)
    assert torch.allclose(out_p.float(), ref.float(), **BF16_TOL), "vs oracle: " + describe(out_p, ref)
    assert torch.allclose(out_p.float(), out_u.float(), **BF16_TOL), "vs single-CTA: " + describe(out_p, out_u)


@pytest.mark.parametrize("case", [c for c in PAIR_CASES if c[1] % 128 == 0 and c[2] % 128 == 0])
def test_umma2_int8(K, case):
    M, I, H, E, topk, style = case
    g = torch.Generator().manual_seed(1111)
    a = (torch.randn(M, H, generator=g) / math.sqrt(H)).bfloat16()
    w1 = torch.randint(-127, 128, (E, 2 * I, H), generator=g, dtype=torch.int8)
    w2 = torch.randint(-127, 128, (E, H, I), generator=g, dtype=torch.int8)
    s1, s2 = torch.rand(E, 2 * I, generator=g) * 1e-2, torch.rand(E, H, generator=g) * 1e-2
    ids = make_ids(M, E, topk, style, g)
    tw = torch.rand(M, topk, generator=g)
    dev = [t.cuda() for t in (a, w1, w2, tw, ids, s1, s2)]
    out_u = K.fused_experts_impl(*dev[:5], False, K.FMT_INT8_W8A8, dev[5], dev[6], None, False, path=K.PATH_UMMA).cpu()
    out_p = K.fused_experts_impl(*dev[:5], False, K.FMT_INT8_W8A8, dev[5], dev[6], None, False, path=K.PATH_UMMA2).cpu()
    assert torch.isfinite(out_p).all(), describe(out_p, out_u)
    # exact integer accumulation and the same f32 epilogue expressions on both kernels
    assert torch.equal(out_p, out_u), "vs single-CTA: " + describe(out_p, out_u)


@pytest.mark.parametrize("block", [[128, 128], [64, 128], [128, 256]])
@pytest.mark.parametrize("case", F8_CASES + [(700, 256, 384, 2, 2, "one"), (1024, 512, 1024, 16, 4, "uniform")])
def test_umma2_fp8(K, case, block):
    """Tensor-native fp8 kernel on a CTA pair (umma2_f8.cu) vs the oracle and the single-CTA kernel."""
    M, I, H, E, topk, style = case
    bn, bk = block
    g = torch.Generator().manual_seed(1111)
    a = (torch.randn(M, H, generator=g) / math.sqrt(H)).bfloat16()
    w1 = (torch.randn(E, 2 * I, H, generator=g) * 400).clamp(-448, 448).to(torch.float8_e4m3fn)
    w2 = (torch.randn(E, H, I, generator=g) * 400).clamp(-448, 448).to(torch.float8_e4m3fn)
    cd = lambda x, y: (x + y - 1) // y
    w1s = torch.randn(E, cd(2 * I, bn), cd(H, bk), generator=g) * 1e-3
    w2s = torch.randn(E, cd(H, bn), cd(I, bk), generator=g) * 1e-3
    ids = make_ids(M, E, topk, style, g)
    tw = torch.rand(M, topk, generator=g) / topk
    ref = O.fused_experts(a, w1, w2, tw, ids, "fp8", w1s, w2s, block)
    dev = [t.cuda() for t in (a, w1, w2, tw, ids, w1s, w2s)]
    out_u = K.fused_experts_impl(*dev[:5], False, K.FMT_FP8_W8A16, dev[5], dev[6], block, False, path=K.PATH_UMMA).cpu()
    out_p = K.fused_experts_impl(*dev[:5], False, K.FMT_FP8_W8A16, dev[5], dev[6], block, False, path=K.PATH_UMMA2).cpu()
    assert torch.isfinite(out_p).all(), describe(out_p, ref)
    assert torch.allclose(out_p.float(), ref.float(), **BF16_TOL), "vs oracle: " + describe(out_p, ref)
    assert torch.allclose(out_p.float(), out_u.float(), **BF16_TOL), "vs single-CTA: " + describe(out_p, out_u)
