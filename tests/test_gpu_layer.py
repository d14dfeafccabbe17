"""MoELayer / GraphedLayer (sgl_cpu_tests_b200/layer.py): router -> fused_experts -> shared_expert as one object.

The layer is the call sequence of the reference's DeepSeek-style block (test_biased_grouped_topk.py:71-80 ->
test_moe_fp8_ext.py:118-124 -> :60-67).  Held here: (1) the layer equals the three drop-in ops called one after the
other, bit for bit; (2) the shared expert RIDING ALONG in the routed experts' launches (sglb_moe_problem.shared_w13)
equals the separate shared_expert call bit for bit, at ragged sizes, with masked ids, eager and captured; (3) both
stay within the reference tolerance of the f32 oracle."""
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


def build(K, fmt, E, topk, H, I, Is, seed, ride, biased=True, path=None):
    import bench
    from sgl_cpu_tests_b200.layer import MoELayer, Router
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(seed)
    w = bench.make_weights(fmt, E, H, I, dev, g)
    sh = bench.make_shared(fmt, H, Is, dev, g) if Is else None
    bias = (torch.randn(E, device=dev, generator=g) * 0.05).bfloat16() if biased else None
    router = Router(topk, True, 4, 2, bias)
    layer = MoELayer(w, bench.FMT_ID[fmt], router, sh, 2.5, path=K.PATH_AUTO if path is None else path, ride=ride)
    return layer, w, sh, bias


@pytest.mark.parametrize("fmt", ["bf16", "int8", "fp8"])
@pytest.mark.parametrize("M", [1, 37, 300])
def test_riding_shared_expert_is_bit_identical(K, fmt, M):
    E, topk, H, I = 16, 4, 1024, 512
    dev = torch.device("cuda", 0)
    ride, w, sh, bias = build(K, fmt, E, topk, H, I, I, 11, True, path=K.PATH_UMMA)
    sep, *_ = build(K, fmt, E, topk, H, I, I, 11, False, path=K.PATH_UMMA)
    assert ride.rides(M) and not sep.rides(M)
    g = torch.Generator(device=dev).manual_seed(100 + M)
    # outputs of O(1), as in the reference's scripts (test_moe_fp8_ext.py:33-35 scales its operands the same way):
    # the op rounds the shared MLP and the sum to bf16, so two ulps at |out| > 2 already exceed rtol = atol = 1e-2
    x = (torch.randn(M, H, device=dev, generator=g) / 16).bfloat16()
    gating = torch.randn(M, E, device=dev, generator=g).bfloat16()
    n0 = K.launch_count()
    a = ride.forward(x, gating)
    n1 = K.launch_count()
    b = sep.forward(x, gating)
    n2 = K.launch_count()
    assert torch.equal(a, b), f"{fmt} M={M}: max diff {(a.float() - b.float()).abs().max()}"
    assert n1 - n0 < n2 - n1, "the riding form must need fewer launches"
    # against the drop-in ops one after the other, and the f32 oracle
    import bench
    fid = bench.FMT_ID[fmt]
    tw, ids = K.biased_grouped_topk(x, gating, bias, topk, True, 4, 2)
    r0 = K.fused_experts_impl(x, w["w1"], w["w2"], tw, ids, False, fid, w["s1"], w["s2"], w["block"], False, path=K.PATH_UMMA)
    r = K.shared_expert_impl(x, sh["w1"], sh["w2"], r0, 2.5, False, fid, sh["s1"], sh["s2"], sh["block"], False)
    assert torch.equal(a, r)
    if fmt != "int8":
        # each op against its oracle on the same inputs, as the reference's scripts do (the shared expert is handed the
        # routed result that the kernel produced: test_moe_fp8_ext.py:60-67 passes one fused_out to both sides)
        c = lambda t: None if t is None else t.cpu()
        fm = {"bf16": "bf16", "fp8": "fp8"}[fmt]
        ref0 = O.fused_experts(x.cpu(), c(w["w1"]), c(w["w2"]), tw.cpu(), ids.cpu(), fm, c(w["s1"]), c(w["s2"]), w["block"])
        assert torch.allclose(r0.float().cpu(), ref0.float(), **BF16_TOL), (r0.float().cpu() - ref0.float()).abs().max()
        ref = O.shared_expert(x.cpu(), c(sh["w1"]), c(sh["w2"]), r0.cpu(), 2.5, fm, c(sh["s1"]), c(sh["s2"]), sh["block"])
        assert torch.allclose(a.float().cpu(), ref.float(), **BF16_TOL), (a.float().cpu() - ref.float()).abs().max()


@pytest.mark.parametrize("fmt", ["bf16", "int8", "fp8"])
def test_riding_with_masked_ids_and_graph(K, fmt):
    """-1 ids (test_moe_offloading_cpu.py:57-68) leave the shared rows untouched; every id masked leaves the shared
    expert alone; the captured layer replays to the same bits."""
    E, topk, H, I, M = 8, 2, 512, 256, 70
    dev = torch.device("cuda", 0)
    import bench
    fid = bench.FMT_ID[fmt]
    ride, w, sh, _ = build(K, fmt, E, topk, H, I, I, 5, True, biased=False, path=K.PATH_UMMA)
    g = torch.Generator(device=dev).manual_seed(9)
    x = (torch.randn(M, H, device=dev, generator=g) / 4).bfloat16()
    gating = torch.randn(M, E, device=dev, generator=g).bfloat16()
    tw, ids = ride.route(gating)
    for masked in (ids.clone(), torch.full_like(ids, -1)):
        masked[::3, 0] = -1
        a = ride.experts(x, tw, masked, with_shared=True)
        r = K.fused_experts_impl(x, w["w1"], w["w2"], tw, masked, False, fid, w["s1"], w["s2"], w["block"], False, path=K.PATH_UMMA)
        r = K.shared_expert_impl(x, sh["w1"], sh["w2"], r, 2.5, False, fid, sh["s1"], sh["s2"], sh["block"], False)
        assert torch.equal(a, r)
    gl = ride.graphed(M)
    assert gl.rides
    eager = ride.forward(x, gating)
    for _ in range(3):
        assert torch.equal(gl(x, gating), eager)


def test_rides_only_where_the_kernels_can(K):
    """Different intermediate size, the streaming path, mxfp4 and the gpt-oss activation keep the separate call."""
    from sgl_cpu_tests_b200 import ops
    L = K
    assert ops.shared_rides(64, 8, 64, 1024, 512, K.FMT_FP8_W8A16, [128, 128], K.PATH_AUTO, shared_I=512)
    assert not ops.shared_rides(64, 8, 64, 1024, 512, K.FMT_FP8_W8A16, [128, 128], K.PATH_AUTO, shared_I=1024)
    assert not ops.shared_rides(64, 8, 64, 1024, 512, K.FMT_BF16, None, K.PATH_STREAM, shared_I=512)
    assert not ops.shared_rides(64, 8, 64, 1024, 512, K.FMT_MXFP4, None, K.PATH_AUTO, shared_I=512)
    assert not ops.shared_rides(0, 8, 64, 1024, 512, K.FMT_BF16, None, K.PATH_UMMA, shared_I=512)
    layer, *_ = build(K, "fp8", 16, 4, 1024, 512, 1024, 3, True)
    assert not layer.rides(64)
    x = torch.zeros(64, 1024, dtype=torch.bfloat16, device="cuda")
    out = layer.forward(x, torch.zeros(64, 16, dtype=torch.bfloat16, device="cuda"))   # separate call, still correct
    assert out.shape == x.shape and torch.isfinite(out.float()).all()
