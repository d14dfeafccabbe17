"""Decode chain fusion (capi.cu, sglb_set_chain_fusion): the single-launch prologue (align + stage-1 operand
preparation; level 1, the default) and, at level 2, the int8 activation quantised inside the stage-2 producer and the
top-k reduce performed inside the stage-2 kernel must give the SAME BITS as the chain of separate launches (align, quant_x, w13, quant_act, w2, topk_reduce),
which in turn is what the oracle tests hold to the reference (tests/test_gpu_parity.py).  Covered: every weight
format, the streaming and the tcgen05 path, ragged token counts, -1 / duplicate ids
(test_moe_offloading_cpu.py:57-68), a token with every id masked, every id masked, the riding shared expert, inplace
output (out aliases the hidden states), graph replay, and the launch count."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    import sgl_cpu_tests_b200 as K
    assert torch.cuda.is_available()
    return K


DEFAULT_LEVEL = 1


def fusion(level):
    """0: separate launches; 1 (default): single-launch prologue; 2: + in-kernel int8 quantisation and top-k reduce."""
    from sgl_cpu_tests_b200 import _lib as L
    L.load().sglb_set_chain_fusion(DEFAULT_LEVEL if level is True else int(level))


def weights(fmt, E, H, I, seed):
    import bench
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(seed)
    return bench.make_weights(fmt, E, H, I, dev, g), bench.FMT_ID[fmt], g


def run(K, x, w, fid, tw, ids, path, inplace=False):
    return K.fused_experts_impl(x.clone(), w["w1"], w["w2"], tw, ids, inplace, fid, w["s1"], w["s2"], w["block"], False, path=path)


def both(K, x, w, fid, tw, ids, path, inplace=False):
    """Separate launches vs level 2; level 1 is checked against the same reference on the way."""
    try:
        fusion(0)
        n0 = K.launch_count()
        ref = run(K, x, w, fid, tw, ids, path, inplace)
        n1 = K.launch_count()
        fusion(1)
        mid = run(K, x, w, fid, tw, ids, path, inplace)
        n1b = K.launch_count()
        assert torch.equal(mid, ref), f"level 1: max diff {(mid.float() - ref.float()).abs().max()}"
        assert n1b - n1 <= n1 - n0
        fusion(2)
        n1c = K.launch_count()
        out = run(K, x, w, fid, tw, ids, path, inplace)
        n2 = K.launch_count()
    finally:
        fusion(True)
    return ref, out, n1 - n0, n2 - n1c


def routing(M, E, topk, g, dev):
    gating = torch.randn(M, E, device=dev, generator=g)
    tw, ids = torch.topk(torch.softmax(gating, -1), topk, dim=-1)
    return tw.float().contiguous(), ids.int().contiguous()


@pytest.mark.parametrize("fmt", ["bf16", "int8", "fp8", "mxfp4"])
@pytest.mark.parametrize("path", ["stream", "umma"])
@pytest.mark.parametrize("M", [1, 5, 37, 130, 300])
def test_fused_chain_is_bit_identical(K, fmt, path, M):
    E, topk, H, I = 16, 4, 1024, 512
    dev = torch.device("cuda", 0)
    w, fid, g = weights(fmt, E, H, I, 3)
    x = (torch.randn(M, H, device=dev, generator=g) / 8).bfloat16()
    tw, ids = routing(M, E, topk, g, dev)
    p = K.PATH_STREAM if path == "stream" else K.PATH_UMMA
    ref, out, n_sep, n_fused = both(K, x, w, fid, tw, ids, p)
    assert torch.isfinite(ref.float()).all() and ref.float().abs().max() > 0
    assert torch.equal(out, ref), f"{fmt} {path} M={M}: max diff {(out.float() - ref.float()).abs().max()}"
    assert n_fused < n_sep, (n_fused, n_sep)
    if fmt in ("bf16", "int8") and path == "umma":
        assert n_fused == 3, n_fused  # prologue, w13, w2 (<= 3 launches per call)


@pytest.mark.parametrize("fmt", ["bf16", "int8", "fp8"])
@pytest.mark.parametrize("path", ["stream", "umma", "auto"])
def test_fused_chain_masked_and_duplicate_ids(K, fmt, path):
    E, topk, H, I, M = 8, 4, 512, 256, 77
    dev = torch.device("cuda", 0)
    w, fid, g = weights(fmt, E, H, I, 5)
    x = (torch.randn(M, H, device=dev, generator=g) / 4).bfloat16()
    tw, ids = routing(M, E, topk, g, dev)
    tw = torch.randn(M, topk, device=dev, generator=g)  # negative weights too (moe_offload fixture)
    p = {"stream": K.PATH_STREAM, "umma": K.PATH_UMMA, "auto": K.PATH_AUTO}[path]
    cases = []
    a = ids.clone(); a[::3, 0] = -1; a[5, :] = -1; a[76, :] = -1; cases.append(a)   # masked slots, two fully masked tokens
    b = ids.clone(); b[:, 1] = b[:, 0]; cases.append(b)                              # duplicate ids inside a token
    cases.append(torch.full_like(ids, -1))                                           # nothing routed at all
    c = torch.zeros_like(ids); c[:, 1:] = -1; cases.append(c)                        # one expert takes every token
    for masked in cases:
        ref, out, _, _ = both(K, x, w, fid, tw, masked, p)
        assert torch.equal(out, ref), f"{fmt} {path}: max diff {(out.float() - ref.float()).abs().max()}"


@pytest.mark.parametrize("fmt", ["bf16", "int8", "fp8"])
def test_fused_chain_inplace_and_qwen_shape(K, fmt):
    """out aliases hidden_states (inplace=True, test_moe.py:79-92): the fused reduce writes rows of `out` while other
    CTAs still run stage 2 -- which never reads the hidden states."""
    E, topk, H, I, M = 128, 8, 2048, 768, 128
    dev = torch.device("cuda", 0)
    w, fid, g = weights(fmt, E, H, I, 7)
    x = (torch.randn(M, H, device=dev, generator=g) / 8).bfloat16()
    tw, ids = routing(M, E, topk, g, dev)
    for M_ in (128, 16, 1):
        ref, out, _, _ = both(K, x[:M_], w, fid, tw[:M_].contiguous(), ids[:M_].contiguous(), K.PATH_AUTO, inplace=True)
        assert torch.equal(out, ref), f"{fmt} M={M_}: max diff {(out.float() - ref.float()).abs().max()}"


@pytest.mark.parametrize("fmt", ["bf16", "int8", "fp8"])
@pytest.mark.parametrize("M", [1, 37, 300])
def test_fused_chain_with_riding_shared_expert(K, fmt, M):
    from test_gpu_layer import build
    E, topk, H, I = 16, 4, 1024, 512
    dev = torch.device("cuda", 0)
    layer, w, sh, bias = build(K, fmt, E, topk, H, I, I, 11, True, path=K.PATH_UMMA)
    assert layer.rides(M)
    g = torch.Generator(device=dev).manual_seed(100 + M)
    x = (torch.randn(M, H, device=dev, generator=g) / 16).bfloat16()
    gating = torch.randn(M, E, device=dev, generator=g).bfloat16()
    try:
        fusion(0)
        ref = layer.forward(x, gating)
        tw, ids = layer.route(gating)
        ids[::2, 1] = -1
        ref_m = layer.experts(x, tw, ids, with_shared=True)
        for level in (1, 2):
            fusion(level)
            out = layer.forward(x, gating)
            out_m = layer.experts(x, tw, ids, with_shared=True)
            assert torch.equal(out, ref), (level, (out.float() - ref.float()).abs().max())
            assert torch.equal(out_m, ref_m), (level, (out_m.float() - ref_m.float()).abs().max())
    finally:
        fusion(True)


def test_fused_chain_graph_replay_and_repeatability(K):
    """Counters and row maxima are re-armed by every prologue: repeated calls and graph replays give the same bits."""
    E, topk, H, I, M = 32, 4, 1024, 512, 64
    dev = torch.device("cuda", 0)
    for fmt in ("int8", "fp8"):
        w, fid, g = weights(fmt, E, H, I, 9)
        x = (torch.randn(M, H, device=dev, generator=g) / 8).bfloat16()
        tw, ids = routing(M, E, topk, g, dev)
        fusion(0)
        try:
            ref = run(K, x, w, fid, tw, ids, K.PATH_AUTO)
            for level in (1, 2):
                fusion(level)
                for _ in range(5):
                    assert torch.equal(run(K, x, w, fid, tw, ids, K.PATH_AUTO), ref)
                ge = K.GraphedExperts(M, w["w1"], w["w2"], topk, fid, w["s1"], w["s2"], w["block"])
                for _ in range(4):
                    assert torch.equal(ge(x, tw, ids), ref)
        finally:
            fusion(True)


def test_many_align_blocks_in_the_prologue(K):
    """M * topk > 1024: the prologue's align blocks count each other's slots themselves (no align_count launch)."""
    E, topk, H, I = 64, 8, 512, 256
    dev = torch.device("cuda", 0)
    for fmt, M in (("bf16", 1000), ("int8", 1000), ("fp8", 2000)):   # 8000 / 16000 slots -> 8 / 16 align blocks
        w, fid, g = weights(fmt, E, H, I, 13)
        x = (torch.randn(M, H, device=dev, generator=g) / 8).bfloat16()
        tw, ids = routing(M, E, topk, g, dev)
        ids[::7, 3] = -1
        ref, out, _, _ = both(K, x, w, fid, tw, ids, K.PATH_UMMA)
        assert torch.equal(out, ref), (out.float() - ref.float()).abs().max()


def test_mxfp4_scale_tiles_are_built_once(K):
    """The e8m0 scale tiles of the block-scaled kernel are a per-weight cache (ops._mx4_tiles): a second call with the
    same weights must not re-tile them (regression: the cache was keyed on a temporary view and re-tiled every call,
    two extra launches = 137 us per gpt-oss step)."""
    E, topk, H, I, M = 16, 4, 1024, 512, 64
    dev = torch.device("cuda", 0)
    w, fid, g = weights("mxfp4", E, H, I, 21)
    x = (torch.randn(M, H, device=dev, generator=g) / 8).bfloat16()
    tw, ids = routing(M, E, topk, g, dev)
    counts = []
    for _ in range(3):
        n0 = K.launch_count()
        run(K, x, w, fid, tw, ids, K.PATH_AUTO)
        counts.append(K.launch_count() - n0)
    assert counts[1] == counts[2] <= 5, counts   # prologue, w13, split, w2, reduce
    assert counts[0] >= counts[1]


def test_graph_outlives_eviction_of_the_mxfp4_scale_tiles(K):
    """A captured graph addresses the registry-owned scale tiles of an mxfp4 weight.  The registry may evict them
    (LRU, clear_weight_cache): the graph object must keep them alive itself, and a replay after the eviction --
    with the freed-looking memory scribbled over by new allocations -- must still give the eager result."""
    from sgl_cpu_tests_b200 import ops
    E, topk, H, I, M = 16, 4, 1024, 512, 64
    dev = torch.device("cuda", 0)
    w, fid, g = weights("mxfp4", E, H, I, 23)
    x = (torch.randn(M, H, device=dev, generator=g) / 8).bfloat16()
    tw, ids = routing(M, E, topk, g, dev)
    ref = run(K, x, w, fid, tw, ids, K.PATH_AUTO).clone()
    gph = K.GraphedExperts(M, w["w1"], w["w2"], topk, fid, w["s1"], w["s2"], w["block"])
    tiles = [t for t in gph._keep if t is not None and t.dtype == torch.uint8 and t.dim() == 1]
    assert len(tiles) >= 2, "the graph does not hold the scale tiles it was captured over"
    ops.clear_weight_cache()
    torch.cuda.synchronize()
    junk = [torch.full((t.numel(),), 0xff, dtype=torch.uint8, device=dev) for t in tiles for _ in range(4)]
    out = gph(x, tw, ids)
    torch.cuda.synchronize()
    assert torch.equal(out, ref)
    del junk


@pytest.mark.parametrize("fmt", ["bf16", "fp8"])
def test_shared_expert_reuses_the_device_twins_of_the_previous_fused_experts_call(K, fmt):
    """Host tensors, the reference's call sequence (test_moe_fp8_ext.py:118 then :60-61): shared_expert_cpu gets the
    hidden_states that fused_experts_cpu just uploaded and the result it just produced.  Both are still in HBM, so
    the second op must not upload them again -- and must give the bits of the plain path.  A hit needs the same
    tensor objects at an unchanged version: an in-place update of either, or an older call's tensors, must miss."""
    from sgl_cpu_tests_b200 import ops
    E, topk, H, I, M = 16, 4, 1024, 512, 48
    dev = torch.device("cuda", 0)
    w, fid, g = weights(fmt, E, H, I, 41)
    sh = {k: (v[0].contiguous() if torch.is_tensor(v) else v) for k, v in weights(fmt, 1, H, I, 43)[0].items()}
    tw_d, ids_d = routing(M, E, topk, g, dev)
    tw, ids = tw_d.cpu(), ids_d.cpu()
    xs = [(torch.randn(M, H, generator=torch.Generator().manual_seed(s)) / 16).bfloat16() for s in (1, 2)]

    def fused(x):
        return K.fused_experts_impl(x, w["w1"], w["w2"], tw, ids, False, fid, w["s1"], w["s2"], w["block"], False)

    def shared(x, fo):
        return K.shared_expert_impl(x, sh["w1"], sh["w2"], fo, 2.5, False, fid, sh["s1"], sh["s2"], sh["block"], False)

    assert ops._SHADOW
    for _ in range(3):                       # the third call replays the cached host-tensor graph (static buffers)
        out1 = fused(xs[0])
        ops._SHADOW = False
        try:
            ref1 = shared(xs[0], out1)       # plain path: both tensors uploaded
        finally:
            ops._SHADOW = True
        h0 = ops.shadow_hits
        got1 = shared(xs[0], out1)
        assert ops.shadow_hits == h0 + 2, "hidden_states and fused_experts_out should both have been found in HBM"
        assert torch.equal(got1, ref1)
    # an older call's tensors: the static buffers now hold call 2's data, call 1's tensors must miss
    out1 = fused(xs[0])
    out2 = fused(xs[1])
    h0 = ops.shadow_hits
    assert torch.equal(shared(xs[0], out1), ref1)
    assert ops.shadow_hits == h0
    # an in-place update through torch bumps the version counter: miss, and the new values are used
    out2.mul_(2)
    xs[1].add_(0.25)
    h0 = ops.shadow_hits
    got = shared(xs[1], out2)
    assert ops.shadow_hits == h0
    ops._SHADOW = False
    try:
        assert torch.equal(got, shared(xs[1], out2))
    finally:
        ops._SHADOW = True


def test_refused_host_registration_leaves_no_cuda_error_behind(K):
    """ops._maybe_pin page-locks large recurring host buffers through sglb_host_register.  A range the driver refuses
    (here: one that is already registered) must be skipped quietly: no exception, no pending CUDA error that the next
    torch launch check would trip over."""
    from sgl_cpu_tests_b200 import _lib as L
    lib = L.load()
    t = torch.empty(8 << 20, dtype=torch.uint8)
    assert lib.sglb_host_register(t.data_ptr(), t.numel()) == 0
    try:
        assert lib.sglb_host_register(t.data_ptr(), t.numel()) != 0      # already registered: refused
        assert b"cudaHostRegister" in lib.sglb_last_error()
        y = torch.ones(1024, device="cuda") * 2                           # a torch launch + its error check
        torch.cuda.synchronize()
        assert float(y.sum()) == 2048.0
    finally:
        assert lib.sglb_host_unregister(t.data_ptr()) == 0
    assert lib.sglb_host_unregister(t.data_ptr()) != 0                    # not registered any more: reported, cleared
    torch.cuda.synchronize()


@pytest.mark.parametrize("M", [1, 16, 40, 100, 250])
def test_fp8_small_tile_deep_ring_is_bit_identical(K, M):
    """fp8 decode takes 32-row tiles on a 5-stage ring (umma_f8.cu F8Cfg<32>; M = 100, 250 here: 80-row tiles on a
    4-stage ring, F8Cfg<80>) when experts get few rows; the per-row
    arithmetic does not depend on the tile a row lands in, so the result equals the 128-row configuration bit for bit
    -- also when one expert receives more than 32 rows and is split into several tiles."""
    from sgl_cpu_tests_b200 import _lib as L
    E, topk, H, I = 16, 4, 1024, 512
    dev = torch.device("cuda", 0)
    w, fid, g = weights("fp8", E, H, I, 31)
    x = (torch.randn(M, H, device=dev, generator=g) / 8).bfloat16()
    tw, ids = routing(M, E, topk, g, dev)
    skew = ids.clone()
    skew[:, 0] = 3                                       # expert 3 takes a row of every token: 40 rows > 32
    skew[:, 1:] = torch.where(skew[:, 1:] == 3, torch.full_like(skew[:, 1:], 4), skew[:, 1:])
    lib = L.load()
    for case in (ids, skew):
        try:
            lib.sglb_set_fp8_native(3)                   # 128-row tiles
            ref = run(K, x, w, fid, tw, case, K.PATH_UMMA)
            lib.sglb_set_fp8_native(5)                   # 32-row tiles (up to 512 slots, <= 12 rows per expert) and 80-row tiles
            out = run(K, x, w, fid, tw, case, K.PATH_UMMA)
        finally:
            lib.sglb_set_fp8_native(1)
        assert torch.isfinite(ref.float()).all() and ref.float().abs().max() > 0
        assert torch.equal(out, ref), (out.float() - ref.float()).abs().max()


def test_fp8_small_tiles_with_riding_shared_expert(K):
    from sgl_cpu_tests_b200 import _lib as L
    from test_gpu_layer import build
    E, topk, H, I, M = 16, 4, 1024, 512, 20
    dev = torch.device("cuda", 0)
    layer, *_ = build(K, "fp8", E, topk, H, I, I, 11, True, path=K.PATH_UMMA)
    g = torch.Generator(device=dev).manual_seed(5)
    x = (torch.randn(M, H, device=dev, generator=g) / 16).bfloat16()
    gating = torch.randn(M, E, device=dev, generator=g).bfloat16()
    lib = L.load()
    try:
        lib.sglb_set_fp8_native(3)
        ref = layer.forward(x, gating)
        lib.sglb_set_fp8_native(1)
        out = layer.forward(x, gating)
    finally:
        lib.sglb_set_fp8_native(1)
    assert torch.equal(out, ref), (out.float() - ref.float()).abs().max()


def test_staged_upload_of_large_pageable_tensors_is_exact(K):
    """ops._upload_staged (first upload of a large pageable host tensor through two pinned staging buffers): every byte
    arrives, also when the size is not a multiple of the staging buffer."""
    from sgl_cpu_tests_b200 import ops
    dev = torch.device("cuda", 0)
    n = ops._UPLOAD_MIN_BYTES + 3 * ops._STAGE_BYTES // 2 + 12345 * 2
    t = torch.randint(-2 ** 15, 2 ** 15 - 1, (n // 2,), dtype=torch.int16).view(torch.bfloat16)
    assert not t.is_pinned()
    d = ops._to_dev(t, dev)
    assert d.dtype == t.dtype and d.shape == t.shape
    assert torch.equal(d.view(torch.int16).cpu(), t.view(torch.int16))
    f8 = torch.randint(0, 255, (ops._UPLOAD_MIN_BYTES + 7,), dtype=torch.uint8).view(torch.float8_e4m3fn)
    assert torch.equal(ops._to_dev(f8, dev).view(torch.uint8).cpu(), f8.view(torch.uint8))


@pytest.mark.parametrize("fmt", ["bf16", "fp8"])
def test_shared_expert_pinned_fused_out_overlapped_upload(K, fmt):
    """A pinned host fused_experts_out is uploaded on a copy stream under the MLP and added by a separate axpy launch
    (ops.shared_expert_impl): same bits as the fused epilogue taken with pageable tensors."""
    import bench
    H, I, M = 1024, 512, 100
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(17)
    sh = bench.make_shared(fmt, H, I, dev, g)
    fid = bench.FMT_ID[fmt]
    x = (torch.randn(M, H, device=dev, generator=g) / 8).bfloat16().cpu()
    fo = (torch.randn(M, H, device=dev, generator=g) / 4).bfloat16().cpu()
    c = lambda t: None if t is None else t.cpu()
    args = (c(sh["w1"]), c(sh["w2"]))
    ref = K.shared_expert_impl(x.clone(), *args, fo.clone(), 2.5, False, fid, c(sh["s1"]), c(sh["s2"]), sh["block"], False)
    out = K.shared_expert_impl(x.clone().pin_memory(), *args, fo.clone().pin_memory(), 2.5, False, fid, c(sh["s1"]), c(sh["s2"]),
                               sh["block"], False)
    assert torch.equal(out, ref), (out.float() - ref.float()).abs().max()
    xin = x.clone().pin_memory()
    out2 = K.shared_expert_impl(xin, *args, fo.clone().pin_memory(), 2.5, True, fid, c(sh["s1"]), c(sh["s2"]), sh["block"], False)
    assert out2 is xin and torch.equal(out2, ref)
