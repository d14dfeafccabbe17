"""CPU oracle for the fused-MoE path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module; nothing under sgl_cpu_tests_b200/ does (tests/test_boundary.py greps for that).

What it restates.  The reference checkout contains no kernel source: the arithmetic of
fused_experts_cpu & co. lives in the sgl-kernel CPU extension of sgl-project/sglang
(sgl-kernel/csrc/cpu/*), version unpinned by the reference (no requirements/lock file) and absent
from /root/reference.  What the reference does hold, and what pins results at the operator boundary,
are the torch-native oracles inside its test scripts.  Each function below restates one of them and
cites it (paths relative to /root/reference).  The restatements are pinned against the reference's
own functions, executed in the build container by tests/golden/make_golden.py, through the
fixtures under tests/golden/ (tests/test_oracle.py).

Three things have no pin in the reference and are defined here -- "parity unpinned"
for exactly these: (1) router tie-break and output order: descending score, ties to the lowest
expert index; (2) the moe_align index layout; (3) mxfp4 *fused MoE* (only the single GEMM is tested
upstream): composed from the mxfp4 dequantiser + the fp32 MoE oracle.
"""
import math

import torch
import torch.nn.functional as F

# ---------------------------------------------------------------------------------------------------
# activation / quantisation helpers


def silu_and_mul(x):
    """test_moe.py:17-19, test_moe_fp8_ext.py:18-20 (dtype-preserving)."""
    d = x.shape[-1] // 2
    return F.silu(x[..., :d]) * x[..., d:]


def silu_and_mul_f32(x):
    """test_moe_int8.py:16-21 (computes in f32, casts back)."""
    dtype = x.dtype
    x = x.to(torch.float32)
    d = x.shape[-1] // 2
    return (F.silu(x[..., :d]) * x[..., d:]).to(dtype)


def per_token_quant_int8(x):
    """test_moe_int8.py:23-31."""
    x = x.float()
    absmax = x.abs().max(dim=-1).values
    absmax = absmax.clamp_min(1e-7).unsqueeze(-1)
    scale_x = absmax / 127
    x_q = x.mul(127 / absmax)
    x_q = torch.round(x_q).to(torch.int8)
    return x_q, scale_x


def native_w8a8_per_token_matmul(A, B, As, Bs, output_dtype=torch.float32):
    """test_moe_int8.py:37-56: C = As * (A @ B^T) * Bs, f32 accumulate."""
    A = A.to(torch.float32)
    B = B.to(torch.float32)
    C = torch.matmul(A, B.t())
    C = As * C * Bs.view(1, -1)
    return C.to(output_dtype)


def scaled_weight(weight, scales, block_n, block_k):
    """test_moe_fp8_ext.py:22-25 generalised to ragged edges (test_moe_fp8.py:15-62 pads the same way):
    w[e, n, k] * scales[e, n // block_n, k // block_k] in f32."""
    squeeze = weight.dim() == 2
    if squeeze:
        weight, scales = weight.unsqueeze(0), scales.unsqueeze(0)
    E, N, K = weight.shape
    s = scales.float().repeat_interleave(block_n, dim=1)[:, :N]
    s = s.repeat_interleave(block_k, dim=2)[:, :, :K]
    out = weight.float() * s
    return out.squeeze(0) if squeeze else out


# ---------------------------------------------------------------------------------------------------
# dense quantised linears; pinned by tests/golden/gemm_*.pt (make_golden_gemm.py)


def linear_bf16(mat1, mat2, bias=None):
    """test_gemm.py:12-20: f32 matmul, bias rounded to bf16 before the add, one rounding of the sum."""
    ref = torch.matmul(mat1.float(), mat2.float().t())
    if bias is not None:
        ref.add_(bias.bfloat16())
    return ref.bfloat16()


def per_token_quant_int8_gemm(x):
    """test_gemm_int8.py:11-20 (as per_token_quant_int8 above, but the clamp is 1e-10)."""
    x = x.float()
    absmax = x.abs().max(dim=-1).values
    absmax = absmax.clamp_min(1e-10).unsqueeze(-1)
    scale_x = absmax / 127
    x_q = x.mul(127 / absmax)
    x_q = torch.round(x_q).to(torch.int8)
    return x_q, scale_x


def linear_int8(Aq, Bq, As, Bs, bias=None, output_dtype=torch.bfloat16):
    """test_gemm_int8.py:23-47 (native_w8a8_per_token_matmul): C = As * (Aq @ Bq^T) * Bs (+ bias), f32."""
    C = torch.matmul(Aq.to(torch.float32), Bq.to(torch.float32).t())
    C = As.reshape(-1, 1) * C * Bs.view(1, -1)
    if bias is not None:
        C.add_(bias.view(1, -1))
    return C.to(output_dtype)


def linear_int8_with_quant(A, Bq, Bs, bias=None, output_dtype=torch.bfloat16):
    """int8_scaled_mm_with_quant (test_gemm_int8.py:74): the quantiser above fused with linear_int8."""
    Aq, As = per_token_quant_int8_gemm(A)
    return linear_int8(Aq, Bq, As, Bs, bias, output_dtype)


def linear_fp8(A, Wq, scales, block_size, bias=None):
    """test_gemm_fp8.py:23-52: the weight is dequantised block-wise in f32 and ROUNDED TO bf16, the matmul
    runs in bf16 (f32 accumulation inside torch), the bias is rounded to bf16 and added in bf16."""
    w = scaled_weight(Wq, scales, block_size[0], block_size[1]).to(A.dtype)
    out = torch.matmul(A, w.t())
    if bias is not None:
        out = out + bias.to(A.dtype)
    return out


def rmsnorm(x, weight, eps=1e-6, residual=None):
    """test_norm.py:13-33 (forward_native): f32 statistics, the normalised value is rounded to the input dtype
    before the weight product; with a residual the f32 sum feeds the variance and is returned rounded."""
    orig_dtype = x.dtype
    x = x.to(torch.float32)
    if residual is not None:
        x = x + residual.to(torch.float32)
        residual = x.to(orig_dtype)
    variance = x.pow(2).mean(dim=-1, keepdim=True)
    x = x * torch.rsqrt(variance + eps)
    x = x.to(orig_dtype) * weight
    return x if residual is None else (x, residual)


# ---------------------------------------------------------------------------------------------------
# MXFP4 (test_mxfp4.py:15-128)

E2M1_VALUES = [0, 0.5, 1, 1.5, 2, 3, 4, 6]
E2M1_BOUNDS = torch.tensor([0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5])
MX_BLOCK = 32


def mxfp4_quantize(x):
    """test_mxfp4.py:21-63.  Returns (packed nibbles u8 [..., K/2], e8m0 u8 [numel/32, 1])."""
    original_shape = x.shape
    x = x.reshape(-1, MX_BLOCK)
    amax = x.abs().max(dim=-1, keepdim=True).values
    descale = amax / 6.0
    e8m0 = torch.ceil(torch.maximum(torch.log2(descale), torch.tensor(-127.0)))
    x = (x / torch.exp2(e8m0)).view(original_shape)
    sign = torch.sign(x)
    sign_bit = (2 - sign) // 2
    ord_ = torch.sum((x.abs().unsqueeze(-1) - E2M1_BOUNDS) > 0, dim=-1)
    q = (sign_bit * 0b1000 + ord_).to(torch.uint8)
    packed = (q[..., 1::2] << 4) + q[..., 0::2]  # low nibble = even index
    return packed, (e8m0 + 127).to(torch.uint8)


def mxfp4_dequantize(packed, dtype, scale):
    """test_mxfp4.py:65-128."""
    lo = packed & 0x0F
    hi = (packed >> 4) & 0x0F
    shape = list(packed.shape)
    shape[-1] *= 2
    q = torch.zeros(shape, dtype=torch.uint8)
    q[..., 0::2] = lo
    q[..., 1::2] = hi
    sign = 1 - 2 * ((q & 0b1000) >> 3).to(torch.float32)
    mag = (q & 0b0111).to(torch.long)
    vals = torch.tensor(E2M1_VALUES, dtype=torch.float32)
    x = sign * vals[mag.reshape(-1)].reshape(mag.shape)
    x = x.reshape(-1, MX_BLOCK) * torch.exp2(scale.float() - 127).reshape(-1, 1)
    return x.reshape(shape).to(dtype)


def mxfp4_scaled_mm(A, Bq, Bs, bias=None):
    """test_mxfp4.py:162-168: bf16(A.f32 @ dequant(B).f32^T) (+ bias added in bf16)."""
    Bdq = mxfp4_dequantize(Bq, A.dtype, Bs)
    out = torch.matmul(A.float(), Bdq.float().t()).bfloat16()
    if bias is not None:
        out.add_(bias.view(1, -1))
    return out


# ---------------------------------------------------------------------------------------------------
# routers


def _topk_lowest_index(scores, k):
    """top-k in descending score order, ties to the lowest index (the build's stated rule)."""
    order = torch.sort(scores, dim=-1, descending=True, stable=True).indices
    ids = order[..., :k]
    return torch.gather(scores, -1, ids), ids


_EXP_C = [float.fromhex(h) for h in (
    "0x1.0000000000000p+0", "0x1.0000000000000p+0", "0x1.0000000000000p-1", "0x1.5555555555555p-3", "0x1.5555555555555p-5",
    "0x1.1111111111111p-7", "0x1.6c16c16c16c17p-10", "0x1.a01a01a01a01ap-13", "0x1.a01a01a01a01ap-16",
    "0x1.71de3a556c734p-19", "0x1.27e4fb7789f5cp-22", "0x1.ae64567f544e4p-26", "0x1.1eed8eff8d898p-29",
    "0x1.6124613a86d09p-33")]                      # 1 / n!, n = 0..13


def sigmoid_f64(x):
    """The build's sigmoid contract (csrc/router.cu:sigmoid_f64, the same sequence operation for operation):
    1 / (1 + exp(-x)) in f64 from correctly rounded IEEE-754 basic operations only -- exp(t) = 2^k P(r) with
    k = rint(t / ln 2), r = t - k ln2_hi - k ln2_lo, P the degree-13 Taylor polynomial in Horner form with separate
    multiplies and adds -- rounded once to f32.  Being a fixed sequence of basic operations it yields the same bits
    on any IEEE machine, which neither the reference's expression (`gating_output.sigmoid()`,
    test_biased_grouped_topk.py:22: torch's f32 CPU sigmoid is a ~1-ulp approximation, 30 % of inputs differ from
    the correctly rounded value) nor a libm exp() (ulp-level differences between libraries flip the f32 rounding
    where sigmoid sits on a half-ulp) can offer.  tests/test_oracle.py counts the rows whose ids differ from the
    reference's expression; tests/test_oracle.py also holds it within one f32 ulp of the exact value."""
    t = (-x.double()).clamp(-50.0, 120.0)
    k = torch.round(t * float.fromhex("0x1.71547652b82fep+0"))            # half to even, like rint()
    r = (t - k * float.fromhex("0x1.62e42fee00000p-1")) - k * float.fromhex("0x1.a39ef35793c76p-33")
    p = torch.full_like(r, _EXP_C[13])
    for n in range(12, -1, -1):
        p = p * r + _EXP_C[n]                                            # two roundings: no fused multiply-add
    e = p * ((k.to(torch.int64) + 1023) << 52).view(torch.float64)       # exact power of two
    return (1.0 / (1.0 + e)).float()


sigmoid_cr = sigmoid_f64   # earlier name


def _seq_sum(w):
    """f32 sum over the last dim in index order (the router kernel adds its picks one by one)."""
    acc = w[..., 0].clone()
    for j in range(1, w.shape[-1]):
        acc = acc + w[..., j]
    return acc.unsqueeze(-1)


def _serving_args(w, ids, E, renormalize, num_fused_shared_experts, routed_scaling_factor, num_token_non_padded):
    """Last three arguments of the 9-argument router call (test_moe.py:61-70 passes 0, None, None).  Semantics of
    sglang's python router ([UPSTREAM-MEMORY] sglang/srt/layers/moe/topk.py) -- PARITY UNPINNED, the reference
    never exercises them: with n fused shared experts the last slot becomes id E + (token % n) (sglang draws the
    replica at random; a fixed rule here) with weight sum(routed) / routed_scaling_factor, renormalisation divides
    by the routed sum only; rows >= num_token_non_padded get ids -1 (and weight 0 here)."""
    n = int(num_fused_shared_experts or 0)
    rsf = 1.0 if routed_scaling_factor is None else float(routed_scaling_factor)
    w = w.to(torch.float32)
    ids = ids.to(torch.int32)
    routed = w[:, :-1] if n > 0 else w
    wsum = _seq_sum(routed)
    if n > 0:
        w = w.clone()
        ids = ids.clone()
        w[:, -1:] = wsum / torch.tensor(rsf, dtype=torch.float32)
        ids[:, -1] = (E + torch.arange(ids.shape[0]) % n).to(torch.int32)
    if renormalize:
        w = w / wsum
    if num_token_non_padded is not None:
        pad = torch.arange(ids.shape[0]) >= int(num_token_non_padded)
        ids = ids.masked_fill(pad.unsqueeze(-1), -1)
        w = w.masked_fill(pad.unsqueeze(-1), 0.0)
    return w, ids


def grouped_topk(gating_output, topk, renormalize, num_expert_group, topk_group, num_fused_shared_experts=0,
                 routed_scaling_factor=None, num_token_non_padded=None):
    """test_grouped_topk.py:9-39 with the tie-break made explicit.  gating_output: any float dtype,
    math in f32 (the script feeds .float() inputs, :49-50)."""
    scores = torch.softmax(gating_output.float(), dim=-1)
    n, E = scores.shape
    G = max(1, num_expert_group)
    group_scores = scores.view(n, G, -1).max(dim=-1).values
    _, group_idx = _topk_lowest_index(group_scores, topk_group if topk_group > 0 else G)
    group_mask = torch.zeros_like(group_scores)
    group_mask.scatter_(1, group_idx, 1)
    score_mask = group_mask.unsqueeze(-1).expand(n, G, E // G).reshape(n, -1)
    tmp = scores.masked_fill(~score_mask.bool(), 0.0)
    w, ids = _topk_lowest_index(tmp, topk)
    return _serving_args(w, ids, E, renormalize, num_fused_shared_experts, routed_scaling_factor, num_token_non_padded)


def biased_grouped_topk(gating_output, correction_bias, topk, renormalize, num_expert_group, topk_group,
                        num_fused_shared_experts=0, routed_scaling_factor=None, num_token_non_padded=None,
                        sigmoid=sigmoid_f64):
    """test_biased_grouped_topk.py:9-47 with the tie-break made explicit and the sigmoid pinned to sigmoid_f64
    (pass sigmoid=torch.sigmoid for the reference's literal expression)."""
    scores = sigmoid(gating_output.float())
    n, E = scores.shape
    G = max(1, num_expert_group)
    choice = scores + correction_bias.float().unsqueeze(0)
    gsz = E // G
    if gsz >= 2:
        group_scores = choice.view(n, G, -1).topk(2, dim=-1)[0].sum(dim=-1)
    else:
        group_scores = choice.view(n, G, -1).sum(dim=-1)
    _, group_idx = _topk_lowest_index(group_scores, topk_group if topk_group > 0 else G)
    group_mask = torch.zeros_like(group_scores)
    group_mask.scatter_(1, group_idx, 1)
    score_mask = group_mask.unsqueeze(-1).expand(n, G, gsz).reshape(n, -1)
    tmp = choice.masked_fill(~score_mask.bool(), float("-inf"))
    _, ids = _topk_lowest_index(tmp, topk)
    w = scores.gather(1, ids)
    return _serving_args(w, ids, E, renormalize, num_fused_shared_experts, routed_scaling_factor, num_token_non_padded)


# ---------------------------------------------------------------------------------------------------
# moe_align (layout of csrc/align.cu; SURVEY.md 3b for where it sits)


def moe_align(topk_ids, num_experts, chunk_rows=8, tile_rows=128):
    ids = topk_ids.reshape(-1).to(torch.int64)
    S = ids.numel()
    valid = (ids >= 0) & (ids < num_experts)
    slots = torch.arange(S)
    key = torch.where(valid, ids, torch.full_like(ids, num_experts))
    order = torch.sort(key, stable=True).indices
    nvalid = int(valid.sum())
    sorted_slot = torch.full((S,), S, dtype=torch.int32)
    sorted_slot[:nvalid] = order[:nvalid].to(torch.int32)
    slot_pos = torch.full((S,), -1, dtype=torch.int32)
    slot_pos[order[:nvalid]] = torch.arange(nvalid, dtype=torch.int32)
    count = torch.bincount(ids[valid], minlength=num_experts).to(torch.int32)
    offset = torch.zeros(num_experts + 1, dtype=torch.int32)
    offset[1:] = torch.cumsum(count, 0)

    def items(rows):
        out = []
        for e in range(num_experts):
            c, o = int(count[e]), int(offset[e])
            for r0 in range(0, c, rows):
                out.append((e, o + r0, min(rows, c - r0)))
        return torch.tensor(out, dtype=torch.int32).reshape(-1, 3)

    chunks, tiles = items(chunk_rows), items(tile_rows)
    counters = torch.tensor([nvalid, chunks.shape[0], tiles.shape[0], int((count > 0).sum())], dtype=torch.int32)
    return dict(expert_count=count, expert_offset=offset, sorted_slot=sorted_slot, slot_pos=slot_pos,
                chunk_list=chunks, tile_list=tiles, counters=counters)


# ---------------------------------------------------------------------------------------------------
# fused MoE oracles


def _masked_rows(topk_ids, i):
    return topk_ids.reshape(-1) == i


def torch_naive_moe(a, w1, w2, topk_weight, topk_ids):
    """test_moe.py:22-54 with the routing passed in (the script recomputes softmax+topk inline):
    bf16 matmuls, bf16 silu_and_mul, weights applied in a.dtype, sum over topk."""
    B, D = a.shape
    topk = topk_ids.shape[1]
    a_rep = a.view(B, -1, D).repeat(1, topk, 1).reshape(-1, D)
    out = torch.zeros(B * topk, w2.shape[1], dtype=a.dtype)
    for i in range(w1.shape[0]):
        mask = _masked_rows(topk_ids, i)
        if mask.sum():
            out[mask] = silu_and_mul(a_rep[mask] @ w1[i].transpose(0, 1)) @ w2[i].transpose(0, 1)
    return (out.view(B, -1, w2.shape[1]) * topk_weight.view(B, -1, 1).to(out.dtype)).sum(dim=1)


def swiglu_oai(x, alpha, limit):
    """gpt-oss clamped swiglu (on-box vllm SwigluOAIAndMul.forward_native, activation.py:388-395) with the gate / up
    halves laid out like every other format of this op (gate = first half, up = second half; the interleaved
    checkpoint layout is a loader concern).  PARITY UNPINNED: no reference call site passes alpha / limit."""
    d = x.shape[-1] // 2
    gate, up = x[..., :d], x[..., d:]
    if limit is not None and limit > 0:
        gate = gate.clamp(min=None, max=limit)
        up = up.clamp(min=-limit, max=limit)
    return gate * torch.sigmoid(gate * (1.702 if alpha is None else alpha)) * (up + 1)


def native_fused_moe(a, w1, w2, topk_weight, topk_ids, w1_bias=None, w2_bias=None, alpha=None, limit=None):
    """test_moe_fp8_ext.py:70-91 / test_moe_offloading_cpu.py:28-52: f32 everywhere, ids == -1 (or any
    id without an expert) contribute zero rows; result cast to a.dtype.  The optional arguments are the
    17-argument vintage (per-expert biases, clamped swiglu) -- parity unpinned, see swiglu_oai."""
    B, D = a.shape
    topk = topk_ids.shape[1]
    old = a.dtype
    a_rep = a.view(B, -1, D).repeat(1, topk, 1).reshape(-1, D).float()
    out = torch.zeros(B * topk, w2.shape[1], dtype=torch.float32)
    oai = alpha is not None or limit is not None
    for i in range(w1.shape[0]):
        mask = _masked_rows(topk_ids, i)
        if mask.sum():
            ic0 = torch.matmul(a_rep[mask], w1[i].float().transpose(0, 1))
            if w1_bias is not None:
                ic0 = ic0 + w1_bias[i].float()
            ic1 = swiglu_oai(ic0, alpha, limit) if oai else silu_and_mul(ic0)
            o = torch.matmul(ic1, w2[i].float().transpose(0, 1))
            if w2_bias is not None:
                o = o + w2_bias[i].float()
            out[mask] = o
    return (out.view(B, -1, w2.shape[1]) * topk_weight.view(B, -1, 1).to(out.dtype)).sum(dim=1).to(old)


def torch_w8a8_per_column_moe(a, w1, w2, w1_s, w2_s, topk_weight, topk_ids):
    """test_moe_int8.py:59-94."""
    B, D = a.shape
    topk = topk_ids.shape[1]
    a_q, a_s = per_token_quant_int8(a)
    a_q = a_q.view(B, -1, D).repeat(1, topk, 1).reshape(-1, D)
    a_s = a_s.view(B, -1, 1).repeat(1, topk, 1).reshape(-1, 1)
    out = torch.zeros(B * topk, w2.shape[1], dtype=torch.float32)
    for i in range(w1.shape[0]):
        mask = _masked_rows(topk_ids, i)
        if mask.sum():
            inter = native_w8a8_per_token_matmul(a_q[mask], w1[i], a_s[mask], w1_s[i])
            act = silu_and_mul_f32(inter)
            act_q, act_s = per_token_quant_int8(act)
            out[mask] = native_w8a8_per_token_matmul(act_q, w2[i], act_s, w2_s[i])
    return (out.view(B, -1, w2.shape[1]) * topk_weight.view(B, -1, 1).to(out.dtype)).sum(dim=1).to(a.dtype)


def fused_experts(a, w1, w2, topk_weight, topk_ids, fmt, w1_scale=None, w2_scale=None, block_size=None, **extra):
    """One entry point per weight format; fmt in {"bf16", "int8", "fp8", "mxfp4"}.
    Weights are row-major (unpacked).  extra: w1_bias, w2_bias, alpha, limit (17-argument vintage; not int8)."""
    if extra and fmt == "int8":
        raise ValueError("the int8 oracle has no bias / clamped-swiglu form")
    if fmt == "bf16":
        return native_fused_moe(a, w1, w2, topk_weight, topk_ids, **extra)
    if fmt == "int8":
        return torch_w8a8_per_column_moe(a, w1, w2, w1_scale, w2_scale, topk_weight, topk_ids)
    if fmt == "fp8":
        bn, bk = block_size
        return native_fused_moe(a, scaled_weight(w1, w1_scale, bn, bk), scaled_weight(w2, w2_scale, bn, bk),
                                topk_weight, topk_ids, **extra)
    if fmt == "mxfp4":
        w1d = mxfp4_dequantize(w1, torch.float32, w1_scale)
        w2d = mxfp4_dequantize(w2, torch.float32, w2_scale)
        return native_fused_moe(a, w1d, w2d, topk_weight, topk_ids, **extra)
    raise ValueError(fmt)


def fused_experts_lean(a, w1, w2, topk_weight, topk_ids, fmt, w1_scale=None, w2_scale=None, block_size=None):
    """fused_experts with the weight-only formats dequantised PER ACTIVE EXPERT inside the loop instead of for the
    whole [E, ...] tensor up front (a DeepSeek-V3 layer is 11 GB of e4m3: 45 GB as f32).  Same arithmetic, same
    results bit for bit (tests/test_oracle.py); used by bench.py's CPU legs."""
    if fmt in ("bf16", "int8"):
        return fused_experts(a, w1, w2, topk_weight, topk_ids, fmt, w1_scale, w2_scale, block_size)
    B, D = a.shape
    topk = topk_ids.shape[1]
    a_rep = a.view(B, -1, D).repeat(1, topk, 1).reshape(-1, D).float()
    out = torch.zeros(B * topk, w2.shape[1], dtype=torch.float32)

    def deq(w, s, i):
        if fmt == "fp8":
            return scaled_weight(w[i], s[i], block_size[0], block_size[1])
        return mxfp4_dequantize(w[i], torch.float32, s[i])
    for i in range(w1.shape[0]):
        mask = _masked_rows(topk_ids, i)
        if mask.sum():
            ic1 = silu_and_mul(torch.matmul(a_rep[mask], deq(w1, w1_scale, i).transpose(0, 1)))
            out[mask] = torch.matmul(ic1, deq(w2, w2_scale, i).transpose(0, 1))
    return (out.view(B, -1, w2.shape[1]) * topk_weight.view(B, -1, 1).to(out.dtype)).sum(dim=1).to(a.dtype)


# ---------------------------------------------------------------------------------------------------
# CPU-baseline variant with the weight-only formats dequantised ONCE (bench.py's "pre-dequantised bf16" figure)
#
# The literal oracles above re-dequantise every active expert to f32 on every call: fine for a checker, unfair as a
# timed CPU baseline (no CPU kernel would do that per call).  This variant hoists the dequantisation out of the
# step -- weights become bf16-resident, which is what test_moe_fp8_ext.py:70-91 itself feeds its GEMMs ("dequantised
# to the activation dtype") -- and runs the reference's bf16 oracle, torch_naive_moe (test_moe.py:22-54: bf16
# matmuls, oneDNN / AMX on a Xeon that has it), on them.  Same algorithm, bf16 GEMM rounding instead of f32.


def dequantise_to_bf16(w, scale, fmt, block_size=None):
    """[E, N, K'] (or 2-D) quantised weights -> bf16 [E, N, K], one expert at a time (bounded temporaries)."""
    if fmt in ("bf16", "int8"):
        raise ValueError("only the weight-only formats (fp8, mxfp4) have a pre-dequantised form")
    squeeze = w.dim() == 2
    if squeeze:
        w, scale = w.unsqueeze(0), scale.unsqueeze(0)
    K = w.shape[2] * (2 if fmt == "mxfp4" else 1)
    out = torch.empty(w.shape[0], w.shape[1], K, dtype=torch.bfloat16)
    for i in range(w.shape[0]):
        if fmt == "fp8":
            out[i] = scaled_weight(w[i], scale[i], block_size[0], block_size[1]).bfloat16()
        else:
            out[i] = mxfp4_dequantize(w[i], torch.bfloat16, scale[i])
    return out[0] if squeeze else out


def shared_expert_resident_bf16(a, w1_bf16, w2_bf16, fused_out, routed_scaling_factor):
    """shared_expert on bf16-resident weights: bf16 GEMMs, f32 epilogue (test_shared_experts.py:34-40 in a.dtype)."""
    ic1 = silu_and_mul(a @ w1_bf16.t())
    return ((ic1 @ w2_bf16.t()).float() + fused_out.float() * routed_scaling_factor).to(a.dtype)


# ---------------------------------------------------------------------------------------------------
# shared expert


def shared_expert(a, w1, w2, fused_out, routed_scaling_factor, fmt, w1_scale=None, w2_scale=None,
                  block_size=None):
    """bf16: test_shared_experts.py:34-40 (f32 inputs, :64); int8: :42-53; fp8: test_moe_fp8_ext.py:51-56."""
    dtype = a.dtype
    if fmt == "int8":
        a_q, a_s = per_token_quant_int8(a.float())
        ic1 = native_w8a8_per_token_matmul(a_q, w1, a_s, w1_scale)
        ic2 = silu_and_mul(ic1)
        q2, s2 = per_token_quant_int8(ic2)
        ic3 = native_w8a8_per_token_matmul(q2, w2, s2, w2_scale)
        return (ic3 + fused_out.float() * routed_scaling_factor).to(dtype)
    if fmt == "fp8":
        bn, bk = block_size
        w1, w2 = scaled_weight(w1, w1_scale, bn, bk), scaled_weight(w2, w2_scale, bn, bk)
    ic1 = silu_and_mul(torch.matmul(a.float(), w1.float().t()))
    out = torch.matmul(ic1, w2.float().t()) + fused_out.float() * routed_scaling_factor
    return out.to(dtype)


# ---------------------------------------------------------------------------------------------------
# prepack layouts


def pack_weight_vnni(w):
    """bf16/fp16: test_gemm.py:38-44, test_moe.py:117-124 -> [E, N/32, K/2, 32, 2], same shape.
    int8: test_gemm_int8.py:84-93 -> per 32-row block [K/4, 32, 4] bytes followed by 32 int32
    compensation values 128*sum_k w[n,k] (behind-the-block placement follows the upstream kernel;
    the reference pins it for N = 32 only); returned as [.., N, K+4] int8.
    fp8 / uint8: same byte tiling as int8 without compensation (not pinned by the reference)."""
    squeeze = w.dim() == 2
    if squeeze:
        w = w.unsqueeze(0)
    E, N, K = w.shape
    if w.dtype in (torch.bfloat16, torch.float16):
        out = w.view(E, N // 32, 32, K // 2, 2).permute(0, 1, 3, 2, 4).contiguous().view(E, N, K)
    elif w.dtype == torch.int8:
        body = w.view(E, N // 32, 32, K // 4, 4).permute(0, 1, 3, 2, 4).contiguous().view(E, N // 32, 32 * K)
        comp = (128 * w.to(torch.int32).sum(dim=-1)).to(torch.int32).view(E, N // 32, 32)
        comp_bytes = comp.contiguous().view(torch.int8).view(E, N // 32, 128)
        out = torch.cat([body, comp_bytes], dim=-1).view(E, N, K + 4)
    else:
        b = w.view(torch.uint8)
        out = b.view(E, N // 32, 32, K // 4, 4).permute(0, 1, 3, 2, 4).contiguous().view(E, N, K).view(w.dtype)
    return out.squeeze(0) if squeeze else out


def pack_scale(s):
    """test_mxfp4.py:185: Bs.view(N/32, 32, K/32).transpose(1, 2), flat order, same shape."""
    squeeze = s.dim() == 2
    if squeeze:
        s = s.unsqueeze(0)
    E, N, G = s.shape
    out = s.view(E, N // 32, 32, G).transpose(2, 3).contiguous().view(E, N, G)
    return out.squeeze(0) if squeeze else out


# ---------------------------------------------------------------------------------------------------
# expert-parallel exchange (no counterpart in the reference; "parity unpinned": the layout is this
# build's own contract, stated identically in csrc/ep.cu.  The only reference hook is that
# fused_experts_cpu ignores ids == -1, test_moe_offloading_cpu.py:57-68.)


def slot_rows(a, w1, w2, topk_weight, topk_ids):
    """One bf16 row per routed slot, [B * topk, H]: bf16(f32(W2[e] . silu_and_mul(W1[e] . x)) * weight) -- what a
    single GPU holds in its ic2 buffer before the top-k reduce.  Rows of masked slots are zero.  f32 weights."""
    B, D = a.shape
    topk = topk_ids.shape[1]
    a_rep = a.view(B, -1, D).repeat(1, topk, 1).reshape(-1, D).float()
    out = torch.zeros(B * topk, w2.shape[1], dtype=torch.float32)
    for i in range(w1.shape[0]):
        mask = _masked_rows(topk_ids, i)
        if mask.sum():
            ic1 = silu_and_mul(torch.matmul(a_rep[mask], w1[i].float().transpose(0, 1)))
            out[mask] = torch.matmul(ic1, w2[i].float().transpose(0, 1))
    return (out * topk_weight.reshape(-1, 1).float()).to(a.dtype)


def reduce_slots(rows, topk_ids, num_experts):
    """csrc/misc.cu topk_reduce: out[m] = bf16(sum_j f32(rows[m, j])) over valid slots, j ascending."""
    B, topk = topk_ids.shape
    acc = torch.zeros(B, rows.shape[-1], dtype=torch.float32)
    r = rows.view(B, topk, rows.shape[-1])
    for j in range(topk):
        ok = (topk_ids[:, j] >= 0) & (topk_ids[:, j] < num_experts)
        acc[ok] += r[ok, j].float()
    return acc.to(rows.dtype)


def ep_plan(topk_ids, num_experts, world):
    """dest_row[m, r] = index of token m in the row segment sent to rank r (token order) or -1;
    send_count[r] = rows in that segment;  slot_index[m, j] = index of slot (m, j) among the slots owned by its
    owner rank, numbered in (token, j) order, -1 for ids outside [0, E);  send_slots[r] = slots owned by r.
    Rank r owns experts [r*E/world, (r+1)*E/world)."""
    M, topk = topk_ids.shape
    e_local = num_experts // world
    dest_row = torch.full((M, world), -1, dtype=torch.int32)
    send_count = torch.zeros(world, dtype=torch.int32)
    slot_index = torch.full((M, topk), -1, dtype=torch.int32)
    send_slots = torch.zeros(world, dtype=torch.int32)
    ids = topk_ids.long()
    for r in range(world):
        mine = (ids >= r * e_local) & (ids < (r + 1) * e_local)
        flag = mine.any(dim=1)
        n = int(flag.sum())
        dest_row[flag, r] = torch.arange(n, dtype=torch.int32)
        send_count[r] = n
        flat = mine.reshape(-1)
        k = int(flat.sum())
        slot_index.view(-1)[flat] = torch.arange(k, dtype=torch.int32)
        send_slots[r] = k
    return dest_row, send_count, slot_index, send_slots


def ep_seg_start(count, world, cap):
    if cap > 0:
        return [r * cap for r in range(world)]
    c = [int(v) for v in count]
    return [sum(c[:r]) for r in range(world)]


def ep_pack(x, topk_ids, topk_weights, dest_row, send_count, num_experts, world, cap):
    """send_x [rows, H], send_meta [rows, 2*topk] int32 (local ids | weight bits); padding rows = -1."""
    M, topk = topk_ids.shape
    e_local = num_experts // world
    rows = cap * world if cap > 0 else int(send_count.sum())
    seg = ep_seg_start(send_count, world, cap)
    send_x = torch.zeros(rows, x.shape[1], dtype=x.dtype)
    meta = torch.full((rows, 2 * topk), -1, dtype=torch.int32)
    wbits = topk_weights.float().contiguous().view(torch.int32)
    for r in range(world):
        sel = dest_row[:, r] >= 0
        pos = seg[r] + dest_row[sel, r].long()
        send_x[pos] = x[sel]
        ids = topk_ids[sel].to(torch.int32)
        mine = (ids >= r * e_local) & (ids < (r + 1) * e_local)
        meta[pos, :topk] = torch.where(mine, ids - r * e_local, torch.full_like(ids, -1))
        meta[pos, topk:] = torch.where(mine, wbits[sel], torch.zeros_like(ids))
    return send_x, meta


def ep_ret_index(recv_meta, topk, world, cap, recv_count=None, recv_slots=None):
    """Owner side: position of received slot (row i, j) in the return buffer, -1 for masked slots.  Valid slots
    of the rows received from source s are numbered in row order from slot_seg(s) = s*cap*topk (fixed capacity)
    or the exclusive scan of recv_slots (exact splits) -- the numbering the source computed in ep_plan."""
    rows = recv_meta.shape[0]
    ids = recv_meta[:, :topk]
    out = torch.full((rows, topk), -1, dtype=torch.int32)
    row_seg = ep_seg_start(recv_count, world, cap) if cap == 0 else [s * cap for s in range(world)]
    slot_seg = ep_seg_start(recv_slots, world, 0) if cap == 0 else [s * cap * topk for s in range(world)]
    for s in range(world):
        n = cap if cap > 0 else int(recv_count[s])
        blk = ids[row_seg[s]:row_seg[s] + n]
        valid = (blk >= 0).reshape(-1)
        o = out[row_seg[s]:row_seg[s] + n].view(-1)
        o[valid] = slot_seg[s] + torch.arange(int(valid.sum()), dtype=torch.int32)
    return out


def ep_combine(ret, topk_ids, slot_index, send_slots, num_experts, world, cap, shared_y=None, rsf=1.0):
    """out[m] = bf16(sum_j f32(ret[slot_seg(owner(m, j)) + slot_index[m, j]])), j ascending over valid slots (the
    single-GPU reduce_slots order); with shared_y: bf16(f32(shared_y) + f32(that bf16) * rsf), fused multiply-add
    (csrc/misc.cu shared_axpy_kernel)."""
    M, topk = topk_ids.shape
    e_local = num_experts // world
    seg = torch.tensor(ep_seg_start(send_slots, world, cap * topk if cap > 0 else 0), dtype=torch.long)
    acc = torch.zeros(M, ret.shape[1], dtype=torch.float32)
    for j in range(topk):
        e = topk_ids[:, j].long()
        ok = (e >= 0) & (e < num_experts)
        pos = seg[(e[ok] // e_local)] + slot_index[ok, j].long()
        acc[ok] += ret[pos].float()
    out = acc.to(ret.dtype)
    if shared_y is not None:
        rsf32 = float(torch.tensor(rsf, dtype=torch.float32))
        out = (shared_y.double() + out.double() * rsf32).float().to(ret.dtype)   # fmaf in f32, then the bf16 rounding
    return out
