"""Dense quantised linears: weight_packed_linear, per_token_quant_int8_cpu,
int8_scaled_mm_cpu, int8_scaled_mm_with_quant, fp8_scaled_mm_cpu.

CPU part: the oracle restatements against fixtures made by the reference's own oracle code
(tests/golden/make_golden_gemm.py).  GPU part (-m gpu): the CUDA path through the C ABI against the same
fixtures, against the oracle on seeded inputs on both kernel paths, and with packed (is_vnni) weights."""
import math

import pytest
import torch

from conftest import load_golden
from oracle import moe_oracle as O

BF16_TOL = dict(rtol=1e-2, atol=1e-2)  # utils.py:3-12 of the reference


def _calls(case):
    return {n: a for n, a in case["calls"]}


# ------------------------------------------------------------------------------------------ oracle (CPU)
def test_oracle_linear_bf16_golden():
    for case in load_golden("gemm_bf16"):
        mat1, mat2, bias, _ = _calls(case)["weight_packed_linear"]
        assert torch.equal(O.linear_bf16(mat1, mat2, bias), case["refs"][0])


def test_oracle_linear_int8_golden():
    for case in load_golden("gemm_int8"):
        c = _calls(case)
        (A,) = c["per_token_quant_int8_cpu"]
        Aq, As = O.per_token_quant_int8_gemm(A)
        _, Bq, Bs, bias, _, _ = c["int8_scaled_mm_with_quant"]
        assert torch.equal(O.linear_int8(Aq, Bq, As, Bs, bias), case["refs"][0])
        assert torch.equal(O.linear_int8_with_quant(A, Bq, Bs, bias), case["refs"][1])


def test_oracle_linear_fp8_golden():
    for case in load_golden("gemm_fp8"):
        A, Wq, scales, block, bias, _, _ = _calls(case)["fp8_scaled_mm_cpu"]
        assert torch.equal(O.linear_fp8(A, Wq, scales, block, bias), case["refs"][0])


# ------------------------------------------------------------------------------------------ CUDA (GPU)
@pytest.fixture(scope="module")
def K():
    import sgl_cpu_tests_b200 as K
    assert torch.cuda.is_available()
    return K


@pytest.mark.gpu
def test_linear_bf16_golden(K):
    for case in load_golden("gemm_bf16"):
        mat1, mat2, bias, _ = _calls(case)["weight_packed_linear"]
        out = K.ops.weight_packed_linear(mat1, mat2, bias, False)
        torch.testing.assert_close(out, case["refs"][0], **BF16_TOL)
        packed = K.ops.convert_weight_packed(mat2)
        out2 = K.ops.weight_packed_linear(mat1, packed, bias, True)
        assert torch.equal(out, out2)


@pytest.mark.gpu
def test_linear_int8_golden(K):
    for case in load_golden("gemm_int8"):
        c = _calls(case)
        (A,) = c["per_token_quant_int8_cpu"]
        _, Bq, Bs, bias, _, _ = c["int8_scaled_mm_with_quant"]
        Aq, As = K.ops.per_token_quant_int8(A)
        rq, rs = O.per_token_quant_int8_gemm(A)
        assert torch.equal(Aq, rq) and torch.equal(As, rs)  # bit-exact quantiser
        out = K.ops.int8_scaled_mm(Aq, Bq, As, Bs, bias, torch.bfloat16, False)
        torch.testing.assert_close(out, case["refs"][0], **BF16_TOL)
        fused = K.ops.int8_scaled_mm_with_quant(A, Bq, Bs, bias, torch.bfloat16, False)
        torch.testing.assert_close(fused, case["refs"][1], **BF16_TOL)
        assert torch.equal(out, fused)
        packed = K.ops.convert_weight_packed(Bq)
        assert torch.equal(K.ops.int8_scaled_mm_with_quant(A, packed, Bs, bias, torch.bfloat16, True), fused)


@pytest.mark.gpu
def test_linear_fp8_golden(K):
    for case in load_golden("gemm_fp8"):
        A, Wq, scales, block, bias, dt, _ = _calls(case)["fp8_scaled_mm_cpu"]
        out = K.ops.fp8_scaled_mm(A, Wq, scales, block, bias, dt, False)
        torch.testing.assert_close(out, case["refs"][0], **BF16_TOL)
        packed = K.ops.convert_weight_packed(Wq)
        assert torch.equal(K.ops.fp8_scaled_mm(A, packed, scales, block, bias, dt, True), out)


def _paths(K):
    return [K.L.PATH_STREAM, K.L.PATH_UMMA]


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K_", [(1, 256, 512), (5, 416, 512), (129, 2816, 1024), (1000, 520, 2048)])
@pytest.mark.parametrize("has_bias", [False, True])
def test_linear_bf16_paths(K, M, N, K_, has_bias):
    torch.manual_seed(M + N)
    a = torch.randn(M, K_, dtype=torch.bfloat16) / 4
    w = torch.randn(N, K_, dtype=torch.bfloat16) / math.sqrt(K_)
    bias = torch.randn(N) if has_bias else None
    ref = O.linear_bf16(a, w, bias)
    for path in _paths(K):
        out = K.ops._linear(a.cuda(), w.cuda(), K.L.FMT_BF16, False, bias=bias, path=path)
        torch.testing.assert_close(out.cpu(), ref, **BF16_TOL)


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K_", [(1, 256, 512), (7, 384, 1024), (300, 2816, 1024)])
def test_linear_int8_paths(K, M, N, K_):
    torch.manual_seed(M + N)
    a = torch.randn(M, K_, dtype=torch.bfloat16) / 10
    bq = ((torch.rand(N, K_) - 0.5) * 2 * 127).clamp(-128, 127).to(torch.int8)
    bs = torch.rand(N) * 1e-2
    bias = torch.randn(N)
    ref = O.linear_int8_with_quant(a, bq, bs, bias)
    for path in _paths(K):
        out = K.ops._linear(a, bq, K.L.FMT_INT8_W8A8, False, w_scale=bs, bias=bias, path=path)
        torch.testing.assert_close(out, ref, **BF16_TOL)


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K_,block", [(1, 2816, 7168, [64, 128]), (4, 2816, 7168, [128, 128]),
                                          (260, 512, 1024, [64, 128]), (1024, 2816, 1024, [128, 128])])
def test_linear_fp8_paths(K, M, N, K_, block):
    torch.manual_seed(M + N)
    a = torch.randn(M, K_, dtype=torch.bfloat16) / math.sqrt(K_)
    wq = (torch.randn(N, K_) * 400).clamp(-400, 400).to(torch.float8_e4m3fn)
    scales = torch.randn(-(-N // block[0]), -(-K_ // block[1])) * 1e-3
    bias = torch.randn(N)
    ref = O.linear_fp8(a, wq, scales, block, bias)
    for path in _paths(K):
        out = K.ops._linear(a, wq, K.L.FMT_FP8_W8A16, False, w_scale=scales, block_size=block, bias=bias,
                            path=path)
        torch.testing.assert_close(out, ref, **BF16_TOL)


@pytest.mark.gpu
def test_linear_errors_and_empty(K):
    a = torch.randn(0, 64, dtype=torch.bfloat16)
    w = torch.randn(32, 64, dtype=torch.bfloat16)
    assert K.ops.weight_packed_linear(a, w, None, False).shape == (0, 32)
    with pytest.raises(RuntimeError):
        K.ops.weight_packed_linear(torch.randn(2, 48, dtype=torch.bfloat16), w, None, False)
    with pytest.raises(RuntimeError):
        K.ops.int8_scaled_mm_with_quant(torch.randn(2, 64, dtype=torch.bfloat16), w.to(torch.int8),
                                        torch.rand(32), None, torch.float16, False)
