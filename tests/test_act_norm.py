"""silu_and_mul_cpu, rmsnorm_cpu, fused_add_rmsnorm_cpu.

CPU part: the oracle restatements against fixtures made by the reference's own functions in test_activation.py /
test_norm.py (tests/golden/make_golden_gemm.py).  GPU part (-m gpu): the CUDA kernels through the C ABI against
the same fixtures (tolerances of bench_silu_and_mul.py:32 and utils.py:3-12) and against the oracle on seeded
inputs at the reference's bench shapes, for device tensors (in place) and host tensors (staged)."""
import pytest
import torch

from conftest import load_golden
from oracle import moe_oracle as O


def tol(dtype):
    return dict(rtol=1e-2, atol=1e-2) if dtype is torch.bfloat16 else dict(rtol=1e-3, atol=1e-3)


def _cases(op):
    return [c for c in load_golden("act_norm") if c["calls"][0][0] == op]


def test_oracle_silu_and_mul_golden():
    for case in _cases("silu_and_mul_cpu"):
        _, x = case["calls"][0][1]
        assert torch.equal(O.silu_and_mul(x), case["refs"][0])


def test_oracle_rmsnorm_golden():
    for case in _cases("rmsnorm_cpu"):
        calls = dict(case["calls"])
        _, x, w, eps = calls["rmsnorm_cpu"]
        assert torch.equal(O.rmsnorm(x, w, eps), case["refs"][0])
        x2, res, w2, eps2 = calls["fused_add_rmsnorm_cpu"]
        y, r = O.rmsnorm(x2, w2, eps2, res)
        assert torch.equal(y, case["refs"][1]) and torch.equal(r, case["refs"][2])


@pytest.fixture(scope="module")
def K():
    import sgl_cpu_tests_b200 as K
    assert torch.cuda.is_available()
    return K


@pytest.mark.gpu
def test_silu_and_mul_golden(K):
    for case in _cases("silu_and_mul_cpu"):
        out, x = case["calls"][0][1]
        got = K.ops.silu_and_mul(x)                       # functional form, host tensor
        torch.testing.assert_close(got, case["refs"][0], **tol(x.dtype))
        out = torch.empty_like(out)
        assert K.ops.silu_and_mul(out, x) is None         # out variant
        assert torch.equal(out, got)
        xd = x.cuda()
        assert torch.equal(K.ops.silu_and_mul(xd).cpu(), got)  # device tensor


@pytest.mark.gpu
def test_rmsnorm_golden(K):
    for case in _cases("rmsnorm_cpu"):
        calls = dict(case["calls"])
        out, x, w, eps = calls["rmsnorm_cpu"]
        out = torch.empty_like(x)
        K.ops.rmsnorm(out, x, w, eps)
        torch.testing.assert_close(out, case["refs"][0], rtol=1e-2, atol=1e-2)
        x2, res, w2, eps2 = calls["fused_add_rmsnorm_cpu"]
        xh, rh = x2.clone(), res.clone()
        K.ops.fused_add_rmsnorm(xh, rh, w2, eps2)         # host tensors, updated in place
        torch.testing.assert_close(xh, case["refs"][1], rtol=1e-2, atol=1e-2)
        assert torch.equal(rh, case["refs"][2])           # the residual sum is one rounding of an exact f32 add
        xd, rd = x2.cuda(), res.cuda()
        K.ops.fused_add_rmsnorm(xd, rd, w2.cuda(), eps2)  # device tensors, in place
        assert torch.equal(xd.cpu(), xh) and torch.equal(rd.cpu(), rh)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(1, 18432), (17, 18432), (1000, 18432), (3, 14)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_silu_and_mul_bench_shapes(K, shape, dtype):
    torch.manual_seed(shape[0])
    x = torch.randn(shape, dtype=dtype)
    torch.testing.assert_close(K.ops.silu_and_mul(x.cuda()).cpu(), O.silu_and_mul(x), **tol(dtype))


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(1, 5120), (1024, 4096), (7, 4109), (2, 16384), (0, 64)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_rmsnorm_shapes(K, shape, dtype):
    torch.manual_seed(shape[1])
    x, res, w = torch.randn(shape, dtype=dtype), torch.randn(shape, dtype=dtype), torch.randn(shape[1], dtype=dtype)
    out = torch.empty_like(x).cuda()
    K.ops.rmsnorm(out, x.cuda(), w.cuda(), 1e-6)
    torch.testing.assert_close(out.cpu(), O.rmsnorm(x, w, 1e-6), rtol=1e-2, atol=1e-2)
    xd, rd = x.cuda(), res.cuda()
    K.ops.fused_add_rmsnorm(xd, rd, w.cuda(), 1e-6)
    y, r = O.rmsnorm(x, w, 1e-6, res)
    torch.testing.assert_close(xd.cpu(), y, rtol=1e-2, atol=1e-2)
    assert torch.equal(rd.cpu(), r)


@pytest.mark.gpu
def test_act_norm_errors(K):
    with pytest.raises(RuntimeError):
        K.ops.silu_and_mul(torch.randn(2, 8))             # f32 is not a supported dtype
    with pytest.raises(RuntimeError):
        K.ops.rmsnorm(torch.empty(2, 8, dtype=torch.bfloat16), torch.randn(2, 8).bfloat16(),
                      torch.randn(4).bfloat16(), 1e-6)    # weight size mismatch
