"""Drop-in `sgl_kernel`: importing it makes torch.ops.sgl_kernel.<op> resolve to the sm_100a
implementation, exactly as `import sgl_kernel` does for the reference scripts (test_moe.py:5-8).

Ops are registered with the PyTorch dispatcher under namespace `sgl_kernel`; fused_experts_cpu and
shared_expert_cpu carry one overload per signature vintage found in the reference,
selected by arity.  The implementation key is CompositeExplicitAutograd so the same op accepts the
CPU tensors the reference scripts build and CUDA tensors.
"""
import os
import sys

import torch

_root = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", "..", ".."))
if _root not in sys.path:
    sys.path.insert(0, _root)

from sgl_cpu_tests_b200 import ops as _ops  # noqa: E402

_ops.L.load()  # fail loudly at import if libsglb200.so is missing

_lib = torch.library.Library("sgl_kernel", "DEF")
_KEY = "CompositeExplicitAutograd"


def _def(schema, fn):
    name = schema.split("(")[0]
    _lib.define(schema)
    _lib.impl(name, fn, _KEY)


_T, _TO = "Tensor", "Tensor?"
_def("convert_weight_packed(Tensor weight) -> Tensor", _ops.convert_weight_packed)
_def("convert_scale_packed(Tensor scale) -> Tensor", _ops.convert_scale_packed)
_def("grouped_topk_cpu(Tensor hidden_states, Tensor gating_output, int topk, bool renormalize, "
     "int num_expert_group, int topk_group, int num_fused_shared_experts, float? routed_scaling_factor, "
     "Tensor? num_token_non_padded) -> (Tensor, Tensor)", _ops.grouped_topk)
_def("biased_grouped_topk_cpu(Tensor hidden_states, Tensor gating_output, Tensor correction_bias, int topk, "
     "bool renormalize, int num_expert_group, int topk_group, int num_fused_shared_experts, "
     "float? routed_scaling_factor, Tensor? num_token_non_padded) -> (Tensor, Tensor)",
     _ops.biased_grouped_topk)

# fused_experts_cpu: 14-arg (default overload), 13-arg and 17-arg vintages
_def("fused_experts_cpu(Tensor hidden_states, Tensor w1, Tensor w2, Tensor topk_weights, Tensor topk_ids, "
     "bool inplace, bool use_int8_w8a8, bool use_fp8_w8a16, Tensor? w1_scale, Tensor? w2_scale, "
     "int[]? block_size, Tensor? a1_scale, Tensor? a2_scale, bool is_vnni) -> Tensor",
     lambda *a: _ops.fused_experts(*a))
_def("fused_experts_cpu.quant_method(Tensor hidden_states, Tensor w1, Tensor w2, Tensor topk_weights, "
     "Tensor topk_ids, bool inplace, int quant_method, Tensor? w1_scale, Tensor? w2_scale, Tensor? w1_zero, "
     "Tensor? w2_zero, int[]? block_size, bool is_vnni) -> Tensor",
     lambda *a: _ops.fused_experts(*a))
_def("fused_experts_cpu.bias(Tensor hidden_states, Tensor w1, Tensor w2, Tensor topk_weights, "
     "Tensor topk_ids, bool inplace, int quant_method, Tensor? w1_scale, Tensor? w2_scale, Tensor? w1_zero, "
     "Tensor? w2_zero, int[]? block_size, Tensor? w1_bias, Tensor? w2_bias, float? alpha, float? limit, "
     "bool is_vnni) -> Tensor",
     lambda *a: _ops.fused_experts(*a))

_def("shared_expert_cpu(Tensor hidden_states, Tensor w1, Tensor w2, Tensor fused_experts_out, "
     "float routed_scaling_factor, bool inplace, bool use_int8_w8a8, bool use_fp8_w8a16, Tensor? w1_scale, "
     "Tensor? w2_scale, int[]? block_size, Tensor? a1_scale, Tensor? a2_scale, bool is_vnni) -> Tensor",
     lambda *a: _ops.shared_expert(*a))
_def("shared_expert_cpu.v12(Tensor hidden_states, Tensor w1, Tensor w2, Tensor fused_experts_out, "
     "float routed_scaling_factor, bool inplace, bool use_int8_w8a8, bool use_fp8_w8a16, Tensor? w1_scale, "
     "Tensor? w2_scale, int[]? block_size, bool is_vnni) -> Tensor",
     lambda *a: _ops.shared_expert(*a))

_def("mxfp4_scaled_mm_cpu(Tensor mat1, Tensor mat2, Tensor scales2, Tensor? bias, bool is_vnni) -> Tensor",
     _ops.mxfp4_scaled_mm)

# dense quantised linears: test_gemm.py:22-25, test_gemm_int8.py:68-74, test_gemm_fp8.py:54-62
_def("weight_packed_linear(Tensor mat1, Tensor mat2, Tensor? bias, bool is_vnni) -> Tensor",
     _ops.weight_packed_linear)
_def("per_token_quant_int8_cpu(Tensor A) -> (Tensor, Tensor)", _ops.per_token_quant_int8)
_def("int8_scaled_mm_cpu(Tensor mat1, Tensor mat2, Tensor scales1, Tensor scales2, Tensor? bias, "
     "ScalarType out_dtype, bool is_vnni) -> Tensor", _ops.int8_scaled_mm)
_def("int8_scaled_mm_with_quant(Tensor mat1, Tensor mat2, Tensor scales2, Tensor? bias, ScalarType out_dtype, "
     "bool is_vnni) -> Tensor", _ops.int8_scaled_mm_with_quant)
_def("fp8_scaled_mm_cpu(Tensor mat1, Tensor mat2, Tensor scales2, int[] block_size, Tensor? bias, "
     "ScalarType out_dtype, bool is_vnni) -> Tensor", _ops.fp8_scaled_mm)

# bench_silu_and_mul.py:31 (functional); the (out, x) form of test_activation.py:26 and the
# norm out-variants of test_norm.py:43-56 are module attributes of common_ops / ops._kernels
_def("silu_and_mul_cpu(Tensor input) -> Tensor", _ops.silu_and_mul)
_def("rmsnorm_cpu(Tensor(a!) output, Tensor input, Tensor weight, float eps) -> ()", _ops.rmsnorm)
_def("fused_add_rmsnorm_cpu(Tensor(a!) input, Tensor(b!) residual, Tensor weight, float eps) -> ()",
     _ops.fused_add_rmsnorm)

class _ArityPacket:
    """Front of an op that exists in several positional vintages.  torch resolves overloads by trying each schema
    in turn and formats an error message -- including the repr of every tensor argument, about 1 ms -- for each
    one that does not match; the 13-argument fused_experts_cpu call paid that on every call.  This object sits on
    the torch.ops.sgl_kernel namespace under the op's name, dispatches on the argument count in Python, and
    forwards everything else (.default, .quant_method, overloads(), ...) to the registered OpOverloadPacket."""

    def __init__(self, name, fn):
        self._packet = getattr(torch.ops.sgl_kernel, name)
        self._fn = fn

    def __call__(self, *args):
        return self._fn(*args)

    def __getattr__(self, item):
        return getattr(self._packet, item)


for _name, _fn in (("fused_experts_cpu", _ops.fused_experts), ("shared_expert_cpu", _ops.shared_expert)):
    setattr(torch.ops.sgl_kernel, _name, _ArityPacket(_name, _fn))


from . import common_ops  # noqa: E402,F401

__all__ = ["common_ops"]
