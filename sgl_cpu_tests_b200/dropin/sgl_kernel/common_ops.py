"""`sgl_kernel.common_ops`: the older import style of the reference scripts
(test_grouped_topk.py:2, test_biased_grouped_topk.py:2, test_moe_fp8.py:4-5).  Routers are the
out-variants those scripts call; a functional call (9 / 10 positional args) is accepted too."""
import torch

from sgl_cpu_tests_b200 import ops as _ops

convert_weight_packed = _ops.convert_weight_packed
convert_scale_packed = _ops.convert_scale_packed
fused_experts_cpu = _ops.fused_experts
shared_expert_cpu = _ops.shared_expert
mxfp4_scaled_mm_cpu = _ops.mxfp4_scaled_mm
weight_packed_linear = _ops.weight_packed_linear
per_token_quant_int8_cpu = _ops.per_token_quant_int8
int8_scaled_mm_cpu = _ops.int8_scaled_mm
int8_scaled_mm_with_quant = _ops.int8_scaled_mm_with_quant
fp8_scaled_mm_cpu = _ops.fp8_scaled_mm
silu_and_mul_cpu = _ops.silu_and_mul
rmsnorm_cpu = _ops.rmsnorm
fused_add_rmsnorm_cpu = _ops.fused_add_rmsnorm


def grouped_topk_cpu(*args):
    if len(args) == 8:
        return _ops.grouped_topk_out(*args)
    return _ops.grouped_topk(*args)


def biased_grouped_topk_cpu(*args):
    if len(args) == 9 and args[0].dtype == torch.float32 and args[1].dtype == torch.int32:
        return _ops.biased_grouped_topk_out(*args)
    return _ops.biased_grouped_topk(*args)
