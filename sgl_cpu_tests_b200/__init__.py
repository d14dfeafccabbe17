"""sgl_cpu_tests_b200 -- B200 (sm_100a) implementation of the quantized fused-MoE expert path that
mingfeima/sgl-cpu-tests drives through torch.ops.sgl_kernel.

  csrc/     hand-written CUDA kernels + the C ABI (include/sglb200.h -> libsglb200.so)
  ops.py    host-side mirror of the reference operator interface (same names / argument lists)
  ep.py     expert-parallel dispatch / combine over NCCL
  dropin/   `sgl_kernel` and `sglang` shim packages: put this directory on PYTHONPATH and the
            reference's test_moe*.py / bench_moe.py run unmodified against this implementation
"""
import os

from . import _lib, ops
from . import _lib as L  # noqa: F401  (K.L.PATH_*, K.L.FMT_*)
from .ops import (CPUQuantMethod, GraphedExperts, biased_grouped_topk, biased_grouped_topk_out, clear_weight_cache,
                  convert_scale_packed, convert_weight_packed, fused_experts, fused_experts_impl,
                  grouped_topk, grouped_topk_out, invalidate, launch_count, moe_align, mxfp4_scaled_mm,
                  shared_expert, shared_expert_impl, split_rows_e4m3, weight_packed_linear, per_token_quant_int8,
                  int8_scaled_mm, int8_scaled_mm_with_quant, fp8_scaled_mm, silu_and_mul, rmsnorm,
                  fused_add_rmsnorm)

from .layer import GraphedLayer, MoELayer, Router  # noqa: E402,F401

DROPIN_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "dropin")
FMT_BF16, FMT_INT8_W8A8, FMT_FP8_W8A16, FMT_MXFP4 = (_lib.FMT_BF16, _lib.FMT_INT8_W8A8,
                                                    _lib.FMT_FP8_W8A16, _lib.FMT_MXFP4)
PATH_AUTO, PATH_STREAM, PATH_UMMA, PATH_UMMA2 = _lib.PATH_AUTO, _lib.PATH_STREAM, _lib.PATH_UMMA, _lib.PATH_UMMA2
