"""`sgl_kernel.ops._kernels` (test_activation.py:10 imports from here).  The ops of the
fused-MoE path, the dense linears and silu_and_mul_cpu / the rmsnorm pair exist in this build."""
from ..common_ops import *  # noqa: F401,F403
