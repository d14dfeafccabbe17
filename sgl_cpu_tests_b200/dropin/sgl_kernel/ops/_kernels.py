"""`sgl_kernel.ops._kernels` (test_activation.py:10 imports from here).  Only the ops on the
fused-MoE path exist in this build; standalone silu_and_mul_cpu is out of scope."""
from ..common_ops import *  # noqa: F401,F403
