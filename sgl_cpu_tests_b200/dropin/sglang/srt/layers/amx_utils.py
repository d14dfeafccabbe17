"""Shim for `from sglang.srt.layers.amx_utils import CPUQuantMethod` (test_moe.py:4)."""
from sgl_cpu_tests_b200.ops import CPUQuantMethod  # noqa: F401
