"""Shim for `from sglang.srt.layers.amx_utils import CPUQuantMethod` (test_moe.py:4).  The reference imports it before
`import sgl_kernel`, so this module finds the package on its own."""
import os
import sys

_root = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", "..", "..", "..", ".."))
if _root not in sys.path:
    sys.path.insert(0, _root)

from sgl_cpu_tests_b200.ops import CPUQuantMethod  # noqa: E402,F401
