"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol the
header declares, the host mirror accepts the reference's signature vintages, the product never
touches the oracle, and it fails loudly without a GPU.  No compute calls here."""
import ctypes
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
PKG = os.path.join(ROOT, "sgl_cpu_tests_b200")


@pytest.fixture(scope="module")
def lib():
    from sgl_cpu_tests_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        from sgl_cpu_tests_b200 import build
        build.build()
    return _lib


def header_symbols():
    text = open(os.path.join(ROOT, "include", "sglb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sglb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(lib):
    dll = ctypes.CDLL(lib.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 14
    for s in syms:
        assert hasattr(dll, s), f"{s} declared in include/sglb200.h but not exported"
    assert sorted(lib.SIGNATURES) == syms, "ctypes table and header disagree"
    assert lib.load().sglb_version() >= 100


def test_library_does_not_link_torch_or_libcuda(lib):
    out = subprocess.run(["ldd", lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "torch" not in out and "libcuda.so" not in out and "libc10" not in out


def test_pure_size_queries(lib):
    L = lib.load()
    assert L.sglb_fused_experts_workspace_bytes(128, 8, 128, 2048, 768, lib.FMT_BF16) > 128 * 8 * (2048 + 768) * 2
    assert L.sglb_fused_experts_workspace_bytes(128, 8, 128, 2048, 768, lib.FMT_INT8_W8A8) > \
        L.sglb_fused_experts_workspace_bytes(128, 8, 128, 2048, 768, lib.FMT_BF16)
    assert L.sglb_moe_align_max_items(128, 8, 128, 8) >= 128 + 128
    assert L.sglb_moe_align_scratch_bytes(8192, 8, 256) == 64 * 256 * 4


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+\.*oracle\b", text, flags=re.M), f
                assert not re.search(r"#include.*oracle|dlopen.*oracle|import_module\(.*oracle", text), f


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_fails_loudly_without_gpu(lib):
    import sgl_cpu_tests_b200 as K
    a = torch.zeros(2, 32, dtype=torch.bfloat16)
    w1 = torch.zeros(4, 64, 32, dtype=torch.bfloat16)
    w2 = torch.zeros(4, 32, 32, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        K.fused_experts(a, w1, w2, torch.zeros(2, 2), torch.zeros(2, 2, dtype=torch.int32), True, False, False, None, None,
                        None, None, None, False)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        K.grouped_topk(a, torch.zeros(2, 8), 2, False, 1, 1, 0, None, None)


def test_signature_vintages_are_validated_before_any_device_work(lib):
    import sgl_cpu_tests_b200 as K
    a = torch.zeros(2, 32, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="13, 14 or 17"):
        K.fused_experts(a, a, a)
    with pytest.raises(RuntimeError, match="12 or 14"):
        K.shared_expert(a, a, a)
    with pytest.raises(RuntimeError, match="mutually exclusive"):
        K.fused_experts(a, a, a, a, a, True, True, True, None, None, None, None, None, False)


def test_dropin_package_registers_ops(lib):
    code = (
        "import torch, sgl_kernel\n"
        "from sglang.srt.layers.amx_utils import CPUQuantMethod\n"
        "from sgl_kernel.common_ops import grouped_topk_cpu, biased_grouped_topk_cpu, shared_expert_cpu, convert_weight_packed\n"
        "ops = torch.ops.sgl_kernel\n"
        "for n in ['convert_weight_packed','convert_scale_packed','grouped_topk_cpu','biased_grouped_topk_cpu',"
        "'fused_experts_cpu','shared_expert_cpu','mxfp4_scaled_mm_cpu']:\n"
        "    getattr(ops, n)\n"
        "assert set(ops.fused_experts_cpu.overloads()) == {'default','quant_method','bias'}\n"
        "assert set(ops.shared_expert_cpu.overloads()) == {'default','v12'}\n"
        "assert int(CPUQuantMethod.UNQUANT) == 0\n"
        "print('ok')\n")
    env = dict(os.environ, PYTHONPATH=os.path.join(PKG, "dropin"))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env)
    assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-2000:]


def test_weight_registry_accounting():
    """Host-side registry logic (no GPU): strides are part of the key, overwriting an entry does not double-count
    its bytes, eviction is least-recently-used under both caps, invalidate() drops the entries of one tensor."""
    import torch
    from sgl_cpu_tests_b200 import ops
    ops.clear_weight_cache()
    w = torch.zeros(8, 8)
    assert ops._registry_key(w, False, "w") != ops._registry_key(w.t(), False, "w")   # same pointer and shape
    dev_copy = torch.zeros(1024, dtype=torch.uint8)
    k1 = ops._registry_key(w, False, "w")
    ops._registry_put(k1, ops._fingerprint(w), dev_copy, w)
    ops._registry_put(k1, ops._fingerprint(w), dev_copy, w)
    assert ops._registry_bytes == 1024 and ops._registry_host_bytes == w.numel() * 4
    old_cap = ops._WEIGHT_CACHE_BYTES
    try:
        ops._WEIGHT_CACHE_BYTES = 3 * 1024
        others = [torch.zeros(4, 4) + i for i in range(3)]
        keys = [ops._registry_key(o, False, "w") for o in others]
        ops._registry_put(keys[0], ops._fingerprint(others[0]), dev_copy.clone(), others[0])
        ops._registry_put(keys[1], ops._fingerprint(others[1]), dev_copy.clone(), others[1])
        assert ops._registry_get(k1, w) is not None            # touch: k1 becomes the most recently used
        ops._registry_put(keys[2], ops._fingerprint(others[2]), dev_copy.clone(), others[2])
        assert k1 in ops._registry and keys[0] not in ops._registry and ops._registry_bytes == 3 * 1024
    finally:
        ops._WEIGHT_CACHE_BYTES = old_cap
    ops.invalidate(w)
    assert k1 not in ops._registry and ops._registry_bytes == 2 * 1024
    w.add_(1)                                                  # in-place through torch: new version, new key
    assert ops._registry_key(w, False, "w") != k1
    ops.clear_weight_cache()
    assert ops._registry_bytes == 0 and ops._registry_host_bytes == 0
