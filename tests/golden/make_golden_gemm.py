"""Golden fixtures for the dense quantised linears, generated in the build container
from the UNMODIFIED reference scripts test_gemm.py, test_gemm_int8.py and test_gemm_fp8.py.

The scripts' own function definitions (their torch oracles and `run_single_test`) are executed as they
are; only their module-level driver calls are skipped (test_gemm_fp8.py's driver clones a 20 MB weight
100 times for its timing loop), and `run_single_test` is then called on the shapes the scripts list
(test_gemm.py:31-34, test_gemm_int8.py:78-80, the sweep at test_gemm_fp8.py:108-109).  The `sgl_kernel`
ops are recording stubs (make_golden.py), `utils.compare` records (reference result, stub result).

Run:  python tests/golden/make_golden_gemm.py      (needs /root/reference; writes tests/golden/gemm_*.pt)
"""
import ast
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden as G  # noqa: E402


def load_defs(name):
    """Executes imports, assignments and function definitions of a reference script, not its driver calls."""
    path = os.path.join(G.REF, name)
    tree = ast.parse(open(path).read(), path)
    keep = (ast.Import, ast.ImportFrom, ast.FunctionDef, ast.Assign)
    body = [n for n in tree.body if isinstance(n, keep)
            or (isinstance(n, ast.Expr) and isinstance(n.value, ast.Call)
                and getattr(n.value.func, "attr", "") == "manual_seed")]
    tree.body = body
    glb = {"__name__": "__ref__", "__file__": path}
    exec(compile(tree, path, "exec"), glb)
    return glb


def install():
    del G.CALLS[:], G.COMPARES[:]
    G._install_stubs()
    ops = {
        "weight_packed_linear": G._record(
            "weight_packed_linear", lambda a, b, *r: torch.zeros(a.shape[0], b.shape[0], dtype=torch.bfloat16)),
        "per_token_quant_int8_cpu": G._record(
            "per_token_quant_int8_cpu",
            lambda a: (torch.zeros(a.shape, dtype=torch.int8), torch.ones(a.shape[0], 1))),
        "int8_scaled_mm_cpu": G._record(
            "int8_scaled_mm_cpu", lambda a, b, *r: torch.zeros(a.shape[0], b.shape[0], dtype=torch.bfloat16)),
        "int8_scaled_mm_with_quant": G._record(
            "int8_scaled_mm_with_quant", lambda a, b, *r: torch.zeros(a.shape[0], b.shape[0], dtype=torch.bfloat16)),
    }
    common = sys.modules["sgl_kernel.common_ops"]
    for k, v in ops.items():
        setattr(common, k, v)
    base = torch.ops.__dict__["sgl_kernel"]

    class _NS:
        def __getattr__(self, name):
            return ops[name] if name in ops else getattr(base, name)

    torch.ops.__dict__["sgl_kernel"] = _NS()


def collect(fn, shapes, ref_second=False):
    """ref_second: the script calls compare(ours, reference) (test_activation.py / test_norm.py) instead of
    compare(reference, ours)."""
    cases = []
    for args in shapes:
        del G.CALLS[:], G.COMPARES[:]
        fn(*args)
        seen, calls = set(), []
        for n, a in G.CALLS:  # the packed-weight repeat of a call carries the same tensors (identity stub)
            if n not in seen and n != "convert_weight_packed":
                seen.add(n)
                calls.append((n, a))
        cases.append({"args": list(args), "calls": calls,
                      "refs": [G._clone(b if ref_second else a) for a, b in G.COMPARES]})
    return cases


def main():
    install()
    g = load_defs("test_gemm.py")
    G.save("gemm_bf16", collect(g["run_single_test"], [(1, 416, 512, True), (11, 416, 512, True),
                                                       (200, 416, 512, True), (37, 256, 1024, False)]))
    install()
    g = load_defs("test_gemm_int8.py")
    G.save("gemm_int8", collect(g["run_single_test"], [
        (128, 384, 544, torch.bfloat16, True), (2, 32, 32, torch.bfloat16, True),
        (128, 384, 544, torch.bfloat16, False), (2, 32, 32, torch.bfloat16, False),
        (70, 256, 1024, torch.bfloat16, True)]))
    install()
    g = load_defs("test_gemm_fp8.py")
    shapes = [(m, n, k, b, False) for m in (1, 11, 111) for n in (128, 224 + 32) for k in (512, 576 + 64)
              for b in (False, True)][:12]
    shapes += [(260, 512, 1024, True, False)]
    G.save("gemm_fp8", collect(g["run_single_test"], shapes))

    # silu_and_mul (test_activation.py) and the rmsnorm pair (test_norm.py)
    install()
    none = lambda *a: None  # noqa: E731
    extra = {"silu_and_mul_cpu": G._record("silu_and_mul_cpu", none), "rmsnorm_cpu": G._record("rmsnorm_cpu", none),
             "fused_add_rmsnorm_cpu": G._record("fused_add_rmsnorm_cpu", none)}
    for k, v in extra.items():
        setattr(sys.modules["sgl_kernel.common_ops"], k, v)
    import types
    opsmod, kern = types.ModuleType("sgl_kernel.ops"), types.ModuleType("sgl_kernel.ops._kernels")
    kern.silu_and_mul_cpu = extra["silu_and_mul_cpu"]
    sys.modules["sgl_kernel.ops"], sys.modules["sgl_kernel.ops._kernels"] = opsmod, kern
    g = load_defs("test_activation.py")
    cases = collect(g["run_single_test"], [([17, 1024], torch.bfloat16, "cpu"), ([5, 2, 344], torch.float16, "cpu"),
                                           ([3, 22016], torch.bfloat16, "cpu"), ([9, 30], torch.float16, "cpu")], True)
    g = load_defs("test_norm.py")
    cases += collect(g["run_single_test"], [([33, 1024], torch.bfloat16, "cpu"), ([9, 4096 + 13], torch.float16, "cpu"),
                                            ([2, 5120], torch.float16, "cpu"), ([4, 8200], torch.bfloat16, "cpu")], True)
    G.save("act_norm", cases)


if __name__ == "__main__":
    main()
