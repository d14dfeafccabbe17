"""Achieved HBM bandwidth of silu_and_mul / rmsnorm / fused_add_rmsnorm at the reference's bench shapes
(bench_silu_and_mul.py:13-20, test_norm.py:83) and at a prefill-sized batch; CUDA events, inputs rotated
over buffers larger than L2 for the big shapes."""
import json
import sys

import torch

sys.path.insert(0, ".")
import sgl_cpu_tests_b200 as K  # noqa: E402

PEAK = 6450.0
try:
    PEAK = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]
except Exception:
    pass


def timeit(fn, n):
    for i in range(5):
        fn(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(n):
        fn(i)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e-3


for dtype in (torch.bfloat16,):
    for rows, width in ((1, 18432), (17, 18432), (1000, 18432), (16384, 18432)):
        copies = max(1, min(8, int(300e6 // (rows * width * 2)) or 1))
        xs = [torch.randn(rows, width, dtype=dtype, device="cuda") for _ in range(copies)]
        t = timeit(lambda i: K.ops.silu_and_mul(xs[i % copies]), 50)
        byt = rows * width * 2 * 1.5
        print(f"silu_and_mul {rows}x{width}: {t*1e6:.1f} us  {byt/t/1e9:.0f} GB/s  {byt/t/1e9/PEAK*100:.0f} % of {PEAK:.0f}")
    for rows, hidden in ((1, 5120), (1024, 4096), (16384, 7168)):
        copies = max(1, min(8, int(300e6 // (rows * hidden * 4)) or 1))
        xs = [torch.randn(rows, hidden, dtype=dtype, device="cuda") for _ in range(copies)]
        rs = [torch.randn(rows, hidden, dtype=dtype, device="cuda") for _ in range(copies)]
        w = torch.randn(hidden, dtype=dtype, device="cuda")
        out = torch.empty_like(xs[0])
        t = timeit(lambda i: K.ops.rmsnorm(out, xs[i % copies], w, 1e-6), 50)
        byt = rows * hidden * 2 * 2
        print(f"rmsnorm {rows}x{hidden}: {t*1e6:.1f} us  {byt/t/1e9:.0f} GB/s  {byt/t/1e9/PEAK*100:.0f} %")
        t = timeit(lambda i: K.ops.fused_add_rmsnorm(xs[i % copies], rs[i % copies], w, 1e-6), 50)
        byt = rows * hidden * 2 * 4
        print(f"fused_add_rmsnorm {rows}x{hidden}: {t*1e6:.1f} us  {byt/t/1e9:.0f} GB/s  {byt/t/1e9/PEAK*100:.0f} %")
