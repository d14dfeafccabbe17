#!/usr/bin/env python
"""Device-resident timing of the dense quantised linears (sglb_linear = stage 2 of the expert kernels
with E = 1) at the reference's bench_gemm.py shapes (bench_gemm.py:147-151): weight_packed_linear (bf16),
fp8_scaled_mm_cpu (block 128 x 128) and int8_scaled_mm_with_quant.  CUDA events, 20 calls after 5 warm-ups, weights
rotated over copies larger than L2."""
import sys

import torch

sys.path.insert(0, ".")
import sgl_cpu_tests_b200 as K  # noqa: E402

SHAPES = [(1000, 18432, 2560), (384, 8192, 7168), (4, 2816, 7168), (4096, 7168, 2816), (1024, 14336, 4096)]


def timed(fn, n=20, w=5):
    for i in range(w):
        fn(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(n):
        fn(i)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def main():
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(1111)
    for M, N, Kd in SHAPES:
        copies = max(2, int(300e6 // (N * Kd)) + 1)
        x = (torch.randn(M, Kd, device=dev, generator=g) / 8).bfloat16()
        wb = [(torch.randn(N, Kd, device=dev, generator=g) / 8).bfloat16() for _ in range(max(2, copies // 2))]
        w8 = [torch.randint(-127, 128, (N, Kd), device=dev, generator=g, dtype=torch.int8) for _ in range(copies)]
        s8 = torch.rand(N, device=dev, generator=g) * 1e-2
        wf = [(torch.randn(N, Kd, device=dev, generator=g) * 50).to(torch.float8_e4m3fn) for _ in range(copies)]
        sf = torch.rand(-(-N // 128), -(-Kd // 128), device=dev, generator=g) * 4e-3
        t_bf16 = timed(lambda i: K.weight_packed_linear(x, wb[i % len(wb)], None, False))
        t_fp8 = timed(lambda i: K.fp8_scaled_mm(x, wf[i % copies], sf, [128, 128], None, torch.bfloat16, False))
        t_int8 = timed(lambda i: K.int8_scaled_mm_with_quant(x, w8[i % copies], s8, None, torch.bfloat16, False))
        fl = 2.0 * M * N * Kd
        line = f"### linear M = {M}, N = {N}, K = {Kd}:"
        for name, t, bw in (("bf16", t_bf16, 2), ("fp8", t_fp8, 1), ("int8", t_int8, 1)):
            line += f" {name}: {t * 1e3:.1f} us = {fl / t / 1e9:.0f} TFLOP/s, {N * Kd * bw / t / 1e6:.0f} GB/s of weights;"
        print(line)
        del wb, w8, wf
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
