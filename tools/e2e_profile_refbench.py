"""Host-side breakdown of the reference's bench_moe.py loop at M = 4 (bench_moe.py:50-82: L rotating clones of the
packed weights and of the input, pageable host tensors) through the drop-in op: cProfile over the rotation."""
import cProfile
import pstats
import sys
import time

import torch

sys.path.insert(0, ".")
import sgl_cpu_tests_b200 as K  # noqa: E402
sys.path.insert(0, K.DROPIN_PATH)
import sgl_kernel  # noqa: E402,F401

L = int(sys.argv[1]) if len(sys.argv) > 1 else 20
M, N, Kd, E, topk = 4, 384, 7168, 256, 8
torch.manual_seed(0)
convert = torch.ops.sgl_kernel.convert_weight_packed
op = torch.ops.sgl_kernel.fused_experts_cpu
w1 = torch.randn(E, 2 * N, Kd, dtype=torch.bfloat16)
w2 = torch.randn(E, Kd, N, dtype=torch.bfloat16)
pw1, pw2 = convert(w1), convert(w2)
x = torch.randn(M, Kd, dtype=torch.bfloat16) / 10
score = torch.softmax(torch.randn(M, E), dim=-1)
tw, ids = torch.topk(score, topk)
ids = ids.to(torch.int32)
inputs = [x.clone() for _ in range(L)]
w1s = [pw1.clone() for _ in range(L)]
w2s = [pw2.clone() for _ in range(L)]


def sweep():
    for i in range(L):
        op(inputs[i], w1s[i], w2s[i], tw, ids, True, False, False, None, None, None, None, None, True)


for _ in range(3):
    sweep()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    sweep()
torch.cuda.synchronize()
print("ms per call", (time.perf_counter() - t0) / (5 * L) * 1e3)
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    sweep()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
