"""Host-side breakdown of the drop-in call with HOST buffers (bench.py's e2e leg): cProfile over 200 calls."""
import cProfile
import pstats
import sys
import time

import torch

sys.path.insert(0, ".")
import sgl_cpu_tests_b200 as K  # noqa: E402
sys.path.insert(0, K.DROPIN_PATH)
import sgl_kernel  # noqa: E402,F401

E, topk, H, I, M = 128, 8, 2048, 768, 128
torch.manual_seed(0)
w1 = torch.randint(-127, 128, (E, 2 * I, H), dtype=torch.int8)
w2 = torch.randint(-127, 128, (E, H, I), dtype=torch.int8)
s1, s2 = torch.rand(E, 2 * I) * 1e-2, torch.rand(E, H) * 1e-2
x = (torch.randn(M, H) / 4).bfloat16().pin_memory()
score = torch.softmax(torch.randn(M, E), dim=-1)
tw, ids = torch.topk(score, topk)
tw, ids = tw.pin_memory(), ids.to(torch.int32).pin_memory()
op = torch.ops.sgl_kernel.fused_experts_cpu


def call():
    return op(x, w1, w2, tw, ids, True, True, False, s1, s2, None, None, None, False)


for _ in range(5):
    call()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(200):
    call()
torch.cuda.synchronize()
print("ms per call", (time.perf_counter() - t0) / 200 * 1e3)
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    call()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
