"""cProfile of the host-buffer call for the mxfp4 layer (gpt-oss shape): where the host time goes."""
import cProfile
import pstats
import sys
import time

import torch

sys.path.insert(0, ".")
import sgl_cpu_tests_b200 as K  # noqa: E402
sys.path.insert(0, K.DROPIN_PATH)
import sgl_kernel  # noqa: E402,F401

E, topk, H, I, M = 128, 4, 2880, 2880, 128
torch.manual_seed(0)
w1 = torch.randint(0, 256, (E, 2 * I, H // 2), dtype=torch.uint8)
w2 = torch.randint(0, 256, (E, H, I // 2), dtype=torch.uint8)
s1 = torch.randint(125, 127, (E, 2 * I, H // 32), dtype=torch.uint8)
s2 = torch.randint(125, 127, (E, H, I // 32), dtype=torch.uint8)
x = (torch.randn(M, H) / 4).bfloat16().pin_memory()
score = torch.softmax(torch.randn(M, E), dim=-1)
tw, ids = torch.topk(score, topk)
tw, ids = tw.pin_memory(), ids.to(torch.int32).pin_memory()
op = torch.ops.sgl_kernel.fused_experts_cpu


def call():
    return op(x, w1, w2, tw, ids, True, 4, s1, s2, None, None, None, False)


for _ in range(3):
    call()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20):
    call()
torch.cuda.synchronize()
print("ms per call", (time.perf_counter() - t0) / 20 * 1e3, "graphs", [type(g).__name__ for g in K.ops._host_graphs.values()])
pr = cProfile.Profile()
pr.enable()
for _ in range(20):
    call()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
