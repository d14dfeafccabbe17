"""CTA-pair kernels (umma2_gemm.cu / umma2_f8.cu, routed rows through TMA gather4): run-to-run determinism, equality
with the single-CTA kernels and per-call time, per format and size.

  python tools/pair_check.py [--quick]
"""
import json
import math
import sys

import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import sgl_cpu_tests_b200 as K  # noqa: E402
import test_gpu_fullsize as T  # noqa: E402

CASES = [("bf16", 128, 8, 2048, 768, 4096), ("bf16", 128, 8, 2048, 768, 8192), ("int8", 128, 8, 2048, 768, 4096),
         ("fp8", 128, 8, 2048, 768, 8192), ("fp8", 256, 8, 7168, 2048, 8192), ("fp8", 256, 8, 7168, 384, 3929),
         ("bf16", 8, 2, 1024, 512, 777)]
if "--quick" in sys.argv:
    CASES = CASES[:1] + CASES[3:4] + CASES[-1:]


def timed(fn, n=5):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


res = []
for fmt, E, topk, H, I, M in CASES:
    fmt_id = {"bf16": K.FMT_BF16, "int8": K.FMT_INT8_W8A8, "fp8": K.FMT_FP8_W8A16}[fmt]
    layer = T.make_layer(fmt, E, H, I)
    g = torch.Generator(device="cuda").manual_seed(9)
    x = torch.randn(M, H, device="cuda", generator=g, dtype=torch.bfloat16) / (4 if fmt != "fp8" else math.sqrt(H))
    tw, ids = T.routing(M, E, topk, 9)
    ids[::5, 0] = -1
    ref = T.call(K, layer, fmt_id, x, tw, ids, path=K.PATH_UMMA)
    outs = [T.call(K, layer, fmt_id, x, tw, ids, path=K.PATH_UMMA2).clone() for _ in range(12)]
    torch.cuda.synchronize()
    bad = sum(0 if torch.equal(o, outs[0]) else 1 for o in outs[1:])
    d = max(float((o.float() - ref.float()).abs().max()) for o in outs)
    t1 = timed(lambda: T.call(K, layer, fmt_id, x, tw, ids, path=K.PATH_UMMA))
    t2 = timed(lambda: T.call(K, layer, fmt_id, x, tw, ids, path=K.PATH_UMMA2))
    fl = 6.0 * M * topk * H * I * 0.8 + 6.0 * M * (topk - 1) * H * I * 0.2   # one slot in five of column 0 is masked
    r = dict(fmt=fmt, E=E, topk=topk, H=H, I=I, M=M, pair_runs_differing=bad, pair_eq_single=bool(torch.equal(outs[0], ref)),
             max_abs_diff=d, single_ms=t1, pair_ms=t2, single_tflops=fl / t1 / 1e9, pair_tflops=fl / t2 / 1e9)
    print(json.dumps(r), flush=True)
    res.append(r)
    del layer
    torch.cuda.empty_cache()
ok = all(r["pair_runs_differing"] == 0 and r["pair_eq_single"] for r in res)
print("RESULT:", "PASS" if ok else "FAIL")
sys.exit(0 if ok else 1)
