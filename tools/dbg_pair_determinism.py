"""Is the CTA-pair kernel bit-reproducible run to run?  12 repeats of the same call per format / size."""
import sys
import torch
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import sgl_cpu_tests_b200 as K
import test_gpu_fullsize as T

for fmt, E, topk, H, I, M in [("bf16", 128, 8, 2048, 768, 4096), ("int8", 128, 8, 2048, 768, 4096), ("bf16", 128, 8, 2048, 768, 8192)]:
    fmt_id = {"bf16": K.FMT_BF16, "int8": K.FMT_INT8_W8A8}[fmt]
    layer = T.make_layer(fmt, E, H, I)
    g = torch.Generator(device="cuda").manual_seed(9)
    x = torch.randn(M, H, device="cuda", generator=g, dtype=torch.bfloat16) / 4
    tw, ids = T.routing(M, E, topk, 9)
    ref = T.call(K, layer, fmt_id, x, tw, ids, path=K.PATH_UMMA)
    outs = [T.call(K, layer, fmt_id, x, tw, ids, path=K.PATH_UMMA2).clone() for _ in range(12)]
    torch.cuda.synchronize()
    bad = sum(0 if torch.equal(o, outs[0]) else 1 for o in outs[1:])
    d = max(float((o.float() - ref.float()).abs().max()) for o in outs)
    print(fmt, M, "pair runs differing from the first:", bad, "of 11; max |pair - single| =", d,
          "pair == single:", torch.equal(outs[0], ref), flush=True)
