"""Debug: is fused_experts bit-exactly equivariant under token permutation, per kernel path?"""
import sys
import torch
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import sgl_cpu_tests_b200 as K
import test_gpu_fullsize as T

for fmt, E, topk, H, I, M in [("bf16", 128, 8, 2048, 768, 4096), ("bf16", 128, 8, 2048, 768, 512), ("int8", 128, 8, 2048, 768, 4096)]:
    fmt_id = {"bf16": K.FMT_BF16, "int8": K.FMT_INT8_W8A8}[fmt]
    layer = T.make_layer(fmt, E, H, I)
    g = torch.Generator(device="cuda").manual_seed(9)
    x = torch.randn(M, H, device="cuda", generator=g, dtype=torch.bfloat16) / 4
    tw, ids = T.routing(M, E, topk, 9)
    perm = torch.randperm(M, device="cuda", generator=g)
    for name, path in (("stream", K.PATH_STREAM), ("umma", K.PATH_UMMA), ("umma2", K.PATH_UMMA2)):
        if path == K.PATH_STREAM and M > 512:
            continue
        a = T.call(K, layer, fmt_id, x, tw, ids, path=path)
        a2 = T.call(K, layer, fmt_id, x, tw, ids, path=path)
        b = T.call(K, layer, fmt_id, x[perm].contiguous(), tw[perm].contiguous(), ids[perm].contiguous(), path=path)
        d = (b.float() - a[perm].float()).abs()
        print(fmt, M, name, "rerun equal", torch.equal(a, a2), "perm equal", torch.equal(b, a[perm]), "max diff", float(d.max()),
              "mismatch frac", float((d > 0).float().mean()), "rows", int((d.amax(1) > 0).sum()), flush=True)
