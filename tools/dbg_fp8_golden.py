import sys, os, torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from conftest import load_golden
import sgl_cpu_tests_b200 as K
from oracle import moe_oracle as O
cases = load_golden("moe_fp8")["cases"] + load_golden("moe_offload")
for i, c in enumerate(cases):
    ref = c["ref_out"].float()
    outs = {}
    for name, path in (("stream", K.PATH_STREAM), ("umma", K.PATH_UMMA)):
        try:
            outs[name] = K.fused_experts_impl(c["a"].cuda(), c["w1"].cuda(), c["w2"].cuda(), c["topk_weight"].cuda(), c["topk_ids"].cuda(),
                                              False, K.FMT_FP8_W8A16, c["w1s"].cuda(), c["w2s"].cuda(), c["block"], False, path=path).float().cpu()
        except RuntimeError as e:
            outs[name] = None; print(i, name, "ERR", str(e)[:100])
    oref = O.fused_experts(c["a"], c["w1"], c["w2"], c["topk_weight"], c["topk_ids"], "fp8", c["w1s"], c["w2s"], c["block"]).float()
    print(i, "shape a", tuple(c["a"].shape), "w1", tuple(c["w1"].shape), "block", c["block"], "topk", c["topk_ids"].shape[1], "|ref|max", float(ref.abs().max()),
          "oracle-vs-golden", float((oref - ref).abs().max()))
    for name, o in outs.items():
        if o is None: continue
        d = (o - ref).abs(); bad = d > 1e-2 + 1e-2 * ref.abs()
        print("   ", name, "max diff", float(d.max()), "bad", int(bad.sum()), "/", bad.numel(), "rows", bad.any(1).nonzero().flatten().tolist()[:8], "cols", bad.any(0).nonzero().flatten().tolist()[:8])
