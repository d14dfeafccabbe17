"""One MoE layer on one GPU: router -> fused_experts -> shared_expert, the call sequence of the reference's
DeepSeek-style layer (test_biased_grouped_topk.py:71-80 -> test_moe_fp8_ext.py:118 -> :60-61), as a single object
over device-resident weights, eager or captured in one CUDA graph.  Every step ends in libsglb200.so kernels; this
module only sequences them (the expert-parallel form of the same layer is ep.EPMoE)."""
import torch

from . import _lib as L
from . import ops


class Router:
    """grouped_topk_cpu (bias=None) or biased_grouped_topk_cpu configuration of a layer."""

    def __init__(self, topk, renormalize=True, num_expert_group=1, topk_group=1, correction_bias=None):
        self.topk, self.renormalize, self.G, self.topk_group, self.bias = topk, renormalize, num_expert_group, topk_group, correction_bias

    def __call__(self, gating, out_w=None, out_ids=None):
        return ops._router(gating, self.bias, self.topk, self.renormalize, self.G, self.topk_group, out_w=out_w, out_ids=out_ids)


class MoELayer:
    """weights: dict(w1 [E,2I,H], w2 [E,H,I], s1, s2, block) on the device, row-major; shared: the same keys with 2-D
    weights or None; forward(x [M,H] bf16, gating [M,E]) -> [M,H] bf16 =
    shared_expert(x) + routed_scaling_factor * fused_experts(x, router(gating))   (shared given)
    fused_experts(x, router(gating))                                             (no shared expert)."""

    def __init__(self, weights, fmt, router, shared=None, routed_scaling_factor=1.0, path=L.PATH_AUTO, ride=True):
        self.w, self.fmt, self.router, self.shared, self.rsf, self.path = weights, fmt, router, shared, float(routed_scaling_factor), path
        self.ride = ride    # let the shared expert ride along in the routed experts' launches where the kernels can

    def rides(self, num_tokens):
        """The shared expert's rows go through the routed experts' two GEMM launches (same result bit for bit, five
        launches and a second pass of under-filled kernels fewer) when it has the experts' own shape and format."""
        if self.shared is None or not self.ride:
            return False
        E, H, I = ops._weight_geometry(self.w["w1"], self.w["w2"], self.fmt, False)
        s = self.shared
        return ops.shared_rides(num_tokens, self.router.topk, E, H, I, self.fmt, self.w.get("block"), self.path,
                                shared_I=s["w2"].shape[-1], shared_fmt=s.get("fmt", self.fmt)) \
            and s.get("block") == self.w.get("block")

    def route(self, gating, out_w=None, out_ids=None):
        return self.router(gating, out_w, out_ids)

    def experts(self, x, tw, ids, out=None, with_shared=False):
        w = self.w
        return ops.fused_experts_impl(x, w["w1"], w["w2"], tw, ids, False, self.fmt, w.get("s1"), w.get("s2"), w.get("block"),
                                      False, path=self.path, out=out, shared=self.shared if with_shared else None,
                                      routed_scaling_factor=self.rsf)

    def add_shared(self, x, routed, out=None):
        if self.shared is None:
            return routed
        s = self.shared
        return ops.shared_expert_impl(x, s["w1"], s["w2"], routed, self.rsf, False, s.get("fmt", self.fmt), s.get("s1"),
                                      s.get("s2"), s.get("block"), False, out=out)

    def forward(self, x, gating):
        tw, ids = self.route(gating)
        if self.rides(x.shape[0]):
            return self.experts(x, tw, ids, with_shared=True)
        return self.add_shared(x, self.experts(x, tw, ids))

    def graphed(self, num_tokens):
        return GraphedLayer(self, num_tokens)


class GraphedLayer:
    """MoELayer.forward for a fixed token count captured in one CUDA graph over static buffers: fill `hidden` and
    `gating` (or pass tensors to __call__), replay, read `out`.  `topk_weights` / `topk_ids` hold the routing of the
    last replay.  At decode sizes the ~14 kernels of a layer call run for a few microseconds each, less than the host
    needs to launch them; a replay is one launch."""

    def __init__(self, layer, num_tokens):
        w = layer.w
        dev = w["w1"].device
        E, H, I = ops._weight_geometry(w["w1"], w["w2"], layer.fmt, False)
        topk = layer.router.topk
        gdt = torch.bfloat16
        self.hidden = torch.zeros(num_tokens, H, dtype=torch.bfloat16, device=dev)
        self.gating = torch.zeros(num_tokens, E, dtype=gdt, device=dev)
        self.topk_weights = torch.zeros(num_tokens, topk, dtype=torch.float32, device=dev)
        self.topk_ids = torch.zeros(num_tokens, topk, dtype=torch.int32, device=dev)
        self.routed = torch.empty_like(self.hidden)
        self.out = torch.empty_like(self.hidden) if layer.shared is not None else self.routed
        lib = L.load()
        nws = lib.sglb_fused_experts_workspace_bytes(num_tokens, topk, E, H, I, layer.fmt)
        if layer.shared is not None:
            Is = layer.shared["w2"].shape[-1]
            nws = max(nws, lib.sglb_fused_experts_workspace_bytes(num_tokens, 1, 1, H, Is, layer.shared.get("fmt", layer.fmt)))
        self._ws = torch.empty(max(nws, 1 << 20), dtype=torch.uint8, device=dev)
        side = torch.cuda.Stream(device=dev)
        key = (dev.index, side.cuda_stream)

        self.rides = layer.rides(num_tokens)

        def call():
            layer.route(self.gating, self.topk_weights, self.topk_ids)
            if self.rides:
                layer.experts(self.hidden, self.topk_weights, self.topk_ids, out=self.out, with_shared=True)
                return
            layer.experts(self.hidden, self.topk_weights, self.topk_ids, out=self.routed)
            if layer.shared is not None:
                layer.add_shared(self.hidden, self.routed, out=self.out)

        with torch.cuda.device(dev):
            side.wait_stream(torch.cuda.current_stream())
            with ops._launch_lock:
                ops._workspaces[key] = self._ws
                try:
                    with torch.cuda.stream(side):  # lazy one-off work (function attributes, tensor maps) outside the capture
                        n0 = ops.launch_count()
                        call()
                        self.kernels_per_replay = ops.launch_count() - n0
                    side.synchronize()
                    self.graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(self.graph, stream=side):
                        call()
                finally:
                    ops._workspaces.pop(key, None)
            torch.cuda.current_stream().wait_stream(side)

    def __call__(self, hidden_states=None, gating=None):
        if hidden_states is not None:
            self.hidden.copy_(hidden_states, non_blocking=True)
        if gating is not None:
            self.gating.copy_(gating, non_blocking=True)
        self.graph.replay()
        return self.out
