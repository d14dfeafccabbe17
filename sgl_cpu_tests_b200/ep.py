"""Expert-parallel fused MoE: experts sharded over the ranks of one NVLink/NVSwitch box, tokens
data-parallel, one exchange each side of the local fused_experts.

Nothing in the reference shards experts.  The hook it offers is that ``fused_experts_cpu`` ignores
``topk_ids == -1`` (test_moe_offloading_cpu.py:57-68) -- exactly what a rank sees for experts it does
not own -- so the local compute is the unchanged single-GPU path.

Per layer call, on every rank (M local tokens, rank r owns experts [r*E/n, (r+1)*E/n)):

  plan      (CUDA, csrc/ep.cu)  which ranks own at least one of each token's experts; stable numbering of the
                                rows that travel out and of the slot rows that come back
  pack      (CUDA)              token rows + (local ids | weights) gathered per destination rank
  dispatch  (NCCL all-to-all)   rows travel once per (token, destination rank), not once per expert
  experts   (CUDA)              ordinary fused_experts over the received rows, ids already local / -1, with
                                slot-addressed output: one bf16 row per routed slot, written by the stage-2
                                epilogue straight into the return buffer
  return    (NCCL all-to-all)   one row per routed slot
  combine   (CUDA)              out[m] = sum of the token's slot rows in slot order, f32, one rounding:
                                the single-GPU reduction, so n-GPU results are BIT-IDENTICAL to one GPU's
                                (optionally + the shared expert's epilogue, whose MLP ran on a side stream
                                while the exchange was in flight)

Three exchange modes, chosen from the group-agreed capacity (never from a rank-local value):
  * p2p   (decode): no collective at all -- stores into / loads from the peers' memory over NVLink
    (csrc/ep_p2p.cu), call counter on the device, CUDA-graph capturable (EPMoE.graphed);
  * fixed (decode without peer access): every segment is padded to `capacity` rows and padding carries
    ids = -1, so the split sizes are static -- no count exchange, no host sync;
  * exact (prefill): row and slot counts are exchanged first and read on the host once.

The device work is behind a small backend object so the host orchestration (split sizes, buffer
offsets, ordering of the collectives) can be exercised on CPU under ``gloo`` by the tests, which plug
in the oracle there; the product backend is `CudaBackend` and nothing else ships.
"""
import os

import torch
import torch.distributed as dist

from . import _lib as L
from . import ops


class CudaBackend:
    """The product backend: every method ends in libsglb200.so kernels."""

    def __init__(self, weights, fmt, e_local, topk, is_vnni=False, path=L.PATH_AUTO, shared=None):
        self.w, self.fmt, self.e_local, self.topk, self.is_vnni, self.path = weights, fmt, e_local, topk, is_vnni, path
        self.shared = shared   # dict(w1, w2, s1, s2, block, fmt) of the replicated shared expert, or None

    @staticmethod
    def _s():
        return torch.cuda.current_stream().cuda_stream

    def plan(self, ids, E, world):
        M, topk = ids.shape
        i32 = dict(dtype=torch.int32, device=ids.device)
        dest_row, send_count = torch.empty(M, world, **i32), torch.empty(world, **i32)
        slot_index, send_slots = torch.empty(M, topk, **i32), torch.empty(world, **i32)
        L.check(L.load().sglb_ep_plan(ids.data_ptr(), M, topk, E, world, dest_row.data_ptr(), send_count.data_ptr(),
                                      slot_index.data_ptr(), send_slots.data_ptr(), self._s()))
        return dest_row, send_count, slot_index, send_slots

    def pack(self, x, ids, tw, dest_row, send_count, E, world, cap, rows):
        M, H = x.shape
        topk = ids.shape[1]
        send_x = torch.empty(rows, H, dtype=x.dtype, device=x.device)
        send_meta = torch.empty(rows, 2 * topk, dtype=torch.int32, device=x.device)
        L.check(L.load().sglb_ep_pack(x.data_ptr(), ids.data_ptr(), tw.data_ptr(), dest_row.data_ptr(),
                                      send_count.data_ptr(), M, topk, E, world, H, cap, rows, send_x.data_ptr(),
                                      send_meta.data_ptr(), self._s()))
        return send_x, send_meta

    def ret_index(self, recv_meta, world, cap, recv_count=None, recv_slots=None):
        rows, topk = recv_meta.shape[0], self.topk
        out = torch.empty(rows, topk, dtype=torch.int32, device=recv_meta.device)
        L.check(L.load().sglb_ep_ret_index(recv_meta.data_ptr(), 2 * topk, rows, topk, world, cap, ops._ptr(recv_count),
                                           ops._ptr(recv_slots), out.data_ptr(), self._s()))
        return out

    def experts_slots(self, x, meta, ret_index, ret_rows, hint=0.0):
        """Local experts over the received rows; the stage-2 epilogue writes slot (i, j) to ret[ret_index[i, j]]."""
        topk = self.topk
        ids = meta[:, :topk].contiguous()
        tw = meta[:, topk:].contiguous().view(torch.float32)
        ret = torch.empty(ret_rows, x.shape[1], dtype=x.dtype, device=x.device)
        if x.shape[0] == 0 or ret_rows == 0:
            return ret
        w = self.w
        return ops.fused_experts_impl(x, w["w1"], w["w2"], tw, ids, False, self.fmt, w.get("s1"), w.get("s2"),
                                      w.get("block"), self.is_vnni, path=self.path, slot_out=ret, slot_map=ret_index,
                                      rows_per_expert_hint=hint)

    def experts_into(self, x, tw, ids, y, hint=0.0):
        """Local experts over received rows with separate ids / weights tensors, slot rows written to `y`."""
        w = self.w
        return ops.fused_experts_impl(x, w["w1"], w["w2"], tw, ids, False, self.fmt, w.get("s1"), w.get("s2"),
                                      w.get("block"), self.is_vnni, path=self.path, slot_out=y,
                                      rows_per_expert_hint=hint)

    def shared_mlp(self, x):
        """silu_and_mul(x W1^T) W2^T of the replicated shared expert on the local tokens (bf16), or None."""
        if self.shared is None or x.shape[0] == 0:
            return None
        s = self.shared
        return ops.shared_expert_impl(x, s["w1"], s["w2"], None, 1.0, False, s.get("fmt", self.fmt), s.get("s1"),
                                      s.get("s2"), s.get("block"), self.is_vnni)

    def combine(self, ret, ids, slot_index, send_slots, E, world, cap, shared_y=None, rsf=1.0):
        M, topk = ids.shape
        H = ret.shape[1]
        out = torch.empty(M, H, dtype=ret.dtype, device=ret.device)
        L.check(L.load().sglb_ep_combine(ret.data_ptr(), ids.data_ptr(), slot_index.data_ptr(), send_slots.data_ptr(),
                                         M, topk, E, world, H, cap, ops._ptr(shared_y), float(rsf), out.data_ptr(),
                                         self._s()))
        return out


class DistComm:
    """torch.distributed plumbing (NCCL over NVLink on the box, gloo in the CPU tests)."""

    def __init__(self, group=None):
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)

    def all_to_all(self, out, inp, out_splits=None, in_splits=None):
        dist.all_to_all_single(out, inp, out_splits, in_splits, group=self.group)

    def exchange_counts(self, send_counts):
        """send_counts [world, c] -> what every peer sends to this rank, same shape."""
        recv = torch.empty_like(send_counts)
        dist.all_to_all_single(recv.view(-1), send_counts.contiguous().view(-1), group=self.group)
        return recv

    def max_int(self, value, device):
        t = torch.tensor([int(value)], dtype=torch.int64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return int(t.item())

    def all_gather_object(self, obj):
        out = [None] * self.world
        dist.all_gather_object(out, obj, group=self.group)
        return out


class _DevMem:
    """Zero-copy torch view of raw device memory through __cuda_array_interface__."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class PeerExchange:
    """Decode-regime exchange over NVLink peer memory (csrc/ep_p2p.cu): one cudaMalloc'd region per rank,
    mapped into every peer through CUDA IPC; rows, ids and weights are stored straight into the owners'
    receive buffers and the per-slot result rows are read back with peer loads.  No NCCL call on the data
    path -- torch.distributed only carries the geometry check and the 64-byte IPC handles once, at construction."""

    def __init__(self, comm, cap, H, topk, device):
        import ctypes as C
        lib = L.load()
        world, rank = comm.world, comm.rank
        self.world, self.rank, self.cap, self.H, self.topk = world, rank, cap, H, topk
        self.comm, self.device = comm, device
        self.base, self._opened, self._closed = None, [], False
        self.pins = 0      # GraphedEP objects whose captured graphs address this region (EPMoE never evicts those)
        # every rank must have built the same geometry: offsets into a peer's region are computed locally
        geo = comm.all_gather_object((world, cap, H, topk))
        if any(g != geo[0] for g in geo):
            raise RuntimeError(f"PeerExchange: ranks disagree on (world, capacity, H, topk): {geo}")
        offs = (C.c_size_t * 8)()
        self.nbytes = lib.sglb_ep_p2p_region_bytes(world, cap, H, topk, offs)
        if self.nbytes == 0:
            raise RuntimeError("PeerExchange: unsupported geometry")
        self.offs = list(offs)
        base = C.c_void_p()
        handle = C.create_string_buffer(64)
        status = "ok"
        try:
            L.check(lib.sglb_p2p_alloc(self.nbytes, C.byref(base), handle))
        except RuntimeError as e:  # keep going: every rank must reach the collective below
            status = str(e)
        gathered = comm.all_gather_object((status, bytes(handle.raw)))
        self.base = base.value
        self.peers = (C.c_void_p * world)()
        if all(g[0] == "ok" for g in gathered):
            for r in range(world):
                if r == rank:
                    self.peers[r] = self.base
                    continue
                p = C.c_void_p()
                try:
                    L.check(lib.sglb_p2p_open(C.create_string_buffer(gathered[r][1], 64), C.byref(p)))
                    self._opened.append(p.value)
                    self.peers[r] = p.value
                except RuntimeError as e:
                    status = str(e)
                    break
        else:
            status = next(g[0] for g in gathered if g[0] != "ok")
        # every rank learns whether every rank mapped every peer: all use the path or none does
        bad = [f for f in comm.all_gather_object(status) if f != "ok"]
        if bad:
            self.close()
            raise RuntimeError(f"PeerExchange: peer mapping failed on some rank: {bad[0]}")
        mem = torch.as_tensor(_DevMem(self.base, self.nbytes), device=device)
        rows = world * cap

        def view(off, dtype, nrows, cols):
            nb = nrows * cols * torch.empty((), dtype=dtype).element_size()
            return mem[off:off + nb].view(dtype).view(nrows, cols)
        self.recv_x = [view(self.offs[p], torch.bfloat16, rows, H) for p in range(2)]
        self.recv_ids = [view(self.offs[2 + p], torch.int32, rows, topk) for p in range(2)]
        self.recv_tw = [view(self.offs[4 + p], torch.float32, rows, topk) for p in range(2)]
        self.y = [view(self.offs[6 + p], torch.bfloat16, rows * topk, H) for p in range(2)]
        self.calls = 0
        self.side = torch.cuda.Stream(device=device)

    def close(self):
        """COLLECTIVE: every rank of the group calls it, at the same point of the program.  A rank's combine waits for
        its peers' experts, not for their combines, so a peer may still be reading this rank's `y` rows over NVLink when
        this rank is done; and CUDA IPC requires every importer to close its mapping before the exporter frees.
        Order: this device idle -> everybody idle -> unmap the peers -> everybody unmapped -> free the own region."""
        if getattr(self, "_closed", True):
            return
        self._closed = True
        lib = L.load()
        torch.cuda.synchronize(self.device)
        self.comm.all_gather_object("idle")          # no kernel of any rank touches any region from here on
        for p in self._opened:
            lib.sglb_p2p_close(p)
        self._opened = []
        self.comm.all_gather_object("unmapped")      # nobody holds a mapping of this rank's region any more
        if self.base:
            lib.sglb_p2p_free(self.base)
            self.base = None

    def __del__(self):
        # Not collective, so the exported region is NOT freed here (a peer may still read it; it goes with the process).
        # Dropping this process's own mappings of the peers is safe once the local device is idle.
        try:
            if not getattr(self, "_closed", True) and self._opened:
                torch.cuda.synchronize(self.device)
                lib = L.load()
                for p in self._opened:
                    lib.sglb_p2p_close(p)
                self._opened = []
        except Exception:  # interpreter shutdown
            pass

    def start_shared(self, x, backend):
        """Forks the replicated shared expert's MLP onto the side stream at THIS point of the current stream.  It needs
        nothing but the local hidden states, so a caller that also routes (GraphedEP) forks it before the router: the
        router, the dispatch kernel and the flag wait keep only a few SMs busy for ~120 us (2 x B200, r2g), which is
        about what the shared MLP of 256 tokens takes; forked after the dispatch (round 2's first version) most of it
        ran behind the persistent expert kernels and showed up as combine time (8 GPUs, r2r: combine 186 us)."""
        if backend.shared is None or x.shape[0] == 0:
            return None
        self.side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.side):
            return backend.shared_mlp(x)

    _NOT_STARTED = object()

    def enqueue(self, x, tw, ids, E, backend, par, rsf=1.0, out=None, hint=0.0, shared_y=_NOT_STARTED):
        """(shared expert on a side stream) -> dispatch -> wait -> local experts into y -> combine, all
        asynchronous on the current stream; `par` selects the buffer set.  shared_y: the result of an earlier
        start_shared() on this stream, else the fork happens here."""
        lib = L.load()
        main = torch.cuda.current_stream()
        s = main.cuda_stream
        M = x.shape[0]
        if shared_y is PeerExchange._NOT_STARTED:
            shared_y = self.start_shared(x, backend)
        L.check(lib.sglb_ep_p2p_dispatch(ops._ptr(x), ops._ptr(ids), ops._ptr(tw), M, self.topk, E, self.world,
                                         self.rank, self.H, self.cap, self.peers, par, s))
        L.check(lib.sglb_ep_p2p_wait_dispatch(self.base, self.world, self.cap, self.H, self.topk, s))
        backend.experts_into(self.recv_x[par], self.recv_tw[par], self.recv_ids[par], self.y[par], hint)
        if out is None:
            out = torch.empty_like(x)
        if shared_y is not None:
            main.wait_stream(self.side)
        L.check(lib.sglb_ep_p2p_combine(ops._ptr(ids), M, self.topk, E, self.world, self.rank, self.H, self.cap,
                                        self.peers, par, ops._ptr(shared_y), float(rsf), ops._ptr(out), s))
        return out

    def run(self, x, tw, ids, E, backend, rsf=1.0, hint=0.0):
        self.calls += 1
        return self.enqueue(x, tw, ids, E, backend, self.calls & 1, rsf, None, hint)


class GraphedEP:
    """The peer-memory layer call captured in CUDA graphs over static buffers (one graph per buffer parity,
    replayed alternately): at decode sizes the ~14 launches of a call cost more host time than the kernels
    run.  Fill `hidden`, `topk_weights`, `topk_ids` -- or, when the layer was given a `router` (layer.Router),
    `hidden` and `gating`: the router kernel then runs inside the graph -- and read `out`.  Every rank must
    replay in lockstep, like any collective."""

    def __init__(self, layer, px, num_tokens, H, device, router=None):
        self.layer, self.px, self.router = layer, px, router
        px.pins += 1       # the graphs below bake in addresses inside px's region: released by close() only
        topk = layer.topk
        self.hidden = torch.zeros(num_tokens, H, dtype=torch.bfloat16, device=device)
        self.gating = torch.zeros(num_tokens, layer.E, dtype=torch.bfloat16, device=device) if router is not None else None
        self.topk_weights = torch.zeros(num_tokens, topk, dtype=torch.float32, device=device)
        self.topk_ids = torch.full((num_tokens, topk), -1, dtype=torch.int32, device=device)
        self.out = torch.empty_like(self.hidden)
        hint = layer._hint(px.cap)
        B = layer.backend
        lib = L.load()
        e_local, _, I = ops._weight_geometry(B.w["w1"], B.w["w2"], B.fmt, B.is_vnni)
        nws = lib.sglb_fused_experts_workspace_bytes(layer.world * px.cap, topk, e_local, H, I, B.fmt)
        self._ws = torch.empty(max(nws, 1 << 20), dtype=torch.uint8, device=device)
        nside = 1 << 20
        if B.shared is not None:
            nside = max(nside, lib.sglb_fused_experts_workspace_bytes(num_tokens, 1, 1, H, B.shared["w2"].shape[-1],
                                                                      B.shared.get("fmt", B.fmt)))
        self._ws_side = torch.empty(nside, dtype=torch.uint8, device=device)
        cap_stream = torch.cuda.Stream(device=device)
        self.graphs = [None, None]

        def enqueue(par):
            shared_y = px.start_shared(self.hidden, B)   # under the router and the exchange
            if router is not None:
                router(self.gating, self.topk_weights, self.topk_ids)
            px.enqueue(self.hidden, self.topk_weights, self.topk_ids, layer.E, B, par, layer.rsf, self.out, hint,
                       shared_y=shared_y)

        with torch.cuda.device(device):
            key = (device.index, cap_stream.cuda_stream)
            skey = (device.index, px.side.cuda_stream)
            cap_stream.wait_stream(torch.cuda.current_stream())
            with ops._launch_lock:
                saved = {k: ops._workspaces.get(k) for k in (key, skey)}
                ops._workspaces[key], ops._workspaces[skey] = self._ws, self._ws_side
                try:
                    with torch.cuda.stream(cap_stream):
                        # two eager calls (both parities): lazy one-off work (function attributes, tensor maps) stays
                        # out of the capture, and every rank's device-side call counter advances identically
                        for _ in range(2):
                            px.calls += 1
                            n0 = ops.launch_count()
                            enqueue(px.calls & 1)
                            self.kernels_per_replay = ops.launch_count() - n0
                    cap_stream.synchronize()
                    for par in (0, 1):
                        g = torch.cuda.CUDAGraph()
                        with torch.cuda.graph(g, stream=cap_stream):
                            enqueue(par)
                        self.graphs[par] = g
                finally:
                    for k, v in saved.items():
                        if v is None:
                            ops._workspaces.pop(k, None)
                        else:
                            ops._workspaces[k] = v
            torch.cuda.current_stream().wait_stream(cap_stream)

    def close(self):
        """Drops the graphs and lets the layer evict / close the peer region again.  Call it on every rank at the same
        point of the program (explicitly: garbage collection runs at different times on different ranks)."""
        if self.graphs is not None:
            self.graphs = None
            self.px.pins -= 1

    def __call__(self, hidden_states=None, topk_weights=None, topk_ids=None, gating=None):
        if self.graphs is None:
            raise RuntimeError("GraphedEP: called after close()")
        if hidden_states is not None:
            self.hidden.copy_(hidden_states, non_blocking=True)
        if gating is not None:
            self.gating.copy_(gating, non_blocking=True)
        if topk_weights is not None:
            self.topk_weights.copy_(topk_weights, non_blocking=True)
        if topk_ids is not None:
            self.topk_ids.copy_(topk_ids, non_blocking=True)
        self.px.calls += 1
        self.graphs[self.px.calls & 1].replay()
        return self.out


class EPMoE:
    """One MoE layer's routed experts (and optionally its shared expert), expert-parallel.

    weights: dict(w1=[E/n, 2I, H], w2=[E/n, H, I], s1, s2, block) -- this rank's experts only.
    shared:  dict(w1=[2Is, H], w2=[H, Is], s1, s2, block[, fmt]) -- the replicated shared expert; with it
             forward returns shared_expert(x) + routed(x) * routed_scaling_factor (shared_expert_cpu's contract).
    capacity: rows reserved per (source, destination) pair.  A GROUP-WIDE constant >= every rank's token count (the
             caller's promise: the check against it is rank-local, a rank that breaks it raises alone).  If omitted
             it is agreed by the group on every call (one all-reduce + host read per call) as the next power of two
             of the largest token count seen so far, and grows on all ranks together.  Ranks may hold different --
             also zero -- token counts.
    close(): collective, like forward.
    forward(x[M,H] bf16, topk_weights[M,k] f32, topk_ids[M,k] i32 with GLOBAL expert ids) -> [M,H] bf16,
    bit-identical to the single-GPU fused_experts (+ shared_expert) on the same tokens when both sides run the
    same GEMM kernel variant (pass `path`; SGLB_PATH_AUTO picks the kernel from the batch size).
    """

    FIXED_CAPACITY_BYTES = 32 << 20   # per-rank dispatch payload up to which padding beats a host sync

    def __init__(self, weights, fmt, num_experts, topk, comm=None, backend=None, is_vnni=False, capacity=None,
                 shared=None, routed_scaling_factor=1.0, path=L.PATH_AUTO):
        self.comm = comm if comm is not None else DistComm()
        world = self.comm.world
        if num_experts % world != 0:
            raise RuntimeError(f"EPMoE: {num_experts} experts do not split evenly over {world} ranks")
        self.E, self.topk, self.world = num_experts, topk, world
        self.backend = backend if backend is not None else CudaBackend(weights, fmt, num_experts // world, topk, is_vnni,
                                                                       path, shared)
        self.rsf = float(routed_scaling_factor)
        self.capacity = None if capacity is None else int(capacity)
        self._auto_capacity = capacity is None
        self.last_mode = None
        self._peer = {}          # (cap, H) -> PeerExchange
        self._peer_failed = False

    def close(self):
        for px in self._peer.values():
            px.close()
        self._peer = {}

    def _hint(self, cap):
        """Expected routed rows per local expert: the received rows are mostly masked (a rank owns ~1/world of
        the slots), so the kernel-path rule of fused_experts gets this estimate instead of rows*topk/E_local."""
        return float(self.world * cap * self.topk) / float(self.E)

    def _agreed_capacity(self, M, device, capacity):
        if capacity is not None:
            return int(capacity)
        if self._auto_capacity:
            # Agreed on EVERY call (one all-reduce + host read): a batch larger than anything seen so far makes the
            # whole group grow together, instead of one rank raising while its peers wait in the exchange for it.
            # Latency-critical callers pass `capacity` (constructor or call) or use graphed(), and skip this.
            mx = self.comm.max_int(M, device)
            if self.capacity is None or mx > self.capacity:
                self.capacity = max(16, 1 << max(0, mx - 1).bit_length())
        return self.capacity

    def _peer_exchange(self, cap, H, device):
        key = (cap, H)
        if key not in self._peer:
            try:
                # bounded: each entry pins an IPC-mapped region on every rank.  Oldest first, but never a region that a
                # live GraphedEP captured graphs over (same decision on every rank: close() is collective)
                while len(self._peer) >= 2:
                    victim = next((k for k, v in self._peer.items() if v.pins == 0), None)
                    if victim is None:
                        break
                    self._peer.pop(victim).close()
                self._peer[key] = PeerExchange(self.comm, cap, H, self.topk, device)
            except RuntimeError:
                self._peer_failed = True     # decided identically on every rank (see PeerExchange.__init__)
                return None
        return self._peer[key]

    def _mode(self, cap, H, x, mode):
        if mode is not None:
            return mode
        # every rank must take the same branch: decide from group-wide constants only
        mode = "fixed" if cap * self.world * H * 2 <= self.FIXED_CAPACITY_BYTES else "exact"
        if (mode == "fixed" and self.world > 1 and x.is_cuda and isinstance(self.backend, CudaBackend) and H % 8 == 0
                and not self._peer_failed and os.environ.get("SGLB_EP_P2P", "1") != "0"):
            mode = "p2p"
        return mode

    def graphed(self, num_tokens, H, device, capacity=None, router=None):
        """CUDA-graph form of the peer-memory path for a fixed token count (decode); None if peer memory is not
        available.  `router` (layer.Router): run the router kernel inside the graph, on a static `gating` buffer.
        Collective: every rank must call it, and replay in lockstep."""
        cap = self._agreed_capacity(num_tokens, device, capacity)
        px = self._peer_exchange(cap, H, device)
        if px is None:
            return None
        self.last_mode = "p2p"
        return GraphedEP(self, px, num_tokens, H, device, router)

    def forward(self, x, topk_weights, topk_ids, mode=None, capacity=None):
        B, comm, world = self.backend, self.comm, self.world
        M, H = x.shape
        if topk_ids.shape != (M, self.topk) or topk_weights.shape != (M, self.topk):
            raise RuntimeError("EPMoE: topk_weights / topk_ids must both be [num_tokens, topk]")
        ids = topk_ids.to(torch.int32).contiguous()
        tw = topk_weights.to(torch.float32).contiguous()
        x = x.contiguous()
        cap = self._agreed_capacity(M, x.device, capacity)
        if cap < M:
            raise RuntimeError(f"EPMoE: the group capacity {cap} is smaller than this rank's {M} tokens; construct the "
                               "layer with capacity >= the largest batch any rank will bring")
        mode = self._mode(cap, H, x, mode)
        if mode == "p2p":
            px = self._peer_exchange(cap, H, x.device)
            if px is None:
                mode = "fixed"
            else:
                self.last_mode = mode
                return px.run(x, tw, ids, self.E, B, self.rsf, self._hint(cap))
        self.last_mode = mode
        shared_y = None
        if getattr(B, "shared", None) is not None and M > 0:
            shared_y = self._shared_async(B, x)
        dest_row, send_count, slot_index, send_slots = B.plan(ids, self.E, world)
        topk = self.topk
        if mode == "fixed":
            rows = recv_rows = cap * world
            ret_rows = send_ret_rows = cap * world * topk
            sc = rc = ss = rs = None
            rcount_dev = rslots_dev = None
            hint = self._hint(cap)
        else:
            counts = torch.stack([send_count, send_slots], dim=1).contiguous()
            recv_counts = comm.exchange_counts(counts)
            both = torch.cat([counts, recv_counts], dim=1).cpu()      # the one host sync of this mode
            sc, ss, rc, rs = (both[:, i].tolist() for i in range(4))
            cap = 0
            rows, recv_rows, ret_rows, send_ret_rows = sum(sc), sum(rc), sum(ss), sum(rs)
            rcount_dev, rslots_dev = recv_counts[:, 0].contiguous(), recv_counts[:, 1].contiguous()
            hint = float(send_ret_rows) / float(self.E // world)
        send_x, send_meta = B.pack(x, ids, tw, dest_row, send_count, self.E, world, cap, rows)
        recv_x = torch.empty(recv_rows, H, dtype=x.dtype, device=x.device)
        recv_meta = torch.empty(recv_rows, 2 * topk, dtype=torch.int32, device=x.device)
        comm.all_to_all(recv_x, send_x, rc, sc)
        comm.all_to_all(recv_meta, send_meta, rc, sc)
        ret_index = B.ret_index(recv_meta, world, cap, rcount_dev, rslots_dev)
        y = B.experts_slots(recv_x, recv_meta, ret_index, send_ret_rows, hint)
        ret = torch.empty(ret_rows, H, dtype=x.dtype, device=x.device)
        comm.all_to_all(ret, y, ss, rs)
        if shared_y is not None:
            shared_y = shared_y()
        return B.combine(ret, ids, slot_index, send_slots, self.E, world, cap, shared_y, self.rsf)

    def _shared_async(self, B, x):
        """Shared expert MLP on a side stream (CUDA) while the exchange runs; returns a join function."""
        if not x.is_cuda:
            y = B.shared_mlp(x)
            return lambda: y
        if not hasattr(self, "_side"):
            self._side = torch.cuda.Stream(device=x.device)
        main = torch.cuda.current_stream()
        self._side.wait_stream(main)
        with torch.cuda.stream(self._side):
            y = B.shared_mlp(x)

        def join():
            main.wait_stream(self._side)
            return y
        return join

    def forward_host(self, x_host, tw_host, ids_host, out_host=None):
        """End-to-end form used by bench.py's e2e leg: pinned host buffers in, host result out."""
        dev = torch.device("cuda", torch.cuda.current_device())
        x = x_host.to(dev, non_blocking=True)
        tw = tw_host.to(dev, non_blocking=True)
        ids = ids_host.to(dev, non_blocking=True)
        out = self.forward(x, tw, ids)
        if out_host is None:
            return out.cpu()
        out_host.copy_(out)
        return out_host
