"""Expert-parallel fused MoE: experts sharded over the ranks of one NVLink/NVSwitch box, tokens
data-parallel, one all-to-all each side of the local fused_experts.

Nothing in the reference shards experts.  The hook it offers is that ``fused_experts_cpu`` ignores
``topk_ids == -1`` (test_moe_offloading_cpu.py:57-68) -- exactly what a rank sees for experts it does
not own -- so the local compute is the unchanged single-GPU path.

Per layer call, on every rank (M local tokens, rank r owns experts [r*E/n, (r+1)*E/n)):

  plan     (CUDA, csrc/ep.cu)  which ranks own at least one of each token's experts; stable row numbering
  pack     (CUDA)              token rows + (local ids | weights) gathered per destination rank
  dispatch (NCCL all-to-all)   rows travel once per (token, destination rank), not once per expert
  experts  (CUDA)              ordinary fused_experts over the received rows, ids already local / -1
  return   (NCCL all-to-all)   one partial row per received row
  combine  (CUDA)              out[m] = sum over ranks in rank order, f32 accumulate -> bf16

Two exchange modes, chosen per call from M:
  * fixed capacity (small M / decode): every segment is padded to M rows and padding carries
    ids = -1, so the split sizes are static -- no count exchange, no host sync, graph-capturable;
  * exact splits (large M / prefill): counts are exchanged first and read on the host once.

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

    def __init__(self, weights, fmt, e_local, topk, is_vnni=False):
        self.w, self.fmt, self.e_local, self.topk, self.is_vnni = weights, fmt, e_local, topk, is_vnni

    @staticmethod
    def _s():
        return torch.cuda.current_stream().cuda_stream

    def plan(self, ids, E, world):
        M, topk = ids.shape
        dest_row = torch.empty(M, world, dtype=torch.int32, device=ids.device)
        send_count = torch.empty(world, dtype=torch.int32, device=ids.device)
        L.check(L.load().sglb_ep_plan(ids.data_ptr(), M, topk, E, world, dest_row.data_ptr(),
                                      send_count.data_ptr(), self._s()))
        return dest_row, send_count

    def pack(self, x, ids, tw, dest_row, send_count, E, world, cap, rows):
        M, H = x.shape
        topk = ids.shape[1]
        send_x = torch.empty(rows, H, dtype=x.dtype, device=x.device)
        send_meta = torch.empty(rows, 2 * topk, dtype=torch.int32, device=x.device)
        L.check(L.load().sglb_ep_pack(x.data_ptr(), ids.data_ptr(), tw.data_ptr(), dest_row.data_ptr(),
                                      send_count.data_ptr(), M, topk, E, world, H, cap, rows, send_x.data_ptr(),
                                      send_meta.data_ptr(), self._s()))
        return send_x, send_meta

    def experts(self, x, meta):
        topk = self.topk
        ids = meta[:, :topk].contiguous()
        tw = meta[:, topk:].contiguous().view(torch.float32)
        w = self.w
        return ops.fused_experts_impl(x, w["w1"], w["w2"], tw, ids, False, self.fmt, w.get("s1"), w.get("s2"),
                                      w.get("block"), self.is_vnni)

    def experts_into(self, x, tw, ids, out):
        """Local experts over received slots with separate ids / weights tensors, result written to `out`."""
        w = self.w
        return ops.fused_experts_impl(x, w["w1"], w["w2"], tw, ids, False, self.fmt, w.get("s1"), w.get("s2"),
                                      w.get("block"), self.is_vnni, out=out)

    def combine(self, ret, dest_row, send_count, world, cap):
        M = dest_row.shape[0]
        H = ret.shape[1]
        out = torch.empty(M, H, dtype=ret.dtype, device=ret.device)
        L.check(L.load().sglb_ep_combine(ret.data_ptr(), dest_row.data_ptr(), send_count.data_ptr(), M, world, H,
                                         cap, out.data_ptr(), self._s()))
        return out


class DistComm:
    """torch.distributed plumbing (NCCL over NVLink on the box, gloo in the CPU tests)."""

    def __init__(self, group=None):
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)

    def all_to_all(self, out, inp, out_splits=None, in_splits=None):
        dist.all_to_all_single(out, inp, out_splits, in_splits, group=self.group)

    def exchange_counts(self, send_count):
        recv = torch.empty_like(send_count)
        dist.all_to_all_single(recv, send_count, group=self.group)
        return recv


class _DevMem:
    """Zero-copy torch view of raw device memory through __cuda_array_interface__."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class PeerExchange:
    """Decode-regime exchange over NVLink peer memory (csrc/ep_p2p.cu): one cudaMalloc'd region per rank,
    mapped into every peer through CUDA IPC; rows, ids and weights are stored straight into the owners'
    receive buffers and partial results are read back with peer loads.  No NCCL call on the data path --
    torch.distributed only carries the 64-byte IPC handles once, at construction."""

    def __init__(self, group, world, rank, cap, H, topk, device):
        import ctypes as C
        lib = L.load()
        self.world, self.rank, self.cap, self.H, self.topk = world, rank, cap, H, topk
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
        gathered = [None] * world
        dist.all_gather_object(gathered, (status, bytes(handle.raw)), group=group)
        self.base = base.value
        self.peers = (C.c_void_p * world)()
        self._opened = []
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
        final = [None] * world
        dist.all_gather_object(final, status, group=group)
        bad = [f for f in final if f != "ok"]
        if bad:
            self.close()
            raise RuntimeError(f"PeerExchange: peer mapping failed on some rank: {bad[0]}")
        mem = torch.as_tensor(_DevMem(self.base, self.nbytes), device=device)
        rows = world * cap

        def view(off, dtype, cols):
            nb = rows * cols * torch.empty((), dtype=dtype).element_size()
            return mem[off:off + nb].view(dtype).view(rows, cols)
        self.recv_x = [view(self.offs[p], torch.bfloat16, H) for p in range(2)]
        self.recv_ids = [view(self.offs[2 + p], torch.int32, topk) for p in range(2)]
        self.recv_tw = [view(self.offs[4 + p], torch.float32, topk) for p in range(2)]
        self.y = [view(self.offs[6 + p], torch.bfloat16, H) for p in range(2)]
        self.seq = 0

    def close(self):
        lib = L.load()
        for p in self._opened:
            lib.sglb_p2p_close(p)
        self._opened = []
        if self.base:
            lib.sglb_p2p_free(self.base)
            self.base = None

    def run(self, x, tw, ids, E, experts_fn):
        """dispatch -> wait -> experts_fn(recv_x, recv_tw, recv_ids, out=y) -> combine; returns [M, H]."""
        lib = L.load()
        s = torch.cuda.current_stream().cuda_stream
        M = x.shape[0]
        self.seq += 1
        seq, par = self.seq & 0xFFFFFFFF, self.seq & 1
        L.check(lib.sglb_ep_p2p_dispatch(x.data_ptr(), ids.data_ptr(), tw.data_ptr(), M, self.topk, E, self.world,
                                         self.rank, self.H, self.cap, self.peers, seq, s))
        L.check(lib.sglb_ep_p2p_wait_dispatch(self.base, self.world, self.cap, self.H, self.topk, seq, s))
        experts_fn(self.recv_x[par], self.recv_tw[par], self.recv_ids[par], self.y[par])
        out = torch.empty_like(x)
        L.check(lib.sglb_ep_p2p_combine(ids.data_ptr(), M, self.topk, E, self.world, self.rank, self.H, self.cap,
                                        self.peers, seq, out.data_ptr(), s))
        return out


class EPMoE:
    """One MoE layer's routed experts, expert-parallel.

    weights: dict(w1=[E/n, 2I, H], w2=[E/n, H, I], s1, s2, block) -- this rank's experts only.
    forward(x[M,H] bf16, topk_weights[M,k] f32, topk_ids[M,k] i32 with GLOBAL expert ids) -> [M,H] bf16.
    """

    FIXED_CAPACITY_BYTES = 32 << 20   # per-rank dispatch payload up to which padding beats a host sync

    def __init__(self, weights, fmt, num_experts, topk, comm=None, backend=None, is_vnni=False):
        self.comm = comm if comm is not None else DistComm()
        world = self.comm.world
        if num_experts % world != 0:
            raise RuntimeError(f"EPMoE: {num_experts} experts do not split evenly over {world} ranks")
        self.E, self.topk, self.world = num_experts, topk, world
        self.backend = backend if backend is not None else CudaBackend(weights, fmt, num_experts // world, topk, is_vnni)
        self.last_mode = None
        self._peer = {}          # (cap, H) -> PeerExchange
        self._peer_failed = False

    def _peer_exchange(self, cap, H, device):
        key = (cap, H)
        if key not in self._peer:
            try:
                self._peer[key] = PeerExchange(self.comm.group, self.world, self.comm.rank, cap, H, self.topk, device)
            except RuntimeError:
                self._peer_failed = True     # decided identically on every rank (see PeerExchange.__init__)
                return None
        return self._peer[key]

    def forward(self, x, topk_weights, topk_ids, mode=None, capacity=None):
        """capacity: rows reserved per destination rank in fixed mode; must be the SAME on every rank and
        >= every rank's token count (default: this rank's M, i.e. all ranks hold the same M)."""
        B, comm, world = self.backend, self.comm, self.world
        M, H = x.shape
        if topk_ids.shape != (M, self.topk) or topk_weights.shape != (M, self.topk):
            raise RuntimeError("EPMoE: topk_weights / topk_ids must both be [num_tokens, topk]")
        ids = topk_ids.to(torch.int32).contiguous()
        tw = topk_weights.to(torch.float32).contiguous()
        x = x.contiguous()
        dest_row, send_count = B.plan(ids, self.E, world)
        cap = M if capacity is None else int(capacity)
        if cap < M:
            raise RuntimeError(f"EPMoE: capacity {cap} is smaller than the {M} local tokens")
        if mode is None:
            # every rank must take the same branch: decide from static shapes only
            mode = "fixed" if cap * world * H * x.element_size() <= self.FIXED_CAPACITY_BYTES else "exact"
            if (mode == "fixed" and world > 1 and x.is_cuda and isinstance(B, CudaBackend) and M >= 1 and H % 8 == 0
                    and not self._peer_failed and os.environ.get("SGLB_EP_P2P", "1") != "0"):
                mode = "p2p"
        if mode == "p2p":
            pcap = max(16, 1 << (cap - 1).bit_length())     # slots per source: next power of two
            px = self._peer_exchange(pcap, H, x.device)
            if px is None:
                mode = "fixed"
            else:
                self.last_mode = mode
                return px.run(x, tw, ids, self.E, B.experts_into)
        self.last_mode = mode
        if mode == "fixed":
            rows, recv_rows, sc, rc = cap * world, cap * world, None, None
        else:
            recv_count = comm.exchange_counts(send_count)
            both = torch.stack([send_count, recv_count]).cpu()      # the one host sync of this mode
            sc, rc = both[0].tolist(), both[1].tolist()
            cap, rows, recv_rows = 0, sum(sc), sum(rc)
        send_x, send_meta = B.pack(x, ids, tw, dest_row, send_count, self.E, world, cap, rows)
        recv_x = torch.empty(recv_rows, H, dtype=x.dtype, device=x.device)
        recv_meta = torch.empty(recv_rows, 2 * self.topk, dtype=torch.int32, device=x.device)
        comm.all_to_all(recv_x, send_x, rc, sc)
        comm.all_to_all(recv_meta, send_meta, rc, sc)
        y = B.experts(recv_x, recv_meta)
        ret = torch.empty(rows, H, dtype=x.dtype, device=x.device)
        comm.all_to_all(ret, y, sc, rc)
        return B.combine(ret, dest_row, send_count, world, cap)

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
