"""Host-side orchestration of the expert-parallel layer (sgl_cpu_tests_b200/ep.py) on CPU:
two gloo ranks, device work replaced by the oracle (tests only -- the product backend is CUDA).
The exchange returns one row per routed slot and the source reduces in slot order, so the n-rank result
must be BIT-IDENTICAL to the single-rank composition reduce_slots(slot_rows(...)) -- torch.equal, no tolerance."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import moe_oracle as O


class OracleBackend:
    """Stands in for ep.CudaBackend in CPU tests: same method contract, oracle arithmetic."""

    def __init__(self, weights, fmt, e_local, topk, shared=None):
        self.w, self.fmt, self.e_local, self.topk, self.shared = weights, fmt, e_local, topk, shared

    def plan(self, ids, E, world):
        return O.ep_plan(ids, E, world)

    def pack(self, x, ids, tw, dest_row, send_count, E, world, cap, rows):
        sx, sm = O.ep_pack(x, ids, tw, dest_row, send_count, E, world, cap)
        assert sx.shape[0] == rows
        return sx, sm

    def ret_index(self, recv_meta, world, cap, recv_count=None, recv_slots=None):
        return O.ep_ret_index(recv_meta, self.topk, world, cap, recv_count, recv_slots)

    def experts_slots(self, x, meta, ret_index, ret_rows, hint=0.0):
        ids = meta[:, :self.topk].contiguous()
        tw = meta[:, self.topk:].contiguous().view(torch.float32)
        ret = torch.zeros(ret_rows, x.shape[1], dtype=x.dtype)
        if x.shape[0] == 0:
            return ret
        rows = slot_rows_rowwise(x, self.w["w1"], self.w["w2"], tw, ids)
        valid = ret_index.view(-1) >= 0
        ret[ret_index.view(-1)[valid].long()] = rows[valid]
        return ret

    def shared_mlp(self, x):
        if self.shared is None or x.shape[0] == 0:
            return None
        return shared_rowwise(x, self.shared["w1"], self.shared["w2"])

    def combine(self, ret, ids, slot_index, send_slots, E, world, cap, shared_y=None, rsf=1.0):
        return O.ep_combine(ret, ids, slot_index, send_slots, E, world, cap, shared_y, rsf)


def slot_rows_rowwise(x, w1, w2, tw, ids):
    """O.slot_rows computed one slot at a time (mat-vec), so a row's bits cannot depend on which other rows share
    the call -- the property the CUDA kernels have per output element and BLAS does not promise across batch sizes."""
    B, topk = ids.shape
    out = torch.zeros(B * topk, w2.shape[1], dtype=x.dtype)
    for m in range(B):
        for j in range(topk):
            e = int(ids[m, j])
            if 0 <= e < w1.shape[0]:
                h = O.silu_and_mul(torch.mv(w1[e].float(), x[m].float()))
                out[m * topk + j] = (torch.mv(w2[e].float(), h) * tw[m, j].float()).to(x.dtype)
    return out


def shared_rowwise(x, w1, w2):
    out = torch.zeros(x.shape[0], w2.shape[0], dtype=x.dtype)
    for m in range(x.shape[0]):
        out[m] = torch.mv(w2.float(), O.silu_and_mul(torch.mv(w1.float(), x[m].float()))).to(x.dtype)
    return out


MS = {"even": (9, 9), "ragged": (9, 12), "empty_rank": (7, 0)}


def make_case(seed, Ms, E, topk, H, I):
    g = torch.Generator().manual_seed(seed)
    w1 = (torch.randn(E, 2 * I, H, generator=g) / 10).bfloat16()
    w2 = (torch.randn(E, H, I, generator=g) / 10).bfloat16()
    sw1 = (torch.randn(2 * I, H, generator=g) / 10).bfloat16()
    sw2 = (torch.randn(H, I, generator=g) / 10).bfloat16()
    xs, tws, idss = [], [], []
    for M in Ms:
        x = (torch.randn(M, H, generator=g) / 10).bfloat16()
        score = torch.softmax(torch.randn(M, E, generator=g), dim=-1)
        tw, ids = torch.topk(score, topk)
        ids = ids.to(torch.int32)
        if M:
            ids[0, 0] = -1                                               # a masked slot (offloading semantics)
            ids[M - 1, 1] = ids[M - 1, 2]                                # a duplicate expert inside a row
        xs.append(x), tws.append(tw), idss.append(ids)
    return w1, w2, sw1, sw2, xs, tws, idss


CASES = [(mode, case, sh) for mode in ("fixed", "exact") for case in ("even", "ragged", "empty_rank") for sh in (False, True)]


def _rank_main(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sgl_cpu_tests_b200 import ep
        E, topk, H, I = 8, 3, 32, 16
        for mode, case, with_shared in CASES:
            w1, w2, sw1, sw2, xs, tws, idss = make_case(7, MS[case], E, topk, H, I)
            el = E // world
            local = dict(w1=w1[rank * el:(rank + 1) * el], w2=w2[rank * el:(rank + 1) * el])
            shared = dict(w1=sw1, w2=sw2) if with_shared else None
            layer = ep.EPMoE(local, "bf16", E, topk, backend=OracleBackend(local, "bf16", el, topk, shared),
                             routed_scaling_factor=2.5)
            x, tw, ids = xs[rank], tws[rank], idss[rank]
            out = layer.forward(x, tw, ids, mode=mode)          # capacity agreed by the group at this first call
            ref = O.reduce_slots(slot_rows_rowwise(x, w1, w2, tw, ids), ids, E)
            if with_shared:      # shared_expert_cpu's epilogue on the rounded routed sum: fmaf in f32, then bf16
                ref = (shared_rowwise(x, sw1, sw2).double() + ref.double() * 2.5).float().to(x.dtype)
            loose = O.fused_experts(x, w1, w2, tw, ids, "bf16") if not with_shared and x.shape[0] else None
            q.put((rank, (mode, case, with_shared), bool(torch.equal(out, ref)), layer.last_mode, tuple(out.shape),
                   layer.capacity, None if loose is None else float((out.float() - loose.float()).abs().max())))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_ep_two_ranks_gloo():
    """Two gloo ranks x {fixed, exact} x {equal, ragged, zero tokens on a rank} x {routed only, + shared expert}."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(2 * len(CASES))]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert len(res) == 2 * len(CASES)
    for rank, (mode, case, with_shared), equal, last_mode, shape, cap, loose_err in res:
        assert last_mode == mode
        assert shape == (MS[case][rank], 32)
        assert cap == 16                     # next power of two of the group maximum, at least 16
        assert equal, (rank, mode, case, with_shared, "n-rank result differs bitwise from the single-rank slot-order reduction")
        if loose_err is not None:            # and it is the reference's MoE within the reference's tolerance
            assert loose_err <= 1e-2


def _regrow_main(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sgl_cpu_tests_b200 import ep
        E, topk, H, I = 8, 3, 32, 16
        el = E // world
        layer = None
        # a warm-up sized batch first, then one that only rank 0 outgrows the agreed capacity with, then a small one
        for step, Ms in enumerate([(3, 2), (20, 5), (1, 1)]):
            w1, w2, _, _, xs, tws, idss = make_case(11 + step, Ms, E, topk, H, I)
            local = dict(w1=w1[rank * el:(rank + 1) * el], w2=w2[rank * el:(rank + 1) * el])
            if layer is None:
                layer = ep.EPMoE(local, "bf16", E, topk, backend=OracleBackend(local, "bf16", el, topk, None))
            else:
                layer.backend = OracleBackend(local, "bf16", el, topk, None)
            x, tw, ids = xs[rank], tws[rank], idss[rank]
            out = layer.forward(x, tw, ids, mode="fixed")
            ref = O.reduce_slots(slot_rows_rowwise(x, w1, w2, tw, ids), ids, E)
            q.put((rank, step, bool(torch.equal(out, ref)), layer.capacity))
    finally:
        dist.destroy_process_group()


def test_ep_auto_capacity_grows_on_every_rank_together():
    """A later batch larger than the capacity agreed at the first call must not raise on one rank while the other
    enters the exchange: without an explicit capacity the group re-agrees per call and grows together."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_regrow_main, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(6)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    caps = {0: 16, 1: 32, 2: 32}
    for rank, step, equal, cap in res:
        assert cap == caps[step], (rank, step, cap)
        assert equal, (rank, step)


def test_ep_oracle_layout_single_process():
    """plan / pack / ret_index / combine contracts, world = 4, both segment layouts: the exchange composed in one
    process reproduces reduce_slots(slot_rows) exactly."""
    torch.manual_seed(3)
    M, E, topk, H, world = 13, 16, 4, 8, 4
    ids = torch.randint(-1, E, (M, topk), dtype=torch.int32)
    tw = torch.rand(M, topk)
    x = torch.randn(M, H).bfloat16()
    dest_row, send_count, slot_index, send_slots = O.ep_plan(ids, E, world)
    el = E // world
    for r in range(world):
        mine = (ids >= r * el) & (ids < (r + 1) * el)
        owns = mine.any(dim=1)
        assert int(send_count[r]) == int(owns.sum()) and int(send_slots[r]) == int(mine.sum())
        assert torch.equal(dest_row[owns, r], torch.arange(int(owns.sum()), dtype=torch.int32))
        assert (dest_row[~owns, r] == -1).all()
        assert torch.equal(slot_index[mine], torch.arange(int(mine.sum()), dtype=torch.int32))   # (token, j) order
    assert (slot_index[(ids < 0)] == -1).all()
    slot_val = (torch.arange(M * topk).view(M, topk) + 1).float()          # a recognisable "row" per slot
    for cap in (0, M):
        sx, meta = O.ep_pack(x, ids, tw, dest_row, send_count, E, world, cap)
        seg = O.ep_seg_start(send_count, world, cap)
        ret_rows = cap * world * topk if cap else int(send_slots.sum())
        ret = torch.zeros(ret_rows, H)
        slot_seg = O.ep_seg_start(send_slots, world, cap * topk)
        for r in range(world):
            n = cap if cap else int(send_count[r])
            recv_meta = meta[seg[r]:seg[r] + n]                         # what rank r receives from this (only) source
            for m in range(M):
                d = int(dest_row[m, r])
                if d >= 0:
                    assert torch.equal(sx[seg[r] + d], x[m])
            # owner r numbers the slots of this source alone: world = 1 from its point of view
            ri = O.ep_ret_index(recv_meta, topk, 1, n if cap else 0, torch.tensor([n]), torch.tensor([int(send_slots[r])]))
            toks = [m for m in range(M) if int(dest_row[m, r]) >= 0]
            for i, m in enumerate(toks):
                for j in range(topk):
                    if int(recv_meta[i, j]) >= 0:
                        assert int(ids[m, j]) == int(recv_meta[i, j]) + r * el
                        ret[slot_seg[r] + int(ri[i, j])] = slot_val[m, j]
                    else:
                        assert int(ri[i, j]) == -1
        out = O.ep_combine(ret.bfloat16(), ids, slot_index, send_slots, E, world, cap)
        want = (slot_val * (ids >= 0)).sum(dim=1, keepdim=True).expand(M, H).bfloat16()
        assert torch.equal(out, want)
