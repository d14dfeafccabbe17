"""Host-side orchestration of the expert-parallel layer (sgl_cpu_tests_b200/ep.py) on CPU:
two gloo ranks, device work replaced by the oracle (tests only -- the product backend is CUDA)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import moe_oracle as O


class OracleBackend:
    """Stands in for ep.CudaBackend in CPU tests: same method contract, oracle arithmetic."""

    def __init__(self, weights, fmt, e_local, topk):
        self.w, self.fmt, self.e_local, self.topk = weights, fmt, e_local, topk

    def plan(self, ids, E, world):
        return O.ep_plan(ids, E, world)

    def pack(self, x, ids, tw, dest_row, send_count, E, world, cap, rows):
        sx, sm = O.ep_pack(x, ids, tw, dest_row, send_count, E, world, cap)
        assert sx.shape[0] == rows
        return sx, sm

    def experts(self, x, meta):
        ids = meta[:, :self.topk].contiguous()
        tw = meta[:, self.topk:].contiguous().view(torch.float32)
        w = self.w
        return O.fused_experts(x, w["w1"], w["w2"], tw, ids, self.fmt, w.get("s1"), w.get("s2"), w.get("block"))

    def combine(self, ret, dest_row, send_count, world, cap):
        return O.ep_combine(ret, dest_row, send_count, world, cap)


def make_case(seed, M, E, topk, H, I):
    g = torch.Generator().manual_seed(seed)
    w1 = (torch.randn(E, 2 * I, H, generator=g) / 10).bfloat16()
    w2 = (torch.randn(E, H, I, generator=g) / 10).bfloat16()
    xs, tws, idss = [], [], []
    for r in range(2):
        x = (torch.randn(M + r, H, generator=g) / 10).bfloat16()        # ragged: ranks hold different M
        score = torch.softmax(torch.randn(M + r, E, generator=g), dim=-1)
        tw, ids = torch.topk(score, topk)
        ids = ids.to(torch.int32)
        ids[0, 0] = -1                                                   # a masked slot (offloading semantics)
        xs.append(x), tws.append(tw), idss.append(ids)
    return w1, w2, xs, tws, idss


def _rank_main(rank, world, port, mode, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sgl_cpu_tests_b200 import ep
        M, E, topk, H, I = 9, 8, 3, 32, 16
        w1, w2, xs, tws, idss = make_case(7, M, E, topk, H, I)
        el = E // world
        local = dict(w1=w1[rank * el:(rank + 1) * el], w2=w2[rank * el:(rank + 1) * el])
        layer = ep.EPMoE(local, "bf16", E, topk, backend=OracleBackend(local, "bf16", el, topk))
        out = layer.forward(xs[rank], tws[rank], idss[rank], mode=mode, capacity=M + world - 1)
        ref = O.fused_experts(xs[rank], w1, w2, tws[rank], idss[rank], "bf16")
        err = float((out.float() - ref.float()).abs().max())
        q.put((rank, err, layer.last_mode, tuple(out.shape)))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("mode", ["fixed", "exact"])
def test_ep_two_ranks_gloo(mode):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, mode, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err, last_mode, shape in res:
        assert last_mode == mode
        assert shape == (9 + rank, 32)
        assert err <= 2e-2, (rank, err)      # partial sums are rounded to bf16 once per rank


def test_ep_oracle_layout_single_process():
    """plan / pack / combine contracts, world = 4, both segment layouts."""
    torch.manual_seed(3)
    M, E, topk, H, world = 13, 16, 4, 8, 4
    ids = torch.randint(-1, E, (M, topk), dtype=torch.int32)
    tw = torch.rand(M, topk)
    x = torch.randn(M, H).bfloat16()
    dest_row, send_count = O.ep_plan(ids, E, world)
    el = E // world
    for r in range(world):
        owns = ((ids >= r * el) & (ids < (r + 1) * el)).any(dim=1)
        assert int(send_count[r]) == int(owns.sum())
        assert torch.equal(dest_row[owns, r], torch.arange(int(owns.sum()), dtype=torch.int32))
        assert (dest_row[~owns, r] == -1).all()
    for cap in (0, M):
        sx, meta = O.ep_pack(x, ids, tw, dest_row, send_count, E, world, cap)
        seg = O.ep_seg_start(send_count, world, cap)
        for r in range(world):
            for m in range(M):
                d = int(dest_row[m, r])
                if d < 0:
                    continue
                row = seg[r] + d
                assert torch.equal(sx[row], x[m])
                for j in range(topk):
                    e = int(ids[m, j])
                    mine = r * el <= e < (r + 1) * el
                    assert int(meta[row, j]) == (e - r * el if mine else -1)
        # identity "experts": returning the sent rows makes combine count the destination ranks
        out = O.ep_combine(sx, dest_row, send_count, world, cap)
        n_dest = (dest_row >= 0).sum(dim=1, keepdim=True).float()
        assert torch.allclose(out.float(), (x.float() * n_dest).bfloat16().float(), rtol=1e-2, atol=1e-2)
