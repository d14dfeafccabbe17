"""GPU tests of the expert-parallel layer: the plan / pack / combine kernels bit-exact against the
oracle, and the whole EPMoE path with `world` ranks emulated as threads of one process on one GPU
(host-side barriers only -- no kernel waits on another) against the single-GPU fused_experts."""
import threading

import pytest
import torch

from oracle import moe_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    import sgl_cpu_tests_b200 as K
    assert torch.cuda.is_available()
    return K


@pytest.mark.parametrize("shape", [(1, 8, 2, 64, 2), (13, 16, 4, 72, 4), (300, 256, 8, 7168, 8), (2500, 128, 8, 512, 8),
                                   (5, 8, 8, 33, 8)])
def test_ep_kernels_bit_exact(K, shape):
    from sgl_cpu_tests_b200 import ep
    M, E, topk, H, world = shape
    torch.manual_seed(M)
    ids = torch.randint(-1, E, (M, topk), dtype=torch.int32)
    tw = torch.rand(M, topk)
    x = torch.randn(M, H).bfloat16()
    B = ep.CudaBackend({}, 0, E // world, topk)
    dr_ref, sc_ref = O.ep_plan(ids, E, world)
    dr, sc = B.plan(ids.cuda(), E, world)
    assert torch.equal(dr.cpu(), dr_ref) and torch.equal(sc.cpu(), sc_ref)
    for cap in (0, M + 3):
        rows = cap * world if cap else int(sc_ref.sum())
        sx_ref, meta_ref = O.ep_pack(x, ids, tw, dr_ref, sc_ref, E, world, cap)
        sx, meta = B.pack(x.cuda(), ids.cuda(), tw.cuda(), dr, sc, E, world, cap, rows)
        assert torch.equal(meta.cpu(), meta_ref)
        live = (meta_ref[:, :topk] >= 0).any(dim=1) | torch.zeros(rows, dtype=torch.bool)
        seg = O.ep_seg_start(sc_ref, world, cap)
        written = torch.zeros(rows, dtype=torch.bool)
        for r in range(world):
            written[seg[r]:seg[r] + int(sc_ref[r])] = True
        assert torch.equal(sx.cpu()[written], sx_ref[written])
        ret = torch.randn(rows, H).bfloat16()
        out = B.combine(ret.cuda(), dr, sc, world, cap)
        assert torch.equal(out.cpu(), O.ep_combine(ret, dr_ref, sc_ref, world, cap))


class ThreadComm:
    """In-process stand-in for the NCCL group: `world` threads, host barriers, same CUDA stream."""

    class Shared:
        def __init__(self, world):
            self.world = world
            self.barrier = threading.Barrier(world)
            self.posted = [None] * world

    def __init__(self, shared, rank):
        self.sh, self.rank, self.world = shared, rank, shared.world

    def all_to_all(self, out, inp, out_splits=None, in_splits=None):
        sh, w = self.sh, self.world
        if in_splits is None:
            in_splits = [inp.shape[0] // w] * w
        if out_splits is None:
            out_splits = [out.shape[0] // w] * w
        sh.posted[self.rank] = (inp, in_splits)
        sh.barrier.wait()
        o = 0
        for q in range(w):
            src, sp = sh.posted[q]
            off = sum(sp[:self.rank])
            n = sp[self.rank]
            assert n == out_splits[q]
            out[o:o + n].copy_(src[off:off + n])
            o += n
        sh.barrier.wait()

    def exchange_counts(self, send_count):
        recv = torch.empty_like(send_count)
        self.all_to_all(recv, send_count)
        return recv


@pytest.mark.parametrize("mode", ["fixed", "exact"])
@pytest.mark.parametrize("fmt", ["bf16", "int8"])
def test_ep_threaded_ranks_match_single_gpu(K, fmt, mode):
    from sgl_cpu_tests_b200 import ep
    world, M, E, topk, H, I = 4, 37, 16, 4, 256, 128
    torch.manual_seed(5)
    dev = torch.device("cuda")
    if fmt == "bf16":
        w1 = (torch.randn(E, 2 * I, H) / 10).bfloat16().to(dev)
        w2 = (torch.randn(E, H, I) / 10).bfloat16().to(dev)
        s1 = s2 = None
        fid = K.FMT_BF16
    else:
        w1 = torch.randint(-127, 128, (E, 2 * I, H), dtype=torch.int8).to(dev)
        w2 = torch.randint(-127, 128, (E, H, I), dtype=torch.int8).to(dev)
        s1, s2 = (torch.rand(E, 2 * I) * 1e-2).to(dev), (torch.rand(E, H) * 1e-2).to(dev)
        fid = K.FMT_INT8_W8A8
    xs, tws, idss = [], [], []
    for r in range(world):
        xs.append((torch.randn(M, H) / 10).bfloat16().to(dev))
        tw, ids = torch.topk(torch.softmax(torch.randn(M, E), dim=-1), topk)
        ids = ids.to(torch.int32)
        ids[r, 0] = -1
        tws.append(tw.to(dev)), idss.append(ids.to(dev))
    refs = [K.fused_experts_impl(xs[r], w1, w2, tws[r], idss[r], False, fid, s1, s2, None, False) for r in range(world)]
    torch.cuda.synchronize()
    shared = ThreadComm.Shared(world)
    outs, errs = [None] * world, []
    el = E // world

    def run(r):
        try:
            torch.cuda.set_device(0)
            local = dict(w1=w1[r * el:(r + 1) * el].contiguous(), w2=w2[r * el:(r + 1) * el].contiguous(),
                         s1=None if s1 is None else s1[r * el:(r + 1) * el].contiguous(),
                         s2=None if s2 is None else s2[r * el:(r + 1) * el].contiguous(), block=None)
            layer = ep.EPMoE(local, fid, E, topk, comm=ThreadComm(shared, r))
            outs[r] = layer.forward(xs[r], tws[r], idss[r], mode=mode)
        except Exception as e:  # noqa: BLE001
            errs.append(e)
            shared.barrier.abort()

    ts = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=120)
    assert not errs, errs
    torch.cuda.synchronize()
    for r in range(world):
        a, b = outs[r].float().cpu(), refs[r].float().cpu()
        assert torch.allclose(a, b, rtol=2e-2, atol=2e-2), float((a - b).abs().max())
