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
                                   (5, 8, 8, 33, 8), (0, 16, 4, 64, 4)])
def test_ep_kernels_bit_exact(K, shape):
    from sgl_cpu_tests_b200 import ep
    M, E, topk, H, world = shape
    torch.manual_seed(M)
    ids = torch.randint(-1, E, (M, topk), dtype=torch.int32)
    tw = torch.rand(M, topk)
    x = torch.randn(M, H).bfloat16()
    B = ep.CudaBackend({}, 0, E // world, topk)
    ref = O.ep_plan(ids, E, world)
    got = B.plan(ids.cuda(), E, world)
    for a, b, name in zip(got, ref, ("dest_row", "send_count", "slot_index", "send_slots")):
        assert torch.equal(a.cpu(), b), name
    dr_ref, sc_ref, si_ref, ss_ref = ref
    dr, sc, si, ss = got
    for cap in (0, M + 3):
        rows = cap * world if cap else int(sc_ref.sum())
        sx_ref, meta_ref = O.ep_pack(x, ids, tw, dr_ref, sc_ref, E, world, cap)
        sx, meta = B.pack(x.cuda(), ids.cuda(), tw.cuda(), dr, sc, E, world, cap, rows)
        assert torch.equal(meta.cpu(), meta_ref)
        seg = O.ep_seg_start(sc_ref, world, cap)
        written = torch.zeros(rows, dtype=torch.bool)
        for r in range(world):
            written[seg[r]:seg[r] + int(sc_ref[r])] = True
        assert torch.equal(sx.cpu()[written], sx_ref[written])
        # owner-side numbering on an arbitrary received buffer: random local ids with -1 holes, `world` sources
        n_src = torch.randint(0, 40, (world,), dtype=torch.int32) if not cap else None
        rrows = cap * world if cap else int(n_src.sum())
        rmeta = torch.randint(-1, 4, (rrows, 2 * topk), dtype=torch.int32)
        rslots = None
        if not cap:
            seg_r = O.ep_seg_start(n_src, world, 0)
            rslots = torch.tensor([int((rmeta[seg_r[s]:seg_r[s] + int(n_src[s]), :topk] >= 0).sum()) for s in range(world)],
                                  dtype=torch.int32)
        ri_ref = O.ep_ret_index(rmeta, topk, world, cap, n_src, rslots)
        ri = B.ret_index(rmeta.cuda(), world, cap, None if n_src is None else n_src.cuda(),
                         None if rslots is None else rslots.cuda())
        assert torch.equal(ri.cpu(), ri_ref)
        ret_rows = cap * world * topk if cap else int(ss_ref.sum())
        ret = torch.randn(max(ret_rows, 1), H).bfloat16()
        for shared_y, rsf in ((None, 1.0), (torch.randn(M, H).bfloat16(), 2.5)):
            out = B.combine(ret.cuda(), ids.cuda(), si, ss, E, world, cap, None if shared_y is None else shared_y.cuda(), rsf)
            assert torch.equal(out.cpu(), O.ep_combine(ret, ids, si_ref, ss_ref, E, world, cap, shared_y, rsf))


def test_slot_addressed_output_matches_reduce(K):
    """fused_experts with slot_out: the per-slot rows reduced in slot order (sglb_topk_reduce's arithmetic, restated by
    O.reduce_slots) reproduce the ordinary call bit for bit, on every GEMM path; with a slot_map the same rows land
    where the map says."""
    torch.manual_seed(11)
    dev = torch.device("cuda")
    M, E, topk, H, I = 70, 8, 4, 256, 128
    x = (torch.randn(M, H) / 10).bfloat16().to(dev)
    w1 = (torch.randn(E, 2 * I, H) / 10).bfloat16().to(dev)
    w2 = (torch.randn(E, H, I) / 10).bfloat16().to(dev)
    tw, ids = torch.topk(torch.softmax(torch.randn(M, E), dim=-1), topk)
    ids = ids.to(torch.int32)
    ids[3, 1] = -1
    ids[5] = torch.tensor([2, 2, 7, -1])
    tw, ids = tw.to(dev), ids.to(dev)
    for path in (K.PATH_STREAM, K.PATH_UMMA):
        ref = K.fused_experts_impl(x, w1, w2, tw, ids, False, K.FMT_BF16, None, None, None, False, path=path)
        rows = torch.zeros(M * topk, H, dtype=torch.bfloat16, device=dev)
        K.fused_experts_impl(x, w1, w2, tw, ids, False, K.FMT_BF16, None, None, None, False, path=path, slot_out=rows)
        assert torch.equal(O.reduce_slots(rows.cpu(), ids.cpu(), E), ref.cpu())
        perm = torch.randperm(M * topk, dtype=torch.int32)
        moved = torch.zeros_like(rows)
        K.fused_experts_impl(x, w1, w2, tw, ids, False, K.FMT_BF16, None, None, None, False, path=path, slot_out=moved,
                             slot_map=perm.to(dev))
        valid = (ids.view(-1) >= 0).cpu()
        assert torch.equal(moved.cpu()[perm.long()][valid], rows.cpu()[valid])


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

    def exchange_counts(self, send_counts):
        recv = torch.empty_like(send_counts)
        self.all_to_all(recv, send_counts.contiguous())
        return recv

    def max_int(self, value, device):
        sh = self.sh
        sh.posted[self.rank] = int(value)
        sh.barrier.wait()
        mx = max(sh.posted)
        sh.barrier.wait()
        return mx

    def all_gather_object(self, obj):
        sh = self.sh
        sh.posted[self.rank] = obj
        sh.barrier.wait()
        out = list(sh.posted)
        sh.barrier.wait()
        return out


@pytest.mark.parametrize("with_shared", [False, True])
@pytest.mark.parametrize("mode", ["fixed", "exact"])
@pytest.mark.parametrize("fmt", ["bf16", "int8", "fp8"])
def test_ep_threaded_ranks_bit_identical_to_single_gpu(K, fmt, mode, with_shared):
    """`world` ranks as threads of one process on one GPU (host barriers stand in for NCCL): every rank's output is
    torch.equal to the single-GPU fused_experts (+ shared_expert) on the same tokens -- ragged token counts, one rank
    without tokens, masked and duplicate ids.  Both sides run the same GEMM kernel variant (explicit path)."""
    from sgl_cpu_tests_b200 import ep
    world, E, topk, H, I = 4, 16, 4, 256, 128
    Ms = [37, 5, 0, 64]
    torch.manual_seed(5)
    dev = torch.device("cuda")
    block = None
    if fmt == "bf16":
        w1 = (torch.randn(E, 2 * I, H) / 10).bfloat16().to(dev)
        w2 = (torch.randn(E, H, I) / 10).bfloat16().to(dev)
        s1 = s2 = None
        sw1, sw2, ss1, ss2 = (torch.randn(2 * I, H) / 10).bfloat16().to(dev), (torch.randn(H, I) / 10).bfloat16().to(dev), None, None
        fid = K.FMT_BF16
    elif fmt == "int8":
        w1 = torch.randint(-127, 128, (E, 2 * I, H), dtype=torch.int8).to(dev)
        w2 = torch.randint(-127, 128, (E, H, I), dtype=torch.int8).to(dev)
        s1, s2 = (torch.rand(E, 2 * I) * 1e-2).to(dev), (torch.rand(E, H) * 1e-2).to(dev)
        sw1 = torch.randint(-127, 128, (2 * I, H), dtype=torch.int8).to(dev)
        sw2 = torch.randint(-127, 128, (H, I), dtype=torch.int8).to(dev)
        ss1, ss2 = (torch.rand(2 * I) * 1e-2).to(dev), (torch.rand(H) * 1e-2).to(dev)
        fid = K.FMT_INT8_W8A8
    else:
        block = [128, 128]
        w1 = (torch.randn(E, 2 * I, H) * 100).clamp(-400, 400).to(torch.float8_e4m3fn).to(dev)
        w2 = (torch.randn(E, H, I) * 100).clamp(-400, 400).to(torch.float8_e4m3fn).to(dev)
        s1, s2 = (torch.randn(E, 2 * I // 128, H // 128) * 1e-3).to(dev), (torch.randn(E, H // 128, I // 128) * 1e-3).to(dev)
        sw1 = (torch.randn(2 * I, H) * 100).clamp(-400, 400).to(torch.float8_e4m3fn).to(dev)
        sw2 = (torch.randn(H, I) * 100).clamp(-400, 400).to(torch.float8_e4m3fn).to(dev)
        ss1, ss2 = (torch.randn(2 * I // 128, H // 128) * 1e-3).to(dev), (torch.randn(H // 128, I // 128) * 1e-3).to(dev)
        fid = K.FMT_FP8_W8A16
    path = K.PATH_UMMA
    xs, tws, idss = [], [], []
    for r in range(world):
        M = Ms[r]
        xs.append((torch.randn(M, H) / 10).bfloat16().to(dev))
        tw, ids = torch.topk(torch.softmax(torch.randn(M, E), dim=-1), topk)
        ids = ids.to(torch.int32)
        if M:
            ids[0, 0] = -1
            ids[M - 1, 1] = ids[M - 1, 2]
        tws.append(tw.to(dev)), idss.append(ids.to(dev))
    refs = []
    for r in range(world):
        ref = K.fused_experts_impl(xs[r], w1, w2, tws[r], idss[r], False, fid, s1, s2, block, False, path=path)
        if with_shared:
            ref = K.shared_expert_impl(xs[r], sw1, sw2, ref, 2.5, False, fid, ss1, ss2, block, False)
        refs.append(ref)
    torch.cuda.synchronize()
    shared = ThreadComm.Shared(world)
    outs, errs = [None] * world, []
    el = E // world

    def run(r):
        try:
            torch.cuda.set_device(0)
            local = dict(w1=w1[r * el:(r + 1) * el].contiguous(), w2=w2[r * el:(r + 1) * el].contiguous(),
                         s1=None if s1 is None else s1[r * el:(r + 1) * el].contiguous(),
                         s2=None if s2 is None else s2[r * el:(r + 1) * el].contiguous(), block=block)
            sh = dict(w1=sw1, w2=sw2, s1=ss1, s2=ss2, block=block) if with_shared else None
            layer = ep.EPMoE(local, fid, E, topk, comm=ThreadComm(shared, r), shared=sh, routed_scaling_factor=2.5, path=path)
            outs[r] = layer.forward(xs[r], tws[r], idss[r], mode=mode)
            assert layer.capacity == 64
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
        assert outs[r].shape == refs[r].shape
        assert torch.equal(outs[r], refs[r]), (r, float((outs[r].float() - refs[r].float()).abs().max()))
