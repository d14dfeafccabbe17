"""sglb_moe_align_static: the expert sort/align pass for callers that cannot allocate device memory.

Who needs it: a serving engine that handed every free byte of HBM to its own KV-cache / weight pools at start-up
(a later cudaMalloc simply fails), a plugin that runs inside somebody else's process and is not allowed to call
cudaMalloc or that process's caching allocator, and latency-critical steps where an allocation's implicit
device synchronisation is not acceptable.  sglb_moe_align makes such a caller bring a scratch buffer for every
batch above 1024 routed slots; this entry point keeps that scratch in a fixed 1 MiB array inside the library's
own device image instead, so the caller passes inputs and outputs only.

The test drives both entry points through the C ABI (ctypes, raw device pointers) on the same routing tables and
requires identical index arrays -- the tolerance of every moe_align test in this repo is bit equality.  It makes
the calls back to back on differently sized batches, because the array is shared by all calls and keeps the
previous call's histograms: a stale entry leaking into the next result is exactly the bug to catch.

This file is named so that pytest collects it last: under `pytest -x` a failure here hides nothing behind it.
It passes alone in a fresh process (the first call then also performs the per-device address look-up) and after
the rest of the suite.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

CHUNK_ROWS, TILE_ROWS = 8, 128
KEYS = ("expert_count", "expert_offset", "sorted_slot", "slot_pos", "chunk_list", "tile_list", "counters")


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from sgl_cpu_tests_b200 import _lib
    torch.cuda.set_device(0)
    torch.zeros(1, device="cuda")  # the device's context exists and is current before the first C-ABI call
    return _lib


def _outputs(L, lib, M, topk, E):
    """Caller-owned output arrays, sized as include/sglb200.h states.  The two work lists start zeroed: the
    kernels write only the items that exist, and the comparison below looks at whole arrays."""
    S = M * topk
    i32 = dict(dtype=torch.int32, device="cuda")
    return {
        "expert_count": torch.full((E,), -7, **i32),
        "expert_offset": torch.full((E + 1,), -7, **i32),
        "sorted_slot": torch.full((max(S, 1),), -7, **i32),
        "slot_pos": torch.full((max(S, 1),), -7, **i32),
        "chunk_list": torch.zeros(3 * max(1, L.sglb_moe_align_max_items(M, topk, E, CHUNK_ROWS)), **i32),
        "tile_list": torch.zeros(3 * max(1, L.sglb_moe_align_max_items(M, topk, E, TILE_ROWS)), **i32),
        "counters": torch.zeros(4, **i32),
    }


def _align(L, lib, ids, E, static):
    """One call through the C ABI on the current torch stream; returns the output arrays."""
    M, topk = ids.shape
    out = _outputs(L, lib, M, topk, E)
    ao = lib.AlignOut(**{k: v.data_ptr() for k, v in out.items()})
    stream = torch.cuda.current_stream().cuda_stream
    if static:
        rc = L.sglb_moe_align_static(ids.data_ptr(), M, topk, E, CHUNK_ROWS, TILE_ROWS, ctypes.byref(ao), stream)
    else:
        # the existing entry point, with the scratch buffer it asks its caller for (poisoned, not zeroed)
        scratch = torch.full((max(1, L.sglb_moe_align_scratch_bytes(M, topk, E)),), 0x5a, dtype=torch.uint8, device="cuda")
        rc = L.sglb_moe_align(ids.data_ptr(), M, topk, E, CHUNK_ROWS, TILE_ROWS, ctypes.byref(ao), scratch.data_ptr(),
                              stream)
    lib.check(rc)
    torch.cuda.synchronize()
    return out


def _routing(M, topk, E, seed):
    g = torch.Generator().manual_seed(seed)
    # -1 = masked slot (test_moe_offloading_cpu.py:57-68 in the reference), repeats inside a row allowed
    return torch.randint(-1, E, (M, topk), dtype=torch.int32, generator=g).cuda()


def test_static_scratch_align_matches_caller_scratch_align_on_consecutive_calls(lib):
    L = lib.load()
    assert L.sglb_moe_align_static_max_bytes() == 1 << 20
    # (tokens, topk, experts): a large batch first, then smaller ones that reuse the front of the array while the
    # large call's histograms are still lying behind it, then the large one again; 3000 x 8 x 8 and 8192 x 8 x 256
    # use several histogram blocks, 100 x 8 x 8 none at all (one block: the scratch is not read)
    calls = [(8192, 8, 256, 1), (3000, 8, 8, 2), (100, 8, 8, 3), (1024, 8, 256, 4), (8192, 8, 256, 5)]
    for M, topk, E, seed in calls:
        ids = _routing(M, topk, E, seed)
        want = _align(L, lib, ids, E, static=False)
        got = _align(L, lib, ids, E, static=True)
        for k in KEYS:
            assert torch.equal(got[k], want[k]), (k, M, topk, E)
    # the same batch twice in a row gives the same answer, and the second call takes no device memory
    ids = _routing(4096, 8, 128, 6)
    first = _align(L, lib, ids, 128, static=True)
    out = _outputs(L, lib, 4096, 8, 128)
    ao = lib.AlignOut(**{k: v.data_ptr() for k, v in out.items()})
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    lib.check(L.sglb_moe_align_static(ids.data_ptr(), 4096, 8, 128, CHUNK_ROWS, TILE_ROWS, ctypes.byref(ao),
                                      torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free_before, "the no-allocation entry point took device memory"
    for k in KEYS:
        assert torch.equal(out[k], first[k]), k


def test_static_scratch_align_refuses_batches_beyond_its_buffer(lib):
    L = lib.load()
    E, topk = 1024, 8
    M = 40000  # ceil(320000 / 1024) * 1024 experts * 4 B = 1.25 MiB of histograms > 1 MiB
    assert L.sglb_moe_align_scratch_bytes(M, topk, E) > L.sglb_moe_align_static_max_bytes()
    ids = torch.zeros(M, topk, dtype=torch.int32, device="cuda")
    out = _outputs(L, lib, M, topk, E)
    ao = lib.AlignOut(**{k: v.data_ptr() for k, v in out.items()})
    rc = L.sglb_moe_align_static(ids.data_ptr(), M, topk, E, CHUNK_ROWS, TILE_ROWS, ctypes.byref(ao),
                                 torch.cuda.current_stream().cuda_stream)
    assert rc == -1  # SGLB_ERR_INVALID
    msg = L.sglb_last_error().decode()
    assert "built-in buffer" in msg and "sglb_moe_align" in msg, msg
    with pytest.raises(RuntimeError, match="built-in buffer"):
        lib.check(rc)
    torch.cuda.synchronize()
    assert int(out["counters"].sum()) == 0  # nothing was launched
