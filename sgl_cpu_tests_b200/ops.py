"""Host-side mirror of the reference operator interface for the fused-MoE path.

Same op names, positional argument lists and error behaviour as the ops the reference scripts
resolve through ``torch.ops.sgl_kernel`` / ``sgl_kernel.common_ops``; every
function here ends in a call into libsglb200.so.  PyTorch is used for memory, streams and dtype
bookkeeping only -- there is no torch compute on this path and no CPU fallback.

Tensors may live on the CPU (what the reference scripts allocate) or on a CUDA device:
  * CUDA tensors are used in place.
  * CPU activations / routing tensors are staged to the current CUDA device for the call and the
    result is copied back (into ``hidden_states`` itself when ``inplace=True``).
  * CPU weights are uploaded once and kept resident in HBM in a registry keyed on the host
    buffer (pointer, shape, dtype, version, sampled fingerprint), so later calls move no weights.
"""
import math
import os
import sys
import threading
import weakref
from enum import IntEnum

import torch

from . import _lib as L


class CPUQuantMethod(IntEnum):  # test_moe.py:4,86 (sglang.srt.layers.amx_utils.CPUQuantMethod)
    UNQUANT = 0
    INT8_W8A8 = 1
    FP8_W8A16 = 2
    INT4_W4A8 = 3
    MXFP4 = 4


_FP8 = torch.float8_e4m3fn
_DT = {torch.bfloat16: L.DT_BF16, torch.float32: L.DT_F32, torch.float16: L.DT_F16}


def _device():
    if not torch.cuda.is_available():
        raise RuntimeError("sgl_cpu_tests_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


# ---- large pageable host buffers that keep coming back are page-locked in place -------------------------
# The reference's scripts allocate ordinary (pageable) CPU tensors and reuse them call after call (bench_moe.py:52
# rotates `inputs[idx]`); a 56 MB activation (M = 3929, H = 7168) crosses PCIe at ~8 GB/s from pageable memory and at
# ~25 GB/s once registered.  A contiguous host tensor of at least _PIN_MIN_BYTES seen for the second time at the same
# address is registered with cudaHostRegister and unregistered when the tensor object dies.  SGLB_PIN_HOST=0 disables.
# The registration goes through libsglb200 (sglb_host_register): a refused range -- a read-only or file-backed mapping
# such as an mmap'ed checkpoint, the locked-memory limit -- is then cleared inside the library instead of staying
# pending in torch's runtime, where the next launch check would report it as an unrelated CUDA error.  At most
# SGLB_PIN_HOST_GB (default 8) of caller memory are page-locked at a time; refused ranges are not tried again.
_PIN_HOST = os.environ.get("SGLB_PIN_HOST", "1") != "0"
_PIN_MIN_BYTES = 4 << 20
_PIN_MAX_BYTES = int(float(os.environ.get("SGLB_PIN_HOST_GB", "8")) * (1 << 30))
_pin_seen = {}
_pinned = {}
_pin_refused = set()


def _unpin(ptr):
    if _pinned.pop(ptr, None) is not None:
        try:
            L.load().sglb_host_unregister(ptr)
        except Exception:  # interpreter / context shutdown
            pass


def _maybe_pin(t):
    if not _PIN_HOST or t is None or t.is_cuda or not t.is_contiguous():
        return
    n = _nbytes(t)
    if n < _PIN_MIN_BYTES:
        return
    ptr = t.data_ptr()
    if ptr in _pinned or (ptr, n) in _pin_refused or sum(_pinned.values()) + n > _PIN_MAX_BYTES:
        return
    seen = _pin_seen.get((ptr, n), 0) + 1
    if seen < 2:
        if len(_pin_seen) > 256:
            _pin_seen.clear()
        _pin_seen[(ptr, n)] = seen
        return
    _pin_seen.pop((ptr, n), None)
    if t.is_pinned():
        return
    if L.load().sglb_host_register(ptr, n) == L.OK:
        _pinned[ptr] = n
        weakref.finalize(t, _unpin, ptr)
    else:
        if len(_pin_refused) > 256:
            _pin_refused.clear()
        _pin_refused.add((ptr, n))


# ---- device shadows of the last host-tensor fused_experts call ---------------------------------------------------
# The reference's API hands tensors from op to op through host memory: shared_expert_cpu(hidden_states, ...,
# fused_experts_out, ...) right after fused_experts_cpu(hidden_states, ...) (test_moe_fp8_ext.py:60-61,118).  Both of
# those tensors were in HBM a moment ago -- hidden_states uploaded, the result produced there -- so the second op need
# not upload them again.  The last host-tensor fused_experts call leaves (host tensor -> device tensor) for its input
# and its result; a hit requires the SAME tensor object (weak reference), an unchanged torch version counter, the
# same Python thread and the same CUDA stream, and every new host-tensor fused_experts call drops both entries
# before it touches the (static, graph-owned) device buffers they point to.  What it cannot see, as for the weight
# registry: writes that bypass the version counter (numpy views, `t.data.copy_()`) between the two ops.
# SGLB_SHADOW=0 disables.
_SHADOW = os.environ.get("SGLB_SHADOW", "1") != "0"
_shadow = {}
shadow_hits = 0   # for tests / bench


def _shadow_put(kind, host_t, dev_t):
    if _SHADOW and host_t is not None and not host_t.is_cuda and host_t.is_contiguous():
        _shadow[kind] = (weakref.ref(host_t), host_t._version, dev_t, threading.get_ident(), _stream())


def _shadow_get(host_t, dev):
    """The device twin of `host_t` left by the last host-tensor fused_experts call, or None.  Call under _launch_lock
    and keep it until the kernels reading the twin are enqueued."""
    global shadow_hits
    if not _shadow or host_t is None or host_t.is_cuda:
        return None
    for ref, ver, d, tid, st in list(_shadow.values()):
        if (ref() is host_t and host_t._version == ver and tid == threading.get_ident() and st == _stream()
                and d.device == dev and d.shape == host_t.shape and d.dtype == host_t.dtype):
            shadow_hits += 1
            return d
    return None


_copy_streams = {}


def _copy_stream(dev):
    st = _copy_streams.get(dev.index)
    if st is None:
        st = _copy_streams[dev.index] = torch.cuda.Stream(device=dev)
    return st


def _to_host(dev_tensor):
    """Device result -> a fresh host tensor; large results land in torch's cached pinned pool (DMA at full rate)."""
    if _PIN_HOST and _nbytes(dev_tensor) >= _PIN_MIN_BYTES:
        h = torch.empty(dev_tensor.shape, dtype=dev_tensor.dtype, pin_memory=True)
        h.copy_(dev_tensor)   # blocking
        return h
    return dev_tensor.cpu()


# ---- first upload of a large pageable host tensor (weights): pipelined through two pinned staging buffers ------
# `tensor.to(device)` from pageable memory runs at ~8 GB/s (the driver stages through a small internal buffer on one
# thread); a multi-threaded memcpy into pinned memory (torch's parallel CPU copy) overlapped with the DMA of the previous
# chunk reaches PCIe rate.  DeepSeek-V3's 11.3 GB layer, or the 84 GB of weight clones bench_moe.py touches once.
_STAGE_BYTES = 64 << 20
_UPLOAD_MIN_BYTES = 256 << 20
_stage = {}


def _upload_staged(t, dev):
    src = t.detach().contiguous().view(torch.uint8).reshape(-1)
    n = src.numel()
    dst = torch.empty(t.shape, dtype=t.dtype, device=dev)
    dflat = dst.view(torch.uint8).reshape(-1)
    if "buf" not in _stage:
        _stage["buf"] = [torch.empty(_STAGE_BYTES, dtype=torch.uint8, pin_memory=True) for _ in range(2)]
        _stage["ev"] = [None, None]
    bufs, evs = _stage["buf"], _stage["ev"]
    with _launch_lock:
        for i, o in enumerate(range(0, n, _STAGE_BYTES)):
            m = min(_STAGE_BYTES, n - o)
            b = i & 1
            if evs[b] is not None:
                evs[b].synchronize()      # the DMA that read this staging buffer has finished
            bufs[b][:m].copy_(src[o:o + m])
            dflat[o:o + m].copy_(bufs[b][:m], non_blocking=True)
            evs[b] = torch.cuda.Event()
            evs[b].record()
        for e in evs:
            if e is not None:
                e.synchronize()
    return dst


def _to_dev(t, dev, dtype=None):
    if t is None:
        return None
    if t.device != dev:
        if (_PIN_HOST and t.device.type == "cpu" and _nbytes(t) >= _UPLOAD_MIN_BYTES and t.is_contiguous()
                and not t.is_pinned()):
            t = _upload_staged(t, dev)
            if dtype is not None and t.dtype != dtype:
                t = t.to(dtype)
            return t
        _maybe_pin(t)
        # pinned host memory: stream-ordered async copy (every entry point ends with a blocking D2H of
        # its result or returns device tensors ordered on the same stream, so the source stays valid)
        t = t.to(dev, non_blocking=(t.device.type == "cpu" and t.is_pinned()))
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


# ---- workspace: one growing buffer per (device, stream) ---------------------------------------------
# Calls that share a stream share its workspace; that is safe because their kernels are stream-ordered,
# provided the launches of one call are not interleaved with another thread's.  _launch_lock is held
# from taking the workspace until the (asynchronous) C call has enqueued everything.
_workspaces = {}
_WORKSPACE_ENTRIES = 16   # streams come and go (side streams, captures): keep the most recently used ones
_launch_lock = threading.RLock()


def _workspace(nbytes, dev):
    key = (dev.index, _stream())
    buf = _workspaces.pop(key, None)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=dev)
    _workspaces[key] = buf            # most recently used last
    while len(_workspaces) > _WORKSPACE_ENTRIES:
        _workspaces.pop(next(iter(_workspaces)))   # stream-ordered frees: torch keeps the memory alive until the stream drains
    return buf


# ---- weight registry ------------------------------------------------------------------------------------
# Host weights (what the reference scripts pass) are uploaded once and served from HBM afterwards.  An entry is
# keyed on (device, pointer, shape, strides, dtype, packing, kind, torch version counter):
#   * host tensors: the entry keeps a reference to the tensor, so the address cannot be recycled under another weight
#     while the entry lives; in-place updates through torch bump the version counter and miss;
#   * CUDA tensors (packed weights returned by convert_weight_packed, scale tiles): nothing is pinned -- a weakref
#     finaliser drops the entry when the keyed tensor dies, so a reallocated address cannot hit a stale entry.
# What the key cannot see: writes that bypass torch's version counter -- `w.data.copy_(...)` (the usual weight-loader /
# weight-sync idiom), numpy views, torch.from_numpy buffers.  For those callers: sgl_cpu_tests_b200.invalidate(w) after
# the update, or SGLB_VERIFY_WEIGHTS=1, which fingerprints 64 sampled cache lines of the host tensor at every lookup
# (~20 us per tensor) and re-uploads on a mismatch.  LRU eviction under two caps: device bytes (SGLB_WEIGHT_CACHE_GB,
# default 140) and the host bytes the entries keep alive (SGLB_WEIGHT_CACHE_HOST_GB, default 160).
_WEIGHT_CACHE_BYTES = int(float(os.environ.get("SGLB_WEIGHT_CACHE_GB", "120")) * (1 << 30))  # of 180 GB: workspaces, graph pools and the caller need the rest
_WEIGHT_CACHE_HOST_BYTES = int(float(os.environ.get("SGLB_WEIGHT_CACHE_HOST_GB", "160")) * (1 << 30))
_VERIFY_WEIGHTS = os.environ.get("SGLB_VERIFY_WEIGHTS", "0") != "0"
_registry = {}          # key -> (keep-alive / fingerprint info, device tensor, host bytes)
_registry_bytes = 0
_registry_host_bytes = 0


def _nbytes(t):
    return t.numel() * t.element_size()


def _sample_hash(t):
    """64 cache lines spread over a contiguous host tensor's bytes."""
    b = t.detach().contiguous().view(torch.uint8).reshape(-1)
    n = b.numel()
    if n <= 64 * 64:
        return hash(b.numpy().tobytes())
    idx = (torch.arange(64, dtype=torch.int64) * ((n - 64) // 63)).unsqueeze(1) + torch.arange(64, dtype=torch.int64)
    return hash(b[idx.reshape(-1)].numpy().tobytes())


def _fingerprint(t):
    """What a registry entry keeps of its key tensor: (the host tensor itself | None for CUDA, sampled hash | None)."""
    if t.is_cuda:
        return (None, None)
    return (t, _sample_hash(t) if _VERIFY_WEIGHTS else None)


def _registry_key(t, is_vnni, kind):
    return (t.device.type, t.device.index, t.data_ptr(), tuple(t.shape), tuple(t.stride()), t.dtype, bool(is_vnni), kind,
            t._version)


def _registry_drop(key):
    global _registry_bytes, _registry_host_bytes
    ent = _registry.pop(key, None)
    if ent is not None:
        _registry_bytes -= _nbytes(ent[1])
        _registry_host_bytes -= ent[2]
        # a cached host-tensor graph keeps the device tensors it was captured over alive: evicting a weight must take
        # its graphs along, or the cap above bounds nothing (bench_moe.py: 60 weight sets, 169 GB held by graphs)
        ptr = ent[1].data_ptr()
        for gk in [gk for gk in _host_graphs if ptr in gk[3:7]]:
            _host_graphs.pop(gk, None)


def _registry_get(key, t=None):
    ent = _registry.get(key)
    if ent is None:
        return None
    if _VERIFY_WEIGHTS and t is not None and ent[0][1] is not None and _sample_hash(t) != ent[0][1]:
        _registry_drop(key)       # rewritten behind torch's version counter
        return None
    _registry[key] = _registry.pop(key)   # most recently used last
    return ent[1]


def _registry_put(key, fp, dev_tensor, key_tensor=None):
    global _registry_bytes, _registry_host_bytes
    _registry_drop(key)           # overwriting: the old entry's bytes must not be counted twice
    nbytes = _nbytes(dev_tensor)
    host = _nbytes(fp[0]) if (isinstance(fp, tuple) and fp[0] is not None) else 0
    while _registry and (_registry_bytes + nbytes > _WEIGHT_CACHE_BYTES or
                         _registry_host_bytes + host > _WEIGHT_CACHE_HOST_BYTES):
        _registry_drop(next(iter(_registry)))   # least recently used first
    _registry[key] = (fp if isinstance(fp, tuple) else (None, None), dev_tensor, host)
    _registry_bytes += nbytes
    _registry_host_bytes += host
    if key_tensor is not None and key_tensor.is_cuda:
        weakref.finalize(key_tensor, _registry_drop, key)   # never serve a freed CUDA buffer's address again


def invalidate(tensor=None):
    """Forget the HBM-resident copies made for `tensor` (a host weight / scale that was rewritten in place through
    `.data`, numpy or another path torch's version counter does not see); None forgets everything."""
    if tensor is None:
        return clear_weight_cache()
    for key in [k for k in _registry if k[0] == tensor.device.type and k[2] == tensor.data_ptr()]:
        _registry_drop(key)
    _host_graphs.clear()


def clear_weight_cache():
    global _registry_bytes, _registry_host_bytes
    _shadow.clear()
    _host_graphs.clear()
    _host_graph_seen.clear()
    _registry.clear()
    _registry_bytes = 0
    _registry_host_bytes = 0


def _rows_k(w, is_int8_packed):
    """[..., N, K] geometry; an int8 packed tensor carries 4 compensation bytes per row."""
    N, K = w.shape[-2], w.shape[-1]
    if is_int8_packed:
        K -= 4
    E = 1
    for d in w.shape[:-2]:
        E *= d
    return E, N, K


def _repack(src, dst, elem_bytes, vnni, E, N, K, comp, to_packed):
    L.check(L.load().sglb_repack_weight(src.data_ptr(), dst.data_ptr(), elem_bytes, vnni, E, N, K,
                                        int(comp), int(to_packed), _stream()))


def _layout_of(dtype):
    """(elem_bytes, vnni, has_compensation) of convert_weight_packed for a weight dtype."""
    if dtype in (torch.bfloat16, torch.float16):
        return 2, 2, False
    if dtype == torch.int8:
        return 1, 4, True
    if dtype in (_FP8, torch.uint8):
        return 1, 4, False
    raise RuntimeError(f"convert_weight_packed: unsupported weight dtype {dtype}")


def _resident_weight(w, is_vnni, dev):
    """Row-major, device-resident view of a weight the caller passed (CPU or CUDA, packed or not)."""
    if w.is_cuda and not is_vnni:
        return w if w.is_contiguous() else w.contiguous()
    key = _registry_key(w, is_vnni, "w")
    hit = _registry_get(key, w)
    if hit is not None:
        return hit
    fp = _fingerprint(w)
    wd = _to_dev(w, dev)
    if is_vnni:
        eb, vnni, comp = _layout_of(w.dtype)
        E, N, K = _rows_k(w, comp)
        out_shape = list(w.shape)
        out_shape[-1] = K
        rm = torch.empty(out_shape, dtype=w.dtype, device=dev)
        _repack(wd, rm, eb, vnni, E, N, K, comp, False)
        wd = rm
    _registry_put(key, fp, wd, w)
    return wd


def _resident_aux(t, dev, dtype, kind="s"):
    """Scales / bias: tiny next to the weights but cached the same way when they come from the CPU."""
    if t is None:
        return None
    if t.is_cuda:
        return _to_dev(t, dev, dtype)
    key = _registry_key(t, False, kind)
    hit = _registry_get(key, t)
    if hit is not None:
        return hit
    td = _to_dev(t, dev, dtype)
    _registry_put(key, _fingerprint(t), td, t)
    return td


# ---- K7: convert_weight_packed / convert_scale_packed --------------------------------------------------
def convert_weight_packed(weight):
    """convert_weight_packed(Tensor) -> Tensor   (test_moe.py:75-76, test_mxfp4.py:159).

    Returns the reference's VNNI tile layout (bit-identical where the reference pins it: bf16
    test_gemm.py:38-44 / test_moe.py:117-124, int8 + compensation test_gemm_int8.py:84-93) on the
    device of the input, and registers the row-major HBM-resident twin the sm_100a kernels read,
    so a later fused_experts(..., is_vnni=True) on the returned tensor moves no weight bytes.
    """
    if weight.dim() not in (2, 3):
        raise RuntimeError(f"convert_weight_packed: expect a 2-D or 3-D weight, got {weight.dim()}-D")
    dev = _device() if not weight.is_cuda else weight.device
    eb, vnni, comp = _layout_of(weight.dtype)
    E, N, K = _rows_k(weight, False)
    if N % 32 != 0:
        raise RuntimeError(f"convert_weight_packed: N={N} must be a multiple of 32")
    if K % vnni != 0:
        raise RuntimeError(f"convert_weight_packed: K={K} must be a multiple of {vnni}")
    with torch.cuda.device(dev):
        wd = _to_dev(weight, dev)
        out_shape = list(weight.shape)
        if comp:
            out_shape[-1] = K + 4
        packed = torch.empty(out_shape, dtype=weight.dtype, device=dev)
        _repack(wd, packed, eb, vnni, E, N, K, comp, True)
        result = packed if weight.is_cuda else packed.cpu()
        _registry_put(_registry_key(result, True, "w"), _fingerprint(result), wd, result)
    return result


def convert_scale_packed(scale):
    """convert_scale_packed(Tensor) -> Tensor   (test_mxfp4.py:160,182-185): e8m0 scales
    [.., N, K/32] u8 -> [.., N/32, K/32, 32] (flat order), same shape and device as the input."""
    if scale.dtype != torch.uint8 or scale.dim() not in (2, 3):
        raise RuntimeError("convert_scale_packed: expect a 2-D or 3-D uint8 (e8m0) tensor")
    dev = _device() if not scale.is_cuda else scale.device
    E, N, G = _rows_k(scale, False)
    if N % 32 != 0:
        raise RuntimeError(f"convert_scale_packed: N={N} must be a multiple of 32")
    with torch.cuda.device(dev):
        sd = _to_dev(scale, dev)
        packed = torch.empty_like(sd)
        L.check(L.load().sglb_repack_scale(sd.data_ptr(), packed.data_ptr(), E, N, G, 1, _stream()))
        result = packed if scale.is_cuda else packed.cpu()
        _registry_put(_registry_key(result, True, "ms"), _fingerprint(result), sd, result)
    return result


def _resident_mx_scale(s, is_vnni, dev):
    """Row-major device copy of an e8m0 scale tensor (packed by convert_scale_packed or not)."""
    if s.is_cuda and not is_vnni:
        return s.contiguous()
    key = _registry_key(s, is_vnni, "ms")
    hit = _registry_get(key, s)
    if hit is not None:
        return hit
    fp = _fingerprint(s)
    sd = _to_dev(s, dev)
    if is_vnni:
        E, N, G = _rows_k(s, False)
        rm = torch.empty_like(sd)
        L.check(L.load().sglb_repack_scale(sd.data_ptr(), rm.data_ptr(), E, N, G, 0, _stream()))
        sd = rm
    _registry_put(key, fp, sd, s)
    return sd


# A CUDA graph bakes in the addresses of everything its kernels touch, including the registry-owned tensors DERIVED
# from a weight (mxfp4 scale tiles, padded mxfp4 weights).  While a graph is being captured the helpers below hand
# every tensor they return to this sink, and the graph object keeps them alive: the registry may evict the entry,
# the memory stays valid for as long as a graph can replay over it.
_capture_sink = None


def _note_captured(t):
    if _capture_sink is not None and t is not None:
        _capture_sink.append(t)
    return t


def _mx4_tiles(s_dev, shape, split):
    """e8m0 scales [E, rows, K/32] on the device -> the 512-byte tiles the block-scaled tensor-core kernel
    copies into TMEM (csrc/umma_mx4.cu); built once per scale tensor and cached.  Keyed on (and dropped with) the
    resident scale tensor itself, not on the reshaped view of it made here: a temporary view dies with the call and
    would take the entry with it (found in the gpt-oss launch list: two re-tiling launches, 137 us, per step)."""
    key = _registry_key(s_dev, False, ("mx4tiles", tuple(shape), split))
    hit = _registry_get(key)
    if hit is not None:
        return _note_captured(hit)
    E, rows, G = shape
    lib = L.load()
    tiles = torch.empty(lib.sglb_mx4_scale_tiles_bytes(E, rows, split, 32 * G), dtype=torch.uint8, device=s_dev.device)
    L.check(lib.sglb_mx4_tile_scales(s_dev.data_ptr(), tiles.data_ptr(), E, rows, split, 32 * G, _stream()))
    _registry_put(key, (None, None), tiles, s_dev)   # dropped when the keyed scale tensor dies
    return _note_captured(tiles)


def _mx4_weight(w_dev):
    """The weight itself when the 4-bit TMA boxes may read it (32-byte aligned; 32 readable bytes past
    the end when K is not a multiple of 128), else a padded device copy (cached); None if unusable."""
    K = 2 * w_dev.shape[-1]
    if K % 64 != 0:
        return None
    if w_dev.data_ptr() % 32 == 0:
        st = w_dev.untyped_storage()
        slack = st.nbytes() - (w_dev.storage_offset() + w_dev.numel())
        if K % 128 == 0 or slack >= 32:
            return w_dev
    key = _registry_key(w_dev, False, "mx4pad")
    hit = _registry_get(key)
    if hit is not None:
        return _note_captured(hit)
    buf = torch.zeros(w_dev.numel() + 64, dtype=torch.uint8, device=w_dev.device)
    padded = buf[:w_dev.numel()].view(w_dev.shape)
    padded.copy_(w_dev)
    _registry_put(key, (None, None), padded, w_dev)
    return _note_captured(padded)


# ---- K1: routers -----------------------------------------------------------------------------------------
def _router(gating_output, correction_bias, topk, renormalize, num_expert_group, topk_group,
            num_fused_shared_experts=0, routed_scaling_factor=None, num_token_non_padded=None,
            out_w=None, out_ids=None):
    if gating_output.dim() != 2:
        raise RuntimeError("grouped_topk: gating_output must be [num_tokens, num_experts]")
    if gating_output.dtype not in _DT:
        raise RuntimeError(f"grouped_topk: unsupported gating dtype {gating_output.dtype}")
    M, E = gating_output.shape
    n_shared = int(num_fused_shared_experts or 0)
    dev = gating_output.device if gating_output.is_cuda else _device()
    with torch.cuda.device(dev):
        g = _to_dev(gating_output, dev)
        b = _to_dev(correction_bias, dev) if correction_bias is not None else None
        if b is not None and (b.dim() != 1 or b.numel() != E or b.dtype not in _DT):
            raise RuntimeError("biased_grouped_topk: correction_bias must be [num_experts] bf16/f16/f32")
        npad, npad64 = None, 0
        if num_token_non_padded is not None:
            npad = num_token_non_padded
            if not torch.is_tensor(npad):
                npad = torch.tensor([int(npad)], dtype=torch.int32)
            if npad.numel() != 1 or npad.dtype not in (torch.int32, torch.int64):
                raise RuntimeError("grouped_topk: num_token_non_padded must hold one int32 / int64 value")
            npad = _to_dev(npad, dev)
            npad64 = int(npad.dtype == torch.int64)
        def _direct(t, dtype):   # an out tensor the kernel can write in place
            return (t is not None and t.is_cuda and t.device == dev and t.dtype == dtype and t.is_contiguous()
                    and tuple(t.shape) == (M, topk))
        direct = _direct(out_w, torch.float32) and _direct(out_ids, torch.int32)
        w = out_w if direct else torch.empty(M, topk, dtype=torch.float32, device=dev)
        ids = out_ids if direct else torch.empty(M, topk, dtype=torch.int32, device=dev)
        L.check(L.load().sglb_grouped_topk(
            g.data_ptr(), _DT[g.dtype], _ptr(b), _DT[b.dtype] if b is not None else 0, M, E, int(topk),
            int(num_expert_group), int(topk_group), int(bool(renormalize)), n_shared,
            float(1.0 if routed_scaling_factor is None else routed_scaling_factor), _ptr(npad), npad64,
            w.data_ptr(), ids.data_ptr(), _stream()))
    if out_w is not None:
        if not direct:
            out_w.copy_(w)
            out_ids.copy_(ids)
        return None
    if not gating_output.is_cuda:
        return w.cpu(), ids.cpu()
    return w, ids


def _check_tokens(hidden_states, gating_output):
    if hidden_states.shape[0] != gating_output.shape[0]:
        raise RuntimeError("Number of tokens mismatch")  # test_grouped_topk.py:18


def grouped_topk(hidden_states, gating_output, topk, renormalize, num_expert_group, topk_group,
                 num_fused_shared_experts=0, routed_scaling_factor=None, num_token_non_padded=None):
    """grouped_topk_cpu, 9-arg functional form (test_moe.py:61-70) -> (topk_weights f32, topk_ids i32).
    The last three arguments follow sglang's python router (csrc/router.cu header); the reference passes
    0, None, None."""
    _check_tokens(hidden_states, gating_output)
    return _router(gating_output, None, topk, renormalize, num_expert_group, topk_group,
                   num_fused_shared_experts, routed_scaling_factor, num_token_non_padded)


def grouped_topk_out(topk_weights, topk_ids, hidden_states, gating_output, topk, renormalize,
                     num_expert_group, topk_group):
    """grouped_topk_cpu, 8-arg out-variant (test_grouped_topk.py:57-69) -> None."""
    _check_tokens(hidden_states, gating_output)
    return _router(gating_output, None, topk, renormalize, num_expert_group, topk_group,
                   out_w=topk_weights, out_ids=topk_ids)


def biased_grouped_topk(hidden_states, gating_output, correction_bias, topk, renormalize,
                        num_expert_group, topk_group, num_fused_shared_experts=0,
                        routed_scaling_factor=None, num_token_non_padded=None):
    """biased_grouped_topk_cpu, functional form (by analogy with grouped_topk_cpu)."""
    _check_tokens(hidden_states, gating_output)
    return _router(gating_output, correction_bias, topk, renormalize, num_expert_group, topk_group,
                   num_fused_shared_experts, routed_scaling_factor, num_token_non_padded)


def biased_grouped_topk_out(topk_weights, topk_ids, hidden_states, gating_output, correction_bias, topk,
                            renormalize, num_expert_group, topk_group):
    """biased_grouped_topk_cpu, 9-arg out-variant (test_biased_grouped_topk.py:67-80) -> None."""
    _check_tokens(hidden_states, gating_output)
    return _router(gating_output, correction_bias, topk, renormalize, num_expert_group, topk_group,
                   out_w=topk_weights, out_ids=topk_ids)


# ---- K2: moe_align (exposed for tests and the expert-parallel layer) -------------------------------------
def moe_align(topk_ids, num_experts, chunk_rows=8, tile_rows=128):
    """Returns a dict of int32 CUDA tensors; layout documented in csrc/align.cu."""
    dev = topk_ids.device if topk_ids.is_cuda else _device()
    with torch.cuda.device(dev):
        ids = _to_dev(topk_ids, dev, torch.int32)
        M, topk = ids.shape
        S = M * topk
        lib = L.load()
        i32 = dict(dtype=torch.int32, device=dev)
        out = {
            "expert_count": torch.empty(num_experts, **i32),
            "expert_offset": torch.empty(num_experts + 1, **i32),
            "sorted_slot": torch.empty(max(S, 1), **i32),
            "slot_pos": torch.empty(max(S, 1), **i32),
            "chunk_list": torch.zeros(3 * max(1, lib.sglb_moe_align_max_items(M, topk, num_experts, chunk_rows)), **i32),
            "tile_list": torch.zeros(3 * max(1, lib.sglb_moe_align_max_items(M, topk, num_experts, tile_rows)), **i32),
            "counters": torch.zeros(4, **i32),
        }
        scratch = torch.empty(max(1, lib.sglb_moe_align_scratch_bytes(M, topk, num_experts)), dtype=torch.uint8, device=dev)
        ao = L.AlignOut(**{k: v.data_ptr() for k, v in out.items()})
        L.check(lib.sglb_moe_align(ids.data_ptr(), M, topk, num_experts, chunk_rows, tile_rows, ao,
                                   scratch.data_ptr(), _stream()))
        out["sorted_slot"] = out["sorted_slot"][:S]
        out["slot_pos"] = out["slot_pos"][:S]
    return out


# ---- K3-K5: fused_experts ---------------------------------------------------------------------------------
def _weight_geometry(w1, w2, fmt, is_vnni):
    if w1.dim() != 3 or w2.dim() != 3:
        raise RuntimeError("fused_experts: w1 and w2 must be 3-D [E, N, K]")
    int8_packed = is_vnni and fmt == L.FMT_INT8_W8A8
    E, N2, K1 = _rows_k(w1, int8_packed)
    E2, H, K2 = _rows_k(w2, int8_packed)
    if fmt == L.FMT_MXFP4:
        K1, K2 = 2 * K1, 2 * K2
    if E != E2 or N2 % 2 != 0 or K1 != H or K2 * 2 != N2:
        raise RuntimeError(f"fused_experts: inconsistent weight shapes w1={tuple(w1.shape)} w2={tuple(w2.shape)}")
    return E, H, N2 // 2


_WDT = {L.FMT_BF16: (torch.bfloat16,), L.FMT_INT8_W8A8: (torch.int8,), L.FMT_FP8_W8A16: (_FP8,),
        L.FMT_MXFP4: (torch.uint8,)}


# ---- host-buffer calls replay a cached CUDA graph ------------------------------------------------------------
# The reference's scripts pass CPU tensors.  For those calls the critical path is host time until the first kernel
# is launched + the kernels + the blocking D2H; a call over static device buffers captured once per (token count,
# weights) needs three H2D copies and ONE launch instead of six (K.GraphedExperts).  SGLB_HOST_GRAPH=0 disables it.
_HOST_GRAPH = os.environ.get("SGLB_HOST_GRAPH", "1") != "0"
_HOST_GRAPH_MAX_TOKENS = 1024
_HOST_GRAPH_ENTRIES = 256          # graphs kept (least recently used first out) ...
_HOST_GRAPH_BYTES = 2 << 30        # ... as long as their static buffers + workspaces stay under this many bytes.
# (Round 2 kept 4: the reference's bench_moe.py rotates 20 weight clones at M = 4, every call missed, re-captured and
# evicted -- 4.5 ms per call instead of 0.2 ms.)
_host_graphs = {}  # key -> GraphedExperts (insertion order = LRU order), or None after a failed capture
_host_graph_seen = {}   # key -> calls seen before a capture: a caller whose batch size changes every call never captures
_host_graph_warned = [False]


def _host_graph(M, topk, fmt, w13, w2d, s13, s2, block_size, path):
    key = (M, topk, fmt, w13.data_ptr(), w2d.data_ptr(), _ptr(s13), _ptr(s2),
           tuple(block_size) if block_size is not None else None, path, w13.device.index)
    if key in _host_graphs:
        g = _host_graphs.pop(key)
        _host_graphs[key] = g  # most recently used last
        return g
    # a capture costs two extra calls and two synchronisations: only for a key that comes back
    seen = _host_graph_seen.get(key, 0) + 1
    if seen < 2:
        if len(_host_graph_seen) > 64:
            _host_graph_seen.clear()
        _host_graph_seen[key] = seen
        return None
    _host_graph_seen.pop(key, None)
    try:
        g = GraphedExperts(M, w13, w2d, topk, fmt, s13, s2, block_size, False, path)
    except Exception as e:  # capture refused: this key stays on the eager path
        torch.cuda.synchronize()
        if not _host_graph_warned[0]:
            _host_graph_warned[0] = True
            sys.stderr.write(f"[sgl_cpu_tests_b200] CUDA-graph capture of a host-tensor fused_experts call failed "
                             f"({type(e).__name__}: {e}); such calls stay on eager launches\n")
        g = None
    _host_graphs[key] = g

    def held(x):
        return 0 if x is None else x._ws.numel() + 2 * _nbytes(x.hidden) + _nbytes(x.topk_weights) + _nbytes(x.topk_ids)

    while len(_host_graphs) > 1 and (len(_host_graphs) > _HOST_GRAPH_ENTRIES or
                                      sum(held(x) for x in _host_graphs.values()) > _HOST_GRAPH_BYTES):
        _host_graphs.pop(next(iter(_host_graphs)))
    return g


def fused_experts_impl(hidden_states, w1, w2, topk_weights, topk_ids, inplace, fmt, w1_scale, w2_scale,
                       block_size, is_vnni, path=L.PATH_AUTO, out=None, w1_bias=None, w2_bias=None, alpha=None,
                       limit=None, slot_out=None, slot_map=None, rows_per_expert_hint=0.0, shared=None,
                       routed_scaling_factor=1.0):
    """Format-explicit core behind every fused_experts_cpu signature vintage.  `out` (CUDA callers only):
    a [M, H] bf16 device tensor that receives the result instead of a fresh allocation.
    w1_bias [E, 2I] / w2_bias [E, H] / alpha / limit: the 17-argument vintage (gpt-oss expert biases and clamped
    swiglu; include/sglb200.h).  slot_out (CUDA callers): bf16 rows addressed by routed slot (through slot_map if
    given) instead of the reduced [M, H] result -- the expert-parallel return path; returns slot_out.
    shared (CUDA callers): dict(w1 [2I, H], w2 [H, I], s1, s2) -- the layer's shared expert riding along in the same
    launches; the result is then shared_expert_cpu(x, ..., fused_experts_cpu(x, ...), routed_scaling_factor), bit for
    bit (include/sglb200.h, sglb_moe_problem.shared_w13).  Only where shared_rides(...) says so."""
    if hidden_states.dim() != 2 or hidden_states.dtype != torch.bfloat16:
        raise RuntimeError("fused_experts: hidden_states must be a 2-D bfloat16 tensor")
    if w1.dtype not in _WDT[fmt] or w2.dtype != w1.dtype:
        raise RuntimeError(f"fused_experts: weight dtype {w1.dtype} does not match the quantization method")
    M, H = hidden_states.shape
    E, Hw, I = _weight_geometry(w1, w2, fmt, is_vnni)
    if Hw != H:
        raise RuntimeError(f"fused_experts: hidden size {H} does not match weights ({Hw})")
    if topk_weights.shape != topk_ids.shape or topk_ids.dim() != 2 or topk_ids.shape[0] != M:
        raise RuntimeError("fused_experts: topk_weights / topk_ids must both be [num_tokens, topk]")
    topk = topk_ids.shape[1]
    bn = bk = 0
    if fmt == L.FMT_FP8_W8A16:
        if block_size is None or len(block_size) != 2:
            raise RuntimeError("fused_experts: fp8 w8a16 needs block_size = [block_n, block_k]")
        bn, bk = int(block_size[0]), int(block_size[1])
    if fmt != L.FMT_BF16 and (w1_scale is None or w2_scale is None):
        raise RuntimeError("fused_experts: quantized weights need w1_scale and w2_scale")

    on_cuda = hidden_states.is_cuda
    dev = hidden_states.device if on_cuda else _device()
    lib = L.load()
    if M == 0:
        return hidden_states if inplace else torch.empty_like(hidden_states)
    with torch.cuda.device(dev):
        w13 = _resident_weight(w1, is_vnni, dev)
        w2d = _resident_weight(w2, is_vnni, dev)
        extras = (w1_bias is not None or w2_bias is not None or alpha is not None or limit is not None or slot_out is not None
                  or shared is not None)
        if not on_cuda and _HOST_GRAPH and M <= _HOST_GRAPH_MAX_TOKENS and out is None and not extras:
            if fmt == L.FMT_MXFP4:
                gs13, gs2 = _resident_mx_scale(w1_scale, is_vnni, dev), _resident_mx_scale(w2_scale, is_vnni, dev)
            elif fmt != L.FMT_BF16:
                gs13, gs2 = _resident_aux(w1_scale, dev, torch.float32), _resident_aux(w2_scale, dev, torch.float32)
            else:
                gs13 = gs2 = None
            with _launch_lock:
                g = _host_graph(M, topk, fmt, w13, w2d, gs13, gs2, block_size, path)
                if g is not None:
                    _shadow.clear()      # the static buffers below are about to be rewritten
                    _maybe_pin(hidden_states)
                    g.hidden.copy_(hidden_states, non_blocking=True)
                    g.topk_weights.copy_(topk_weights, non_blocking=True)
                    g.topk_ids.copy_(topk_ids, non_blocking=True)
                    g.graph.replay()
                    if inplace:
                        hidden_states.copy_(g.out)  # blocking D2H into the caller's buffer
                        _shadow_put("out", hidden_states, g.out)   # hidden_states now holds the result
                        return hidden_states
                    h = _to_host(g.out)
                    _shadow_put("x", hidden_states, g.hidden)
                    _shadow_put("out", h, g.out)
                    return h
        if not on_cuda:
            _shadow.clear()
        x = _to_dev(hidden_states, dev)
        tw = _to_dev(topk_weights, dev, torch.float32)
        ids = _to_dev(topk_ids, dev, torch.int32)
        s13 = s2 = t13 = t2 = None
        if fmt == L.FMT_MXFP4:
            s13 = _resident_mx_scale(w1_scale, is_vnni, dev)
            s2 = _resident_mx_scale(w2_scale, is_vnni, dev)
            if H % 64 == 0 and I % 64 == 0:  # block-scaled tensor-core path (umma_mx4.cu)
                w13p, w2p = _mx4_weight(w13), _mx4_weight(w2d)
                if w13p is not None and w2p is not None:
                    w13, w2d = w13p, w2p
                    t13 = _mx4_tiles(s13, (E, 2 * I, H // 32), I)
                    t2 = _mx4_tiles(s2, (E, H, I // 32), 0)
        elif fmt != L.FMT_BF16:
            s13 = _resident_aux(w1_scale, dev, torch.float32)
            s2 = _resident_aux(w2_scale, dev, torch.float32)
        b13 = b2 = None
        if w1_bias is not None or w2_bias is not None:
            if w1_bias is None or w2_bias is None or tuple(w1_bias.shape) != (E, 2 * I) or tuple(w2_bias.shape) != (E, H):
                raise RuntimeError("fused_experts: w1_bias / w2_bias must be [E, 2I] and [E, H]")
            b13 = _resident_aux(w1_bias, dev, torch.float32, "b")
            b2 = _resident_aux(w2_bias, dev, torch.float32, "b")
        act = L.ACT_SILU
        if alpha is not None or limit is not None:
            act = L.ACT_SWIGLU_OAI
        if slot_out is not None:
            if not on_cuda or not slot_out.is_cuda or slot_out.dtype != torch.bfloat16 or not slot_out.is_contiguous() \
                    or slot_out.shape[-1] != H:
                raise RuntimeError("fused_experts: slot_out needs CUDA inputs and a contiguous bf16 [rows, H] CUDA tensor")
            if slot_map is None and slot_out.numel() < M * topk * H:
                raise RuntimeError("fused_experts: slot_out must hold M * topk rows")
            if slot_map is not None and (slot_map.dtype != torch.int32 or slot_map.numel() != M * topk or not slot_map.is_cuda):
                raise RuntimeError("fused_experts: slot_map must be an int32 CUDA tensor of M * topk entries")
        elif out is not None:
            if not on_cuda or inplace or out.shape != x.shape or out.dtype != x.dtype or not out.is_contiguous() \
                    or out.device != x.device:
                raise RuntimeError("fused_experts: out= needs CUDA inputs, inplace=False and a contiguous tensor like hidden_states")
        elif on_cuda and inplace and x.data_ptr() == hidden_states.data_ptr():
            out = x
        else:
            out = torch.empty_like(x)
        sh = [None] * 4
        if shared is not None:
            sw1, sw2 = shared["w1"], shared["w2"]
            if not (on_cuda and sw1.is_cuda and sw2.is_cuda) or tuple(sw1.shape) != (2 * I, H) or tuple(sw2.shape) != (H, I) \
                    or sw1.dtype != w13.dtype or not (sw1.is_contiguous() and sw2.is_contiguous()):
                raise RuntimeError("fused_experts: a riding shared expert needs CUDA tensors w1 [2I, H], w2 [H, I] of the experts' "
                                   "own I and dtype")
            sh = [sw1, sw2, shared.get("s1"), shared.get("s2")]
            if fmt != L.FMT_BF16 and (sh[2] is None or sh[3] is None or sh[2].dtype != torch.float32 or not sh[2].is_cuda):
                raise RuntimeError("fused_experts: a riding shared expert needs its f32 CUDA scales")
        nws = lib.sglb_fused_experts_workspace_bytes(M, topk, E, H, I, fmt)
        with _launch_lock:
            ws = _workspace(nws, dev)
            p = L.MoeProblem(M=M, topk=topk, E=E, H=H, I=I, fmt=fmt, path=path, hidden=x.data_ptr(),
                             w13=w13.data_ptr(), w2=w2d.data_ptr(), topk_weights=tw.data_ptr(),
                             topk_ids=ids.data_ptr(), w13_scale=_ptr(s13), w2_scale=_ptr(s2), block_n=bn,
                             block_k=bk, out=_ptr(out), workspace=ws.data_ptr(),
                             workspace_bytes=ws.numel(), w13_scale_tiles=_ptr(t13), w2_scale_tiles=_ptr(t2),
                             w13_bias=_ptr(b13), w2_bias=_ptr(b2), act=act,
                             alpha=float(1.702 if alpha is None else alpha), limit=float(0.0 if limit is None else limit),
                             slot_out=_ptr(slot_out), slot_map=_ptr(slot_map),
                             rows_per_expert_hint=float(rows_per_expert_hint),
                             shared_w13=_ptr(sh[0]), shared_w2=_ptr(sh[1]), shared_w13_scale=_ptr(sh[2]),
                             shared_w2_scale=_ptr(sh[3]), shared_rsf=float(routed_scaling_factor))
            L.check(lib.sglb_fused_experts(p, _stream()))
        if slot_out is not None:
            return slot_out
        if on_cuda:
            if inplace and out is not hidden_states:
                hidden_states.copy_(out)
                return hidden_states
            return out
        if inplace:
            hidden_states.copy_(out)  # D2H into the caller's buffer ("fused moe mutates content in hs")
            _shadow_put("out", hidden_states, out)
            return hidden_states
        h = _to_host(out)
        _shadow_put("x", hidden_states, x)
        _shadow_put("out", h, out)
        return h


def shared_rides(M, topk, E, H, I, fmt, block_size=None, path=L.PATH_AUTO, shared_I=None, shared_fmt=None):
    """True where fused_experts_impl(shared=...) can carry the layer's shared expert inside the routed experts'
    launches (sglb_fused_experts_shared_rides + the shape conditions: same I and format as the routed experts)."""
    if M <= 0 or (shared_I is not None and shared_I != I) or (shared_fmt is not None and shared_fmt != fmt):
        return False
    bn, bk = (int(block_size[0]), int(block_size[1])) if (fmt == L.FMT_FP8_W8A16 and block_size is not None) else (0, 0)
    p = L.MoeProblem(M=M, topk=topk, E=E, H=H, I=I, fmt=fmt, path=path, block_n=bn, block_k=bk, act=L.ACT_SILU)
    return bool(L.load().sglb_fused_experts_shared_rides(p))


class GraphedExperts:
    """One fused_experts call captured in a CUDA graph over static device buffers and replayed per step.

    At decode batch sizes the six kernels of a call run for a few microseconds each, less than it takes the
    host to launch them (ctypes + cudaLaunchKernelEx per kernel); a replay is one launch.  Fill `hidden`,
    `topk_weights` and `topk_ids` in place (or pass tensors to __call__, which copies them on the current
    stream), then read `out`.  Weights must be device tensors: their addresses are baked into the graph
    (kernel arguments and TMA descriptors)."""

    def __init__(self, num_tokens, w1, w2, topk, fmt, w1_scale=None, w2_scale=None, block_size=None,
                 is_vnni=False, path=L.PATH_AUTO):
        if not (w1.is_cuda and w2.is_cuda):
            raise RuntimeError("GraphedExperts: weights must be CUDA tensors")
        dev = w1.device
        E, H, I = _weight_geometry(w1, w2, fmt, is_vnni)
        self.hidden = torch.zeros(num_tokens, H, dtype=torch.bfloat16, device=dev)
        self.topk_weights = torch.zeros(num_tokens, topk, dtype=torch.float32, device=dev)
        self.topk_ids = torch.zeros(num_tokens, topk, dtype=torch.int32, device=dev)
        self.out = torch.empty_like(self.hidden)
        self._keep = (w1, w2, w1_scale, w2_scale)
        nws = L.load().sglb_fused_experts_workspace_bytes(num_tokens, topk, E, H, I, fmt)
        self._ws = torch.empty(max(nws, 1 << 20), dtype=torch.uint8, device=dev)
        side = torch.cuda.Stream(device=dev)
        key = (dev.index, side.cuda_stream)

        def call():
            fused_experts_impl(self.hidden, w1, w2, self.topk_weights, self.topk_ids, False, fmt, w1_scale,
                               w2_scale, block_size, is_vnni, path=path, out=self.out)

        with torch.cuda.device(dev):
            side.wait_stream(torch.cuda.current_stream())
            with _launch_lock:
                _workspaces[key] = self._ws
                try:
                    with torch.cuda.stream(side):  # lazy one-off work (function attributes, scale tiles) outside the capture
                        n0 = launch_count()
                        call()
                        self.kernels_per_replay = launch_count() - n0
                    side.synchronize()
                    self.graph = torch.cuda.CUDAGraph()
                    global _capture_sink
                    derived = _capture_sink = []
                    try:
                        with torch.cuda.graph(self.graph, stream=side):
                            call()
                    finally:
                        _capture_sink = None
                    self._keep = self._keep + tuple(derived)   # scale tiles / padded weights the graph addresses
                finally:
                    _workspaces.pop(key, None)
            torch.cuda.current_stream().wait_stream(side)

    def __call__(self, hidden_states=None, topk_weights=None, topk_ids=None):
        if hidden_states is not None:
            self.hidden.copy_(hidden_states, non_blocking=True)
        if topk_weights is not None:
            self.topk_weights.copy_(topk_weights, non_blocking=True)
        if topk_ids is not None:
            self.topk_ids.copy_(topk_ids, non_blocking=True)
        self.graph.replay()
        return self.out


def fused_experts(*args):
    """fused_experts_cpu with the reference's three positional signature vintages:
      14 args (bench_moe.py:65-82, test_moe_int8.py:129, test_moe_fp8_ext.py:118):
         hs, w1, w2, topk_w, topk_ids, inplace, use_int8_w8a8, use_fp8_w8a16, w1_scale, w2_scale,
         block_size, a1_scale, a2_scale, is_vnni
      13 args (test_moe.py:79-92):
         hs, w1, w2, topk_w, topk_ids, inplace, quant_method, w1_scale, w2_scale, w1_zero, w2_zero,
         block_size, is_vnni
      17 args (vllm cpu_moe.py:308-326): the 13-arg list + w1_bias, w2_bias, alpha, limit before is_vnni
    """
    n = len(args)
    if n == 14:
        hs, w1, w2, tw, ids, inplace, use_int8, use_fp8, s1, s2, block, a1s, a2s, is_vnni = args
        if use_int8 and use_fp8:
            raise RuntimeError("fused_experts: use_int8_w8a8 and use_fp8_w8a16 are mutually exclusive")
        if a1s is not None or a2s is not None:
            raise RuntimeError("fused_experts: static activation scales (a1_scale/a2_scale) are not supported")
        fmt = L.FMT_INT8_W8A8 if use_int8 else (L.FMT_FP8_W8A16 if use_fp8 else L.FMT_BF16)
    elif n in (13, 17):
        hs, w1, w2, tw, ids, inplace, method, s1, s2, z1, z2, block = args[:12]
        is_vnni = args[-1]
        fmt = int(method)
        if fmt == int(CPUQuantMethod.INT4_W4A8) or z1 is not None or z2 is not None:
            raise RuntimeError("fused_experts: int4 w4a8 (zero points) is not supported")
        if fmt not in (L.FMT_BF16, L.FMT_INT8_W8A8, L.FMT_FP8_W8A16, L.FMT_MXFP4):
            raise RuntimeError(f"fused_experts: unknown quant method {fmt}")
        if n == 17:
            b1, b2, alpha, limit = args[12:16]
            return fused_experts_impl(hs, w1, w2, tw, ids, bool(inplace), fmt, s1, s2, block, bool(is_vnni),
                                      w1_bias=b1, w2_bias=b2, alpha=alpha, limit=limit)
    else:
        raise RuntimeError(f"fused_experts_cpu: expected 13, 14 or 17 positional arguments, got {n}")
    return fused_experts_impl(hs, w1, w2, tw, ids, bool(inplace), fmt, s1, s2, block, bool(is_vnni))


# ---- K6: shared_expert -------------------------------------------------------------------------------------
def shared_expert(*args):
    """shared_expert_cpu: 14 args (test_moe_fp8_ext.py:60-61, test_moe_fp8.py:87-88)
         hs, w1, w2, fused_experts_out, routed_scaling_factor, inplace, use_int8_w8a8, use_fp8_w8a16,
         w1_scale, w2_scale, block_size, a1_scale, a2_scale, is_vnni
       or 12 args (test_shared_experts.py:68,78): ... w1_scale, w2_scale, block_size, is_vnni."""
    n = len(args)
    if n == 14:
        hs, w1, w2, fused_out, rsf, inplace, use_int8, use_fp8, s1, s2, block, a1s, a2s, is_vnni = args
        if a1s is not None or a2s is not None:
            raise RuntimeError("shared_expert: static activation scales are not supported")
    elif n == 12:
        hs, w1, w2, fused_out, rsf, inplace, use_int8, use_fp8, s1, s2, block, is_vnni = args
    else:
        raise RuntimeError(f"shared_expert_cpu: expected 12 or 14 positional arguments, got {n}")
    if use_int8 and use_fp8:
        raise RuntimeError("shared_expert: use_int8_w8a8 and use_fp8_w8a16 are mutually exclusive")
    fmt = L.FMT_INT8_W8A8 if use_int8 else (L.FMT_FP8_W8A16 if use_fp8 else L.FMT_BF16)
    if fused_out is None:
        raise RuntimeError("shared_expert: fused_experts_out must be a tensor")
    return shared_expert_impl(hs, w1, w2, fused_out, rsf, bool(inplace), fmt, s1, s2, block, bool(is_vnni))


def shared_expert_impl(hs, w1, w2, fused_out, rsf, inplace, fmt, s1, s2, block, is_vnni, out=None):
    """Format-explicit core of shared_expert_cpu.  fused_out=None (CUDA callers): the MLP result alone,
    silu_and_mul(x W1^T) W2^T in bf16 -- the expert-parallel layer adds `combined * rsf` in its combine kernel
    (the same arithmetic as the fused epilogue).  `out`: CUDA tensor that receives the result."""
    if hs.dim() != 2 or hs.dtype != torch.bfloat16:
        raise RuntimeError("shared_expert: hidden_states must be a 2-D bfloat16 tensor")
    if w1.dim() != 2 or w2.dim() != 2 or w1.dtype not in _WDT[fmt] or w2.dtype != w1.dtype:
        raise RuntimeError("shared_expert: w1/w2 must be 2-D and match the quantization method")
    if fused_out is not None and (fused_out.shape != hs.shape or fused_out.dtype != torch.bfloat16):
        raise RuntimeError("shared_expert: fused_experts_out must match hidden_states")
    M, H = hs.shape
    packed8 = bool(is_vnni) and fmt == L.FMT_INT8_W8A8
    _, N2, K1 = _rows_k(w1, packed8)
    _, Hw, I = _rows_k(w2, packed8)
    if K1 != H or Hw != H or N2 != 2 * I:
        raise RuntimeError(f"shared_expert: inconsistent shapes hs={tuple(hs.shape)} w1={tuple(w1.shape)} w2={tuple(w2.shape)}")
    bn = bk = 0
    if fmt == L.FMT_FP8_W8A16:
        if block is None or len(block) != 2:
            raise RuntimeError("shared_expert: fp8 w8a16 needs block_size = [block_n, block_k]")
        bn, bk = int(block[0]), int(block[1])
    if fmt != L.FMT_BF16 and (s1 is None or s2 is None):
        raise RuntimeError("shared_expert: quantized weights need w1_scale and w2_scale")
    on_cuda = hs.is_cuda
    dev = hs.device if on_cuda else _device()
    lib = L.load()
    with torch.cuda.device(dev), _launch_lock:     # the lock spans shadow look-up .. enqueue (see _shadow_get)
        x = _shadow_get(hs, dev) if not on_cuda else None
        if x is None:
            x = _to_dev(hs, dev)
        fo_shadow = _shadow_get(fused_out, dev) if (not on_cuda and out is None) else None
        # A pinned host `fused_out` is only needed by the final axpy: its upload runs on a copy stream UNDER the MLP's
        # GEMMs (the reference's API hands the routed result from op to op through host memory; at DeepSeek-V3's 256
        # tokens that is 3.7 MB = about as long as the shared MLP itself).
        fo_event = None
        if fo_shadow is not None:       # still in HBM from the fused_experts call that produced it: no upload at all
            fo = fo_shadow
        elif (fused_out is not None and not on_cuda and out is None and fused_out.is_contiguous()
                and (fused_out.is_pinned() or fused_out.data_ptr() in _pinned)):
            side = _copy_stream(dev)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                fo = torch.empty(fused_out.shape, dtype=fused_out.dtype, device=dev)
                fo.copy_(fused_out, non_blocking=True)
                fo_event = torch.cuda.Event()
                fo_event.record(side)
            fo.record_stream(torch.cuda.current_stream())
        else:
            fo = _to_dev(fused_out, dev)
        w1d = _resident_weight(w1, bool(is_vnni), dev)
        w2d = _resident_weight(w2, bool(is_vnni), dev)
        s1d = _resident_aux(s1, dev, torch.float32)
        s2d = _resident_aux(s2, dev, torch.float32)
        if out is not None:
            if not on_cuda or inplace or out.shape != x.shape or out.dtype != x.dtype or not out.is_contiguous():
                raise RuntimeError("shared_expert: out= needs CUDA inputs, inplace=False and a contiguous tensor like hidden_states")
        else:
            out = x if (on_cuda and inplace and x.data_ptr() == hs.data_ptr()) else torch.empty_like(x)
        nws = lib.sglb_fused_experts_workspace_bytes(M, 1, 1, H, I, fmt)
        with _launch_lock:
            ws = _workspace(nws, dev)
            p = L.SharedProblem(M=M, H=H, I=I, fmt=fmt, path=L.PATH_AUTO, hidden=x.data_ptr(), w1=w1d.data_ptr(),
                                w2=w2d.data_ptr(), w1_scale=_ptr(s1d), w2_scale=_ptr(s2d), block_n=bn, block_k=bk,
                                fused_out=None if fo_event is not None else _ptr(fo), routed_scaling_factor=float(rsf),
                                out=out.data_ptr(), workspace=ws.data_ptr(), workspace_bytes=ws.numel())
            L.check(lib.sglb_shared_expert(p, _stream()))
            if fo_event is not None:   # out = bf16(fma(fused_out, rsf, mlp)): the epilogue sglb_shared_expert fuses
                torch.cuda.current_stream().wait_event(fo_event)
                L.check(lib.sglb_shared_axpy(out.data_ptr(), fo.data_ptr(), float(rsf), out.numel(), out.data_ptr(), _stream()))
        if on_cuda:
            if inplace and out is not hs:
                hs.copy_(out)
                return hs
            return out
        if inplace:
            hs.copy_(out)
            return hs
        return out.cpu()


# ---- K9: mxfp4 linear --------------------------------------------------------------------------------------
def mxfp4_scaled_mm(mat1, mat2, scales2, bias, is_vnni):
    """mxfp4_scaled_mm_cpu(A, B_packed, Bs_packed, bias?, is_vnni) -> Tensor (test_mxfp4.py:167,196)."""
    if mat1.dim() != 2 or mat1.dtype != torch.bfloat16:
        raise RuntimeError("mxfp4_scaled_mm: mat1 must be a 2-D bfloat16 tensor")
    if mat2.dim() != 2 or mat2.dtype != torch.uint8 or scales2.dtype != torch.uint8:
        raise RuntimeError("mxfp4_scaled_mm: mat2 / scales2 must be 2-D uint8 tensors")
    M, K = mat1.shape
    N = mat2.shape[0]
    if mat2.shape[1] * 2 != K or K % 32 != 0 or scales2.numel() != N * (K // 32):
        raise RuntimeError(f"mxfp4_scaled_mm: inconsistent shapes A={tuple(mat1.shape)} B={tuple(mat2.shape)} Bs={tuple(scales2.shape)}")
    scale_vnni = bool(is_vnni)
    if tuple(scales2.shape) != (N, K // 32):
        # MXFP4QuantizeUtil.quantize (test_mxfp4.py:69-88) returns the scales flattened to
        # [N*K/32, 1]; convert_scale_packed of that shape re-tiles with G = 1, i.e. not at all
        # (test_mxfp4.py:200-201), so the bytes are still row-major [N, K/32].
        if scales2.dim() != 2 or scales2.shape[1] != 1:
            raise RuntimeError(f"mxfp4_scaled_mm: scales must be [N, K/32] or [N*K/32, 1], got {tuple(scales2.shape)}")
        scales2 = scales2.reshape(N, K // 32)
        scale_vnni = False
    on_cuda = mat1.is_cuda
    dev = mat1.device if on_cuda else _device()
    lib = L.load()
    with torch.cuda.device(dev):
        a = _to_dev(mat1, dev)
        bq = _resident_weight(mat2, bool(is_vnni), dev)
        bs = _resident_mx_scale(scales2, scale_vnni, dev)
        bd = _resident_aux(bias, dev, torch.float32, "b")
        out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        nws = lib.sglb_fused_experts_workspace_bytes(M, 1, 1, K, N, L.FMT_MXFP4)
        with _launch_lock:
            ws = _workspace(nws, dev)
            L.check(lib.sglb_mxfp4_mm(a.data_ptr(), bq.data_ptr(), bs.data_ptr(), _ptr(bd), M, N, K, out.data_ptr(),
                                      ws.data_ptr(), ws.numel(), _stream()))
    return out if on_cuda else out.cpu()


# ---- K10: dense quantised linears ---------------------------------------------------
def _linear(A, W, fmt, is_vnni, w_scale=None, a_scale=None, block_size=None, bias=None, out_dtype=None,
            path=L.PATH_AUTO, who="linear"):
    """out[M,N] bf16 = A[M,K] . dequant(W[N,K])^T + bias through sglb_linear (stage 2 of the expert kernels
    with E = 1).  W may be the packed tensor convert_weight_packed returned (is_vnni=True)."""
    if out_dtype not in (None, torch.bfloat16):
        raise RuntimeError(f"{who}: only bfloat16 outputs are supported (got {out_dtype})")
    if A.dim() != 2 or W.dim() != 2:
        raise RuntimeError(f"{who}: mat1 and mat2 must be 2-D")
    M, K = A.shape
    comp = fmt == L.FMT_INT8_W8A8 and bool(is_vnni)
    _, N, Kw = _rows_k(W, comp)
    if Kw != K:
        raise RuntimeError(f"{who}: inner dimensions differ (mat1 K={K}, mat2 K={Kw})")
    bn = bk = 0
    if fmt == L.FMT_FP8_W8A16:
        if block_size is None or len(block_size) != 2:
            raise RuntimeError(f"{who}: fp8 needs block_size=[block_n, block_k]")
        bn, bk = int(block_size[0]), int(block_size[1])
        if w_scale is None or w_scale.numel() != -(-N // bn) * -(-K // bk):
            raise RuntimeError(f"{who}: fp8 scales must be [ceil(N/{bn}), ceil(K/{bk})]")
    if fmt == L.FMT_INT8_W8A8 and (w_scale is None or w_scale.numel() != N):
        raise RuntimeError(f"{who}: int8 needs a per-channel scale of {N} elements")
    if a_scale is not None and a_scale.numel() != M:
        raise RuntimeError(f"{who}: activation scale must hold one value per row ({M})")
    on_cuda = A.is_cuda
    dev = A.device if on_cuda else _device()
    lib = L.load()
    with torch.cuda.device(dev):
        a = _to_dev(A, dev)
        wd = _resident_weight(W, bool(is_vnni), dev)
        sd = _resident_aux(w_scale, dev, torch.float32)
        asd = _to_dev(a_scale.reshape(-1), dev, torch.float32) if a_scale is not None else None
        bd = _resident_aux(bias, dev, torch.float32, "b")
        out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        nws = lib.sglb_linear_workspace_bytes(M, N, K, fmt)
        with _launch_lock:
            ws = _workspace(nws, dev)
            p = L.LinearProblem(M=M, N=N, K=K, fmt=fmt, path=path, A=_ptr(a), a_scale=_ptr(asd), W=_ptr(wd),
                                w_scale=_ptr(sd), block_n=bn, block_k=bk, bias=_ptr(bd), out=_ptr(out),
                                workspace=ws.data_ptr(), workspace_bytes=ws.numel())
            L.check(lib.sglb_linear(p, _stream()))
    return out if on_cuda else out.cpu()


def weight_packed_linear(mat1, mat2, bias, is_vnni):
    """weight_packed_linear(mat1, mat2, bias?, is_vnni) -> Tensor   (test_gemm.py:22-25)."""
    if mat1.dtype != torch.bfloat16 or mat2.dtype != torch.bfloat16:
        raise RuntimeError("weight_packed_linear: mat1 / mat2 must be bfloat16")
    return _linear(mat1, mat2, L.FMT_BF16, is_vnni, bias=bias, who="weight_packed_linear")


def per_token_quant_int8(A):
    """per_token_quant_int8_cpu(A) -> (Aq int8 [M,K], As f32 [M,1])   (test_gemm_int8.py:68, oracle :11-20)."""
    if A.dim() != 2 or A.dtype != torch.bfloat16:
        raise RuntimeError("per_token_quant_int8: A must be a 2-D bfloat16 tensor")
    on_cuda = A.is_cuda
    dev = A.device if on_cuda else _device()
    M, K = A.shape
    with torch.cuda.device(dev):
        a = _to_dev(A, dev)
        q = torch.empty(M, K, dtype=torch.int8, device=dev)
        sc = torch.empty(M, 1, dtype=torch.float32, device=dev)
        L.check(L.load().sglb_quant_rows_int8(_ptr(a), M, K, _ptr(q), _ptr(sc), _stream()))
    return (q, sc) if on_cuda else (q.cpu(), sc.cpu())


def int8_scaled_mm(mat1, mat2, scales1, scales2, bias, out_dtype, is_vnni):
    """int8_scaled_mm_cpu(Aq, Bq, As, Bs, bias?, out_dtype, is_vnni) -> Tensor   (test_gemm_int8.py:69)."""
    if mat1.dtype != torch.int8 or mat2.dtype != torch.int8:
        raise RuntimeError("int8_scaled_mm: mat1 / mat2 must be int8")
    return _linear(mat1, mat2, L.FMT_INT8_W8A8, is_vnni, w_scale=scales2, a_scale=scales1, bias=bias,
                   out_dtype=out_dtype, who="int8_scaled_mm")


def int8_scaled_mm_with_quant(mat1, mat2, scales2, bias, out_dtype, is_vnni):
    """int8_scaled_mm_with_quant(A, Bq, Bs, bias?, out_dtype, is_vnni) -> Tensor   (test_gemm_int8.py:74):
    per-token quantisation of A fused into the call."""
    if mat1.dtype != torch.bfloat16 or mat2.dtype != torch.int8:
        raise RuntimeError("int8_scaled_mm_with_quant: mat1 must be bfloat16 and mat2 int8")
    return _linear(mat1, mat2, L.FMT_INT8_W8A8, is_vnni, w_scale=scales2, bias=bias, out_dtype=out_dtype,
                   who="int8_scaled_mm_with_quant")


def fp8_scaled_mm(mat1, mat2, scales2, block_size, bias, out_dtype, is_vnni):
    """fp8_scaled_mm_cpu(A, Bq, Bs, block_size, bias?, out_dtype, is_vnni) -> Tensor   (test_gemm_fp8.py:54-62)."""
    if mat1.dtype != torch.bfloat16 or mat2.dtype != torch.float8_e4m3fn:
        raise RuntimeError("fp8_scaled_mm: mat1 must be bfloat16 and mat2 float8_e4m3fn")
    return _linear(mat1, mat2, L.FMT_FP8_W8A16, is_vnni, w_scale=scales2, block_size=block_size, bias=bias,
                   out_dtype=out_dtype, who="fp8_scaled_mm")


# ---- K11: silu_and_mul / rmsnorm / fused_add_rmsnorm -------------------------------
def _act_dtype(t, who):
    if t.dtype not in (torch.bfloat16, torch.float16):
        raise RuntimeError(f"{who}: bfloat16 or float16 tensors only (got {t.dtype})")
    return _DT[t.dtype]


def silu_and_mul(*args):
    """silu_and_mul_cpu(x) -> Tensor (bench_silu_and_mul.py:31) or silu_and_mul_cpu(out, x) -> None
    (test_activation.py:26)."""
    if len(args) == 1:
        out, x = None, args[0]
    elif len(args) == 2:
        out, x = args
    else:
        raise RuntimeError("silu_and_mul: expected (x) or (out, x)")
    dt = _act_dtype(x, "silu_and_mul")
    if x.shape[-1] % 2:
        raise RuntimeError("silu_and_mul: the last dimension must be even")
    d = x.shape[-1] // 2
    oshape = tuple(x.shape[:-1]) + (d,)
    if out is not None and (tuple(out.shape) != oshape or out.dtype != x.dtype):
        raise RuntimeError("silu_and_mul: out must be [..., d] of the input dtype")
    on_cuda = x.is_cuda
    dev = x.device if on_cuda else _device()
    with torch.cuda.device(dev):
        xd = _to_dev(x, dev)
        od = out if (out is not None and out.is_cuda and out.is_contiguous()) else torch.empty(oshape, dtype=x.dtype, device=dev)
        rows = xd.numel() // (2 * d) if d else 0
        L.check(L.load().sglb_silu_and_mul(_ptr(xd), _ptr(od), rows, d, dt, _stream()))
    if out is None:
        return od if on_cuda else od.cpu()
    if od is not out:
        out.copy_(od)
    return None


def rmsnorm(out, x, weight, eps):
    """rmsnorm_cpu(out, x, weight, eps) -> None   (test_norm.py:43-45)."""
    dt = _act_dtype(x, "rmsnorm")
    if weight.dtype != x.dtype or out.dtype != x.dtype or out.shape != x.shape or weight.numel() != x.shape[-1]:
        raise RuntimeError("rmsnorm: out / weight must match the input's dtype, shape and hidden size")
    dev = x.device if x.is_cuda else _device()
    hidden = x.shape[-1]
    with torch.cuda.device(dev):
        xd, wd = _to_dev(x, dev), _to_dev(weight, dev)
        od = out if (out.is_cuda and out.is_contiguous()) else torch.empty_like(xd)
        L.check(L.load().sglb_rmsnorm(_ptr(xd), _ptr(wd), float(eps), _ptr(od), xd.numel() // hidden, hidden, dt,
                                      _stream()))
    if od is not out:
        out.copy_(od)
    return None


def fused_add_rmsnorm(x, residual, weight, eps):
    """fused_add_rmsnorm_cpu(x, residual, weight, eps) -> None, both updated in place   (test_norm.py:56)."""
    dt = _act_dtype(x, "fused_add_rmsnorm")
    if residual.dtype != x.dtype or residual.shape != x.shape or weight.dtype != x.dtype or \
            weight.numel() != x.shape[-1]:
        raise RuntimeError("fused_add_rmsnorm: residual / weight must match the input's dtype, shape and hidden size")
    dev = x.device if x.is_cuda else _device()
    hidden = x.shape[-1]
    with torch.cuda.device(dev):
        inplace = x.is_cuda and x.is_contiguous() and residual.is_cuda and residual.is_contiguous()
        xd = x if inplace else _to_dev(x, dev).clone()
        rd = residual if inplace else _to_dev(residual, dev).clone()
        wd = _to_dev(weight, dev)
        L.check(L.load().sglb_fused_add_rmsnorm(_ptr(xd), _ptr(rd), _ptr(wd), float(eps), xd.numel() // hidden,
                                                hidden, dt, _stream()))
    if not inplace:
        x.copy_(xd)
        residual.copy_(rd)
    return None


def split_rows_e4m3(x):
    """bf16 rows -> (hi, lo, rexp): two e4m3 pieces (uint8 storage) and the per-row power-of-two factor with
    x ~= (hi + lo) * rexp -- the activation format of the tensor-native fp8 path (csrc/umma_f8.cu)."""
    dev = _device()
    x = _to_dev(x, dev, torch.bfloat16)
    rows, K = x.shape
    hi = torch.empty(rows, K, dtype=torch.uint8, device=dev)
    lo = torch.empty_like(hi)
    rexp = torch.empty(rows, dtype=torch.float32, device=dev)
    L.check(L.load().sglb_split_rows_e4m3(_ptr(x), rows, K, _ptr(hi), _ptr(lo), _ptr(rexp), _stream()))
    return hi, lo, rexp


def launch_count():
    return int(L.load().sglb_launch_count())
