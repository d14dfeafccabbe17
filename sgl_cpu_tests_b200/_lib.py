"""ctypes binding of libsglb200.so (include/sglb200.h).  No torch types cross this boundary."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsglb200.so")

OK = 0
DT_BF16, DT_F32, DT_F16 = 0, 1, 2
FMT_BF16, FMT_INT8_W8A8, FMT_FP8_W8A16, FMT_MXFP4 = 0, 1, 2, 4
PATH_AUTO, PATH_STREAM, PATH_UMMA, PATH_UMMA2 = 0, 1, 2, 3
ACT_SILU, ACT_SWIGLU_OAI = 0, 1

_vp, _i32, _i64, _f32, _sz = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_size_t


class AlignOut(C.Structure):
    _fields_ = [(n, _vp) for n in ("expert_count", "expert_offset", "sorted_slot", "slot_pos",
                                   "chunk_list", "tile_list", "counters")]


class MoeProblem(C.Structure):
    _fields_ = [
        ("M", _i64), ("topk", _i32), ("E", _i32), ("H", _i32), ("I", _i32), ("fmt", _i32), ("path", _i32),
        ("hidden", _vp), ("w13", _vp), ("w2", _vp), ("topk_weights", _vp), ("topk_ids", _vp),
        ("w13_scale", _vp), ("w2_scale", _vp), ("block_n", _i32), ("block_k", _i32),
        ("out", _vp), ("workspace", _vp), ("workspace_bytes", _sz),
        ("w13_scale_tiles", _vp), ("w2_scale_tiles", _vp),
        ("w13_bias", _vp), ("w2_bias", _vp), ("act", _i32), ("alpha", _f32), ("limit", _f32),
        ("slot_out", _vp), ("slot_map", _vp), ("rows_per_expert_hint", _f32),
        ("shared_w13", _vp), ("shared_w2", _vp), ("shared_w13_scale", _vp), ("shared_w2_scale", _vp), ("shared_rsf", _f32),
    ]


class SharedProblem(C.Structure):
    _fields_ = [
        ("M", _i64), ("H", _i32), ("I", _i32), ("fmt", _i32), ("path", _i32),
        ("hidden", _vp), ("w1", _vp), ("w2", _vp), ("w1_scale", _vp), ("w2_scale", _vp),
        ("block_n", _i32), ("block_k", _i32), ("fused_out", _vp), ("routed_scaling_factor", _f32),
        ("out", _vp), ("workspace", _vp), ("workspace_bytes", _sz),
    ]


class LinearProblem(C.Structure):
    _fields_ = [
        ("M", _i64), ("N", _i32), ("K", _i32), ("fmt", _i32), ("path", _i32),
        ("A", _vp), ("a_scale", _vp), ("W", _vp), ("w_scale", _vp),
        ("block_n", _i32), ("block_k", _i32), ("bias", _vp), ("out", _vp),
        ("workspace", _vp), ("workspace_bytes", _sz),
    ]


# name -> (restype, argtypes); also the list tests check against include/sglb200.h
SIGNATURES = {
    "sglb_version": (C.c_int, []),
    "sglb_last_error": (C.c_char_p, []),
    "sglb_launch_count": (C.c_uint64, []),
    "sglb_host_register": (C.c_int, [_vp, _sz]),
    "sglb_host_unregister": (C.c_int, [_vp]),
    "sglb_profile_enable": (None, [C.c_int]),
    "sglb_set_fp8_native": (None, [C.c_int]),
    "sglb_set_chain_fusion": (None, [C.c_int]),
    "sglb_set_mx4_native": (None, [C.c_int]),
    "sglb_mx4_scale_tiles_bytes": (_sz, [_i64, C.c_int, C.c_int, C.c_int]),
    "sglb_mx4_tile_scales": (C.c_int, [_vp, _vp, _i64, C.c_int, C.c_int, C.c_int, _vp]),
    "sglb_profile_read": (C.c_int, [C.POINTER(C.c_float), C.POINTER(C.c_int), C.c_int]),
    "sglb_grouped_topk": (C.c_int, [_vp, C.c_int, _vp, C.c_int, _i64, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.c_int, C.c_int, _f32, _vp, C.c_int, _vp, _vp, _vp]),
    "sglb_moe_align_scratch_bytes": (_sz, [_i64, C.c_int, C.c_int]),
    "sglb_moe_align_max_items": (_i64, [_i64, C.c_int, C.c_int, C.c_int]),
    "sglb_moe_align": (C.c_int, [_vp, _i64, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(AlignOut), _vp, _vp]),
    "sglb_moe_align_static_max_bytes": (_sz, []),
    "sglb_moe_align_static": (C.c_int, [_vp, _i64, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(AlignOut), _vp]),
    "sglb_fused_experts_workspace_bytes": (_sz, [_i64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "sglb_fused_experts": (C.c_int, [C.POINTER(MoeProblem), _vp]),
    "sglb_fused_experts_shared_rides": (C.c_int, [C.POINTER(MoeProblem)]),
    "sglb_shared_expert": (C.c_int, [C.POINTER(SharedProblem), _vp]),
    "sglb_linear_workspace_bytes": (_sz, [_i64, C.c_int, C.c_int, C.c_int]),
    "sglb_linear": (C.c_int, [C.POINTER(LinearProblem), _vp]),
    "sglb_silu_and_mul": (C.c_int, [_vp, _vp, _i64, C.c_int, C.c_int, _vp]),
    "sglb_rmsnorm": (C.c_int, [_vp, _vp, _f32, _vp, _i64, C.c_int, C.c_int, _vp]),
    "sglb_fused_add_rmsnorm": (C.c_int, [_vp, _vp, _vp, _f32, _i64, C.c_int, C.c_int, _vp]),
    "sglb_mxfp4_mm": (C.c_int, [_vp, _vp, _vp, _vp, _i64, C.c_int, C.c_int, _vp, _vp, _sz, _vp]),
    "sglb_repack_weight": (C.c_int, [_vp, _vp, C.c_int, C.c_int, _i64, _i64, _i64, C.c_int, C.c_int, _vp]),
    "sglb_repack_scale": (C.c_int, [_vp, _vp, _i64, _i64, _i64, C.c_int, _vp]),
    "sglb_quant_rows_int8": (C.c_int, [_vp, _i64, C.c_int, _vp, _vp, _vp]),
    "sglb_split_rows_e4m3": (C.c_int, [_vp, _i64, C.c_int, _vp, _vp, _vp, _vp]),
    "sglb_ep_plan": (C.c_int, [_vp, _i64, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp]),
    "sglb_ep_ret_index": (C.c_int, [_vp, C.c_int, _i64, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, _vp]),
    "sglb_shared_axpy": (C.c_int, [_vp, _vp, _f32, _i64, _vp, _vp]),
    "sglb_ep_pack": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _i64, _vp,
                               _vp, _vp]),
    "sglb_ep_combine": (C.c_int, [_vp, _vp, _vp, _vp, _i64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _vp, _f32,
                                  _vp, _vp]),
    "sglb_p2p_alloc": (C.c_int, [_sz, C.POINTER(_vp), _vp]),
    "sglb_p2p_open": (C.c_int, [_vp, C.POINTER(_vp)]),
    "sglb_p2p_close": (C.c_int, [_vp]),
    "sglb_p2p_free": (C.c_int, [_vp]),
    "sglb_ep_p2p_region_bytes": (_sz, [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(_sz)]),
    "sglb_ep_p2p_dispatch": (C.c_int, [_vp, _vp, _vp, _i64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.POINTER(_vp), C.c_int, _vp]),
    "sglb_ep_p2p_wait_dispatch": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, _vp]),
    "sglb_ep_p2p_combine": (C.c_int, [_vp, _i64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(_vp),
                                      C.c_int, _vp, _f32, _vp, _vp]),
    "sglb_topk_reduce": (C.c_int, [_vp, _vp, _i64, C.c_int, C.c_int, _vp, _vp]),
}

_lib = None


def load():
    """Loads the CUDA library.  There is no fallback: a missing build is an error."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m sgl_cpu_tests_b200.build` "
            "(or __graft_entry__.build()).  There is no CPU or PyTorch fallback for this path.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here == header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != OK:
        msg = load().sglb_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"sglb200: {msg} (status {rc})")
