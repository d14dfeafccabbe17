// Shared device helpers for the sm_100a fused-MoE path.  Everything in csrc/ is written for
// one target only (-gencode arch=compute_100a,code=sm_100a); there is no fallback.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/sglb200.h"

namespace sglb {

// ---- error plumbing (thread-local message behind sglb_last_error) -------------------------
void set_error(const char* fmt, ...);
void count_launch(int n = 1);

#define SGLB_CHECK(cond, ...)                 \
  do {                                        \
    if (!(cond)) {                            \
      ::sglb::set_error(__VA_ARGS__);         \
      return SGLB_ERR_INVALID;                \
    }                                         \
  } while (0)

#define SGLB_CUDA(expr)                                                              \
  do {                                                                               \
    cudaError_t _e = (expr);                                                         \
    if (_e != cudaSuccess) {                                                         \
      ::sglb::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),      \
                        __FILE__, __LINE__);                                         \
      return SGLB_ERR_CUDA;                                                          \
    }                                                                                \
  } while (0)

#define SGLB_LAUNCH_CHECK()                                                          \
  do {                                                                               \
    cudaError_t _e = cudaGetLastError();                                             \
    if (_e != cudaSuccess) {                                                         \
      ::sglb::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e),  \
                        __FILE__, __LINE__);                                         \
      return SGLB_ERR_CUDA;                                                          \
    }                                                                                \
  } while (0)

constexpr int kWarp = 32;

// ---- programmatic dependent launch -----------------------------------------------------------------
// Every kernel of the fused_experts chain is launched with programmatic stream serialization: the
// next kernel's CTAs may start (and run their prologue) while this grid drains, and block in
// pdl_wait() until the previous grid has completed and its writes are visible.  pdl_wait() must
// precede the first access to any buffer another kernel of the chain reads or writes.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

#ifdef __CUDACC__
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
__host__ __device__ inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
__host__ __device__ inline size_t align_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

// ---- loads ---------------------------------------------------------------------------------
// Weight stream: read once, keep out of L1 so the (re-used) activation rows stay there.
__device__ __forceinline__ int4 ldg_stream16(const void* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
// Activation rows: small, re-read by every warp of the CTA -> normal caching.
__device__ __forceinline__ int4 ldg_cached16(const void* p) {
  int4 r;
  asm volatile("ld.global.nc.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// ---- bf16 helpers --------------------------------------------------------------------------
__device__ __forceinline__ float bf16_bits_to_float(uint32_t b16) { return __uint_as_float(b16 << 16); }
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

template <typename T>
__device__ __forceinline__ float load_as_float(const T* p);
template <>
__device__ __forceinline__ float load_as_float<float>(const float* p) { return *p; }
template <>
__device__ __forceinline__ float load_as_float<__nv_bfloat16>(const __nv_bfloat16* p) {
  return __bfloat162float(*p);
}
template <>
__device__ __forceinline__ float load_as_float<__half>(const __half* p) { return __half2float(*p); }

__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + __expf(-x)); }
// Stage-1 activation.  act 0: silu(gate) * up (test_moe.py:17-19).  act 1: gpt-oss clamped swiglu of the 17-argument
// fused_experts_cpu vintage (alpha, limit): g = min(gate, limit), u = clamp(up, -limit, limit),
// out = g * sigmoid(alpha * g) * (u + 1); limit <= 0 means "no clamp".
enum { SGLB_ACT_SILU = 0, SGLB_ACT_SWIGLU_OAI = 1 };
__device__ __forceinline__ float glu_f(float gate, float up, int act, float alpha, float limit) {
  if (act == SGLB_ACT_SILU) return silu_f(gate) * up;
  if (limit > 0.f) {
    gate = fminf(gate, limit);
    up = fminf(fmaxf(up, -limit), limit);
  }
  return gate / (1.0f + __expf(-alpha * gate)) * (up + 1.0f);
}
// Where stage 2 writes routed row p (position in expert-sorted order): row p of `out`, or -- when the caller
// asked for slot-addressed output (expert-parallel return path) -- row sorted_slot[p] (= token * topk + slot),
// optionally sent through one more table (the compact return buffer of the NCCL exchange).
__device__ __forceinline__ int64_t out_row_of(int p, const int32_t* sorted_slot, int by_slot, const int32_t* out_map) {
  if (!by_slot) return p;
  const int s = sorted_slot[p];
  return out_map != nullptr ? out_map[s] : s;
}

// ---- fused top-k reduce (decode chain fusion, capi.cu) ---------------------------------------------------------
// The stage-2 GEMM kernels finish a token themselves: every (routed row, output-column block) a CTA / warp completes is
// counted on red_cnt[token, block]; whoever brings the count to `need` (= topk, + 1 with a riding shared expert; the
// prologue pre-counts masked slots) adds the token's slot rows of that block IN SLOT ORDER j = 0..topk-1 in f32 and
// rounds once -- topk_reduce_kernel's arithmetic (misc.cu), so the result is bit-identical to the separate launch.
// Rows written by other SMs are read through L2 (ld.global.cg) after a device-scope fence.
struct FusedReduce {
  int32_t* cnt;              // [M, blocks]; nullptr = no fusion (the caller launches topk_reduce)
  const int32_t* slot_pos;   // [M * topk]: sorted position of each slot or -1
  const int32_t* tok_valid;  // [M]: valid slots per token (0 and no shared rider -> the row is zero-filled)
  void* out;                 // [M, N] bf16
  int M;
  int blocks;                // counters per token
  int need;                  // arrivals that complete a (token, block)
  int shared_row0;           // riding shared expert: its stage-2 rows sit at ic2[shared_row0 + token]; -1 = none
  float rsf;
};
// 8 consecutive columns [col0, col0 + 8) of token t.  ic2: the stage-2 output rows (bf16, row stride N).
__device__ __forceinline__ void fused_reduce_cols8(const FusedReduce& r, const __nv_bfloat16* ic2, int topk, int N, int t,
                                                   int col0) {
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  int4 sv = make_int4(0, 0, 0, 0);  // the riding shared expert's row travels with the first pass
  if (r.shared_row0 >= 0) sv = __ldcg(reinterpret_cast<const int4*>(ic2 + (size_t)(r.shared_row0 + t) * N + col0));
  // eight slots per pass: the positions, then the rows, are fetched together (two L2 round trips per pass instead of
  // two per slot); the additions keep the slot order
  for (int j0 = 0; j0 < topk; j0 += 8) {
    int p[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) p[i] = j0 + i < topk ? r.slot_pos[(size_t)t * topk + j0 + i] : -1;
    int4 v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
      v[i] = p[i] >= 0 ? __ldcg(reinterpret_cast<const int4*>(ic2 + (size_t)p[i] * N + col0)) : make_int4(0, 0, 0, 0);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (p[i] < 0) continue;
      const uint32_t* w = reinterpret_cast<const uint32_t*>(&v[i]);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        acc[2 * q] += bf16_bits_to_float(w[q] & 0xffffu);
        acc[2 * q + 1] += bf16_bits_to_float(w[q] >> 16);
      }
    }
  }
  if (r.shared_row0 >= 0) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(&sv);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      acc[2 * i] = fmaf(bf16_round(acc[2 * i]), r.rsf, bf16_bits_to_float(w[i] & 0xffffu));
      acc[2 * i + 1] = fmaf(bf16_round(acc[2 * i + 1]), r.rsf, bf16_bits_to_float(w[i] >> 16));
    }
  }
  uint32_t o[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) o[i] = pack_bf16x2(acc[2 * i], acc[2 * i + 1]);
  *reinterpret_cast<int4*>((__nv_bfloat16*)r.out + (size_t)t * N + col0) = make_int4(o[0], o[1], o[2], o[3]);
}
// rows of tokens without any valid slot (and no shared rider) are zero: nobody ever arrives for them
__device__ __forceinline__ void fused_reduce_zero_rows(const FusedReduce& r, int N, int worker, int workers, int lane) {
  if (r.cnt == nullptr || r.shared_row0 >= 0) return;
  for (int t = worker; t < r.M; t += workers) {
    if (r.tok_valid[t] != 0) continue;
    for (int c = lane * 8; c < N; c += 32 * 8)
      *reinterpret_cast<int4*>((__nv_bfloat16*)r.out + (size_t)t * N + c) = make_int4(0, 0, 0, 0);
  }
}

// ---- legacy warp MMA (decode / weight-streaming path) ----------------------------------------
// D(16x8,f32) += A(16x16,bf16,row) * B(16x8,bf16,col)
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                               uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// D(16x8,s32) += A(16x32,s8,row) * B(32x8,s8,col)
__device__ __forceinline__ void mma_s8_16832(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                             uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// ---- in-register dequantisation ----------------------------------------------------------------
// Two e4m3 bytes (already moved to bits [7:0] and [23:16] of p, other bytes zero) -> bf16x2.
// s.eeee.mmm lands in bf16 as s | eeeemmm<<4, i.e. the value scaled by 2^-120 (bias 7 vs 127);
// one packed bf16 multiply by 2^120 restores it, e4m3 subnormals included (bf16 mul keeps
// subnormal inputs).  Exact: e4m3 has 3 mantissa bits, bf16 has 7.
__device__ __forceinline__ uint32_t e4m3x2_to_bf16x2(uint32_t p) {
  uint32_t sign = p & 0x00800080u;
  uint32_t raw = p * 16u + sign * 240u;  // (mag << 4) | (sign << 8)
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&raw);
  const uint32_t kScale = 0x7B807B80u;  // bf16 2^120 in both halves
  __nv_bfloat162 s = *reinterpret_cast<const __nv_bfloat162*>(&kScale);
  v = __hmul2(v, s);
  return *reinterpret_cast<uint32_t*>(&v);
}
// word with 4 e4m3 bytes b0..b3 -> (b0,b1) and (b2,b3) as bf16x2
__device__ __forceinline__ void e4m3x4_to_bf16x4(uint32_t w, uint32_t& lo, uint32_t& hi) {
  lo = e4m3x2_to_bf16x2(__byte_perm(w, 0u, 0x4140));  // b0 -> byte0, b1 -> byte2
  hi = e4m3x2_to_bf16x2(__byte_perm(w, 0u, 0x4342));  // b2 -> byte0, b3 -> byte2
}

// e2m1 nibbles: s.ee.m -> bf16 bits s<<15 | eem<<6 equals value * 2^-126 for all 16 codes
// (0.5 = 0b001 becomes a bf16 subnormal).  Multiply by 2^(e8m0-1) = 2^126 * 2^(e8m0-127) afterwards.
// For word w (nibbles n0..n7, n0 lowest) returns r[i] = (n_i, n_{i+4}) as raw (unscaled) bf16x2.
__device__ __forceinline__ void e2m1x8_to_raw_bf16(uint32_t w, uint32_t (&r)[4]) {
#pragma unroll
  for (int i = 3; i >= 0; --i) {
    r[i] = (w & 0x80008000u) | ((w & 0x70007000u) >> 6);
    w <<= 4;
  }
}
__device__ __forceinline__ uint32_t bf16x2_mul(uint32_t a, uint32_t b) {
  __nv_bfloat162 x = *reinterpret_cast<__nv_bfloat162*>(&a);
  __nv_bfloat162 y = *reinterpret_cast<__nv_bfloat162*>(&b);
  x = __hmul2(x, y);
  return *reinterpret_cast<uint32_t*>(&x);
}
// e8m0 byte s -> bf16 2^(s-127) in both halves (s = 0 is the bf16 subnormal 2^-127).
__device__ __forceinline__ uint32_t e8m0_to_bf16x2(uint32_t s) {
  uint32_t b = s ? (s << 7) : 0x0040u;
  return b | (b << 16);
}
// raw e2m1 pair (see e2m1x8_to_raw_bf16) * 2^126 * scale.  Two exact power-of-two multiplies so no
// intermediate exponent can wrap for any e8m0 code whose result is representable.
__device__ __forceinline__ uint32_t mxfp4_finish(uint32_t raw, uint32_t scale2) {
  const uint32_t k126 = 0x7E807E80u;  // bf16 2^126 in both halves
  return bf16x2_mul(bf16x2_mul(raw, k126), scale2);
}

}  // namespace sglb
