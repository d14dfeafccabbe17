// Per-row activation preparation shared by the stand-alone kernels (misc.cu quant_rows_kernel, umma_f8.cu
// split_rows_e4m3_kernel) and the single-launch prologue of the fused decode chain (align.cu): 256 threads work on one
// row and meet on a named barrier, so a 1024-thread block can host four independent rows.
#pragma once
#include <cuda_fp8.h>

#include "common.cuh"

namespace sglb {

constexpr int kPrologueMaxAlignBlocks = 16;  // align blocks (1024 slots each) the single-launch prologue handles

struct PrologueRows {
  int mode;                 // 0: nothing to prepare (bf16), 1: int8 per-row quantisation, 2: e4m3 hi / lo split
  const __nv_bfloat16* x;   // [M, K] hidden states
  int M, K;
  void *q, *q2;             // mode 1: s8 [M, K]; mode 2: hi, lo e4m3 [M, K]
  float* xs;                // [M]: int8 scale / 2^e factor
  int32_t* tok_valid;       // [M] or nullptr (no fused reduce)
  int32_t* red_cnt;         // [M, red_blocks]
  int red_blocks;
  uint32_t* row_amax;       // [amax_rows] zeroed, or nullptr
  int amax_rows;
};

// barrier among the 256 threads of one row (id 0 == __syncthreads of a 256-thread block)
__device__ __forceinline__ void row_bar(int id) { asm volatile("bar.sync %0, 256;" ::"r"(id) : "memory"); }

template <typename TIN>
__device__ __forceinline__ void load8(const TIN* p, float (&v)[8]);
template <>
__device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[8]) {
  int4 r = *reinterpret_cast<const int4*>(p);
  const uint32_t* w = reinterpret_cast<const uint32_t*>(&r);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    v[2 * j] = bf16_bits_to_float(w[j] & 0xffffu);
    v[2 * j + 1] = bf16_bits_to_float(w[j] >> 16);
  }
}
template <>
__device__ __forceinline__ void load8<float>(const float* p, float (&v)[8]) {
  float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w, v[4] = b.x, v[5] = b.y, v[6] = b.z, v[7] = b.w;
}

// max over the 256 threads of a row group; warp_max_s: 8 floats of shared memory owned by the group
__device__ __forceinline__ float row_max256(float amax, int t, float* warp_max_s, int bar_id) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, off));
  if ((t & 31) == 0) warp_max_s[t >> 5] = amax;
  row_bar(bar_id);
  amax = warp_max_s[0];
#pragma unroll
  for (int w = 1; w < 8; ++w) amax = fmaxf(amax, warp_max_s[w]);
  return amax;
}

// ---- per-row dynamic int8 quantisation (test_moe_int8.py:23-31) ---------------------------------
// absmax over the row, scale = absmax.clamp_min(1e-7) / 127, q = rint(x * (127 / absmax)) -- the same two f32
// expressions the reference oracle evaluates.  t = 0..255.
template <typename TIN>
__device__ __forceinline__ void quant_row_int8(const TIN* __restrict__ xr, int K, int8_t* __restrict__ qr,
                                               float* __restrict__ scale, int t, float* warp_max_s, int bar_id) {
  float amax = 0.f;
  const bool vec = (K % 8 == 0);
  if (vec) {
    for (int i = t * 8; i < K; i += 256 * 8) {
      float v[8];
      load8<TIN>(xr + i, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fabsf(v[j]));
    }
  } else {
    for (int i = t; i < K; i += 256) amax = fmaxf(amax, fabsf(load_as_float<TIN>(xr + i)));
  }
  amax = fmaxf(row_max256(amax, t, warp_max_s, bar_id), 1e-7f);
  // torch evaluates `127 / absmax` (Tensor.__rtruediv__) as reciprocal(absmax) * 127: two roundings
  const float inv = __fmul_rn(__frcp_rn(amax), 127.0f);
  if (t == 0) *scale = amax / 127.0f;
  if (vec) {
    for (int i = t * 8; i < K; i += 256 * 8) {
      float v[8];
      load8<TIN>(xr + i, v);
      uint32_t o[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        int q0 = __float2int_rn(v[4 * j] * inv), q1 = __float2int_rn(v[4 * j + 1] * inv);
        int q2 = __float2int_rn(v[4 * j + 2] * inv), q3 = __float2int_rn(v[4 * j + 3] * inv);
        o[j] = (uint32_t)(q0 & 0xff) | ((uint32_t)(q1 & 0xff) << 8) | ((uint32_t)(q2 & 0xff) << 16) |
               ((uint32_t)(q3 & 0xff) << 24);
      }
      *reinterpret_cast<uint2*>(qr + i) = make_uint2(o[0], o[1]);
    }
  } else {
    for (int i = t; i < K; i += 256) qr[i] = (int8_t)__float2int_rn(load_as_float<TIN>(xr + i) * inv);
  }
}

// ---- bf16 row -> (hi, lo) e4m3 pieces + the row's power-of-two factor (umma_f8.cu) -------------------
// x' = x * 2^(7 - e), e = floor(log2(max|x|)) so max|x'| is in [128, 256) (e4m3 max 448);
// hi = RN_e4m3(x'), lo = RN_e4m3(x' - hi), rexp = 2^(e - 7):  x ~= (hi + lo) * rexp.  K % 8 == 0.
__device__ __forceinline__ void split_row_e4m3(const __nv_bfloat16* __restrict__ xr, int K, uint8_t* __restrict__ hi,
                                               uint8_t* __restrict__ lo, float* __restrict__ rexp, int t,
                                               float* warp_max_s, int bar_id) {
  float amax = 0.f;
  for (int i = t * 8; i < K; i += 256 * 8) {
    const int4 r = *reinterpret_cast<const int4*>(xr + i);
    const uint32_t* wv = reinterpret_cast<const uint32_t*>(&r);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      amax = fmaxf(amax, fabsf(bf16_bits_to_float(wv[j] & 0xffffu)));
      amax = fmaxf(amax, fabsf(bf16_bits_to_float(wv[j] >> 16)));
    }
  }
  amax = row_max256(amax, t, warp_max_s, bar_id);
  int e = (int)((__float_as_uint(amax) >> 23) & 0xff) - 127;  // floor(log2) of a normal float
  if (!(amax > 0.f) || e < -100) e = -100;                   // all-zero / denormal rows: any factor works
  if (e > 120) e = 120;
  const float up = __uint_as_float((uint32_t)(127 + 7 - e) << 23);
  if (t == 0) *rexp = __uint_as_float((uint32_t)(127 + e - 7) << 23);
  for (int i = t * 8; i < K; i += 256 * 8) {
    const int4 r = *reinterpret_cast<const int4*>(xr + i);
    const uint32_t* wv = reinterpret_cast<const uint32_t*>(&r);
    uint32_t oh[2] = {0, 0}, ol[2] = {0, 0};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 v = make_float2(bf16_bits_to_float(wv[j] & 0xffffu) * up, bf16_bits_to_float(wv[j] >> 16) * up);
      const __nv_fp8x2_storage_t h = __nv_cvt_float2_to_fp8x2(v, __NV_SATFINITE, __NV_E4M3);
      const __half2_raw hr = __nv_cvt_fp8x2_to_halfraw2(h, __NV_E4M3);
      const float2 hf = __half22float2(*reinterpret_cast<const __half2*>(&hr));
      const __nv_fp8x2_storage_t l =
          __nv_cvt_float2_to_fp8x2(make_float2(v.x - hf.x, v.y - hf.y), __NV_SATFINITE, __NV_E4M3);
      oh[j >> 1] |= (uint32_t)h << (16 * (j & 1));
      ol[j >> 1] |= (uint32_t)l << (16 * (j & 1));
    }
    *reinterpret_cast<uint2*>(hi + i) = make_uint2(oh[0], oh[1]);
    *reinterpret_cast<uint2*>(lo + i) = make_uint2(ol[0], ol[1]);
  }
}

}  // namespace sglb
