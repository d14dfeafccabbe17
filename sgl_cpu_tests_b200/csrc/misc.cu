// Small HBM-bound kernels around the GEMMs: per-row int8 quantisation, the topk reduce (K5), the
// shared-expert axpy epilogue and the prepack re-tiling kernels (K7).
#include "common.cuh"
#include "rowops.cuh"

namespace sglb {



// ---- per-row dynamic int8 quantisation: one CTA per row (the row function lives in rowops.cuh) ----------
template <typename TIN>
__global__ void __launch_bounds__(256) quant_rows_kernel(const TIN* __restrict__ x, int K,
                                                         int8_t* __restrict__ q,
                                                         float* __restrict__ scale, const int32_t* __restrict__ n_valid,
                                                         int always_from, int64_t rows) {
  pdl_wait();
  pdl_trigger();
  __shared__ float warp_max_s[8];
  const int nv = n_valid != nullptr ? *n_valid : 0;
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    // act rows in expert-sorted order: [*n_valid, always_from) hold nothing (masked slots), see split_rows_e4m3_kernel
    if (n_valid != nullptr && row >= nv && row < always_from) continue;
    quant_row_int8<TIN>(x + row * K, K, q + row * K, scale + row, threadIdx.x, warp_max_s, 0);
    __syncthreads();  // warp_max_s is reused by the next row
  }
}

// ---- K5: out[m,:] = sum over the token's valid slots, in slot order j = 0..topk-1 ----------------
// f32 accumulation, one rounding to bf16 at the end.  Fixed order -> the result does not depend on
// which GPU produced which slot.
__global__ void __launch_bounds__(256)
    topk_reduce_kernel(const __nv_bfloat16* __restrict__ ic2, const int32_t* __restrict__ slot_pos,
                       int topk, int H, const __nv_bfloat16* __restrict__ shared_y, float rsf,
                       __nv_bfloat16* __restrict__ out) {
  pdl_wait();
  pdl_trigger();
  const int64_t m = blockIdx.y;
  const int h0 = (blockIdx.x * 256 + threadIdx.x) * 8;
  if (h0 >= H) return;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const bool full = (h0 + 8 <= H) && (H % 8 == 0);
  if (full) {
    // eight slots per pass: the positions, then the rows, are requested together (two dependent round trips per pass
    // instead of two per slot -- at decode this kernel is nothing but latency); the additions keep the slot order
    for (int j0 = 0; j0 < topk; j0 += 8) {
      int p[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) p[i] = j0 + i < topk ? slot_pos[m * topk + j0 + i] : -1;
      int4 v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i)
        v[i] = p[i] >= 0 ? *reinterpret_cast<const int4*>(ic2 + (size_t)p[i] * H + h0) : make_int4(0, 0, 0, 0);
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
  } else {
    for (int j = 0; j < topk; ++j) {
      const int p = slot_pos[m * topk + j];
      if (p < 0) continue;
      const __nv_bfloat16* src = ic2 + (size_t)p * H + h0;
      for (int i = 0; i < 8 && h0 + i < H; ++i) acc[i] += __bfloat162float(src[i]);
    }
  }
  if (shared_y != nullptr) {  // shared_expert_cpu's epilogue on the rounded sum (shared_axpy_kernel's arithmetic)
    const __nv_bfloat16* y = shared_y + (size_t)m * H + h0;
    for (int i = 0; i < 8 && h0 + i < H; ++i) acc[i] = fmaf(bf16_round(acc[i]), rsf, __bfloat162float(y[i]));
  }
  __nv_bfloat16* dst = out + (size_t)m * H + h0;
  if (full) {
    uint32_t o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) o[i] = pack_bf16x2(acc[2 * i], acc[2 * i + 1]);
    *reinterpret_cast<int4*>(dst) = make_int4(o[0], o[1], o[2], o[3]);
  } else {
    for (int i = 0; i < 8 && h0 + i < H; ++i) dst[i] = __float2bfloat16_rn(acc[i]);
  }
}

// ---- routed rows into expert-sorted order (CTA-pair kernels read them as plain TMA boxes) -------------------------
// y[p] = x[sorted_slot[p] / topk] for the valid positions p (sentinel entries = S are skipped); x2 / y2: a second
// array gathered the same way (the lo pieces of the fp8 split) or nullptr.
__global__ void __launch_bounds__(128)
    gather_rows_kernel(const int32_t* __restrict__ sorted_slot, int S, int topk, int row_bytes, const char* __restrict__ x,
                       const char* __restrict__ x2, char* __restrict__ y, char* __restrict__ y2) {
  pdl_wait();
  pdl_trigger();
  const int p = blockIdx.x;
  const int slot = sorted_slot[p];
  if (slot >= S) return;
  const size_t src = (size_t)(slot / topk) * row_bytes, dst = (size_t)p * row_bytes;
  for (int i = threadIdx.x * 16; i < row_bytes; i += 128 * 16) {
    *reinterpret_cast<int4*>(y + dst + i) = *reinterpret_cast<const int4*>(x + src + i);
    if (x2 != nullptr) *reinterpret_cast<int4*>(y2 + dst + i) = *reinterpret_cast<const int4*>(x2 + src + i);
  }
}

// ---- shared expert epilogue: out = y + fused_out * rsf ------------------------------------------
__global__ void __launch_bounds__(256)
    shared_axpy_kernel(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ fused,
                       float rsf, int64_t n, __nv_bfloat16* __restrict__ out) {
  pdl_wait();
  pdl_trigger();
  int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i < n) out[i] = __float2bfloat16_rn(fmaf(__bfloat162float(fused[i]), rsf, __bfloat162float(y[i])));
}

// ---- K7 prepack: row-major [N,K] <-> [N/32, K/V, 32, V] (element size EB bytes) -------------------
// One thread moves one V-element group (V*EB = 4 bytes).  Threads are ordered along the *packed*
// layout so the packed side is fully coalesced and the row-major side is read in 4-byte pieces.
__global__ void __launch_bounds__(256)
    repack_kernel(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst, int64_t E, int64_t N,
                  int64_t KG /* 4-byte groups per row */, int64_t packed_block_words,
                  int to_packed) {
  const int64_t per_e = N * KG;
  const int64_t total = E * per_e;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * 256) {
    const int64_t e = i / per_e;
    int64_t r = i % per_e;
    const int64_t nb = r / (KG * 32);
    r %= KG * 32;
    const int64_t kg = r / 32, nn = r % 32;
    const int64_t rm = e * per_e + (nb * 32 + nn) * KG + kg;
    const int64_t pk = (e * (N / 32) + nb) * packed_block_words + kg * 32 + nn;
    if (to_packed)
      dst[pk] = src[rm];
    else
      dst[rm] = src[pk];
  }
}

// comp[e,n] = 128 * sum_k w[e,n,k]; written behind each packed 32-row block when dst_packed != NULL.
__global__ void __launch_bounds__(128)
    int8_comp_kernel(const int8_t* __restrict__ w, int64_t N, int64_t K, int32_t* __restrict__ comp,
                     int64_t comp_block_stride_words, int64_t comp_block_offset_words) {
  const int64_t row = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);  // global row over E*N
  const int lane = threadIdx.x & 31;
  const int8_t* wr = w + row * K;
  int s = 0;
  for (int64_t k = lane; k < K; k += 32) s += wr[k];
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
  if (lane == 0) {
    const int64_t blk = row / 32, nn = row % 32;
    comp[blk * comp_block_stride_words + comp_block_offset_words + nn] = 128 * s;
  }
}

// e8m0 scales [E*N, G] <-> [E*N/32, G, 32]
__global__ void __launch_bounds__(256)
    repack_scale_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int64_t rows,
                        int64_t G, int to_packed) {
  const int64_t total = rows * G;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * 256) {
    const int64_t nb = i / (G * 32);
    const int64_t r = i % (G * 32);
    const int64_t gg = r / 32, nn = r % 32;
    const int64_t rm = (nb * 32 + nn) * G + gg;
    if (to_packed)
      dst[i] = src[rm];
    else
      dst[rm] = src[i];
  }
}

}  // namespace sglb

using namespace sglb;

extern "C" int sglb_quant_rows_int8(const void* x_bf16, int64_t rows, int K, int8_t* q,
                                    float* scale, sglb_stream_t stream) {
  SGLB_CHECK(x_bf16 && q && scale, "quant_rows: null pointer");
  SGLB_CHECK(K > 0 && rows >= 0, "quant_rows: bad shape");
  if (rows == 0) return SGLB_OK;
  SGLB_CUDA(launch_pdl(quant_rows_kernel<__nv_bfloat16>, dim3((unsigned)(rows < 16 * 148 ? rows : 16 * 148)), dim3(256), 0,
                       (cudaStream_t)stream, (const __nv_bfloat16*)x_bf16, K, q, scale, (const int32_t*)nullptr, 0, rows));
  count_launch();
  return SGLB_OK;
}

namespace sglb {
// f32 rows (the int8 path keeps silu_and_mul in f32 until it is re-quantised, test_moe_int8.py:84-90)
int launch_quant_rows_f32(const float* x, int64_t rows, int K, int8_t* q, float* scale,
                          cudaStream_t stream, const int32_t* n_valid, int always_from) {
  if (rows == 0) return SGLB_OK;
  SGLB_CUDA(launch_pdl(quant_rows_kernel<float>, dim3((unsigned)(rows < 16 * 148 ? rows : 16 * 148)), dim3(256), 0, stream, x,
                       K, q, scale, n_valid, always_from, rows));
  count_launch();
  return SGLB_OK;
}
}  // namespace sglb

namespace sglb {
int launch_topk_reduce(const void* ic2, const int32_t* slot_pos, int64_t M, int topk, int H, const void* shared_y,
                       float rsf, void* out, cudaStream_t stream) {
  SGLB_CHECK(ic2 && slot_pos && out, "topk_reduce: null pointer");
  SGLB_CHECK(M < 65536ll * 32768ll, "topk_reduce: M too large");
  if (M == 0) return SGLB_OK;
  // grid.y carries M (<= 65535 per launch)
  for (int64_t m0 = 0; m0 < M; m0 += 65535) {
    int64_t mm = M - m0 < 65535 ? M - m0 : 65535;
    dim3 grid((unsigned)ceil_div(ceil_div(H, 8), 256), (unsigned)mm);
    SGLB_CUDA(launch_pdl(topk_reduce_kernel, grid, dim3(256), 0, stream, (const __nv_bfloat16*)ic2,
                         slot_pos + m0 * topk, topk, H,
                         shared_y ? (const __nv_bfloat16*)shared_y + m0 * H : (const __nv_bfloat16*)nullptr, rsf,
                         (__nv_bfloat16*)out + m0 * H));
    count_launch();
  }
  return SGLB_OK;
}
}  // namespace sglb

extern "C" int sglb_topk_reduce(const void* ic2, const int32_t* slot_pos, int64_t M, int topk,
                                int H, void* out, sglb_stream_t stream) {
  return sglb::launch_topk_reduce(ic2, slot_pos, M, topk, H, nullptr, 1.f, out, (cudaStream_t)stream);
}

namespace sglb {
int launch_gather_rows(const int32_t* sorted_slot, int64_t S, int topk, int row_bytes, const void* x, const void* x2, void* y,
                       void* y2, cudaStream_t stream) {
  SGLB_CHECK(row_bytes % 16 == 0, "gather_rows: rows must be whole 16-byte pieces (%d bytes)", row_bytes);
  if (S == 0) return SGLB_OK;
  SGLB_CUDA(launch_pdl(gather_rows_kernel, dim3((unsigned)S), dim3(128), 0, stream, sorted_slot, (int)S, topk, row_bytes,
                       (const char*)x, (const char*)x2, (char*)y, (char*)y2));
  count_launch();
  return SGLB_OK;
}
int launch_shared_axpy(const void* y, const void* fused, float rsf, int64_t n, void* out,
                       cudaStream_t stream) {
  if (n == 0) return SGLB_OK;
  SGLB_CUDA(launch_pdl(shared_axpy_kernel, dim3((unsigned)ceil_div64(n, 256)), dim3(256), 0, stream,
                       (const __nv_bfloat16*)y, (const __nv_bfloat16*)fused, rsf, n, (__nv_bfloat16*)out));
  count_launch();
  return SGLB_OK;
}
}  // namespace sglb

extern "C" int sglb_repack_weight(const void* src, void* dst, int elem_bytes, int vnni, int64_t E,
                                  int64_t N, int64_t K, int int8_comp, int to_packed,
                                  sglb_stream_t stream) {
  SGLB_CHECK(src && dst, "repack_weight: null pointer");
  SGLB_CHECK(elem_bytes * vnni == 4, "repack_weight: elem_bytes*vnni must be 4 (got %d*%d)",
             elem_bytes, vnni);
  SGLB_CHECK(N % 32 == 0, "repack_weight: N=%lld must be a multiple of 32", (long long)N);
  SGLB_CHECK(K % vnni == 0, "repack_weight: K=%lld must be a multiple of %d", (long long)K, vnni);
  SGLB_CHECK(!int8_comp || elem_bytes == 1, "repack_weight: compensation is int8 only");
  const int64_t KG = K / vnni;
  const int64_t block_words = KG * 32 + (int8_comp ? 32 : 0);
  const int64_t total = E * N * KG;
  if (total == 0) return SGLB_OK;
  unsigned grid = (unsigned)(ceil_div64(total, 256) < 148 * 16 ? ceil_div64(total, 256) : 148 * 16);
  repack_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const uint32_t*)src, (uint32_t*)dst, E, N,
                                                       KG, block_words, to_packed);
  SGLB_LAUNCH_CHECK();
  count_launch();
  if (int8_comp && to_packed) {
    SGLB_CHECK((E * N) % 4 == 0, "repack_weight: bad row count");
    int8_comp_kernel<<<(unsigned)(E * N / 4), 128, 0, (cudaStream_t)stream>>>(
        (const int8_t*)src, N, K, (int32_t*)dst, block_words, KG * 32);
    SGLB_LAUNCH_CHECK();
    count_launch();
  }
  return SGLB_OK;
}

extern "C" int sglb_repack_scale(const void* src, void* dst, int64_t E, int64_t N, int64_t G,
                                 int to_packed, sglb_stream_t stream) {
  SGLB_CHECK(src && dst, "repack_scale: null pointer");
  SGLB_CHECK(N % 32 == 0, "repack_scale: N=%lld must be a multiple of 32", (long long)N);
  const int64_t total = E * N * G;
  if (total == 0) return SGLB_OK;
  unsigned grid = (unsigned)(ceil_div64(total, 256) < 148 * 16 ? ceil_div64(total, 256) : 148 * 16);
  repack_scale_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const uint8_t*)src, (uint8_t*)dst,
                                                             E * N, G, to_packed);
  SGLB_LAUNCH_CHECK();
  count_launch();
  return SGLB_OK;
}
