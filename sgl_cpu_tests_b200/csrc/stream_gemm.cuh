// K3/K4 (decode regime): weight-streaming grouped GEMM on warp-level MMA.
//
// Each expert that received rows is visited in chunks of 8*NG rows ("chunk_list" from moe_align).
// The weight matrix is the m16 operand (so one 16-B global load per lane feeds the tensor op
// directly, no shared-memory staging); the few routed rows are the n8 operand.  Every weight byte
// is read from HBM exactly once per chunk with 16-B vector loads marked L1::no_allocate; the routed
// rows are re-read from L1/L2.  For M_e <= 8*NG rows per expert the kernel is bound by HBM.
//
// Which k index sits in which MMA k-slot is irrelevant for a dot product as long as both operands
// agree, so lane (g,t) simply takes the 16 contiguous bytes at k0 + 16*t of weight row g (and the
// same k positions of routed row g) and feeds them to consecutive MMAs.
//
//   STAGE 1: out[p, n]  = silu(x_p . W13[e, n]) * (x_p . W13[e, I+n])        (fused silu_and_mul)
//   STAGE 2: out[p, h]  = (a_p . W2[e, h]) * topk_weight[slot(p)]            (slot-major ic2)
//            (row_scale == nullptr and an optional f32 bias turn stage 2 into a plain linear)
// p = position in expert-sorted order (sorted_slot), e = expert of the chunk.
#pragma once
#include "common.cuh"

namespace sglb {

struct StreamArgs {
  const void* W;           // [E, rows_per_expert, K] (mxfp4: K/2 bytes per row)
  const void* Wscale;      // int8: f32 [E, rows]; fp8: f32 [E, rows/bn, K/bk]; mxfp4: u8 [E, rows, K/32]
  const void* X;           // routed rows: bf16 (or s8 for int8) with row stride K
  const float* Xscale;     // int8 only: per X row
  const int32_t* chunk_list;
  const int32_t* counters;
  const int32_t* sorted_slot;
  const float* row_scale;  // stage 2: topk_weights flat [S] (indexed by slot), else nullptr
  const float* bias;       // stage 2 optional f32 [N]
  void* out;               // bf16 [*, N]  (stage 1: N = I; f32 for int8 stage 1)
  int K;
  int N;                   // outputs per row: I (stage 1), H (stage 2)
  int rows_per_expert;     // 2I (stage 1), H (stage 2)
  int topk;
  int gather;              // 1: X row = sorted_slot[p] / topk ; 0: X row = p
  int block_n, block_k;    // fp8 scale block
  int scale_cols;          // fp8: ceil(K/bk); mxfp4: K/32
  const void* X2;          // tcgen05 fp8 path only (see UmmaArgs)
  const float* ebias;      // per-expert f32 bias [E, rows_per_expert] or nullptr (gpt-oss w1_bias / w2_bias): added to
                           // the dequantised accumulator before the activation (stage 1) / before row_scale (stage 2)
  int act;                 // stage 1: SGLB_ACT_SILU or SGLB_ACT_SWIGLU_OAI (common.cuh glu_f)
  float act_alpha, act_limit;
  int out_by_slot;         // stage 2: write routed row p to row sorted_slot[p] (through out_map if set) instead of p
  const int32_t* out_map;
  const void* Wsh;         // tcgen05 paths only: the shared expert riding along (see UmmaArgs)
  const void* Wsh_scale;
  int shared_row0;
  FusedReduce red;         // stage 2: fused top-k reduce, one counter per (token, 32-column warp block); cnt == nullptr = off
  uint32_t* row_amax;      // tcgen05 int8 path only (see UmmaArgs)
  const float* Xraw;
};

template <int FMT>
struct FmtTraits;
template <>
struct FmtTraits<SGLB_FMT_BF16> {
  static constexpr int PE = 8;       // weight elements per 16-B piece
  static constexpr int XV = 1;       // int4 loads of X per piece
  static constexpr int WBITS = 16;
  using acc_t = float;
};
template <>
struct FmtTraits<SGLB_FMT_INT8_W8A8> {
  static constexpr int PE = 16;
  static constexpr int XV = 1;
  static constexpr int WBITS = 8;
  using acc_t = int;
};
template <>
struct FmtTraits<SGLB_FMT_FP8_W8A16> {
  static constexpr int PE = 16;
  static constexpr int XV = 2;
  static constexpr int WBITS = 8;
  using acc_t = float;
};
template <>
struct FmtTraits<SGLB_FMT_MXFP4> {
  static constexpr int PE = 32;
  static constexpr int XV = 4;
  static constexpr int WBITS = 4;
  using acc_t = float;
};

template <int FMT, int NG>
struct UnitFrag {
  int4 w[2][2];                        // [block][row g / g+8]
  int4 x[NG][FmtTraits<FMT>::XV];
  float sc[2][2];                      // fp8: f32 block scales; mxfp4: e8m0 byte (as float bits)
};

template <int FMT, int STAGE, int NG, int WN, int WK, int PD>
__global__ void __launch_bounds__(WN* WK * 32)
    stream_gemm_kernel(const StreamArgs a) {
  using T = FmtTraits<FMT>;
  using acc_t = typename T::acc_t;
  constexpr int PE = T::PE;
  constexpr int U = 4 * PE;  // k elements per unit (one 16-B load per lane and weight row)
  constexpr int XV = T::XV;

  pdl_wait();
  pdl_trigger();
  if constexpr (STAGE == 2)
    fused_reduce_zero_rows(a.red, a.N, (blockIdx.y * gridDim.x + blockIdx.x) * (WN * WK) + (threadIdx.x >> 5),
                           gridDim.x * gridDim.y * (WN * WK), threadIdx.x & 31);
  const int chunk = blockIdx.y;
  if (chunk >= a.counters[1]) return;
  const int e = a.chunk_list[3 * chunk + 0];
  const int row_start = a.chunk_list[3 * chunk + 1];
  const int rows = a.chunk_list[3 * chunk + 2];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int wn = warp % WN, wk = warp / WN;
  const int K = a.K;

  // ---- weight rows of this warp: two 16-row blocks ----
  int nbase[2];
  if (STAGE == 1) {
    nbase[0] = (blockIdx.x * WN + wn) * 16;  // gate rows
    nbase[1] = a.N + nbase[0];               // matching up rows (N = I)
  } else {
    nbase[0] = (blockIdx.x * WN + wn) * 32;
    nbase[1] = nbase[0] + 16;
  }
  const bool warp_active = (STAGE == 1) ? (nbase[0] < a.N) : (nbase[0] < a.N);
  const int row_limit = a.rows_per_expert;

  const size_t w_row_bytes = (size_t)K * T::WBITS / 8;
  const char* wptr[2][2];
  int wrow[2][2];
#pragma unroll
  for (int b = 0; b < 2; ++b)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      int r = nbase[b] + g + 8 * h;
      r = min(r, row_limit - 1);
      wrow[b][h] = r;
      wptr[b][h] = (const char*)a.W + ((size_t)e * row_limit + r) * w_row_bytes + t * 16;
    }

  // ---- routed rows of this lane (n8 operand): row g of each group ----
  const size_t x_elem_bytes = (FMT == SGLB_FMT_INT8_W8A8) ? 1 : 2;
  const char* xptr[NG];
  int xrow_idx[NG];
#pragma unroll
  for (int q = 0; q < NG; ++q) {
    int r = min(q * 8 + g, rows - 1);
    int p = row_start + r;
    int xr = a.gather ? (a.sorted_slot[p] / a.topk) : p;
    xrow_idx[q] = xr;
    xptr[q] = (const char*)a.X + (size_t)xr * K * x_elem_bytes + (size_t)t * PE * x_elem_bytes;
  }

  // ---- k slice of this warp ----
  int kslice = ((K + WK - 1) / WK + U - 1) / U * U;
  const int kbeg = wk * kslice;
  const int kend = min(K, kbeg + kslice);

  acc_t acc[2][NG][4];
#pragma unroll
  for (int b = 0; b < 2; ++b)
#pragma unroll
    for (int q = 0; q < NG; ++q)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[b][q][c] = 0;

  auto load_unit = [&](UnitFrag<FMT, NG>& f, int k) {
    const bool ok = warp_active && (k + t * PE < kend);
    const size_t wb = (size_t)k * T::WBITS / 8;
#pragma unroll
    for (int b = 0; b < 2; ++b)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        f.w[b][h] = ok ? ldg_stream16(wptr[b][h] + wb) : make_int4(0, 0, 0, 0);
      }
#pragma unroll
    for (int q = 0; q < NG; ++q)
#pragma unroll
      for (int v = 0; v < XV; ++v) {
        f.x[q][v] = ok ? ldg_cached16(xptr[q] + (size_t)k * x_elem_bytes + v * 16)
                       : make_int4(0, 0, 0, 0);
      }
    if (FMT == SGLB_FMT_FP8_W8A16) {
      const float* sc = (const float*)a.Wscale;
      const int kb = k / a.block_k;  // the whole unit lies in one scale block (block_k % 64 == 0)
      const bool uok = warp_active && (k < kend);
      const int nblk = (row_limit + a.block_n - 1) / a.block_n;
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int h = 0; h < 2; ++h)
          f.sc[b][h] = uok ? __ldg(sc + ((size_t)e * nblk + wrow[b][h] / a.block_n) * a.scale_cols + kb)
                           : 0.f;
    } else if (FMT == SGLB_FMT_MXFP4) {
      const uint8_t* sc = (const uint8_t*)a.Wscale;
      const int kg = (k + t * PE) / 32;
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int h = 0; h < 2; ++h)
          f.sc[b][h] = ok ? __uint_as_float((uint32_t)__ldg(
                                sc + ((size_t)e * row_limit + wrow[b][h]) * a.scale_cols + kg))
                          : 0.f;
    }
  };

  auto compute_unit = [&](const UnitFrag<FMT, NG>& f) {
    if constexpr (FMT == SGLB_FMT_BF16) {
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int q = 0; q < NG; ++q) {
          mma_bf16_16816(acc[b][q], f.w[b][0].x, f.w[b][1].x, f.w[b][0].y, f.w[b][1].y,
                         f.x[q][0].x, f.x[q][0].y);
          mma_bf16_16816(acc[b][q], f.w[b][0].z, f.w[b][1].z, f.w[b][0].w, f.w[b][1].w,
                         f.x[q][0].z, f.x[q][0].w);
        }
    } else if constexpr (FMT == SGLB_FMT_INT8_W8A8) {
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int q = 0; q < NG; ++q) {
          mma_s8_16832(acc[b][q], f.w[b][0].x, f.w[b][1].x, f.w[b][0].y, f.w[b][1].y, f.x[q][0].x,
                       f.x[q][0].y);
          mma_s8_16832(acc[b][q], f.w[b][0].z, f.w[b][1].z, f.w[b][0].w, f.w[b][1].w, f.x[q][0].z,
                       f.x[q][0].w);
        }
    } else if constexpr (FMT == SGLB_FMT_FP8_W8A16) {
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        uint32_t wg[8], wh[8];
        const uint32_t* pg = reinterpret_cast<const uint32_t*>(&f.w[b][0]);
        const uint32_t* ph = reinterpret_cast<const uint32_t*>(&f.w[b][1]);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          e4m3x4_to_bf16x4(pg[i], wg[2 * i], wg[2 * i + 1]);
          e4m3x4_to_bf16x4(ph[i], wh[2 * i], wh[2 * i + 1]);
        }
#pragma unroll
        for (int q = 0; q < NG; ++q) {
          float part[4] = {0.f, 0.f, 0.f, 0.f};
          const uint32_t* px = reinterpret_cast<const uint32_t*>(&f.x[q][0]);
#pragma unroll
          for (int i = 0; i < 4; ++i)
            mma_bf16_16816(part, wg[2 * i], wh[2 * i], wg[2 * i + 1], wh[2 * i + 1], px[2 * i],
                           px[2 * i + 1]);
          acc[b][q][0] = fmaf(part[0], f.sc[b][0], acc[b][q][0]);
          acc[b][q][1] = fmaf(part[1], f.sc[b][0], acc[b][q][1]);
          acc[b][q][2] = fmaf(part[2], f.sc[b][1], acc[b][q][2]);
          acc[b][q][3] = fmaf(part[3], f.sc[b][1], acc[b][q][3]);
        }
      }
    } else {  // MXFP4
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const uint32_t sg = e8m0_to_bf16x2(__float_as_uint(f.sc[b][0]));
        const uint32_t sh = e8m0_to_bf16x2(__float_as_uint(f.sc[b][1]));
        const uint32_t* pg = reinterpret_cast<const uint32_t*>(&f.w[b][0]);
        const uint32_t* ph = reinterpret_cast<const uint32_t*>(&f.w[b][1]);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint32_t rg[4], rh[4];
          e2m1x8_to_raw_bf16(pg[i], rg);
          e2m1x8_to_raw_bf16(ph[i], rh);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            rg[j] = mxfp4_finish(rg[j], sg);
            rh[j] = mxfp4_finish(rh[j], sh);
          }
#pragma unroll
          for (int q = 0; q < NG; ++q) {
            const uint32_t* px = reinterpret_cast<const uint32_t*>(&f.x[q][0]) + 4 * i;
            uint32_t b00 = __byte_perm(px[0], px[2], 0x5410), b01 = __byte_perm(px[0], px[2], 0x7632);
            uint32_t b10 = __byte_perm(px[1], px[3], 0x5410), b11 = __byte_perm(px[1], px[3], 0x7632);
            mma_bf16_16816(acc[b][q], rg[0], rh[0], rg[1], rh[1], b00, b01);
            mma_bf16_16816(acc[b][q], rg[2], rh[2], rg[3], rh[3], b10, b11);
          }
        }
      }
    }
  };

  // ---- main loop: PD units in flight per lane ----
  UnitFrag<FMT, NG> f[PD];
#pragma unroll
  for (int p = 0; p < PD; ++p) load_unit(f[p], kbeg + p * U);
  for (int k = kbeg; k < kend; k += PD * U) {
#pragma unroll
    for (int p = 0; p < PD; ++p) {
      if (k + p * U < kend) compute_unit(f[p]);
      load_unit(f[p], k + (PD + p) * U);
    }
  }

  // ---- split-K reduction across the WK warps that share the same rows ----
  if constexpr (WK > 1) {
    __shared__ acc_t red[(WK - 1) * WN * 32 * 2 * NG * 4];
    if (wk > 0) {
      acc_t* dst = red + (((wk - 1) * WN + wn) * 32 + lane) * (2 * NG * 4);
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int q = 0; q < NG; ++q)
#pragma unroll
          for (int c = 0; c < 4; ++c) dst[(b * NG + q) * 4 + c] = acc[b][q][c];
    }
    __syncthreads();
    if (wk > 0) return;
#pragma unroll
    for (int s = 0; s < WK - 1; ++s) {
      const acc_t* src = red + ((s * WN + wn) * 32 + lane) * (2 * NG * 4);
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int q = 0; q < NG; ++q)
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[b][q][c] += src[(b * NG + q) * 4 + c];
    }
  }
  if (!warp_active) return;

  // ---- epilogue ----
  __nv_bfloat16* out = (__nv_bfloat16*)a.out;
#pragma unroll
  for (int q = 0; q < NG; ++q)
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int r = q * 8 + 2 * t + (c & 1);
      if (r >= rows) continue;
      const int p = row_start + r;
      const int rofs = g + ((c & 2) ? 8 : 0);
      float xs = 1.f;
      if (FMT == SGLB_FMT_INT8_W8A8) {
        int xr = a.gather ? (a.sorted_slot[p] / a.topk) : p;
        xs = a.Xscale[xr];
      }
      if (STAGE == 1) {
        const int n = nbase[0] + rofs;
        if (n >= a.N) continue;
        float gate = (float)acc[0][q][c], up = (float)acc[1][q][c];
        if (FMT == SGLB_FMT_INT8_W8A8) {
          const float* ws = (const float*)a.Wscale + (size_t)e * row_limit;
          gate = xs * gate * ws[n];
          up = xs * up * ws[a.N + n];
        }
        if (a.ebias != nullptr) {
          const float* eb = a.ebias + (size_t)e * a.rows_per_expert;
          gate += eb[n];
          up += eb[a.N + n];
        }
        const float actv = glu_f(gate, up, a.act, a.act_alpha, a.act_limit);
        if (FMT == SGLB_FMT_INT8_W8A8)
          reinterpret_cast<float*>(a.out)[(size_t)p * a.N + n] = actv;
        else
          out[(size_t)p * a.N + n] = __float2bfloat16_rn(actv);
      } else {
        float rs = 1.f;
        if (a.row_scale != nullptr) rs = a.row_scale[a.sorted_slot[p]];
        const int64_t orow = out_row_of(p, a.sorted_slot, a.out_by_slot, a.out_map);
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          const int n = nbase[b] + rofs;
          if (n >= a.N) continue;
          float v = (float)acc[b][q][c];
          if (FMT == SGLB_FMT_INT8_W8A8) {
            const float* ws = (const float*)a.Wscale + (size_t)e * row_limit;
            v = xs * v * ws[n];
          }
          if (a.bias != nullptr) v += a.bias[n];
          if (a.ebias != nullptr) v += a.ebias[(size_t)e * a.rows_per_expert + n];
          out[(size_t)orow * a.N + n] = __float2bfloat16_rn(v * rs);
        }
      }
    }
  // ---- fused top-k reduce (decode chain fusion): this warp finished 32 columns of the chunk's rows ----
  if constexpr (STAGE == 2) {
    if (a.red.cnt != nullptr) {  // kernel-uniform
      __threadfence();
      __syncwarp();
      bool done = false;
      int tok = 0;
      if (lane < rows) {
        tok = a.sorted_slot[row_start + lane] / a.topk;
        __threadfence();
        done = atomicAdd(a.red.cnt + (size_t)tok * a.red.blocks + (blockIdx.x * WN + wn), 1) + 1 == a.red.need;
        if (done) __threadfence();
      }
      unsigned mask = __ballot_sync(0xffffffffu, done);
      const int col0 = nbase[0] + lane * 8;
      while (mask) {
        const int srcl = __ffs(mask) - 1;
        mask &= mask - 1;
        const int tkn = __shfl_sync(0xffffffffu, tok, srcl);
        if (lane < 4 && col0 < a.N) fused_reduce_cols8(a.red, (const __nv_bfloat16*)a.out, a.topk, a.N, tkn, col0);
      }
    }
  }
}

// Host-side launcher shared by the per-format translation units.
struct StreamPlan {
  int ng;         // 1, 2 or 4 row groups (chunk rows = 8*ng)
  int cfg;        // 0: WN=8,WK=1   1: WN=2,WK=4   2: WN=1,WK=8
  int max_chunks; // grid.y
};

template <int FMT, int STAGE>
int launch_stream_gemm(const StreamArgs& a, const StreamPlan& plan, cudaStream_t stream);

}  // namespace sglb

// ---- launcher (included by the per-format translation units) -------------------------------
namespace sglb {

template <int FMT, int NG>
constexpr int stream_prefetch_depth() {
  if (FMT == SGLB_FMT_BF16 || FMT == SGLB_FMT_INT8_W8A8) return NG <= 2 ? 4 : 2;
  if (FMT == SGLB_FMT_FP8_W8A16) return NG <= 2 ? 2 : 1;
  return NG <= 1 ? 2 : 1;
}

template <int FMT, int STAGE, int NG, int WN, int WK>
static int stream_launch_one(const StreamArgs& a, int max_chunks, cudaStream_t stream) {
  constexpr int PD = stream_prefetch_depth<FMT, NG>();
  const int rows_per_cta = (STAGE == 1 ? 16 : 32) * WN;
  dim3 grid((unsigned)ceil_div(a.N, rows_per_cta), (unsigned)max_chunks);
  SGLB_CUDA(launch_pdl(stream_gemm_kernel<FMT, STAGE, NG, WN, WK, PD>, grid, dim3(WN * WK * 32), 0, stream, a));
  return SGLB_OK;
}

template <int FMT, int STAGE, int NG>
static int stream_launch_cfg(const StreamArgs& a, const StreamPlan& plan, cudaStream_t stream) {
  switch (plan.cfg) {
    case 0: return stream_launch_one<FMT, STAGE, NG, 8, 1>(a, plan.max_chunks, stream);
    case 1: return stream_launch_one<FMT, STAGE, NG, 2, 4>(a, plan.max_chunks, stream);
    default: return stream_launch_one<FMT, STAGE, NG, 1, 8>(a, plan.max_chunks, stream);
  }
}

#define SGLB_INSTANTIATE_STREAM(FMT)                                                              \
  template <>                                                                                     \
  int launch_stream_gemm<FMT, 1>(const StreamArgs& a, const StreamPlan& plan, cudaStream_t s) {   \
    switch (plan.ng) {                                                                            \
      case 1: return stream_launch_cfg<FMT, 1, 1>(a, plan, s);                                    \
      case 2: return stream_launch_cfg<FMT, 1, 2>(a, plan, s);                                    \
      default: return stream_launch_cfg<FMT, 1, 4>(a, plan, s);                                   \
    }                                                                                             \
  }                                                                                               \
  template <>                                                                                     \
  int launch_stream_gemm<FMT, 2>(const StreamArgs& a, const StreamPlan& plan, cudaStream_t s) {   \
    switch (plan.ng) {                                                                            \
      case 1: return stream_launch_cfg<FMT, 2, 1>(a, plan, s);                                    \
      case 2: return stream_launch_cfg<FMT, 2, 2>(a, plan, s);                                    \
      default: return stream_launch_cfg<FMT, 2, 4>(a, plan, s);                                   \
    }                                                                                             \
  }

}  // namespace sglb
