// Shared pieces of the tcgen05 grouped-GEMM kernels (umma_gemm.cu: operands straight from HBM;
// umma_dq.cu: weights dequantised to bf16 on the way): PTX wrappers, descriptors, the work-item
// decode, the routed-row gather producers and the epilogue warps.
#pragma once
#include <cuda.h>  // CUtensorMap and enums only; the encoder is resolved through cudart, no -lcuda

#include "umma_gemm.cuh"

namespace sglb {
namespace ummak {


constexpr int kRowBytes = 128;                       // bytes of K per operand row per stage
constexpr int kWSub = 128;                           // weight rows per accumulator (UMMA M)
constexpr int kWTileBytes = kWSub * kRowBytes;       // 16 KB
constexpr int kXTileBytes = kUmmaTileRows * kRowBytes;
constexpr int kGatherThreads = 128;
constexpr int kEpiWarps = 8;                         // two per TMEM lane quadrant (one per 16-column parity)
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kTmemCols = 512;
constexpr int kAccCols = 256;                        // per accumulator buffer: 2 sub-tiles x 128
constexpr int kEpiBufBytes = 16 * 32 * 4;            // per epilogue warp: 16 tokens x 32 outputs, f32 worst case

// ---- PTX wrappers ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!ok);
}
// Wait of a role that is not on the critical path (producers waiting for a free slot, the MMA
// issuer and the epilogue waiting for data in the HBM-bound regime).  A failed poll sleeps so the
// spinning warp stops competing for issue slots with the warps doing ALU work (ncu, r1d: try_wait
// spin loops were 49 % of all issued instructions of the fp8 kernel); the try_wait itself carries a
// suspend-time hint so the hardware may park the warp until the phase completes.
__device__ __forceinline__ void mbar_wait_relaxed(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity), "r"(100000u)
        : "memory");
    if (ok) break;
    __nanosleep(32);
  }
}
// one lane of a converged warp; ptxas keeps tcgen05 / TMA instructions guarded by it on the uniform
// datapath without the per-lane "waterfall" loop it emits under an `if (lane == 0)` branch
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, "
      "%4, %5}], [%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
// mbarrier arrive (counted in the barrier's expected arrivals) once all cp.async issued so far by this
// thread have landed: the gather producers never block on their own loads
__device__ __forceinline__ void cp_async_arrive_noinc(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; issued by one thread on behalf of the CTA
template <int FMT>
__device__ __forceinline__ void umma(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                     uint32_t accumulate) {
  if constexpr (FMT == SGLB_FMT_INT8_W8A8) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// mbarrier arrive once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
               : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, "
      "%13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]),
        "=r"(v[15])
      : "r"(taddr));
}
// The loaded registers are threaded through the wait as "+r" operands so the compiler cannot
// schedule a use of them above it.
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&v)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]),
                 "+r"(v[7]), "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]),
                 "+r"(v[14]), "+r"(v[15])
               :
               : "memory");
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor (sm_100 format): start address >> 4,
// LBO = 1 (unused for swizzled K-major), SBO = 1024 B between 8-row groups, version 1, layout 2.
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t addr) {
  return (uint64_t)((addr >> 4) & 0x3FFFu) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) |
         (1ull << 46) | (2ull << 61);
}
// instruction descriptor: D f32 (or s32), A and B K-major, M = 128, N = n
template <int FMT>
__device__ __forceinline__ uint32_t instr_desc(int n) {
  const uint32_t cfmt = FMT == SGLB_FMT_INT8_W8A8 ? 2u : 1u;  // S32 : F32
  const uint32_t abfmt = 1u;                                  // kind::f16: BF16; kind::i8: signed
  return (cfmt << 4) | (abfmt << 7) | (abfmt << 10) | ((uint32_t)(n >> 3) << 17) |
         ((uint32_t)(kWSub >> 4) << 24);
}

__device__ __forceinline__ float sel4(const float (&v)[4], int i) {
  return i == 0 ? v[0] : (i == 1 ? v[1] : (i == 2 ? v[2] : v[3]));
}

struct WorkItem {
  int e, row_start, rows, nb;
};

template <int STAGE>
__device__ __forceinline__ int num_weight_blocks(const UmmaArgs& a) {
  return STAGE == 1 ? ceil_div(a.N, kWSub) : ceil_div(a.N, 2 * kWSub);
}
__device__ __forceinline__ WorkItem get_item(const UmmaArgs& a, int w, int n_blocks) {
  WorkItem it;
  const int t = w / n_blocks;
  it.nb = w - t * n_blocks;
  it.e = a.tile_list[3 * t + 0];
  it.row_start = a.tile_list[3 * t + 1];
  it.rows = a.tile_list[3 * t + 2];
  return it;
}


// ---- gather producers: routed rows -> 128-B-swizzled K-major smem tiles ------------------------------
// 128 threads (g = 0..127).  Thread g owns the 16-B chunk c = g & 7 of rows rbase + 16 j.  Stage s of the
// ring lives at x_tile0 + s * x_stride; full/empty barriers at xfull0 + 8 s / xempty0 + 8 s.
template <bool kRelaxed>
__device__ __forceinline__ void gather_role(const UmmaArgs& a, int total, int n_blocks, int num_kb,
                                            int k_bytes, int g, uint32_t x_tile0, uint32_t x_stride,
                                            uint32_t xfull0, uint32_t xempty0, int nstages) {
  auto x_tile = [&](int s) { return x_tile0 + s * x_stride; };
  auto full_bar = [&](int s) { return xfull0 + 8 * s; };
  auto empty_bar = [&](int s) { return xempty0 + 8 * s; };
  const int kStages = nstages;
    const int c = g & 7, rbase = g >> 3;  // 16-B chunk of the 128-B row; rows rbase + 16*j
    const uint32_t dst0 = rbase * kRowBytes + ((c ^ (rbase & 7)) << 4);
    uint32_t it = 0;
    // Two-level lookahead over the dependent loads tile_list -> sorted_slot: the descriptor of item i + 2 and
    // the row indices of item i + 1 are fetched while item i is copied.
    const int step = gridDim.x;
    auto rows_of = [&](const WorkItem& wi, int (&xr)[8]) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int p = wi.row_start + min(rbase + 16 * j, wi.rows - 1);
        xr[j] = a.gather ? a.sorted_slot[p] : p;  // raw slot: the division waits until the value is used
      }
    };
    const int w0 = blockIdx.x;
    // no item for this CTA (fewer items than CTAs, or every id masked: then tile_list holds nothing at all and the
    // lookahead below would chase uninitialised row indices)
    if (w0 >= total) return;
    WorkItem cur = get_item(a, w0, n_blocks);
    WorkItem nxt = get_item(a, w0 + step < total ? w0 + step : 0, n_blocks);
    int xr_cur[8];
    rows_of(cur, xr_cur);
    for (int w = w0; w < total; w += step) {
      const WorkItem wi = cur;
      const int n_eff = (wi.rows + 15) & ~15;
      const char* src[8];
#pragma unroll
      for (int j = 0; j < 8; ++j)
        src[j] = (const char*)a.X + (size_t)(a.gather ? xr_cur[j] / a.topk : xr_cur[j]) * k_bytes + c * 16;
      cur = nxt;
      if (w + step < total) rows_of(cur, xr_cur);  // consumed by the next iteration
      if (w + 2 * step < total) nxt = get_item(a, w + 2 * step, n_blocks);
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        const int s = it % kStages;
        if (kRelaxed) mbar_wait_relaxed(empty_bar(s), ((it / kStages) & 1) ^ 1);
        else mbar_wait(empty_bar(s), ((it / kStages) & 1) ^ 1);
        const uint32_t dst = x_tile(s) + dst0;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (rbase + 16 * j < n_eff) cp_async16(dst + j * 16 * kRowBytes, src[j] + kb * kRowBytes);
        cp_async_arrive_noinc(full_bar(s));
      }
    }
}

// ---- int8 stage 2 in the fused decode chain: the gather warps quantise f32 act rows on the way in ----------
// Same thread -> (row, 16-byte chunk) map as gather_role.  A chunk is 16 s8 of K = 64 bytes of f32 act; the row's
// scale comes from a.row_amax (maxed by the stage-1 epilogue): inv = rcp(max(amax, 1e-7)) * 127 and
// q = rint(v * inv), the two f32 expressions of quant_rows_kernel (misc.cu), so the operand bytes are the ones the
// separate quantisation launch would have written.
__device__ __forceinline__ void gather_quant_role(const UmmaArgs& a, int total, int n_blocks, int num_kb, int g,
                                                  uint32_t x_tile0, uint32_t x_stride, uint32_t xfull0, uint32_t xempty0,
                                                  int nstages) {
  const int c = g & 7, rbase = g >> 3;
  const uint32_t dst0 = rbase * kRowBytes + ((c ^ (rbase & 7)) << 4);
  uint32_t it = 0;
  for (int w = blockIdx.x; w < total; w += gridDim.x) {
    const WorkItem wi = get_item(a, w, n_blocks);
    const int n_eff = (wi.rows + 15) & ~15;
    const float* src[8];
    float inv[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int p = wi.row_start + min(rbase + 16 * j, wi.rows - 1);
      src[j] = a.Xraw + (size_t)p * a.K + c * 16;
      inv[j] = 0.f;
      if (rbase + 16 * j < n_eff) inv[j] = __fmul_rn(__frcp_rn(fmaxf(__uint_as_float(a.row_amax[p]), 1e-7f)), 127.0f);
    }
    // rows j = 0, 1 (all there is at decode: tiles of up to 32 rows): the 64 bytes of k block kb + 1 are requested
    // before k block kb is converted, so an L2 round trip never sits between two stages
    auto load_pair = [&](float4 (&v)[2][4], int j0, int kb) {
#pragma unroll
      for (int jj = 0; jj < 2; ++jj)
        if (rbase + 16 * (j0 + jj) < n_eff) {
          const float4* p4 = reinterpret_cast<const float4*>(src[j0 + jj] + kb * kRowBytes);
#pragma unroll
          for (int i = 0; i < 4; ++i) v[jj][i] = __ldcg(p4 + i);
        }
    };
    auto store_pair = [&](const float4 (&v)[2][4], int j0, uint32_t dst) {
#pragma unroll
      for (int jj = 0; jj < 2; ++jj)
        if (rbase + 16 * (j0 + jj) < n_eff) {
          const float sc = inv[j0 + jj];
          uint32_t o[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int q0 = __float2int_rn(v[jj][i].x * sc), q1 = __float2int_rn(v[jj][i].y * sc);
            const int q2 = __float2int_rn(v[jj][i].z * sc), q3 = __float2int_rn(v[jj][i].w * sc);
            o[i] = (uint32_t)(q0 & 0xff) | ((uint32_t)(q1 & 0xff) << 8) | ((uint32_t)(q2 & 0xff) << 16) |
                   ((uint32_t)(q3 & 0xff) << 24);
          }
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + (j0 + jj) * 16 * kRowBytes), "r"(o[0]),
                       "r"(o[1]), "r"(o[2]), "r"(o[3])
                       : "memory");
        }
    };
    float4 nxt[2][4];
    load_pair(nxt, 0, 0);
    for (int kb = 0; kb < num_kb; ++kb, ++it) {
      const int s = it % nstages;
      float4 cur[2][4];
#pragma unroll
      for (int jj = 0; jj < 2; ++jj)
#pragma unroll
        for (int i = 0; i < 4; ++i) cur[jj][i] = nxt[jj][i];
      if (kb + 1 < num_kb) load_pair(nxt, 0, kb + 1);
      mbar_wait(xempty0 + 8 * s, ((it / nstages) & 1) ^ 1);
      const uint32_t dst = x_tile0 + s * x_stride + dst0;
      store_pair(cur, 0, dst);
#pragma unroll
      for (int j0 = 2; j0 < 8; j0 += 2) {
        if (rbase + 16 * j0 < n_eff) {  // larger tiles (skewed routing): plain load -> convert -> store
          float4 v[2][4];
          load_pair(v, j0, kb);
          store_pair(v, j0, dst);
        }
      }
      fence_proxy_async();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
      mbar_arrive(xfull0 + 8 * s);
    }
  }
}

// ---- fused top-k reduce: a dedicated reducer warp (stage 2 of umma_gemm.cu in the fused decode chain) -------
// The epilogue warps only ARRIVE on red_full[b] after the stores of a work item (and wait on red_empty[b] before they
// reuse the slot two items later), so the L2 round trips of the reduce -- fence, arrival atomics, the slot rows --
// never sit on the path that frees TMEM.  The reducer counts each routed row of the tile on red.cnt[token, nb];
// tokens it completes are reduced by the whole warp, 8 columns per lane (256 per block).  Barriers at red_bar0:
// full[0], full[1], empty[0], empty[1].
__device__ __forceinline__ void reducer_role(const UmmaArgs& a, int total, int n_blocks, uint32_t red_bar0, int lane) {
  uint32_t li = 0;
  for (int w = blockIdx.x; w < total; w += gridDim.x, ++li) {
    const WorkItem wi = get_item(a, w, n_blocks);
    const int b = li & 1;
    int tok[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {  // before the wait: the dependent tile_list -> sorted_slot loads hide under the item
      const int r = lane + 32 * i;
      tok[i] = r < wi.rows ? a.sorted_slot[wi.row_start + r] / a.topk : -1;
    }
    mbar_wait_relaxed(red_bar0 + 8 * b, (li >> 1) & 1);  // every epilogue thread has issued its stores of the item
    __threadfence();  // cumulative over the stores observed through the barrier (the grid-sync pattern)
    int old[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
      old[i] = tok[i] >= 0 ? atomicAdd(a.red.cnt + (size_t)tok[i] * a.red.blocks + wi.nb, 1) : -2;
    bool any = false;
#pragma unroll
    for (int i = 0; i < 4; ++i) any |= old[i] + 1 == a.red.need;
    if (any) __threadfence();
    const int col0 = wi.nb * 2 * kWSub + lane * 8;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      unsigned mask = __ballot_sync(0xffffffffu, old[i] + 1 == a.red.need);
      while (mask) {
        const int srcl = __ffs(mask) - 1;
        mask &= mask - 1;
        const int t = __shfl_sync(0xffffffffu, tok[i], srcl);
        if (col0 < a.N) fused_reduce_cols8(a.red, (const __nv_bfloat16*)a.out, a.topk, a.N, t, col0);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(red_bar0 + 8 * (2 + b));
  }
}

// ---- epilogue warps -------------------------------------------------------------------------------------
// kEpiWarps warps; `ew` = index among them, `warp` = warp index in the CTA (its TMEM lane quadrant is
// warp & 3).  INT8: s32 accumulators with activation / weight scales; otherwise f32 accumulators.
// kTokScale: f32 accumulators times a per-routed-row factor a.Xscale (the 2^e of the e4m3 hi/lo split).
// kAccStride / kSubStride: TMEM columns between the two accumulator buffers / the two sub-tiles.
// kConcat: the accumulator of a tile holds the hi-piece products in columns [0, n_eff) and the lo-piece
// products in [n_eff, 2 n_eff) (one MMA over the concatenated pieces); the epilogue adds them.
template <bool kInt8, int STAGE, bool kRelaxed, bool kTokScale = false, int kAccStride = kAccCols,
          int kSubStride = kWSub, bool kConcat = false>
__device__ __forceinline__ void epilogue_role(const UmmaArgs& a, int total, int n_blocks, uint8_t* smem_raw,
                                              uint32_t epi_base, uint32_t tfull0, uint32_t tempty0,
                                              uint32_t tmem_base, int warp, int ew, int lane, uint32_t red_bar0 = 0) {
  auto tfull_bar = [&](int b) { return tfull0 + 8 * b; };
  auto tempty_bar = [&](int b) { return tempty0 + 8 * b; };
    // ================= epilogue =================
    // Eight warps: warp (q, par) owns TMEM lanes [32q, 32q+32) (= 32 consecutive outputs n) and the
    // 16-column (= 16-token) chunks of parity par.  A chunk is transposed through a private smem
    // buffer so that global stores are 16 B per lane and whole 64/128-B row segments per token.
    const int q = warp & 3, par = ew >> 2;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    constexpr int OB = (kInt8 && STAGE == 1) ? 4 : 2;  // output element bytes (f32 act for int8 stage 1)
    constexpr int ROWB = 32 * OB;                      // bytes per token row in the transpose buffer
    constexpr int SEGS = ROWB / 16;                    // lanes per token row on the read side
    constexpr int TPI = 32 / SEGS;                     // token rows per read instruction
    constexpr int NRD = 16 / TPI;                      // read instructions per chunk
    uint8_t* ebuf = smem_raw + (epi_base + ew * kEpiBufBytes - smem_u32(smem_raw));
    const int rd_tok = lane / SEGS, rd_seg = lane % SEGS;
    if constexpr (STAGE == 2) fused_reduce_zero_rows(a.red, a.N, blockIdx.x * kEpiWarps + ew, gridDim.x * kEpiWarps, lane);
    uint32_t li = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x, ++li) {
      const WorkItem wi = get_item(a, w, n_blocks);
      const int n_eff = (wi.rows + 15) & ~15;
      const int buf = li & 1;
      if constexpr (STAGE == 2) {  // fused reduce: the reducer warp must be done with the item two back
        if (red_bar0 != 0) mbar_wait_relaxed(red_bar0 + 8 * (2 + buf), ((li >> 1) & 1) ^ 1);
      }
      // Per-token scalars of this tile (activation scale, topk weight): lane l keeps tokens
      // l, l+32, l+64, l+96.  Fetched before the accumulator wait so the dependent loads
      // (sorted_slot -> scale) hide under the MMAs; broadcast with shuffles in the column loop.
      float xs_r[4], rs_r[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        xs_r[i] = 1.f;
        rs_r[i] = 1.f;
        const int p = wi.row_start + min(lane + 32 * i, wi.rows - 1);
        if (32 * i < n_eff) {
          if constexpr (STAGE == 1) {
            if (kInt8 || kTokScale) xs_r[i] = a.Xscale[a.sorted_slot[p] / a.topk];
          } else {
            if constexpr (kInt8) {  // fused chain: the row's scale is quant_rows_kernel's amax / 127
              xs_r[i] = a.Xraw != nullptr ? fmaxf(__uint_as_float(a.row_amax[p]), 1e-7f) / 127.0f : a.Xscale[p];
            } else if (kTokScale) {
              xs_r[i] = a.Xscale[p];
            }
            if (a.row_scale != nullptr && wi.e < a.E) rs_r[i] = a.row_scale[a.sorted_slot[p]];
          }
        }
      }
      if (kRelaxed) mbar_wait_relaxed(tfull_bar(buf), (li >> 1) & 1);
      else mbar_wait(tfull_bar(buf), (li >> 1) & 1);
      tc_fence_after();
      const uint32_t acc = tmem_base + lane_off + buf * kAccStride;
      constexpr int kSubs = STAGE == 1 ? 1 : 2;
#pragma unroll 1
      for (int sub = 0; sub < kSubs; ++sub) {
        const int nbase = STAGE == 1 ? wi.nb * kWSub + q * 32 : wi.nb * 2 * kWSub + sub * kWSub + q * 32;
        const int n = nbase + lane;
        float ws0 = 1.f, ws1 = 1.f;
        if (kInt8 && n < a.N) {
          const float* ws = wi.e < a.E ? (const float*)a.Wscale + (size_t)wi.e * a.rows_per_expert : (const float*)a.Wsh_scale;
          ws0 = ws[n];
          if (STAGE == 1) ws1 = ws[a.N + n];
        }
        float bias_n = 0.f;  // dense linears: f32 bias of output n (this lane's TMEM lane)
        if (STAGE == 2 && a.bias != nullptr && n < a.N) bias_n = a.bias[n];
        float eb0 = 0.f, eb1 = 0.f;  // per-expert bias of this lane's output(s)
        if (a.ebias != nullptr && n < a.N && wi.e < a.E) {
          const float* eb = a.ebias + (size_t)wi.e * a.rows_per_expert;
          eb0 = eb[n];
          if (STAGE == 1) eb1 = eb[a.N + n];
        }
#pragma unroll 1
        for (int c0 = 16 * par; c0 < n_eff; c0 += 32) {  // n_eff is warp-uniform
          const float xsel = sel4(xs_r, c0 >> 5), rsel = sel4(rs_r, c0 >> 5);
          if constexpr (STAGE == 1) {
            uint32_t gv[16], uv[16];
            tmem_ld16(acc + c0, gv);
            tmem_ld16(acc + kSubStride + c0, uv);
            tmem_ld_wait(gv);
            tmem_ld_wait(uv);
            if constexpr (kConcat) {
              uint32_t gl[16], ul[16];
              tmem_ld16(acc + n_eff + c0, gl);
              tmem_ld16(acc + kSubStride + n_eff + c0, ul);
              tmem_ld_wait(gl);
              tmem_ld_wait(ul);
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                gv[j] = __float_as_uint(__uint_as_float(gv[j]) + __uint_as_float(gl[j]));
                uv[j] = __float_as_uint(__uint_as_float(uv[j]) + __uint_as_float(ul[j]));
              }
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              if constexpr (kInt8) {
                const float xs = __shfl_sync(0xffffffffu, xsel, 16 * par + j);
                const float gate = xs * (float)(int)gv[j] * ws0 + eb0;
                const float up = xs * (float)(int)uv[j] * ws1 + eb1;
                *reinterpret_cast<float*>(ebuf + j * ROWB + lane * 4) = glu_f(gate, up, a.act, a.act_alpha, a.act_limit);
              } else {
                float gate = __uint_as_float(gv[j]), up = __uint_as_float(uv[j]);
                if constexpr (kTokScale) {
                  const float xs = __shfl_sync(0xffffffffu, xsel, 16 * par + j);
                  gate *= xs;
                  up *= xs;
                }
                *reinterpret_cast<__nv_bfloat16*>(ebuf + j * ROWB + lane * 2) =
                    __float2bfloat16_rn(glu_f(gate + eb0, up + eb1, a.act, a.act_alpha, a.act_limit));
              }
            }
          } else {
            uint32_t v[16];
            tmem_ld16(acc + sub * kSubStride + c0, v);
            tmem_ld_wait(v);
            if constexpr (kConcat) {
              uint32_t vl[16];
              tmem_ld16(acc + sub * kSubStride + n_eff + c0, vl);
              tmem_ld_wait(vl);
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(vl[j]));
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float rs = __shfl_sync(0xffffffffu, rsel, 16 * par + j);
              float val;
              if constexpr (kInt8) {
                const float xs = __shfl_sync(0xffffffffu, xsel, 16 * par + j);
                val = xs * (float)(int)v[j] * ws0;
              } else {
                val = __uint_as_float(v[j]);
                if constexpr (kTokScale) val *= __shfl_sync(0xffffffffu, xsel, 16 * par + j);
              }
              *reinterpret_cast<__nv_bfloat16*>(ebuf + j * ROWB + lane * 2) = __float2bfloat16_rn(fmaf(val + eb0, rs, bias_n));
            }
          }
          __syncwarp();
          const int nn = nbase + rd_seg * (16 / OB);  // first output held by this lane's 16 bytes
#pragma unroll
          for (int i = 0; i < NRD; ++i) {
            const int tok = rd_tok + TPI * i;
            const int4 piece = *reinterpret_cast<const int4*>(ebuf + tok * ROWB + rd_seg * 16);
            const bool live = c0 + tok < wi.rows && nn < a.N;
            if (live) {
              const int64_t orow = STAGE == 2 ? out_row_of(wi.row_start + c0 + tok, a.sorted_slot, a.out_by_slot, a.out_map)
                                              : (int64_t)(wi.row_start + c0 + tok);
              *reinterpret_cast<int4*>((char*)a.out + ((size_t)orow * a.N + nn) * OB) = piece;
            }
            if constexpr (kInt8 && STAGE == 1) {
              // fused chain: max |act| of the row, for the quantisation the stage-2 gather warps perform
              if (a.row_amax != nullptr) {  // warp-uniform
                const float4 f = *reinterpret_cast<const float4*>(&piece);
                float m = live ? fmaxf(fmaxf(fabsf(f.x), fabsf(f.y)), fmaxf(fabsf(f.z), fabsf(f.w))) : 0.f;
#pragma unroll
                for (int off = 1; off < SEGS; off <<= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
                if (rd_seg == 0 && c0 + tok < wi.rows) atomicMax(a.row_amax + wi.row_start + c0 + tok, __float_as_uint(m));
              }
            }
          }
          __syncwarp();
        }
      }
      tc_fence_before();
      mbar_arrive(tempty_bar(buf));
      if constexpr (STAGE == 2) {
        if (red_bar0 != 0) mbar_arrive(red_bar0 + 8 * buf);  // this thread's stores of the item are issued (release)
      }
    }
}

}  // namespace ummak

// host helpers shared by the two tcgen05 translation units (defined in umma_gemm.cu)
int weight_tensor_map(const void* W, uint64_t kbytes, uint64_t rows, uint64_t E, uint32_t box_bytes,
                      uint32_t box_rows, int swizzle128, CUtensorMap* out);
// Routed-row map (rows already in expert-sorted order): bytes {K_bytes, rows}, box {128, box_rows}, 128-B swizzle
int rows_tensor_map(const void* X, uint64_t kbytes, uint64_t rows, uint32_t box_rows, CUtensorMap* out);
int sm_count();

// Persistent grid of the tcgen05 kernels: one CTA per SM, at most one per work item.  In the weight-streaming regime
// (capi.cu sets est_items only there) every item costs the same, so the grid is sized to WHOLE WAVES of the expected
// item count -- Qwen3 w13 at M = 128 is 768 items: 128 CTAs x 6 instead of 148 CTAs x 5.2; an expert-parallel rank of
// 8 streams DeepSeek-V3's w13 as 512 items: 128 CTAs x 4 instead of 148 x 3.46.  Measured (SGLB_GRID experiment,
// DESIGN 4a): 96..148 CTAs stream at the same rate, the partial last wave costs its full length.
inline int persistent_grid(int64_t max_items, int64_t est_items) {
  const int sms = sm_count();
  int grid = (int)(max_items < sms ? (max_items > 0 ? max_items : 1) : sms);
  if (est_items > 0) {
    const int64_t waves = (est_items + sms - 1) / sms;
    const int64_t even = (est_items + waves - 1) / waves;
    if (even < grid && even >= (sms * 3) / 4) grid = (int)even;
  }
  return grid;
}

}  // namespace sglb
