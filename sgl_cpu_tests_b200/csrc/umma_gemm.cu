// K3/K4 (prefill regime): grouped GEMM on the 5th-generation tensor cores.
//
//   D^T[w, t] = sum_k W[e, w, k] * X[row(t), k]        (weights are the UMMA "A"/M side,
//                                                       routed rows the "B"/N side)
//
// Putting the weight rows on the 128-lane M side and the routed rows on the N side lets the token
// count of a tile be any multiple of 16 (tcgen05.mma takes N at run time through the instruction
// descriptor), so ragged experts do not pay for a fixed 128-row M tile.  One CTA per SM, persistent
// over (tile, weight-block) work items, warp-specialised:
//
//   warp 0      TMA producer: weight boxes [128 rows x 128 B], SWIZZLE_128B, 3-D map {K, rows, E} so
//               rows past an expert's end are zero-filled, never the next expert's
//   warp 1      TMEM allocator + the single thread that issues tcgen05.mma / tcgen05.commit
//   warps 2-5   gather producers: routed rows come through sorted_slot (stage 1) so they cannot be
//               a TMA box; 16-B cp.async into the same 128-B-swizzled K-major layout the UMMA
//               descriptor expects (chunk' = chunk ^ (row & 7)), then fence.proxy.async + mbarrier
//   warps 6-9   epilogue: tcgen05.ld the f32/s32 accumulators, fuse silu_and_mul (stage 1) or the
//               topk weight (stage 2) and the int8 dequant scales, store
//
// Pipelines: a 4-deep smem ring (full/empty mbarriers; empty is signalled by tcgen05.commit) and a
// 2-deep TMEM accumulator ring (2 x 256 columns) so the epilogue of item i overlaps the MMAs of i+1.
// All operand tiles are [rows x 128 bytes]: 64 bf16 or 128 s8 of K per stage, four UMMA K-steps
// (32 B each) per stage for either type, so the kernel is written in bytes and only the instruction
// descriptor (kind::f16 / kind::i8) and the epilogue differ per format.
#include <cuda.h>  // CUtensorMap and enums only; the encoder is resolved through cudart, no -lcuda
#include <mutex>
#include <unordered_map>

#include "umma_gemm.cuh"

namespace sglb {

namespace {

constexpr int kStages = 4;
constexpr int kRowBytes = 128;                       // bytes of K per operand row per stage
constexpr int kWSub = 128;                           // weight rows per accumulator (UMMA M)
constexpr int kWTileBytes = kWSub * kRowBytes;       // 16 KB
constexpr int kXTileBytes = kUmmaTileRows * kRowBytes;
constexpr int kStageBytes = 2 * kWTileBytes + kXTileBytes;  // 48 KB
constexpr int kGatherThreads = 128;
constexpr int kEpiWarps = 8;                         // two per TMEM lane quadrant (one per 16-column parity)
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kFirstEpiWarp = 6;
constexpr int kThreads = (kFirstEpiWarp + kEpiWarps) * 32;  // 448
constexpr int kTmemCols = 512;
constexpr int kAccCols = 256;                        // per accumulator buffer: 2 sub-tiles x 128
constexpr int kEpiBufBytes = 16 * 32 * 4;            // per epilogue warp: 16 tokens x 32 outputs, f32 worst case
constexpr int kSmemBytes = kStages * kStageBytes + kEpiWarps * kEpiBufBytes + 1024 /*alignment slack*/ +
                           256 /*barriers*/;

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

template <int FMT, int STAGE>
__global__ void __launch_bounds__(kThreads, 1)
    umma_moe_kernel(const __grid_constant__ CUtensorMap tmap_w, const UmmaArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B: 1 KB aligned
  const uint32_t epi_base = smem_base + kStages * kStageBytes;
  const uint32_t bar_base = epi_base + kEpiWarps * kEpiBufBytes;
  auto w_tile = [&](int s) { return smem_base + s * kStageBytes; };
  auto x_tile = [&](int s) { return smem_base + s * kStageBytes + 2 * kWTileBytes; };
  auto full_bar = [&](int s) { return bar_base + 8 * s; };
  auto empty_bar = [&](int s) { return bar_base + 8 * (kStages + s); };
  auto tfull_bar = [&](int b) { return bar_base + 8 * (2 * kStages + b); };
  auto tempty_bar = [&](int b) { return bar_base + 8 * (2 * kStages + 2 + b); };
  const uint32_t tmem_slot = bar_base + 8 * (2 * kStages + 4);
  uint32_t* tmem_slot_ptr =
      reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int EB = FMT == SGLB_FMT_INT8_W8A8 ? 1 : 2;  // operand element bytes
  const int k_bytes = a.K * EB;
  const int num_kb = k_bytes / kRowBytes;
  const int n_blocks = num_weight_blocks<STAGE>(a);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w)) : "memory");
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1 + kGatherThreads);
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tfull_bar(b), 1);
      mbar_init(tempty_bar(b), kEpiThreads);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                 "n"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  // Everything above overlapped the previous kernel's tail; its outputs are read from here on.
  pdl_wait();
  pdl_trigger();
  const int total = a.counters[2] * n_blocks;

  if (warp == 0) {
    // ================= TMA producer: weights =================
    if (lane == 0) {
      uint32_t it = 0;
      for (int w = blockIdx.x; w < total; w += gridDim.x) {
        const WorkItem wi = get_item(a, w, n_blocks);
        const int wrow0 = STAGE == 1 ? wi.nb * kWSub : wi.nb * 2 * kWSub;
        const int wrow1 = STAGE == 1 ? a.N + wi.nb * kWSub : wrow0 + kWSub;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % kStages;
          mbar_wait(empty_bar(s), ((it / kStages) & 1) ^ 1);
          mbar_arrive_expect_tx(full_bar(s), 2 * kWTileBytes);
          tma_load_3d(w_tile(s), &tmap_w, full_bar(s), kb * kRowBytes, wrow0, wi.e);
          tma_load_3d(w_tile(s) + kWTileBytes, &tmap_w, full_bar(s), kb * kRowBytes, wrow1, wi.e);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      uint32_t it = 0, li = 0;
      for (int w = blockIdx.x; w < total; w += gridDim.x, ++li) {
        const WorkItem wi = get_item(a, w, n_blocks);
        const int n_eff = (wi.rows + 15) & ~15;
        const uint32_t idesc = instr_desc<FMT>(n_eff);
        const int buf = li & 1;
        mbar_wait(tempty_bar(buf), ((li >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t acc = tmem_base + buf * kAccCols;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % kStages;
          mbar_wait(full_bar(s), (it / kStages) & 1);
          fence_proxy_async();  // routed rows were written by cp.async (generic proxy)
          tc_fence_after();
#pragma unroll
          for (int sub = 0; sub < 2; ++sub) {
#pragma unroll
            for (int j = 0; j < kRowBytes / 32; ++j) {
              const uint64_t ad = smem_desc_sw128(w_tile(s) + sub * kWTileBytes + j * 32);
              const uint64_t bd = smem_desc_sw128(x_tile(s) + j * 32);
              umma<FMT>(acc + sub * kWSub, ad, bd, idesc, (kb > 0 || j > 0) ? 1u : 0u);
            }
          }
          umma_commit(empty_bar(s));  // frees the smem stage once these MMAs have read it
        }
        umma_commit(tfull_bar(buf));  // accumulators of this item are complete
      }
    }
    __syncwarp();
  } else if (warp < 6) {
    // ================= gather producers: routed rows =================
    const int g = threadIdx.x - 64;
    const int c = g & 7, rbase = g >> 3;  // 16-B chunk of the 128-B row; rows rbase + 16*j
    const uint32_t dst0 = rbase * kRowBytes + ((c ^ (rbase & 7)) << 4);
    uint32_t it = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x) {
      const WorkItem wi = get_item(a, w, n_blocks);
      const int n_eff = (wi.rows + 15) & ~15;
      const char* src[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int r = rbase + 16 * j;
        const int p = wi.row_start + min(r, wi.rows - 1);
        const int xr = a.gather ? (a.sorted_slot[p] / a.topk) : p;
        src[j] = (const char*)a.X + (size_t)xr * k_bytes + c * 16;
      }
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        const int s = it % kStages;
        mbar_wait(empty_bar(s), ((it / kStages) & 1) ^ 1);
        const uint32_t dst = x_tile(s) + dst0;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (rbase + 16 * j < n_eff) cp_async16(dst + j * 16 * kRowBytes, src[j] + kb * kRowBytes);
        cp_async_arrive_noinc(full_bar(s));
      }
    }
  } else {
    // ================= epilogue =================
    // Eight warps: warp (q, par) owns TMEM lanes [32q, 32q+32) (= 32 consecutive outputs n) and the
    // 16-column (= 16-token) chunks of parity par.  A chunk is transposed through a private smem
    // buffer so that global stores are 16 B per lane and whole 64/128-B row segments per token.
    const int ew = warp - kFirstEpiWarp;
    const int q = warp & 3, par = ew >> 2;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    constexpr bool kInt8 = FMT == SGLB_FMT_INT8_W8A8;
    constexpr int OB = (kInt8 && STAGE == 1) ? 4 : 2;  // output element bytes (f32 act for int8 stage 1)
    constexpr int ROWB = 32 * OB;                      // bytes per token row in the transpose buffer
    constexpr int SEGS = ROWB / 16;                    // lanes per token row on the read side
    constexpr int TPI = 32 / SEGS;                     // token rows per read instruction
    constexpr int NRD = 16 / TPI;                      // read instructions per chunk
    uint8_t* ebuf = smem_raw + (epi_base + ew * kEpiBufBytes - smem_u32(smem_raw));
    const int rd_tok = lane / SEGS, rd_seg = lane % SEGS;
    uint32_t li = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x, ++li) {
      const WorkItem wi = get_item(a, w, n_blocks);
      const int n_eff = (wi.rows + 15) & ~15;
      const int buf = li & 1;
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
            if (kInt8) xs_r[i] = a.Xscale[a.sorted_slot[p] / a.topk];
          } else {
            if (kInt8) xs_r[i] = a.Xscale[p];
            if (a.row_scale != nullptr) rs_r[i] = a.row_scale[a.sorted_slot[p]];
          }
        }
      }
      mbar_wait(tfull_bar(buf), (li >> 1) & 1);
      tc_fence_after();
      const uint32_t acc = tmem_base + lane_off + buf * kAccCols;
      constexpr int kSubs = STAGE == 1 ? 1 : 2;
#pragma unroll 1
      for (int sub = 0; sub < kSubs; ++sub) {
        const int nbase = STAGE == 1 ? wi.nb * kWSub + q * 32 : wi.nb * 2 * kWSub + sub * kWSub + q * 32;
        const int n = nbase + lane;
        float ws0 = 1.f, ws1 = 1.f;
        if (kInt8 && n < a.N) {
          const float* ws = (const float*)a.Wscale + (size_t)wi.e * a.rows_per_expert;
          ws0 = ws[n];
          if (STAGE == 1) ws1 = ws[a.N + n];
        }
#pragma unroll 1
        for (int c0 = 16 * par; c0 < n_eff; c0 += 32) {  // n_eff is warp-uniform
          const float xsel = sel4(xs_r, c0 >> 5), rsel = sel4(rs_r, c0 >> 5);
          if constexpr (STAGE == 1) {
            uint32_t gv[16], uv[16];
            tmem_ld16(acc + c0, gv);
            tmem_ld16(acc + kWSub + c0, uv);
            tmem_ld_wait(gv);
            tmem_ld_wait(uv);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              if constexpr (kInt8) {
                const float xs = __shfl_sync(0xffffffffu, xsel, 16 * par + j);
                const float gate = xs * (float)(int)gv[j] * ws0;
                const float up = xs * (float)(int)uv[j] * ws1;
                *reinterpret_cast<float*>(ebuf + j * ROWB + lane * 4) = silu_f(gate) * up;
              } else {
                const float gate = __uint_as_float(gv[j]), up = __uint_as_float(uv[j]);
                *reinterpret_cast<__nv_bfloat16*>(ebuf + j * ROWB + lane * 2) =
                    __float2bfloat16_rn(silu_f(gate) * up);
              }
            }
          } else {
            uint32_t v[16];
            tmem_ld16(acc + sub * kWSub + c0, v);
            tmem_ld_wait(v);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float rs = __shfl_sync(0xffffffffu, rsel, 16 * par + j);
              float val;
              if constexpr (kInt8) {
                const float xs = __shfl_sync(0xffffffffu, xsel, 16 * par + j);
                val = xs * (float)(int)v[j] * ws0;
              } else {
                val = __uint_as_float(v[j]);
              }
              *reinterpret_cast<__nv_bfloat16*>(ebuf + j * ROWB + lane * 2) = __float2bfloat16_rn(val * rs);
            }
          }
          __syncwarp();
          const int nn = nbase + rd_seg * (16 / OB);  // first output held by this lane's 16 bytes
#pragma unroll
          for (int i = 0; i < NRD; ++i) {
            const int tok = rd_tok + TPI * i;
            const int4 piece = *reinterpret_cast<const int4*>(ebuf + tok * ROWB + rd_seg * 16);
            if (c0 + tok < wi.rows && nn < a.N)
              *reinterpret_cast<int4*>((char*)a.out +
                                       ((size_t)(wi.row_start + c0 + tok) * a.N + nn) * OB) = piece;
          }
          __syncwarp();
        }
      }
      tc_fence_before();
      mbar_arrive(tempty_bar(buf));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "n"(kTmemCols)
                 : "memory");
  }
}

// ---- host: tensor maps ------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

struct MapKey {
  const void* ptr;
  uint64_t kbytes, rows, E;
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && kbytes == o.kbytes && rows == o.rows && E == o.E;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    return std::hash<const void*>()(k.ptr) ^ (k.kbytes * 0x9E3779B97F4A7C15ull) ^ (k.rows << 20) ^ (k.E << 44);
  }
};

// Weight map: bytes {K_bytes, rows_per_expert, E}, box {128 B, 128 rows, 1}, 128-B swizzle.  The
// expert dimension is separate so that a box reaching past an expert's last row is zero-filled.
int weight_tensor_map(const void* W, uint64_t kbytes, uint64_t rows, uint64_t E, CUtensorMap* out) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  const MapKey key{W, kbytes, rows, E};
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(key);
  if (it != cache.end()) {
    *out = it->second;
    return SGLB_OK;
  }
  EncodeTiledFn enc = encode_tiled_fn();
  SGLB_CHECK(enc != nullptr, "tcgen05 path: cuTensorMapEncodeTiled is not available from this driver");
  SGLB_CHECK(((uintptr_t)W & 15) == 0 && kbytes % 16 == 0, "tcgen05 path: weights must be 16-byte aligned");
  const cuuint64_t gdim[3] = {kbytes, rows, E};
  const cuuint64_t gstride[2] = {kbytes, kbytes * rows};
  const cuuint32_t box[3] = {(cuuint32_t)kRowBytes, (cuuint32_t)kWSub, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(W), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SGLB_CHECK(r == CUDA_SUCCESS, "tcgen05 path: cuTensorMapEncodeTiled failed (CUresult %d)", (int)r);
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, m);
  *out = m;
  return SGLB_OK;
}

int sm_count() {
  static int n = [] {
    int dev = 0, v = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    return v > 0 ? v : 148;
  }();
  return n;
}

template <int FMT, int STAGE>
int launch_one(const UmmaArgs& a, cudaStream_t stream) {
  constexpr int EB = FMT == SGLB_FMT_INT8_W8A8 ? 1 : 2;
  CUtensorMap map;
  int rc = weight_tensor_map(a.W, (uint64_t)a.K * EB, (uint64_t)a.rows_per_expert, (uint64_t)a.E, &map);
  if (rc) return rc;
  static bool attr_set = false;  // per instantiation
  if (!attr_set) {
    SGLB_CUDA(cudaFuncSetAttribute(umma_moe_kernel<FMT, STAGE>,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_set = true;
  }
  const int n_blocks = STAGE == 1 ? ceil_div(a.N, kWSub) : ceil_div(a.N, 2 * kWSub);
  const int64_t max_items = (int64_t)a.max_tiles * n_blocks;
  const int grid = (int)(max_items < sm_count() ? (max_items > 0 ? max_items : 1) : sm_count());
  SGLB_CUDA(launch_pdl(umma_moe_kernel<FMT, STAGE>, dim3(grid), dim3(kThreads), kSmemBytes, stream, map, a));
  return SGLB_OK;
}

}  // namespace

int umma_supported(int fmt, int64_t M, int topk, int E, int H, int I) {
  (void)M;
  (void)topk;
  (void)E;
  if (fmt != SGLB_FMT_BF16 && fmt != SGLB_FMT_INT8_W8A8) return 0;
  const int eb = fmt == SGLB_FMT_INT8_W8A8 ? 1 : 2;
  // whole 128-byte K slabs per stage; 16-B vector stores need nothing beyond that
  if ((H * eb) % kRowBytes != 0 || (I * eb) % kRowBytes != 0) return 0;
  return 1;
}

int launch_umma_gemm(int fmt, int stage, const UmmaArgs& a, cudaStream_t stream) {
  int rc;
  if (fmt == SGLB_FMT_BF16)
    rc = stage == 1 ? launch_one<SGLB_FMT_BF16, 1>(a, stream) : launch_one<SGLB_FMT_BF16, 2>(a, stream);
  else if (fmt == SGLB_FMT_INT8_W8A8)
    rc = stage == 1 ? launch_one<SGLB_FMT_INT8_W8A8, 1>(a, stream)
                    : launch_one<SGLB_FMT_INT8_W8A8, 2>(a, stream);
  else {
    set_error("tcgen05 path: unsupported weight format %d", fmt);
    return SGLB_ERR_INVALID;
  }
  if (rc == SGLB_OK) count_launch();
  return rc;
}

}  // namespace sglb
