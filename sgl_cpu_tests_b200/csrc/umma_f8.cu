// K3/K4 for fp8 e4m3 w8a16 block-scaled weights on the fp8 tensor cores: no weight dequantisation.
//
// The dequantising kernel (umma_dq.cu) spends ~3 ALU instructions per weight element to widen e4m3
// to bf16 and is bound by the ALU pipes (ncu r1d: 55 % of HBM at decode, 36 % of the bf16 tensor
// peak at prefill).  Here the e4m3 bytes go HBM --TMA--> smem --tcgen05.mma kind::f8f6f4--> TMEM
// untouched, and the two things the weight-only format needs are moved to where they are cheap:
//
//  * bf16 activations on an fp8 MMA: every bf16 value (8 significant bits) is split once per call
//    into hi + lo, two e4m3 numbers (4 significant bits each) after a per-row power-of-two
//    normalisation (split_rows_e4m3 below).  W.x = W.hi + W.lo: two MMAs at twice the bf16 rate.
//    The split is exact for every element within 2^-9 of its row maximum and has an absolute error
//    below 2^-18 of the row maximum otherwise -- finer than the bf16 rounding of the operands.
//  * the f32 scale of each [block_n, block_k] weight block multiplies the *partial accumulator* of
//    that k block: the MMA issuer accumulates one k block (block_k / 128 smem stages) into one of two
//    TMEM partial buffers, the epilogue warps pull it with tcgen05.ld and do acc += scale * partial
//    in registers while the tensor core fills the other buffer.
//
// Numerics: sum_k e4m3(w) * x accumulated in f32 and scaled in f32 is the exact-arithmetic value
// of what the reference oracle computes with bf16(w * scale) (test_moe_fp8_ext.py:22-25); the
// difference is the oracle's own bf16 rounding of the dequantised weight.  Outputs are held to the
// reference tolerance (rtol = atol = 1e-2) in tests/test_gpu_umma.py.
//
// Roles (12 warps; setmaxnreg moves registers from the producer warpgroup to the epilogue):
//   warp 0        TMA producer: weight boxes [128 rows x 128 B] x 2 (gate/up or two row blocks)
//   warp 1        MMA issuer (one thread)
//   warps 2-3     gather producers: hi and lo rows of the routed tokens, 16-B cp.async, 128-B swizzle
//   warps 4-11    epilogue: per-k-block promotion into 128 f32 registers per thread, then the fused
//                 silu_and_mul (stage 1) / topk-weight (stage 2) epilogue
#include <cuda_fp8.h>

#include <type_traits>

#include "rowops.cuh"
#include "umma_common.cuh"

namespace sglb {

using namespace ummak;

namespace {

// Two shared-memory configurations (template parameter XROWS = routed rows an operand slot holds = row capacity of the
// tiles moe_align hands out):
//   XROWS = 128: stage = W gate/up 32 KB + X hi + X lo 2 x 16 KB = 64 KB, 3 stages (prefill and general decode)
//   XROWS = 80:  stage = 32 KB + 2 x 10 KB = 52 KB, 4 stages: 128 KB of weight bytes in flight per SM
//   XROWS = 32:  stage = 32 KB + 2 x 4 KB = 40 KB, 5 stages: 160 KB instead of 96 KB in flight.
// At ~2.2 us of loaded HBM latency 96 KB in flight is ~43 GB/s per SM -- enough when all 148 SMs stream, the cap when a
// decode call has fewer work items than SMs (DeepSeek-V3 at M = 1: 144 items) or sizes its grid to whole waves (an
// expert-parallel rank of 8: 128 CTAs, 33 GB/s per SM measured); capi.cu picks the smallest capacity the expected rows
// per expert fit with two standard deviations to spare (a longer expert is split and streams its weights twice).
template <int XROWS>
struct F8Cfg {
  static constexpr int kXBytes = XROWS * kRowBytes;
  static constexpr int kStages = XROWS == 128 ? 3 : (XROWS == 80 ? 4 : 5);
  static_assert(XROWS == 128 || XROWS == 80 || XROWS == 32, "operand slot rows");
  static constexpr int kStageBytes = 2 * kWTileBytes + 2 * kXBytes;
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiWarps * kEpiBufBytes + 1024 + 256;
};
constexpr int kGatherThreadsF8 = 64;
constexpr int kFirstEpiWarp = 4;
constexpr int kThreads = (kFirstEpiWarp + kEpiWarps) * 32;  // 384
constexpr int kProducerRegs = 56, kEpilogueRegs = 216;  // 128*56 + 256*216 <= 384*168

__device__ __forceinline__ void umma_f8(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                        uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D f32, A and B e4m3 (format 0), both K-major, M = 128, N = n
__device__ __forceinline__ uint32_t instr_desc_f8(int n) {
  return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kWSub >> 4) << 24);
}

template <int STAGE, int XROWS>
__global__ void __launch_bounds__(kThreads, 1)
    umma_f8_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_sh,
                   const UmmaArgs a) {
  constexpr int kStages = F8Cfg<XROWS>::kStages, kStageBytes = F8Cfg<XROWS>::kStageBytes;
  constexpr int kXTileBytes = F8Cfg<XROWS>::kXBytes;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t epi_base = smem_base + kStages * kStageBytes;
  const uint32_t bar_base = epi_base + kEpiWarps * kEpiBufBytes;
  auto w_tile = [&](int s) { return smem_base + s * kStageBytes; };
  auto xh_tile = [&](int s) { return smem_base + s * kStageBytes + 2 * kWTileBytes; };
  auto xl_tile = [&](int s) { return smem_base + s * kStageBytes + 2 * kWTileBytes + kXTileBytes; };
  auto full_bar = [&](int s) { return bar_base + 8 * s; };
  auto empty_bar = [&](int s) { return bar_base + 8 * (kStages + s); };
  auto pfull_bar = [&](int b) { return bar_base + 8 * (2 * kStages + b); };
  auto pempty_bar = [&](int b) { return bar_base + 8 * (2 * kStages + 2 + b); };
  const uint32_t tmem_slot = bar_base + 8 * (2 * kStages + 4);
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_kb = a.K / kRowBytes;           // 128 e4m3 of K per stage
  const int G = a.block_k / kRowBytes;          // stages per scale block
  const int n_blocks = num_weight_blocks<STAGE>(a);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w)) : "memory");
    if (a.Wsh != nullptr) asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_sh)) : "memory");
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1 + kGatherThreadsF8);
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(pfull_bar(b), 1);
      mbar_init(pempty_bar(b), kEpiThreads);
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
  pdl_wait();
  pdl_trigger();
  const int total = a.counters[2] * n_blocks;

  if (warp < kFirstEpiWarp) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kProducerRegs));
    if (warp == 0) {
      // ================= TMA producer: e4m3 weights, untouched =================
      {
        uint32_t it = 0;
        for (int w = blockIdx.x; w < total; w += gridDim.x) {
          const WorkItem wi = get_item(a, w, n_blocks);
          const int wrow0 = STAGE == 1 ? wi.nb * kWSub : wi.nb * 2 * kWSub;
          const int wrow1 = STAGE == 1 ? a.N + wi.nb * kWSub : wrow0 + kWSub;
          const CUtensorMap* tm = wi.e < a.E ? &tmap_w : &tmap_sh;  // the shared expert rides along as expert id E
          const int we = wi.e < a.E ? wi.e : 0;
          for (int kb = 0; kb < num_kb; ++kb, ++it) {
            const int s = it % kStages;
            mbar_wait(empty_bar(s), ((it / kStages) & 1) ^ 1);
            if (elect_one()) {
              mbar_arrive_expect_tx(full_bar(s), 2 * kWTileBytes);
              tma_load_3d(w_tile(s), tm, full_bar(s), kb * kRowBytes, wrow0, we);
              tma_load_3d(w_tile(s) + kWTileBytes, tm, full_bar(s), kb * kRowBytes, wrow1, we);
            }
            __syncwarp();
          }
        }
      }
      __syncwarp();
    } else if (warp == 1) {
      // ================= MMA issuer =================
      // The whole warp runs the loop converged; one elected lane issues (see elect_one()).
      uint32_t it = 0, gi = 0;
      for (int w = blockIdx.x; w < total; w += gridDim.x) {
        const WorkItem wi = get_item(a, w, n_blocks);
        const int n_eff = (wi.rows + 15) & ~15;
        const uint32_t idesc = instr_desc_f8(n_eff);
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const bool first = (kb % G) == 0;
          const int pb = gi & 1;
          if (first) {
            mbar_wait(pempty_bar(pb), ((gi >> 1) & 1) ^ 1);
            tc_fence_after();
          }
          const int s = it % kStages;
          mbar_wait(full_bar(s), (it / kStages) & 1);
          fence_proxy_async();  // routed rows were written by cp.async (generic proxy)
          tc_fence_after();
          const uint32_t acc = tmem_base + pb * kAccCols;
          const bool last = (kb % G) == G - 1 || kb == num_kb - 1;
          if (elect_one()) {
#pragma unroll
            for (int sub = 0; sub < 2; ++sub) {
#pragma unroll
              for (int j = 0; j < kRowBytes / 32; ++j) {
                const uint64_t ad = smem_desc_sw128(w_tile(s) + sub * kWTileBytes + j * 32);
                const uint64_t bh = smem_desc_sw128(xh_tile(s) + j * 32);
                const uint64_t bl = smem_desc_sw128(xl_tile(s) + j * 32);
                umma_f8(acc + sub * kWSub, ad, bh, idesc, (first && j == 0) ? 0u : 1u);
                umma_f8(acc + sub * kWSub, ad, bl, idesc, 1u);
              }
            }
            umma_commit(empty_bar(s));
            if (last) umma_commit(pfull_bar(pb));  // this k block's partial sums are complete
          }
          __syncwarp();
          if (last) ++gi;
        }
      }
    } else {
      // ================= gather producers: hi and lo rows =================
      // 64 threads: thread g owns 16-B chunk c = g & 7 of rows (g >> 3) + 8 j, j = 0..15
      const int g = threadIdx.x - 64;
      const int c = g & 7, rbase = g >> 3;
      const uint32_t dst0 = rbase * kRowBytes + ((c ^ (rbase & 7)) << 4);  // (rbase + 8j) & 7 == rbase & 7
      const char* xh = (const char*)a.X + c * 16;
      const char* xl = (const char*)a.X2 + c * 16;
      uint32_t it = 0;
      for (int w = blockIdx.x; w < total; w += gridDim.x) {
        const WorkItem wi = get_item(a, w, n_blocks);
        const int n_eff = (wi.rows + 15) & ~15;
        uint32_t off[XROWS / 8];
#pragma unroll
        for (int j = 0; j < XROWS / 8; ++j) {
          const int r = rbase + 8 * j;
          const int p = wi.row_start + min(r, wi.rows - 1);
          const int xr = a.gather ? (a.sorted_slot[p] / a.topk) : p;
          off[j] = (uint32_t)xr * (uint32_t)a.K;
        }
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % kStages;
          mbar_wait(empty_bar(s), ((it / kStages) & 1) ^ 1);
          const uint32_t dh = xh_tile(s) + dst0, dl = xl_tile(s) + dst0;
#pragma unroll
          for (int j = 0; j < XROWS / 8; ++j) {
            if (rbase + 8 * j < n_eff) {
              cp_async16(dh + j * 8 * kRowBytes, xh + off[j] + kb * kRowBytes);
              cp_async16(dl + j * 8 * kRowBytes, xl + off[j] + kb * kRowBytes);
            }
          }
          cp_async_arrive_noinc(full_bar(s));
        }
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kEpilogueRegs));
    // ================= epilogue: promotion + fused output =================
    // warp (q, par): TMEM lanes [32q, 32q+32) = 32 consecutive weight rows; token chunks
    // c0 = 32 c + 16 par, c = 0..3.  acc[sub][c][j]: sub-tile, chunk, token within the chunk.
    const int ew = warp - kFirstEpiWarp;
    const int q = warp & 3, par = ew >> 2;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    constexpr int ROWB = 64, SEGS = 4, TPI = 8, NRD = 2;  // bf16 transpose buffer: 16 tokens x 32 outputs
    uint8_t* ebuf = smem_raw + (epi_base + ew * kEpiBufBytes - smem_u32(smem_raw));
    const int rd_tok = lane / SEGS, rd_seg = lane % SEGS;
    const int nblk_rows = ceil_div(a.rows_per_expert, a.block_n);
    uint32_t gi = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x) {
      const WorkItem wi = get_item(a, w, n_blocks);
      const int n_eff = (wi.rows + 15) & ~15;
      const int wrow0 = STAGE == 1 ? wi.nb * kWSub : wi.nb * 2 * kWSub;
      const int wrow1 = STAGE == 1 ? a.N + wi.nb * kWSub : wrow0 + kWSub;
      // scale rows of this warp's 32 weight rows (block_n is a multiple of 32): one scalar per k block
      const float* sbase = wi.e < a.E ? (const float*)a.Wscale + (size_t)wi.e * nblk_rows * a.scale_cols : (const float*)a.Wsh_scale;
      const float* srow0 = sbase + (size_t)(min(wrow0 + q * 32, a.rows_per_expert - 1) / a.block_n) * a.scale_cols;
      const float* srow1 = sbase + (size_t)(min(wrow1 + q * 32, a.rows_per_expert - 1) / a.block_n) * a.scale_cols;
      // per-token scalars: lane l keeps tokens l, l+32, l+64, l+96
      float xs_r[4], rs_r[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        xs_r[i] = 1.f;
        rs_r[i] = 1.f;
        const int p = wi.row_start + min(lane + 32 * i, wi.rows - 1);
        if (32 * i < n_eff) {
          if constexpr (STAGE == 1) {
            xs_r[i] = a.Xscale[a.sorted_slot[p] / a.topk];
          } else {
            xs_r[i] = a.Xscale[p];
            if (a.row_scale != nullptr && wi.e < a.E) rs_r[i] = a.row_scale[a.sorted_slot[p]];
          }
        }
      }
      // acc[sub][c][j/2]: two tokens per register pair, accumulated with the packed f32x2 FMA
      float2 acc[2][4][8];
#pragma unroll
      for (int s = 0; s < 2; ++s)
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[s][c][j] = make_float2(0.f, 0.f);

      const int ngroups = ceil_div(num_kb, G);
      float s0n = __ldg(srow0), s1n = __ldg(srow1);
      for (int g = 0; g < ngroups; ++g, ++gi) {
        const float2 s0 = make_float2(s0n, s0n), s1 = make_float2(s1n, s1n);
        if (g + 1 < ngroups) {
          s0n = __ldg(srow0 + g + 1);
          s1n = __ldg(srow1 + g + 1);
        }
        const int pb = gi & 1;
        mbar_wait(pfull_bar(pb), (gi >> 1) & 1);
        tc_fence_after();
        const uint32_t part = tmem_base + lane_off + pb * kAccCols + 16 * par;
        // Software pipeline over the four 16-token chunks: the TMEM loads of chunk c + 1 are in flight
        // while chunk c is accumulated (ncu r1f: with load -> wait -> FMA per chunk the promotion took
        // ~750 clk per k block against ~900 clk of MMAs, and the two serialised on the two partial buffers).
        // Eight half-steps h = (chunk c, sub-tile): v[sub] is reloaded for step h + 1 as soon as the wait of
        // step h has returned, so one 16-column load is always in flight under the 8 packed FMAs of a step.
        // Full tiles (the prefill case) take a branch-free copy so the loads keep one uniform base register.
        uint32_t v[2][16];
        auto promote = [&](auto full_tag) {
          constexpr bool kFull = decltype(full_tag)::value;
          if (kFull || 16 * par < n_eff) tmem_ld16(part, v[0]);
#pragma unroll
          for (int h = 0; h < 8; ++h) {
            const int c = h >> 1, sub = h & 1;
            if (kFull || 32 * c + 16 * par < n_eff) {  // warp-uniform, monotone in c
              tmem_ld_wait(v[sub]);
              if (h + 1 < 8) {
                const int c2 = (h + 1) >> 1, sub2 = (h + 1) & 1;
                if (kFull || 32 * c2 + 16 * par < n_eff) tmem_ld16(part + sub2 * kWSub + 32 * c2, v[sub2]);
              }
              const float2 sc = sub ? s1 : s0;
#pragma unroll
              for (int j = 0; j < 8; ++j)
                acc[sub][c][j] = __ffma2_rn(
                    sc, make_float2(__uint_as_float(v[sub][2 * j]), __uint_as_float(v[sub][2 * j + 1])),
                    acc[sub][c][j]);
            }
          }
        };
        if (n_eff == kUmmaTileRows) promote(std::true_type{});
        else promote(std::false_type{});
        tc_fence_before();
        mbar_arrive(pempty_bar(pb));
      }

      // ---- fused output of this work item ----
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int c0 = 32 * c + 16 * par;
        if (c0 >= n_eff) continue;
        const float xsel = sel4(xs_r, c), rsel = sel4(rs_r, c);
        constexpr int kSubs = STAGE == 1 ? 1 : 2;
#pragma unroll
        for (int sub = 0; sub < kSubs; ++sub) {
          const int nbase = STAGE == 1 ? wi.nb * kWSub + q * 32 : wi.nb * 2 * kWSub + sub * kWSub + q * 32;
          float bias_n = 0.f;
          if (STAGE == 2 && a.bias != nullptr && nbase + lane < a.N) bias_n = a.bias[nbase + lane];
          float eb0 = 0.f, eb1 = 0.f;  // per-expert bias of this lane's output(s)
          if (a.ebias != nullptr && nbase + lane < a.N && wi.e < a.E) {
            const float* eb = a.ebias + (size_t)wi.e * a.rows_per_expert;
            eb0 = eb[nbase + lane];
            if (STAGE == 1) eb1 = eb[a.N + nbase + lane];
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float xs = __shfl_sync(0xffffffffu, xsel, 16 * par + j);
            float val;
            if constexpr (STAGE == 1) {
              const float g_ = (j & 1) ? acc[0][c][j >> 1].y : acc[0][c][j >> 1].x;
              const float u_ = (j & 1) ? acc[1][c][j >> 1].y : acc[1][c][j >> 1].x;
              val = glu_f(g_ * xs + eb0, u_ * xs + eb1, a.act, a.act_alpha, a.act_limit);
            } else {
              const float rs = __shfl_sync(0xffffffffu, rsel, 16 * par + j);
              const float a_ = (j & 1) ? acc[sub][c][j >> 1].y : acc[sub][c][j >> 1].x;
              val = fmaf(a_ * xs + eb0, rs, bias_n);
            }
            *reinterpret_cast<__nv_bfloat16*>(ebuf + j * ROWB + lane * 2) = __float2bfloat16_rn(val);
          }
          __syncwarp();
          const int nn = nbase + rd_seg * 8;
#pragma unroll
          for (int i = 0; i < NRD; ++i) {
            const int tok = rd_tok + TPI * i;
            const int4 piece = *reinterpret_cast<const int4*>(ebuf + tok * ROWB + rd_seg * 16);
            if (c0 + tok < wi.rows && nn < a.N) {
              const int64_t orow = STAGE == 2 ? out_row_of(wi.row_start + c0 + tok, a.sorted_slot, a.out_by_slot, a.out_map)
                                              : (int64_t)(wi.row_start + c0 + tok);
              *reinterpret_cast<int4*>((char*)a.out + ((size_t)orow * a.N + nn) * 2) = piece;
            }
          }
          __syncwarp();
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols)
                 : "memory");
  }
}

// ---- bf16 rows -> (hi, lo) e4m3 pieces + per-row power-of-two factor: one CTA per row (rowops.cuh) ----
// n_valid / always_from (act rows in expert-sorted order): rows [*n_valid, always_from) hold nothing -- masked slots
// sort behind the valid ones (an expert-parallel rank sees world x cap slots of which ~1/world are its own) -- and are
// skipped; rows from always_from on are the riding shared expert's.
__global__ void __launch_bounds__(256) split_rows_e4m3_kernel(const __nv_bfloat16* __restrict__ x, int K,
                                                              uint8_t* __restrict__ hi, uint8_t* __restrict__ lo,
                                                              float* __restrict__ rexp, const int32_t* __restrict__ n_valid,
                                                              int always_from, int64_t rows) {
  pdl_wait();
  pdl_trigger();
  __shared__ float warp_max_s[8];
  const int nv = n_valid != nullptr ? *n_valid : 0;
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    if (n_valid != nullptr && row >= nv && row < always_from) continue;
    split_row_e4m3(x + row * K, K, hi + row * K, lo + row * K, rexp + row, threadIdx.x, warp_max_s, 0);
    __syncthreads();  // warp_max_s is reused by the next row
  }
}

template <int STAGE, int XROWS>
int launch_one(const UmmaArgs& a, cudaStream_t stream) {
  constexpr int kSmemBytes = F8Cfg<XROWS>::kSmemBytes;
  CUtensorMap map;
  int rc = weight_tensor_map(a.W, (uint64_t)a.K, (uint64_t)a.rows_per_expert, (uint64_t)a.E, kRowBytes, kWSub, 1, &map);
  if (rc) return rc;
  CUtensorMap map_sh = map;
  if (a.Wsh != nullptr) {
    rc = weight_tensor_map(a.Wsh, (uint64_t)a.K, (uint64_t)a.rows_per_expert, 1, kRowBytes, kWSub, 1, &map_sh);
    if (rc) return rc;
  }
  static bool attr_set = false;  // per instantiation
  if (!attr_set) {
    SGLB_CUDA(cudaFuncSetAttribute(umma_f8_kernel<STAGE, XROWS>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_set = true;
  }
  const int n_blocks = STAGE == 1 ? ceil_div(a.N, kWSub) : ceil_div(a.N, 2 * kWSub);
  const int64_t max_items = (int64_t)a.max_tiles * n_blocks;
  const int grid = persistent_grid(max_items, (int64_t)a.est_tiles * n_blocks);
  SGLB_CUDA(launch_pdl(umma_f8_kernel<STAGE, XROWS>, dim3(grid), dim3(kThreads), kSmemBytes, stream, map, map_sh, a));
  return SGLB_OK;
}

}  // namespace

int umma_f8_supported(int H, int I, int block_n, int block_k) {
  return H % kRowBytes == 0 && I % kRowBytes == 0 && block_k % kRowBytes == 0 && block_n % 32 == 0;
}

int umma_f8_small_tile_rows() { return 32; }
int umma_f8_mid_tile_rows() { return 80; }

// a.tile_cap: row capacity of the tiles in tile_list -- 0 / 128 (general) or umma_f8_small_tile_rows() (deep ring)
int launch_umma_f8(int stage, const UmmaArgs& a, cudaStream_t stream) {
  int rc;
  if (a.tile_cap == 32) rc = stage == 1 ? launch_one<1, 32>(a, stream) : launch_one<2, 32>(a, stream);
  else if (a.tile_cap == 80) rc = stage == 1 ? launch_one<1, 80>(a, stream) : launch_one<2, 80>(a, stream);
  else rc = stage == 1 ? launch_one<1, 128>(a, stream) : launch_one<2, 128>(a, stream);
  if (rc == SGLB_OK) count_launch();
  return rc;
}

int launch_split_rows_e4m3(const void* x_bf16, int64_t rows, int K, void* hi, void* lo, float* rexp,
                           cudaStream_t stream, const int32_t* n_valid, int always_from) {
  SGLB_CHECK(K % 8 == 0, "split_rows_e4m3: K must be a multiple of 8 (K=%d)", K);
  if (rows == 0) return SGLB_OK;
  // one CTA per row up to 16 waves of the machine, then rows are strided (an expert-parallel rank launches over
  // world x cap x topk slots of which ~1/world hold a row)
  const int64_t grid = rows < 16 * 148 ? rows : 16 * 148;
  SGLB_CUDA(launch_pdl(split_rows_e4m3_kernel, dim3((unsigned)grid), dim3(256), 0, stream,
                       (const __nv_bfloat16*)x_bf16, K, (uint8_t*)hi, (uint8_t*)lo, rexp, n_valid, always_from, rows));
  count_launch();
  return SGLB_OK;
}

}  // namespace sglb
