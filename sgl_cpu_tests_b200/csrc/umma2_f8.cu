// K3/K4 for fp8 e4m3 w8a16 weights, prefill regime: the tensor-native fp8 kernel of umma_f8.cu on a CTA
// pair (tcgen05.mma.cta_group::2.kind::f8f6f4, M = 256).
//
// Same arithmetic as umma_f8.cu -- e4m3 weights untouched from HBM to the tensor core, activations as
// exact hi + lo e4m3 pieces, the f32 block scale applied to per-k-block TMEM partials by the epilogue
// warps -- with the operand layout of umma2_gemm.cu: each CTA of the pair feeds its 128 weight rows
// and only its half of the (up to 256) routed rows, which halves the shared-memory operand traffic
// per MMA and doubles the work per issued instruction.  Stage 1 gives each CTA 64 gate + 64 up rows;
// the warps holding `up` pass their finished accumulators to the `gate` warps through shared memory.
//
// Roles (12 warps; setmaxnreg: 56 registers for the producer warpgroup, 216 for the epilogue):
//   warp 0      TMA producer, weights (cta_group::2 loads, bytes counted on the leader's full barrier)
//   warp 1      leader: MMA issuer; peer: idle
//   warps 2-3   idle (the routed rows arrive in expert-sorted order: warp 0 loads this CTA's hi and lo halves as two
//               2-D TMA boxes counted on the leader's full barrier, umma2_gemm.cu header)
//   warps 4-11  epilogue: acc += scale * partial per k block (128 f32 registers per thread), fused output
#include <type_traits>

#include "umma_pair.cuh"

namespace sglb {

using namespace ummak;

namespace {

constexpr int kStagesP = 4;
constexpr int kStageBytesP = kWTileBytes + 2 * kXTileBytes;  // 48 KB: A half + X hi + X lo (this CTA's rows)
constexpr int kFirstEpiWarpP = 4;
constexpr int kThreadsP = (kFirstEpiWarpP + kEpiWarps) * 32;  // 384
constexpr int kXchgBytesP = 4 * 2 * 2048;
constexpr int kSmemBytesP = kStagesP * kStageBytesP + kEpiWarps * kEpiBufBytes + kXchgBytesP + 1024 + 256;
constexpr int kProducerRegsP = 56, kEpilogueRegsP = 216;

__device__ __forceinline__ void umma_pair_f8(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D f32, A and B e4m3, K-major, M = 256 (the pair), N = n
__device__ __forceinline__ uint32_t instr_desc_pair_f8(int n) {
  return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
}

template <int STAGE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreadsP, 1)
    umma2_f8_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_xh,
                    const __grid_constant__ CUtensorMap tmap_xl, const UmmaArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t epi_base = smem_base + kStagesP * kStageBytesP;
  const uint32_t xchg_base = epi_base + kEpiWarps * kEpiBufBytes;
  const uint32_t bar_base = xchg_base + kXchgBytesP;
  auto w_tile = [&](int s) { return smem_base + s * kStageBytesP; };
  auto xh_tile = [&](int s) { return smem_base + s * kStageBytesP + kWTileBytes; };
  auto xl_tile = [&](int s) { return smem_base + s * kStageBytesP + kWTileBytes + kXTileBytes; };
  auto full_bar = [&](int s) { return bar_base + 8 * s; };
  auto empty_bar = [&](int s) { return bar_base + 8 * (2 * kStagesP + s); };
  auto pfull_bar = [&](int b) { return bar_base + 8 * (3 * kStagesP + b); };
  auto pempty_bar = [&](int b) { return bar_base + 8 * (3 * kStagesP + 2 + b); };
  const uint32_t tmem_slot = bar_base + 8 * (3 * kStagesP + 4);
  static_assert(8 * (3 * kStagesP + 4) + 4 <= 256, "barrier region overflow");
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int num_kb = a.K / kRowBytes;   // 128 e4m3 of K per stage
  const int G = a.block_k / kRowBytes;  // stages per scale block
  const int n_blocks = num_weight_blocks<STAGE>(a);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_xh)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_xl)) : "memory");
    for (int s = 0; s < kStagesP; ++s) {
      mbar_init(full_bar(s), 1);  // the leader's expect_tx arrive; everything else is TMA bytes
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(pfull_bar(b), 1);
      mbar_init(pempty_bar(b), 2 * kEpiWarps);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                 "n"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  pdl_wait();
  pdl_trigger();
  const int total = a.counters[2] * n_blocks;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

  if (warp < kFirstEpiWarpP) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kProducerRegsP));
    if (warp == 0) {
      // ================= TMA producer =================
      const bool l2_hints = a.l2_hints != 0;
      uint32_t it = 0;
      for (int w = cluster_id; w < total; w += num_clusters) {
        const PairItem wi = get_pair_item(a, w, n_blocks);
        const int n_eff = (wi.rows + 15) & ~15;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % kStagesP;
          mbar_wait(empty_bar(s), ((it / kStagesP) & 1) ^ 1);
          if (elect_one()) {
            const uint32_t lbar = mapa(full_bar(s), 0);
            // both CTAs' weight halves + both CTAs' hi and lo routed-row boxes (whole 128-row boxes)
            if (rank == 0) mbar_arrive_expect_tx(full_bar(s), 2 * kWTileBytes + 4 * kXTileBytes);
            const int r0 = wi.row_start + (int)rank * (n_eff >> 1);
            if (l2_hints) {
              tma_load_2d_pair_hint(xh_tile(s), &tmap_xh, lbar, kb * kRowBytes, r0, kL2EvictLast);
              tma_load_2d_pair_hint(xl_tile(s), &tmap_xl, lbar, kb * kRowBytes, r0, kL2EvictLast);
              if constexpr (STAGE == 1) {
                const int g0 = wi.nb * kWSub + (int)rank * 64;
                tma_load_3d_pair_hint(w_tile(s), &tmap_w, lbar, kb * kRowBytes, g0, wi.e, kL2EvictFirst);
                tma_load_3d_pair_hint(w_tile(s) + kWTileBytes / 2, &tmap_w, lbar, kb * kRowBytes, a.N + g0, wi.e, kL2EvictFirst);
              } else {
                tma_load_3d_pair_hint(w_tile(s), &tmap_w, lbar, kb * kRowBytes, wi.nb * 2 * kWSub + (int)rank * kWSub, wi.e,
                                      kL2EvictFirst);
              }
            } else {
            tma_load_2d_pair(xh_tile(s), &tmap_xh, lbar, kb * kRowBytes, r0);
            tma_load_2d_pair(xl_tile(s), &tmap_xl, lbar, kb * kRowBytes, r0);
            if constexpr (STAGE == 1) {
              const int g0 = wi.nb * kWSub + (int)rank * 64;
              tma_load_3d_pair(w_tile(s), &tmap_w, lbar, kb * kRowBytes, g0, wi.e);
              tma_load_3d_pair(w_tile(s) + kWTileBytes / 2, &tmap_w, lbar, kb * kRowBytes, a.N + g0, wi.e);
            } else {
              tma_load_3d_pair(w_tile(s), &tmap_w, lbar, kb * kRowBytes, wi.nb * 2 * kWSub + (int)rank * kWSub, wi.e);
            }
            }
          }
          __syncwarp();
        }
      }
    } else if (warp == 1) {
      if (rank == 0) {
        // ================= MMA issuer (leader) =================
        uint32_t it = 0, gi = 0;
        for (int w = cluster_id; w < total; w += num_clusters) {
          const PairItem wi = get_pair_item(a, w, n_blocks);
          const int n_eff = (wi.rows + 15) & ~15;
          const uint32_t idesc = instr_desc_pair_f8(n_eff);
          for (int kb = 0; kb < num_kb; ++kb, ++it) {
            const bool first = (kb % G) == 0;
            const int pb = gi & 1;
            if (first) {
              mbar_wait(pempty_bar(pb), ((gi >> 1) & 1) ^ 1);
              tc_fence_after();
            }
            const int s = it % kStagesP;
            mbar_wait(full_bar(s), (it / kStagesP) & 1);
            tc_fence_after();
            const uint32_t acc = tmem_base + pb * kAccCols;
            const bool last = (kb % G) == G - 1 || kb == num_kb - 1;
            if (elect_one()) {
#pragma unroll
              for (int j = 0; j < kRowBytes / 32; ++j) {
                const uint64_t ad = smem_desc_sw128(w_tile(s) + j * 32);
                const uint64_t bh = smem_desc_sw128(xh_tile(s) + j * 32);
                const uint64_t bl = smem_desc_sw128(xl_tile(s) + j * 32);
                umma_pair_f8(acc, ad, bh, idesc, (first && j == 0) ? 0u : 1u);
                umma_pair_f8(acc, ad, bl, idesc, 1u);
              }
              umma_commit_pair(empty_bar(s));
              if (last) umma_commit_pair(pfull_bar(pb));
            }
            __syncwarp();
            if (last) ++gi;
          }
        }
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kEpilogueRegsP));
    // ================= epilogue: promotion + fused output =================
    // warp (q, par): this CTA's TMEM lanes [32q, 32q+32); token chunks c0 = 32 c + 16 par, c = 0..7.
    const int ew = warp - kFirstEpiWarpP;
    const int q = warp & 3, par = ew >> 2;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    constexpr int ROWB = 64, SEGS = 4, TPI = 8, NRD = 2;  // bf16 transpose buffer: 16 tokens x 32 outputs
    uint8_t* ebuf = smem_raw + (epi_base + ew * kEpiBufBytes - smem_u32(smem_raw));
    const int rd_tok = lane / SEGS, rd_seg = lane % SEGS;
    const bool is_up = STAGE == 1 && q >= 2;  // lanes 64..127 hold `up` of the outputs lanes 0..63 gate
    const int pair_idx = (q & 1) + 2 * par;
    float* xbuf = reinterpret_cast<float*>(smem_raw + (xchg_base + pair_idx * 4096 - smem_u32(smem_raw)));
    const uint32_t pempty_l0 = mapa(pempty_bar(0), 0), pempty_l1 = mapa(pempty_bar(1), 0);
    const int nblk_rows = ceil_div(a.rows_per_expert, a.block_n);
    uint32_t gi = 0;
    uint32_t xi = 0;  // exchange-buffer parity over all chunks of this warp pair, across items (umma2_gemm.cu)
    for (int w = cluster_id; w < total; w += num_clusters) {
      const PairItem wi = get_pair_item(a, w, n_blocks);
      const int n_eff = (wi.rows + 15) & ~15;
      int nbase;
      if constexpr (STAGE == 1) nbase = wi.nb * kWSub + (int)rank * 64 + (q & 1) * 32;
      else nbase = wi.nb * 2 * kWSub + (int)rank * kWSub + q * 32;
      const int wrow = (STAGE == 1 && is_up) ? a.N + nbase : nbase;  // first weight row of this warp
      const float* srow = (const float*)a.Wscale +
                          ((size_t)wi.e * nblk_rows + min(wrow, a.rows_per_expert - 1) / a.block_n) * a.scale_cols;
      float xs_r[8], rs_r[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        xs_r[i] = 1.f;
        rs_r[i] = 1.f;
        const int p = wi.row_start + min(lane + 32 * i, wi.rows - 1);
        if (32 * i < n_eff) {
          if constexpr (STAGE == 1) {
            xs_r[i] = a.Xscale[a.sorted_slot[p] / a.topk];
          } else {
            xs_r[i] = a.Xscale[p];
            if (a.row_scale != nullptr) rs_r[i] = a.row_scale[a.sorted_slot[p]];
          }
        }
      }
      // acc[c][j/2]: two tokens per register pair, accumulated with the packed f32x2 FMA
      float2 acc[8][8];
#pragma unroll
      for (int c = 0; c < 8; ++c)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[c][j] = make_float2(0.f, 0.f);
      auto accv = [&](int c, int j) { return (j & 1) ? acc[c][j >> 1].y : acc[c][j >> 1].x; };

      const int ngroups = ceil_div(num_kb, G);
      float sn = __ldg(srow);
      for (int g = 0; g < ngroups; ++g, ++gi) {
        const float2 sc = make_float2(sn, sn);
        if (g + 1 < ngroups) sn = __ldg(srow + g + 1);
        const int pb = gi & 1;
        mbar_wait(pfull_bar(pb), (gi >> 1) & 1);
        tc_fence_after();
        const uint32_t part = tmem_base + lane_off + pb * kAccCols + 16 * par;
        // Software pipeline over the eight 16-token chunks (same scheme as umma_f8.cu): the load of chunk
        // c + 1 is in flight under the 8 packed FMAs of chunk c; full tiles take a branch-free copy.
        uint32_t v[2][16];
        auto promote = [&](auto full_tag) {
          constexpr bool kFull = decltype(full_tag)::value;
          if (kFull || 16 * par < n_eff) tmem_ld16(part, v[0]);
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            if (kFull || 32 * c + 16 * par < n_eff) {  // warp-uniform, monotone in c
              tmem_ld_wait(v[c & 1]);
              if (c + 1 < 8 && (kFull || 32 * (c + 1) + 16 * par < n_eff))
                tmem_ld16(part + 32 * (c + 1), v[(c + 1) & 1]);
#pragma unroll
              for (int j = 0; j < 8; ++j)
                acc[c][j] = __ffma2_rn(
                    sc, make_float2(__uint_as_float(v[c & 1][2 * j]), __uint_as_float(v[c & 1][2 * j + 1])),
                    acc[c][j]);
            }
          }
        };
        if (n_eff == 256) promote(std::true_type{});
        else promote(std::false_type{});
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(pb ? pempty_l1 : pempty_l0);
      }

      // ---- fused output of this work item ----
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const int c0 = 32 * c + 16 * par;
        if (c0 >= n_eff) continue;
        const float xsel = sel8(xs_r, c), rsel = sel8(rs_r, c);
        if constexpr (STAGE == 1) {
          float* xb = xbuf + (xi++ & 1) * 512;
          if (is_up) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float xs = __shfl_sync(0xffffffffu, xsel, 16 * par + j);
              xb[j * 32 + lane] = accv(c, j) * xs;
            }
            named_bar_sync(1 + pair_idx, 64);
            continue;
          }
          named_bar_sync(1 + pair_idx, 64);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float xs = __shfl_sync(0xffffffffu, xsel, 16 * par + j);
            const float r = silu_f(accv(c, j) * xs) * xb[j * 32 + lane];
            *reinterpret_cast<__nv_bfloat16*>(ebuf + j * ROWB + lane * 2) = __float2bfloat16_rn(r);
          }
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float xs = __shfl_sync(0xffffffffu, xsel, 16 * par + j);
            const float rs = __shfl_sync(0xffffffffu, rsel, 16 * par + j);
            *reinterpret_cast<__nv_bfloat16*>(ebuf + j * ROWB + lane * 2) = __float2bfloat16_rn(accv(c, j) * xs * rs);
          }
        }
        __syncwarp();
        const int nn = nbase + rd_seg * 8;
#pragma unroll
        for (int i = 0; i < NRD; ++i) {
          const int tok = rd_tok + TPI * i;
          const int4 piece = *reinterpret_cast<const int4*>(ebuf + tok * ROWB + rd_seg * 16);
          if (c0 + tok < wi.rows && nn < a.N)
            *reinterpret_cast<int4*>((char*)a.out + ((size_t)(wi.row_start + c0 + tok) * a.N + nn) * 2) = piece;
        }
        __syncwarp();
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols)
                 : "memory");
  }
}

template <int STAGE>
int launch_pair_f8(const UmmaArgs& a, cudaStream_t stream) {
  CUtensorMap map;
  int rc = weight_tensor_map(a.W, (uint64_t)a.K, (uint64_t)a.rows_per_expert, (uint64_t)a.E, kRowBytes,
                             STAGE == 1 ? 64 : kWSub, 1, &map);
  if (rc) return rc;
  CUtensorMap map_xh, map_xl;
  SGLB_CHECK(a.gather == 0, "CTA-pair tcgen05 path: routed rows must be in expert-sorted order (gather = 0)");
  rc = rows_tensor_map(a.X, (uint64_t)a.K, (uint64_t)a.x_rows, kUmmaTileRows, &map_xh);
  if (rc) return rc;
  rc = rows_tensor_map(a.X2, (uint64_t)a.K, (uint64_t)a.x_rows, kUmmaTileRows, &map_xl);
  if (rc) return rc;
  static bool attr_set = false;  // per instantiation
  if (!attr_set) {
    SGLB_CUDA(cudaFuncSetAttribute(umma2_f8_kernel<STAGE>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytesP));
    attr_set = true;
  }
  const int n_blocks = STAGE == 1 ? ceil_div(a.N, kWSub) : ceil_div(a.N, 2 * kWSub);
  const int64_t max_items = (int64_t)a.max_tiles * n_blocks;
  const int max_clusters = sm_count() / 2;
  const int clusters = (int)(max_items < max_clusters ? (max_items > 0 ? max_items : 1) : max_clusters);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * clusters);
  cfg.blockDim = dim3(kThreadsP);
  cfg.dynamicSmemBytes = kSmemBytesP;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  SGLB_CUDA(cudaLaunchKernelEx(&cfg, umma2_f8_kernel<STAGE>, map, map_xh, map_xl, a));
  return SGLB_OK;
}

}  // namespace

int launch_umma2_f8(int stage, const UmmaArgs& a, cudaStream_t stream) {
  int rc = stage == 1 ? launch_pair_f8<1>(a, stream) : launch_pair_f8<2>(a, stream);
  if (rc == SGLB_OK) count_launch();
  return rc;
}

}  // namespace sglb
