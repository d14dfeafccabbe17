// K3/K4 for weight-only quantised formats (fp8 e4m3 w8a16 block-scaled, mxfp4): the tcgen05 grouped
// GEMM of umma_gemm.cu with one more pipeline stage in front of the tensor cores.
//
//   HBM --TMA--> raw ring (e4m3 bytes / e2m1 nibbles, dense rows)
//       --8 dequant warps--> bf16 operand ring (K-major, 128-B swizzle), scale applied on the way
//       --tcgen05.mma kind::f16--> TMEM --epilogue warps--> HBM
//
// Weights cross HBM in their packed form (1 or 0.5 byte per element) and are widened once per CTA
// tile, not once per routed row.  Numerics follow the reference oracles: the dequantised weight is
// bf16(f32(w) * scale) (test_moe_fp8_ext.py:22-25 `scaled_weight`; test_mxfp4.py:90-128 dequantize),
// then a bf16 x bf16 -> f32 GEMM.  e4m3 -> f32 is a shift (s.eeee.mmm -> f32 bits << 20, times 2^120
// folded into the block scale, subnormals included); e2m1 * 2^(e8m0-127) is exact in bf16.
//
// A stage is 64 K-elements: 128 B per routed row (bf16), 64 B (fp8) or 32 B (mxfp4) per weight row.
// Rings: raw weights RAWS x (256 rows x RB), bf16 weight operand 2 x 32 KB, routed rows 4 x 16 KB.
#include "umma_common.cuh"

namespace sglb {

using namespace ummak;

namespace {

constexpr int kXStages = 4;
constexpr int kWopStages = 2;
constexpr int kWopBytes = 2 * kWTileBytes;  // 32 KB: two 128-row sub-tiles
constexpr int kFirstEpiWarp = 6;
constexpr int kFirstDqWarp = kFirstEpiWarp + kEpiWarps;  // 14
constexpr int kDqWarps = 8;
constexpr int kDqThreads = kDqWarps * 32;
constexpr int kThreads = (kFirstDqWarp + kDqWarps) * 32;  // 704
constexpr int kKPerStage = 64;

template <int FMT>
struct DqTraits;
template <>
struct DqTraits<SGLB_FMT_FP8_W8A16> {
  static constexpr int RB = 64;    // raw bytes per weight row per stage
  static constexpr int RAWS = 4;   // raw ring depth
};
template <>
struct DqTraits<SGLB_FMT_MXFP4> {
  static constexpr int RB = 32;
  static constexpr int RAWS = 8;
};

template <int FMT>
constexpr int smem_bytes() {
  return DqTraits<FMT>::RAWS * 256 * DqTraits<FMT>::RB + kWopStages * kWopBytes + kXStages * kXTileBytes +
         kEpiWarps * kEpiBufBytes + 1024 /*alignment slack*/ + 512 /*barriers*/;
}

// four e4m3 bytes of a word -> four bf16(f32(w) * scale) packed as two bf16x2.  The hardware converter
// (cvt.rn.f16x2.e4m3x2, exact: f16 holds every e4m3 value) replaces the shift/mask arithmetic; the
// product is the same single-rounded f32 multiply the reference oracle does, then one RN to bf16.
__device__ __forceinline__ void e4m3x4_scale_to_bf16x4(uint32_t w, float s, uint32_t& lo, uint32_t& hi) {
  uint32_t h01, h23;
  asm("{\n\t.reg .b16 l, h;\n\tmov.b32 {l, h}, %2;\n\t"
      "cvt.rn.f16x2.e4m3x2 %0, l;\n\tcvt.rn.f16x2.e4m3x2 %1, h;\n\t}"
      : "=r"(h01), "=r"(h23)
      : "r"(w));
  const float2 f01 = __half22float2(*reinterpret_cast<const __half2*>(&h01));
  const float2 f23 = __half22float2(*reinterpret_cast<const __half2*>(&h23));
  lo = pack_bf16x2(f01.x * s, f01.y * s);
  hi = pack_bf16x2(f23.x * s, f23.y * s);
}
// word w of e2m1 nibbles n0..n7 (n0 lowest = lowest k) -> raw bf16x2 words (value * 2^-126) in k
// order: r[0] = (n0,n1) ... r[3] = (n6,n7).  Pairs (n_i, n_{i+4}) come out of one mask/shift each
// (see e2m1x8_to_raw_bf16 in common.cuh); two byte permutes per pair of words restore the k order.
__device__ __forceinline__ void e2m1x8_to_raw_bf16_korder(uint32_t w, uint32_t (&r)[4]) {
  uint32_t p[4];
  e2m1x8_to_raw_bf16(w, p);                  // p[i] = (n_i, n_{i+4})
  r[0] = __byte_perm(p[0], p[1], 0x5410);    // (n0, n1)
  r[2] = __byte_perm(p[0], p[1], 0x7632);    // (n4, n5)
  r[1] = __byte_perm(p[2], p[3], 0x5410);    // (n2, n3)
  r[3] = __byte_perm(p[2], p[3], 0x7632);    // (n6, n7)
}
// A raw e2m1 pair (value * 2^-126) times 2^(e-1) is value * 2^(e8m0-127).  The bf16 code of 2^(e-1)
// has biased exponent e + 126, valid for e <= 128 (e = 0 included: 2^-1); larger scales (> 2) take a
// second exact factor 2^(e-128).  bf16x2_pow2(b) = the bf16 number with biased exponent b, both halves.
__device__ __forceinline__ uint32_t bf16x2_pow2(int biased) {  // 1 <= biased <= 254
  const uint32_t b = (uint32_t)biased << 7;
  return b | (b << 16);
}

template <int FMT, int STAGE>
__global__ void __launch_bounds__(kThreads, 1)
    umma_dq_kernel(const __grid_constant__ CUtensorMap tmap_w, const UmmaArgs a) {
  using T = DqTraits<FMT>;
  constexpr int RB = T::RB, RAWS = T::RAWS;
  constexpr int kRawStageBytes = 256 * RB;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t wop_base = smem_base;                                  // 1 KB aligned (swizzled)
  const uint32_t x_base = wop_base + kWopStages * kWopBytes;            // 1 KB aligned (swizzled)
  const uint32_t raw_base = x_base + kXStages * kXTileBytes;
  const uint32_t epi_base = raw_base + RAWS * kRawStageBytes;
  const uint32_t bar_base = epi_base + kEpiWarps * kEpiBufBytes;
  // barriers: x_full[4] x_empty[4] wop_full[2] wop_empty[2] raw_full[RAWS] raw_empty[RAWS] tfull[2] tempty[2]
  auto x_full = [&](int s) { return bar_base + 8 * s; };
  auto x_empty = [&](int s) { return bar_base + 8 * (kXStages + s); };
  auto wop_full = [&](int s) { return bar_base + 8 * (2 * kXStages + s); };
  auto wop_empty = [&](int s) { return bar_base + 8 * (2 * kXStages + kWopStages + s); };
  constexpr int kRawBar0 = 2 * kXStages + 2 * kWopStages;
  auto raw_full = [&](int s) { return bar_base + 8 * (kRawBar0 + s); };
  auto raw_empty = [&](int s) { return bar_base + 8 * (kRawBar0 + RAWS + s); };
  constexpr int kTBar0 = kRawBar0 + 2 * RAWS;
  auto tfull_bar = [&](int b) { return bar_base + 8 * (kTBar0 + b); };
  auto tempty_bar = [&](int b) { return bar_base + 8 * (kTBar0 + 2 + b); };
  const uint32_t tmem_slot = bar_base + 8 * (kTBar0 + 4);
  static_assert(8 * (kTBar0 + 4) + 4 <= 512, "barrier region overflow");
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_kb = a.K / kKPerStage;
  const int n_blocks = num_weight_blocks<STAGE>(a);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w)) : "memory");
    for (int s = 0; s < kXStages; ++s) {
      mbar_init(x_full(s), kGatherThreads);
      mbar_init(x_empty(s), 1);
    }
    for (int s = 0; s < kWopStages; ++s) {
      mbar_init(wop_full(s), kDqThreads);
      mbar_init(wop_empty(s), 1);
    }
    for (int s = 0; s < RAWS; ++s) {
      mbar_init(raw_full(s), 1);
      mbar_init(raw_empty(s), kDqThreads);
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
  pdl_wait();
  pdl_trigger();
  const int total = a.counters[2] * n_blocks;

  if (warp == 0) {
    // ================= TMA producer: raw weights =================
    {
      uint32_t it = 0;
      for (int w = blockIdx.x; w < total; w += gridDim.x) {
        const WorkItem wi = get_item(a, w, n_blocks);
        const int wrow0 = STAGE == 1 ? wi.nb * kWSub : wi.nb * 2 * kWSub;
        const int wrow1 = STAGE == 1 ? a.N + wi.nb * kWSub : wrow0 + kWSub;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % RAWS;
          mbar_wait_relaxed(raw_empty(s), ((it / RAWS) & 1) ^ 1);
          if (elect_one()) {
            mbar_arrive_expect_tx(raw_full(s), kRawStageBytes);
            const uint32_t dst = raw_base + s * kRawStageBytes;
            tma_load_3d(dst, &tmap_w, raw_full(s), kb * RB, wrow0, wi.e);
            tma_load_3d(dst + 128 * RB, &tmap_w, raw_full(s), kb * RB, wrow1, wi.e);
          }
          __syncwarp();
        }
      }
    }
      __syncwarp();
  } else if (warp == 1) {
    // ================= MMA issuer (converged warp, one elected lane issues) =================
    uint32_t it = 0, li = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x, ++li) {
      const WorkItem wi = get_item(a, w, n_blocks);
      const int n_eff = (wi.rows + 15) & ~15;
      const uint32_t idesc = instr_desc<SGLB_FMT_BF16>(n_eff);
      const int buf = li & 1;
      mbar_wait_relaxed(tempty_bar(buf), ((li >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t acc = tmem_base + buf * kAccCols;
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        const int sx = it % kXStages, sw = it % kWopStages;
        mbar_wait_relaxed(x_full(sx), (it / kXStages) & 1);
        mbar_wait_relaxed(wop_full(sw), (it / kWopStages) & 1);
        fence_proxy_async();
        tc_fence_after();
        const uint32_t wt = wop_base + sw * kWopBytes, xt = x_base + sx * kXTileBytes;
        if (elect_one()) {
#pragma unroll
          for (int sub = 0; sub < 2; ++sub) {
#pragma unroll
            for (int j = 0; j < kRowBytes / 32; ++j) {
              const uint64_t ad = smem_desc_sw128(wt + sub * kWTileBytes + j * 32);
              const uint64_t bd = smem_desc_sw128(xt + j * 32);
              umma<SGLB_FMT_BF16>(acc + sub * kWSub, ad, bd, idesc, (kb > 0 || j > 0) ? 1u : 0u);
            }
          }
          umma_commit(x_empty(sx));
          umma_commit(wop_empty(sw));
          if (kb == num_kb - 1) umma_commit(tfull_bar(buf));
        }
        __syncwarp();
      }
    }
  } else if (warp < kFirstEpiWarp) {
    gather_role<true>(a, total, n_blocks, num_kb, a.K * 2, threadIdx.x - 64, x_base, kXTileBytes, x_full(0),
                x_empty(0), kXStages);
  } else if (warp < kFirstDqWarp) {
    epilogue_role<false, STAGE, true>(a, total, n_blocks, smem_raw, epi_base, tfull_bar(0), tempty_bar(0), tmem_base,
                                warp, warp - kFirstEpiWarp, lane);
  } else {
    // ================= dequant warps: raw ring -> bf16 operand ring =================
    // Task t = i * 256 + d (d = thread index among the 256): weight row t / CPR of the 256-row window,
    // raw 16-B chunk t % CPR, so consecutive threads read consecutive 16-B pieces (conflict free).
    constexpr int CPR = RB / 16;            // raw chunks per row: 4 (fp8), 2 (mxfp4)
    constexpr int TASKS = 256 * CPR / kDqThreads;
    constexpr int OCH = 8 / CPR;            // bf16 16-B chunks produced per raw chunk: 2 (fp8), 4 (mxfp4)
    const int d = threadIdx.x - kFirstDqWarp * 32;
    const int chunk = d % CPR;
    uint32_t it = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x) {
      const WorkItem wi = get_item(a, w, n_blocks);
      const int wrow0 = STAGE == 1 ? wi.nb * kWSub : wi.nb * 2 * kWSub;
      const int wrow1 = STAGE == 1 ? a.N + wi.nb * kWSub : wrow0 + kWSub;
      // per-task scale row pointers (rows past the expert's end read zero weights: clamp the index)
      const void* srow[TASKS];
#pragma unroll
      for (int i = 0; i < TASKS; ++i) {
        const int r = (i * kDqThreads + d) / CPR;  // 0..255
        int wr = (r < kWSub ? wrow0 + r : wrow1 + (r - kWSub));
        wr = min(wr, a.rows_per_expert - 1);
        if constexpr (FMT == SGLB_FMT_FP8_W8A16) {
          const int nblk = ceil_div(a.rows_per_expert, a.block_n);
          srow[i] = (const float*)a.Wscale + ((size_t)wi.e * nblk + wr / a.block_n) * a.scale_cols;
        } else {
          srow[i] = (const uint8_t*)a.Wscale + ((size_t)wi.e * a.rows_per_expert + wr) * a.scale_cols;
        }
      }
      // scales are fetched one k block ahead (registers), so their HBM/L2 latency overlaps the
      // conversion of the current block instead of stalling every stage
      auto load_scale = [&](int i, int kb) -> uint32_t {
        if constexpr (FMT == SGLB_FMT_FP8_W8A16)
          return __float_as_uint(__ldg((const float*)srow[i] + (kb * kKPerStage) / a.block_k));
        else
          return (uint32_t)__ldg((const uint8_t*)srow[i] + kb * 2 + chunk);
      };
      uint32_t sc_next[TASKS];
#pragma unroll
      for (int i = 0; i < TASKS; ++i) sc_next[i] = load_scale(i, 0);
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        uint32_t sc[TASKS];
#pragma unroll
        for (int i = 0; i < TASKS; ++i) sc[i] = sc_next[i];
        if (kb + 1 < num_kb) {
#pragma unroll
          for (int i = 0; i < TASKS; ++i) sc_next[i] = load_scale(i, kb + 1);
        }
        const int sr = it % RAWS, sw = it % kWopStages;
        mbar_wait(raw_full(sr), (it / RAWS) & 1);
        mbar_wait(wop_empty(sw), ((it / kWopStages) & 1) ^ 1);
        const uint8_t* raw = smem_raw + (raw_base + sr * kRawStageBytes - smem_u32(smem_raw));
        uint8_t* wop = smem_raw + (wop_base + sw * kWopBytes - smem_u32(smem_raw));
#pragma unroll
        for (int i = 0; i < TASKS; ++i) {
          const int t = i * kDqThreads + d;
          const int r = t / CPR;
          const int4 q = *reinterpret_cast<const int4*>(raw + t * 16);
          const uint32_t* qw = reinterpret_cast<const uint32_t*>(&q);
          uint32_t o[16 / CPR * 2];  // 8 (fp8) or 16 (mxfp4) bf16x2 words
          if constexpr (FMT == SGLB_FMT_FP8_W8A16) {
            const float sf = __uint_as_float(sc[i]);
#pragma unroll
            for (int k = 0; k < 4; ++k) e4m3x4_scale_to_bf16x4(qw[k], sf, o[2 * k], o[2 * k + 1]);
          } else {
            const int e = (int)sc[i];
            const uint32_t m1 = bf16x2_pow2((e < 128 ? e : 128) + 126);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              uint32_t rr[4];
              e2m1x8_to_raw_bf16_korder(qw[k], rr);
#pragma unroll
              for (int b = 0; b < 4; ++b) o[4 * k + b] = bf16x2_mul(rr[b], m1);
            }
            if (e > 128) {  // rare: scales above 2 need a second exact factor 2^(e-128)
              const uint32_t m2 = bf16x2_pow2(e - 1);
#pragma unroll
              for (int k = 0; k < 16; ++k) o[k] = bf16x2_mul(o[k], m2);
            }
          }
          uint8_t* orow = wop + (r >> 7) * kWTileBytes + (r & 127) * kRowBytes;
#pragma unroll
          for (int h = 0; h < OCH; ++h) {
            const int oc = (chunk * OCH + h) ^ (r & 7);
            *reinterpret_cast<int4*>(orow + oc * 16) =
                make_int4((int)o[4 * h], (int)o[4 * h + 1], (int)o[4 * h + 2], (int)o[4 * h + 3]);
          }
        }
        fence_proxy_async();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
        mbar_arrive(wop_full(sw));
        mbar_arrive(raw_empty(sr));
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

template <int FMT, int STAGE>
int launch_one(const UmmaArgs& a, cudaStream_t stream) {
  using T = DqTraits<FMT>;
  const uint64_t kbytes = FMT == SGLB_FMT_FP8_W8A16 ? (uint64_t)a.K : (uint64_t)a.K / 2;
  CUtensorMap map;
  int rc = weight_tensor_map(a.W, kbytes, (uint64_t)a.rows_per_expert, (uint64_t)a.E, T::RB, kWSub, 0, &map);
  if (rc) return rc;
  static bool attr_set = false;  // per instantiation
  if (!attr_set) {
    SGLB_CUDA(cudaFuncSetAttribute(umma_dq_kernel<FMT, STAGE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   smem_bytes<FMT>()));
    attr_set = true;
  }
  const int n_blocks = STAGE == 1 ? ceil_div(a.N, kWSub) : ceil_div(a.N, 2 * kWSub);
  const int64_t max_items = (int64_t)a.max_tiles * n_blocks;
  const int grid = persistent_grid(max_items, (int64_t)a.est_tiles * n_blocks);
  SGLB_CUDA(launch_pdl(umma_dq_kernel<FMT, STAGE>, dim3(grid), dim3(kThreads), smem_bytes<FMT>(), stream, map, a));
  return SGLB_OK;
}

}  // namespace

int launch_umma_dq(int fmt, int stage, const UmmaArgs& a, cudaStream_t stream) {
  if (fmt == SGLB_FMT_FP8_W8A16)
    return stage == 1 ? launch_one<SGLB_FMT_FP8_W8A16, 1>(a, stream) : launch_one<SGLB_FMT_FP8_W8A16, 2>(a, stream);
  return stage == 1 ? launch_one<SGLB_FMT_MXFP4, 1>(a, stream) : launch_one<SGLB_FMT_MXFP4, 2>(a, stream);
}

}  // namespace sglb
