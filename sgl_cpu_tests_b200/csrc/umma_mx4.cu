// K3/K4 for mxfp4 weights on the block-scaled tensor cores: no dequantisation instructions at all.
//
// The dequantising kernel (umma_dq.cu) widens every e2m1 nibble to bf16 with ALU instructions and is
// bound by them (gpt-oss shape, M = 128: 30 % of the HBM roofline; M = 4096: 42 % of the bf16 tensor
// peak).  tcgen05.mma kind::mxf8f6f4.block_scale consumes the format natively:
//
//   e2m1 nibbles  HBM --TMA (CU_TENSOR_MAP_DATA_TYPE_16U4_ALIGN16B: 16 nibbles -> a 16-byte slot, one
//                 byte of K per element, 128-B swizzle)--> smem --A operand, format E2M1--> tensor core
//   e8m0 scales   one per 32 k per weight row (test_mxfp4.py:69-88) = exactly the scale_vec::1X block of
//                 the instruction.  They are re-tiled once per weight (mx4_tile_scales_kernel) into the
//                 512-byte chunks tcgen05.cp wants -- 128 rows x 4 k-groups, byte (r%32)*16 + (r/32)*4 + g
//                 -- bulk-copied next to the weight stage and copied smem -> TMEM by the MMA thread.
//   activations   bf16 rows split once per call into hi + lo e4m3 pieces with a per-row power-of-two
//                 factor (split_rows_e4m3, umma_f8.cu); B operand format E4M3, its scale tile is the
//                 constant 2^0.  W.x = W.hi + W.lo: two MMAs at twice the bf16 rate.
//
// Numerics: e2m1 * 2^(e8m0-127) is exact, the products are exact in f32 and accumulate in the tensor
// core's f32 accumulator: the result equals the reference oracle's bf16-dequantised GEMM
// (test_mxfp4.py:90-128) up to the activation split (relative 2^-9 of the row maximum at worst).
//
// K need not be a multiple of the 128-element stage (gpt-oss: 2880 = 22.5 stages): the weight box of
// the last stage runs into the next row's first bytes (the tensor map declares K rounded up, the row
// stride stays K/2), and those columns meet zero activations (cp.async zero-fill) and the scale 2^0.
//
// TMEM (512 columns): two accumulator buffers of 2 sub-tiles x 112 token columns (tiles hold at most
// 112 routed rows so that the scale tiles fit), 6 x 8 columns of weight scales (one slot per smem
// stage), 4 columns of activation scales.
//
// Roles (12 warps): warp 0 TMA producer (weights + scale chunks), warp 1 MMA issuer, warps 2-3 gather
// producers (hi/lo rows, cp.async), warps 4-11 epilogue (shared with the other tcgen05 kernels).
#include "umma_common.cuh"

namespace sglb {

using namespace ummak;

namespace {

constexpr int kMaxStages = 6;
constexpr int kTileCapMax = 112;                  // routed rows per tile: 2 x 2 x 112 = 448 accumulator columns
constexpr int kAccStrideMx = 2 * kTileCapMax;     // columns per accumulator buffer
constexpr int kSfaCol0 = 2 * kAccStrideMx;        // 448: weight-scale slots, 8 columns per smem stage
constexpr int kSfbCol = kSfaCol0 + 8 * kMaxStages;  // 496: activation scales (constant 2^0)
constexpr int kSfChunk = 512;                     // bytes of one 128-row x 4-group scale tile
constexpr int kGatherThreadsMx = 64;
constexpr int kFirstEpiWarp = 4;
constexpr int kThreads = (kFirstEpiWarp + kEpiWarps) * 32;  // 384
constexpr int kBarBytes = 512;
constexpr int kSmemLimit = 232448;                // 227 KB per CTA on sm_100

__host__ __device__ constexpr int stage_bytes(int cap) { return 2 * kWTileBytes + 2 * cap * kRowBytes + 2 * kSfChunk; }
__host__ __device__ constexpr int fixed_bytes() { return kEpiWarps * kEpiBufBytes + kSfChunk + 1024 + kBarBytes; }

// Bounded wait: a wrong transaction count or descriptor must end in a trap, not in a hung GPU.
__device__ __forceinline__ void mbar_wait_bounded(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  for (uint32_t spin = 0;; ++spin) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity), "r"(1000u)
        : "memory");
    if (ok) return;
    if (spin > (1u << 21)) __trap();
  }
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}
// 16-byte cp.async that writes zeros when `valid` is false (src-size 0)
__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void* src, bool valid) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(valid ? 16u : 0u)
               : "memory");
}
// K-major, no swizzle: 8-row x 16-byte core matrices, 128 B between 8-row groups (the scale chunk is
// 32 rows x 16 bytes, row stride 16)
__device__ __forceinline__ uint64_t smem_desc_sf(uint32_t addr) {
  return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
}
// smem -> TMEM: 32 rows x 128 bits, replicated to the four 32-lane quadrants
__device__ __forceinline__ void tmem_cp_sf(uint32_t taddr, uint64_t desc) {
  asm volatile("tcgen05.cp.cta_group::1.32x128b.warpx4 [%0], %1;" ::"r"(taddr), "l"(desc) : "memory");
}
// block-scaled instruction descriptor: A e2m1 (5), B e4m3 (0), both K-major, e8m0 scales, M = 128, N = n
__device__ __forceinline__ uint32_t instr_desc_mx4(int n) {
  return (5u << 7) | (0u << 10) | ((uint32_t)(n >> 3) << 17) | (1u << 23) | ((uint32_t)(kWSub >> 4) << 24);
}
// D[tmem] (+)= (A * SFA) (B * SFB); `idesc` carries the byte (k-group) of the scale columns to use
__device__ __forceinline__ void umma_mx4(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                         uint32_t tsfa, uint32_t tsfb, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf8f6f4.block_scale [%0], %1, %2, %3, [%5], [%6], p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(tsfa), "r"(tsfb)
      : "memory");
}

// CONCAT (tiles of at most 48 routed rows, the decode regime): the lo pieces sit right behind the hi
// pieces in shared memory, one MMA of N = 2 n_eff covers both and the epilogue adds the two column
// ranges.  An MMA costs at least the 4 KB read of its weight operand (32 cycles) whatever N is (ncu:
// with separate hi / lo MMAs the tensor pipe was busy 57 % of the time at 4 rows per expert and bounded
// the kernel at 55 % of the HBM roofline), so halving their number matters more than their size.
template <int STAGE, bool CONCAT>
__global__ void __launch_bounds__(kThreads, 1)
    umma_mx4_kernel(const __grid_constant__ CUtensorMap tmap_w, const UmmaArgs a, const int nstages,
                    const int w_tx_bytes) {
  extern __shared__ uint8_t smem_raw[];
  const int cap = a.tile_cap;
  const int x_bytes = cap * kRowBytes;
  const int st_bytes = stage_bytes(cap);
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t epi_base = smem_base + nstages * st_bytes;
  const uint32_t sfb_smem = epi_base + kEpiWarps * kEpiBufBytes;
  const uint32_t bar_base = sfb_smem + kSfChunk;
  auto w_tile = [&](int s) { return smem_base + s * st_bytes; };
  auto xh_tile = [&](int s) { return smem_base + s * st_bytes + 2 * kWTileBytes; };
  auto xl_tile = [&](int s) { return smem_base + s * st_bytes + 2 * kWTileBytes + x_bytes; };
  auto sf_tile = [&](int s) { return smem_base + s * st_bytes + 2 * kWTileBytes + 2 * x_bytes; };
  auto full_bar = [&](int s) { return bar_base + 8 * s; };
  auto empty_bar = [&](int s) { return bar_base + 8 * (kMaxStages + s); };
  auto tfull_bar = [&](int b) { return bar_base + 8 * (2 * kMaxStages + b); };
  auto tempty_bar = [&](int b) { return bar_base + 8 * (2 * kMaxStages + 2 + b); };
  const uint32_t tmem_slot = bar_base + 8 * (2 * kMaxStages + 4);
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_kb = ceil_div(a.K, kRowBytes);  // 128 elements of K per stage
  const int n_blocks = num_weight_blocks<STAGE>(a);
  // scale tiles: [E][blocks][num_kb][512]; stage 1 keeps the gate rows' blocks before the up rows'
  const int sf_blk0 = STAGE == 1 ? ceil_div(a.N, kWSub) : 0;
  const int sf_blocks = STAGE == 1 ? 2 * sf_blk0 : ceil_div(a.rows_per_expert, kWSub);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w)) : "memory");
    for (int s = 0; s < nstages; ++s) {
      mbar_init(full_bar(s), 1 + kGatherThreadsMx);
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
    // activation scale tile: e8m0 127 = 2^0 for every token and k group
    uint32_t* sfb = reinterpret_cast<uint32_t*>(smem_raw + (sfb_smem - smem_u32(smem_raw)));
    for (int i = lane; i < kSfChunk / 4; i += 32) sfb[i] = 0x7F7F7F7Fu;
    fence_proxy_async();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  pdl_wait();
  pdl_trigger();
  const int total = a.counters[2] * n_blocks;

  if (warp == 0) {
    // ================= TMA producer: e2m1 weights (unpacked by TMA) + their scale chunks =================
    uint32_t it = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x) {
      const WorkItem wi = get_item(a, w, n_blocks);
      const int wrow0 = STAGE == 1 ? wi.nb * kWSub : wi.nb * 2 * kWSub;
      const int wrow1 = STAGE == 1 ? a.N + wi.nb * kWSub : wrow0 + kWSub;
      const int blk0 = STAGE == 1 ? wi.nb : 2 * wi.nb;
      const int blk1 = STAGE == 1 ? sf_blk0 + wi.nb : min(2 * wi.nb + 1, sf_blocks - 1);
      const uint8_t* sf0 = (const uint8_t*)a.Wsf + ((size_t)wi.e * sf_blocks + blk0) * num_kb * kSfChunk;
      const uint8_t* sf1 = (const uint8_t*)a.Wsf + ((size_t)wi.e * sf_blocks + blk1) * num_kb * kSfChunk;
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        const int s = it % nstages;
        mbar_wait_bounded(empty_bar(s), ((it / nstages) & 1) ^ 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(full_bar(s), 2 * w_tx_bytes + 2 * kSfChunk);
          tma_load_3d(w_tile(s), &tmap_w, full_bar(s), kb * kRowBytes, wrow0, wi.e);
          tma_load_3d(w_tile(s) + kWTileBytes, &tmap_w, full_bar(s), kb * kRowBytes, wrow1, wi.e);
          bulk_load(sf_tile(s), sf0 + (size_t)kb * kSfChunk, kSfChunk, full_bar(s));
          bulk_load(sf_tile(s) + kSfChunk, sf1 + (size_t)kb * kSfChunk, kSfChunk, full_bar(s));
        }
        __syncwarp();
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================= MMA issuer (converged warp, one elected lane issues) =================
    const uint32_t tsfb = tmem_base + kSfbCol;
    if (elect_one()) tmem_cp_sf(tsfb, smem_desc_sf(sfb_smem));
    __syncwarp();
    uint32_t it = 0, li = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x, ++li) {
      const WorkItem wi = get_item(a, w, n_blocks);
      const int n_eff = (wi.rows + 15) & ~15;
      const uint32_t idesc = instr_desc_mx4(CONCAT ? 2 * n_eff : n_eff);
      const int buf = li & 1;
      mbar_wait_bounded(tempty_bar(buf), ((li >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t acc = tmem_base + buf * kAccStrideMx;
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        const int s = it % nstages;
        mbar_wait_bounded(full_bar(s), (it / nstages) & 1);
        fence_proxy_async();  // routed rows were written by cp.async (generic proxy)
        tc_fence_after();
        if (elect_one()) {
          const uint32_t tsfa = tmem_base + kSfaCol0 + 8 * s;
          tmem_cp_sf(tsfa, smem_desc_sf(sf_tile(s)));
          tmem_cp_sf(tsfa + 4, smem_desc_sf(sf_tile(s) + kSfChunk));
#pragma unroll
          for (int sub = 0; sub < 2; ++sub) {
#pragma unroll
            for (int j = 0; j < kRowBytes / 32; ++j) {
              const uint64_t ad = smem_desc_sw128(w_tile(s) + sub * kWTileBytes + j * 32);
              const uint64_t bh = smem_desc_sw128(xh_tile(s) + j * 32);
              const uint64_t bl = smem_desc_sw128(xl_tile(s) + j * 32);
              const uint32_t id = idesc | ((uint32_t)j << 29);  // weight scale of k group j
              umma_mx4(acc + sub * kTileCapMax, ad, bh, id, tsfa + 4 * sub, tsfb, (kb > 0 || j > 0) ? 1u : 0u);
              if constexpr (!CONCAT) umma_mx4(acc + sub * kTileCapMax, ad, bl, id, tsfa + 4 * sub, tsfb, 1u);
            }
          }
          umma_commit(empty_bar(s));
          if (kb == num_kb - 1) umma_commit(tfull_bar(buf));
        }
        __syncwarp();
      }
    }
  } else if (warp < kFirstEpiWarp) {
    // ================= gather producers: hi and lo rows =================
    // 64 threads: thread g owns 16-B chunk c = g & 7 of rows (g >> 3) + 8 j; chunks past K read as zero
    const int g = threadIdx.x - 64;
    const int c = g & 7, rbase = g >> 3;
    const uint32_t dst0 = rbase * kRowBytes + ((c ^ (rbase & 7)) << 4);
    const char* xh = (const char*)a.X + c * 16;
    const char* xl = (const char*)a.X2 + c * 16;
    uint32_t it = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x) {
      const WorkItem wi = get_item(a, w, n_blocks);
      const int n_eff = (wi.rows + 15) & ~15;
      uint32_t off[kTileCapMax / 8];
#pragma unroll
      for (int j = 0; j < kTileCapMax / 8; ++j) {
        const int r = rbase + 8 * j;
        const int p = wi.row_start + min(r, wi.rows - 1);
        const int xr = a.gather ? (a.sorted_slot[p] / a.topk) : p;
        off[j] = (uint32_t)xr * (uint32_t)a.K;
      }
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        const int s = it % nstages;
        mbar_wait_bounded(empty_bar(s), ((it / nstages) & 1) ^ 1);
        const uint32_t dh = xh_tile(s) + dst0;
        const uint32_t dl = (CONCAT ? xh_tile(s) + n_eff * kRowBytes : xl_tile(s)) + dst0;
        const bool valid = kb * kRowBytes + c * 16 < a.K;
        const int koff = valid ? kb * kRowBytes : 0;
#pragma unroll
        for (int j = 0; j < kTileCapMax / 8; ++j) {
          if (rbase + 8 * j < n_eff) {
            cp_async16_zfill(dh + j * 8 * kRowBytes, xh + off[j] + koff, valid);
            cp_async16_zfill(dl + j * 8 * kRowBytes, xl + off[j] + koff, valid);
          }
        }
        cp_async_arrive_noinc(full_bar(s));
      }
    }
  } else {
    epilogue_role<false, STAGE, true, true, kAccStrideMx, kTileCapMax, CONCAT>(
        a, total, n_blocks, smem_raw, epi_base, tfull_bar(0), tempty_bar(0), tmem_base, warp, warp - kFirstEpiWarp,
        lane);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols)
                 : "memory");
  }
}

// ---- e8m0 scales [E, rows, K/32] -> tcgen05.cp tiles -------------------------------------------------
// Block b of an expert covers 128 weight rows: blocks 0..nblk0-1 the rows [0, split) (or all rows when
// split == 0), the rest the rows [split, rows).  One thread per (block, k stage, row): 4 bytes.
__global__ void __launch_bounds__(128) mx4_tile_scales_kernel(const uint8_t* __restrict__ src,
                                                              uint8_t* __restrict__ dst, int rows, int split,
                                                              int K, int nblk0, int nblk, int num_kb) {
  const int kb = blockIdx.x, b = blockIdx.y;
  const int64_t e = blockIdx.z;
  const int r = threadIdx.x;
  const int seg_start = b < nblk0 ? 0 : split;
  const int seg_end = b < nblk0 ? (split > 0 ? split : rows) : rows;
  const int row = seg_start + (b < nblk0 ? b : b - nblk0) * 128 + r;
  const int G = K / 32;
  uint32_t v = 0;
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    const int kg = kb * 4 + g;
    uint32_t s = 127;  // 2^0 for padding rows / k groups (they meet zero weights or zero activations)
    if (row < seg_end && kg < G) s = src[(e * rows + row) * G + kg];
    v |= s << (8 * g);
  }
  uint8_t* chunk = dst + ((e * nblk + b) * num_kb + kb) * (int64_t)kSfChunk;
  *reinterpret_cast<uint32_t*>(chunk + (r & 31) * 16 + (r >> 5) * 4) = v;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// e2m1 weights [E, rows, K] (K/2 bytes per row): 4-bit elements, box {128 elements, 128 rows, 1}
int mx4_weight_map(const void* W, int K, int rows, int E, CUtensorMap* out) {
  static EncodeTiledFn enc = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  SGLB_CHECK(enc != nullptr, "mxfp4 tensor path: cuTensorMapEncodeTiled is not available from this driver");
  SGLB_CHECK(((uintptr_t)W & 31) == 0, "mxfp4 tensor path: weights must be 32-byte aligned");
  const cuuint64_t gdim[3] = {(cuuint64_t)ceil_div(K, 128) * 128, (cuuint64_t)rows, (cuuint64_t)E};
  const cuuint64_t gstride[2] = {(cuuint64_t)K / 2, (cuuint64_t)K / 2 * rows};
  const cuuint32_t box[3] = {128, (cuuint32_t)kWSub, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_16U4_ALIGN16B, 3, const_cast<void*>(W), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SGLB_CHECK(r == CUDA_SUCCESS, "mxfp4 tensor path: cuTensorMapEncodeTiled(16U4_ALIGN16B) failed (CUresult %d)", (int)r);
  return SGLB_OK;
}

constexpr int kConcatCapMax = 48;  // 2 x 48 columns per sub-tile fit the 112-column accumulator slot

template <int STAGE, bool CONCAT>
int launch_one(const UmmaArgs& a, cudaStream_t stream) {
  CUtensorMap map;
  int rc = mx4_weight_map(a.W, a.K, a.rows_per_expert, a.E, &map);
  if (rc) return rc;
  const int cap = a.tile_cap;
  SGLB_CHECK(cap >= 16 && cap <= kTileCapMax && cap % 16 == 0, "mxfp4 tensor path: bad tile capacity %d", cap);
  int nstages = (kSmemLimit - fixed_bytes()) / stage_bytes(cap);
  if (nstages > kMaxStages) nstages = kMaxStages;
  static const int forced_stages = [] {
    const char* e = getenv("SGLB_MX4_STAGES");  // experiments
    return e ? atoi(e) : 0;
  }();
  if (forced_stages >= 2 && forced_stages < nstages) nstages = forced_stages;
  const int smem = nstages * stage_bytes(cap) + fixed_bytes();
  static int attr_smem = 0;  // per instantiation
  if (attr_smem < smem) {
    SGLB_CUDA(cudaFuncSetAttribute(umma_mx4_kernel<STAGE, CONCAT>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    attr_smem = kSmemLimit;
  }
  // bytes one weight box adds to the mbarrier transaction count: the packed nibbles (the 8-byte gaps of
  // the 16-byte slots are not written).  SGLB_MX4_TX overrides for bring-up.
  static const int tx = [] {
    const char* s = getenv("SGLB_MX4_TX");
    return s ? atoi(s) : kWTileBytes / 2;
  }();
  const int n_blocks = STAGE == 1 ? ceil_div(a.N, kWSub) : ceil_div(a.N, 2 * kWSub);
  const int64_t max_items = (int64_t)a.max_tiles * n_blocks;
  const int grid = persistent_grid(max_items, (int64_t)a.est_tiles * n_blocks);
  SGLB_CUDA(launch_pdl(umma_mx4_kernel<STAGE, CONCAT>, dim3(grid), dim3(kThreads), (size_t)smem, stream, map, a, nstages, tx));
  return SGLB_OK;
}

}  // namespace

// K: whole 32-element scale groups and a 32-byte row stride (TMA rule for 4-bit data); I: 16-B stores
int umma_mx4_supported(int H, int I) { return H % 64 == 0 && I % 64 == 0; }

// Tile row capacity handed to moe_align: small tiles leave shared memory for more weight stages in
// flight (decode), 112 is the most the TMEM budget allows (prefill).
int umma_mx4_tile_rows(double rows_per_expert) {
  static const int forced = [] {
    const char* e = getenv("SGLB_MX4_CAP");  // experiments
    return e ? atoi(e) : 0;
  }();
  if (forced >= 16 && forced <= kTileCapMax && forced % 16 == 0) return forced;
  return rows_per_expert <= 8.0 ? 32 : (rows_per_expert <= 40.0 ? kConcatCapMax : kTileCapMax);
}

int launch_umma_mx4(int stage, const UmmaArgs& a, cudaStream_t stream) {
  SGLB_CHECK(a.Wsf != nullptr && a.X2 != nullptr && a.Xscale != nullptr, "mxfp4 tensor path: missing operand");
  int rc;
  if (a.tile_cap <= kConcatCapMax)
    rc = stage == 1 ? launch_one<1, true>(a, stream) : launch_one<2, true>(a, stream);
  else
    rc = stage == 1 ? launch_one<1, false>(a, stream) : launch_one<2, false>(a, stream);
  if (rc == SGLB_OK) count_launch();
  return rc;
}

static void mx4_blocks(int rows, int split, int* nblk0, int* nblk) {
  *nblk0 = ceil_div(split > 0 ? split : rows, 128);
  *nblk = *nblk0 + (split > 0 ? ceil_div(rows - split, 128) : 0);
}

size_t mx4_scale_tiles_bytes(int64_t E, int rows, int split, int K) {
  int nblk0, nblk;
  mx4_blocks(rows, split, &nblk0, &nblk);
  return (size_t)E * nblk * ceil_div(K, 128) * kSfChunk;
}

int launch_mx4_tile_scales(const void* src, void* dst, int64_t E, int rows, int split, int K, cudaStream_t stream) {
  SGLB_CHECK(K % 32 == 0 && rows > 0 && split >= 0 && split < rows, "mx4_tile_scales: bad shape rows=%d split=%d K=%d",
             rows, split, K);
  if (E == 0) return SGLB_OK;
  int nblk0, nblk;
  mx4_blocks(rows, split, &nblk0, &nblk);
  const int num_kb = ceil_div(K, 128);
  mx4_tile_scales_kernel<<<dim3(num_kb, nblk, (unsigned)E), 128, 0, stream>>>((const uint8_t*)src, (uint8_t*)dst, rows,
                                                                              split, K, nblk0, nblk, num_kb);
  SGLB_CUDA(cudaGetLastError());
  count_launch();
  return SGLB_OK;
}

}  // namespace sglb
