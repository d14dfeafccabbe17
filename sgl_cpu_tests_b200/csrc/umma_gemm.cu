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
#include <algorithm>
#include <iterator>
#include <mutex>
#include <unordered_map>
#include <vector>

#include <stdlib.h>

#include "umma_common.cuh"

namespace sglb {

using namespace ummak;

namespace {

constexpr int kStages = 4;
constexpr int kStageBytes = 2 * kWTileBytes + kXTileBytes;  // 48 KB
constexpr int kFirstEpiWarp = 6;
constexpr int kThreads = (kFirstEpiWarp + kEpiWarps) * 32;  // 448
constexpr int kReducerWarp = kFirstEpiWarp + kEpiWarps;     // 14: present only in the fused decode chain (stage 2)
constexpr int kThreadsFused = kThreads + 32;
constexpr int kSmemBytes = kStages * kStageBytes + kEpiWarps * kEpiBufBytes + 1024 /*alignment slack*/ +
                           256 /*barriers*/;

// RED: the level-2 chain-fusion variant of stage 2 with the reducer warp (its own instantiation, so that the default
// kernels keep the register budget of a 448-thread block)
template <int FMT, int STAGE, bool RED = false>
__global__ void __launch_bounds__(RED ? kThreadsFused : kThreads, 1)
    umma_moe_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_sh,
                    const UmmaArgs a) {
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
  const uint32_t red_bar0 = bar_base + 128;  // fused reduce: full[2], empty[2] between the epilogue and the reducer warp
  constexpr bool fused_reduce = RED;  // then the block has kThreadsFused threads
  uint32_t* tmem_slot_ptr =
      reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int EB = FMT == SGLB_FMT_INT8_W8A8 ? 1 : 2;  // operand element bytes
  const int k_bytes = a.K * EB;
  const int num_kb = k_bytes / kRowBytes;
  const int n_blocks = num_weight_blocks<STAGE>(a);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w)) : "memory");
    if (a.Wsh != nullptr) asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_sh)) : "memory");
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1 + kGatherThreads);
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tfull_bar(b), 1);
      mbar_init(tempty_bar(b), kEpiThreads);
      mbar_init(red_bar0 + 8 * b, kEpiThreads);
      mbar_init(red_bar0 + 8 * (2 + b), 1);
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
    {
      uint32_t it = 0;
      // The descriptor of the next item is fetched while this item streams: a dependent tile_list load at
      // every item start leaves the ring's freed slots unissued for its latency (one item is only ~11 us of
      // HBM time per SM for int8 at decode).
      WorkItem nxt = {0, 0, 0, 0};
      if ((int)blockIdx.x < total) nxt = get_item(a, blockIdx.x, n_blocks);
      for (int w = blockIdx.x; w < total; w += gridDim.x) {
        const WorkItem wi = nxt;
        if (w + (int)gridDim.x < total) nxt = get_item(a, w + gridDim.x, n_blocks);
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
    // The whole warp runs the loop converged and one elected lane issues: under `if (lane == 0)` ptxas
    // wraps every tcgen05.mma in a per-lane waterfall loop (~65 issue cycles per MMA, ncu r1e), which
    // made the issuing thread -- not the tensor pipe -- the bottleneck at prefill.
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
        if (elect_one()) {
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
          if (kb == num_kb - 1) umma_commit(tfull_bar(buf));  // accumulators of this item are complete
        }
        __syncwarp();
      }
    }
  } else if (warp < 6) {
    if (RED && FMT == SGLB_FMT_INT8_W8A8 && a.Xraw != nullptr)  // fused decode chain: quantise on the way in
      gather_quant_role(a, total, n_blocks, num_kb, threadIdx.x - 64, x_tile(0), kStageBytes, full_bar(0), empty_bar(0),
                        kStages);
    else
      gather_role<false>(a, total, n_blocks, num_kb, k_bytes, threadIdx.x - 64, x_tile(0), kStageBytes, full_bar(0),
                         empty_bar(0), kStages);
  } else if (warp < kReducerWarp) {
    epilogue_role<FMT == SGLB_FMT_INT8_W8A8, STAGE, false>(a, total, n_blocks, smem_raw, epi_base, tfull_bar(0),
                                                    tempty_bar(0), tmem_base, warp, warp - kFirstEpiWarp, lane,
                                                    fused_reduce ? red_bar0 : 0u);
  } else {
    if constexpr (RED) reducer_role(a, total, n_blocks, red_bar0, lane);
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
  uint32_t box_bytes, box_rows, swizzle;
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && kbytes == o.kbytes && rows == o.rows && E == o.E && box_bytes == o.box_bytes &&
           box_rows == o.box_rows && swizzle == o.swizzle;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    return std::hash<const void*>()(k.ptr) ^ (k.kbytes * 0x9E3779B97F4A7C15ull) ^ (k.rows << 20) ^ (k.E << 44) ^
           ((uint64_t)k.box_bytes << 7) ^ k.swizzle;
  }
};

}  // namespace

// Weight map: bytes {K_bytes, rows_per_expert, E}, box {box_bytes, box_rows, 1}, 128-B swizzle or none.
// The expert dimension is separate so that a box reaching past an expert's last row is zero-filled.
int weight_tensor_map(const void* W, uint64_t kbytes, uint64_t rows, uint64_t E, uint32_t box_bytes,
                      uint32_t box_rows, int swizzle128, CUtensorMap* out) {
  // descriptor cache with least-recently-used eviction: every entry carries the tick of its last use and, when the
  // table is full, the older half is dropped (a descriptor is 128 bytes and ~1 us to rebuild, so this is cheap)
  struct Entry {
    CUtensorMap map;
    uint64_t tick;
  };
  constexpr size_t kMaxEntries = 4096;
  static std::mutex mu;
  static std::unordered_map<MapKey, Entry, MapKeyHash> cache;
  static uint64_t tick = 0;
  const MapKey key{W, kbytes, rows, E, box_bytes, box_rows, (uint32_t)swizzle128};
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(key);
  if (it != cache.end()) {
    it->second.tick = ++tick;
    *out = it->second.map;
    return SGLB_OK;
  }
  EncodeTiledFn enc = encode_tiled_fn();
  SGLB_CHECK(enc != nullptr, "tcgen05 path: cuTensorMapEncodeTiled is not available from this driver");
  SGLB_CHECK(((uintptr_t)W & 15) == 0 && kbytes % 16 == 0, "tcgen05 path: weights must be 16-byte aligned");
  const cuuint64_t gdim[3] = {kbytes, rows, E};
  const cuuint64_t gstride[2] = {kbytes, kbytes * rows};
  const cuuint32_t box[3] = {box_bytes, box_rows, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(W), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SGLB_CHECK(r == CUDA_SUCCESS, "tcgen05 path: cuTensorMapEncodeTiled failed (CUresult %d)", (int)r);
  if (cache.size() >= kMaxEntries) {
    std::vector<uint64_t> ticks;
    ticks.reserve(cache.size());
    for (const auto& kv : cache) ticks.push_back(kv.second.tick);
    std::nth_element(ticks.begin(), ticks.begin() + ticks.size() / 2, ticks.end());
    const uint64_t cut = ticks[ticks.size() / 2];
    for (auto i = cache.begin(); i != cache.end();) i = i->second.tick < cut ? cache.erase(i) : std::next(i);
  }
  cache.emplace(key, Entry{m, ++tick});
  *out = m;
  return SGLB_OK;
}

int rows_tensor_map(const void* X, uint64_t kbytes, uint64_t rows, uint32_t box_rows, CUtensorMap* out) {
  EncodeTiledFn enc = encode_tiled_fn();
  SGLB_CHECK(enc != nullptr, "tcgen05 path: cuTensorMapEncodeTiled is not available from this driver");
  SGLB_CHECK(((uintptr_t)X & 15) == 0 && kbytes % 16 == 0, "tcgen05 path: routed rows must be 16-byte aligned");
  const cuuint64_t gdim[2] = {kbytes, rows};
  const cuuint64_t gstride[1] = {kbytes};
  const cuuint32_t box[2] = {kRowBytes, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(X), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SGLB_CHECK(r == CUDA_SUCCESS, "tcgen05 path: cuTensorMapEncodeTiled (routed rows) failed (CUresult %d)", (int)r);
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

namespace {

template <int FMT, int STAGE, bool RED = false>
int launch_one(const UmmaArgs& a, cudaStream_t stream) {
  if constexpr (STAGE == 2 && !RED) {
    if (a.red.cnt != nullptr || a.Xraw != nullptr) return launch_one<FMT, 2, true>(a, stream);
  }
  constexpr int EB = FMT == SGLB_FMT_INT8_W8A8 ? 1 : 2;
  CUtensorMap map;
  int rc = weight_tensor_map(a.W, (uint64_t)a.K * EB, (uint64_t)a.rows_per_expert, (uint64_t)a.E, kRowBytes, kWSub, 1,
                             &map);
  if (rc) return rc;
  CUtensorMap map_sh = map;
  if (a.Wsh != nullptr) {
    rc = weight_tensor_map(a.Wsh, (uint64_t)a.K * EB, (uint64_t)a.rows_per_expert, 1, kRowBytes, kWSub, 1, &map_sh);
    if (rc) return rc;
  }
  static bool attr_set = false;  // per instantiation
  if (!attr_set) {
    SGLB_CUDA(cudaFuncSetAttribute(umma_moe_kernel<FMT, STAGE, RED>,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_set = true;
  }
  const int n_blocks = STAGE == 1 ? ceil_div(a.N, kWSub) : ceil_div(a.N, 2 * kWSub);
  const int64_t max_items = (int64_t)a.max_tiles * n_blocks;
  int grid = persistent_grid(max_items, (int64_t)a.est_tiles * n_blocks);  // whole waves in the streaming regime
  static const int forced_grid = [] {
    const char* e = getenv("SGLB_GRID");  // experiments: persistent grid size
    return e ? atoi(e) : 0;
  }();
  if (forced_grid > 0 && forced_grid < grid) grid = forced_grid;
  const int threads = RED ? kThreadsFused : kThreads;  // + the reducer warp
  SGLB_CUDA(launch_pdl(umma_moe_kernel<FMT, STAGE, RED>, dim3(grid), dim3(threads), kSmemBytes, stream, map, map_sh, a));
  return SGLB_OK;
}

}  // namespace

int umma_supported(int fmt, int64_t M, int topk, int E, int H, int I) {
  (void)M;
  (void)topk;
  (void)E;
  if (fmt == SGLB_FMT_FP8_W8A16 || fmt == SGLB_FMT_MXFP4) return H % 64 == 0 && I % 64 == 0;
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
  else if (fmt == SGLB_FMT_FP8_W8A16 || fmt == SGLB_FMT_MXFP4)
    rc = launch_umma_dq(fmt, stage, a, stream);
  else {
    set_error("tcgen05 path: unsupported weight format %d", fmt);
    return SGLB_ERR_INVALID;
  }
  if (rc == SGLB_OK) count_launch();
  return rc;
}

}  // namespace sglb
