// C-ABI entry points that orchestrate the kernels: fused_experts, shared_expert, mxfp4_mm.
#include <atomic>
#include <vector>
#include <math.h>
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>

#include "common.cuh"
#include "rowops.cuh"
#include "stream_gemm.cuh"
#include "umma_gemm.cuh"

namespace sglb {

static thread_local char g_err[1024] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

// ---- optional per-stage CUDA-event timing (bench.py's roofline leg) ------------------------------
// When enabled, fused_experts brackets each stage with events on the launching stream; the events
// are resolved by sglb_profile_read() after the caller has synchronised.
enum { ST_ALIGN = 0, ST_QUANT_X, ST_GEMM1, ST_QUANT_ACT, ST_GEMM2, ST_REDUCE, ST_COUNT };
struct StageEvents {
  cudaEvent_t a, b;
  int stage;
};
static bool g_profile = false;
static bool g_mx4_native = true;  // mxfp4: block-scaled tensor-native kernel (umma_mx4.cu) when scale tiles are given
static bool g_f8_native = true;  // fp8: tensor-native kernel (umma_f8.cu) vs dequantising kernel (umma_dq.cu)
static bool g_f8_small_tiles = true;  // fp8 decode: 32-row tiles on the deep-ring configuration (SGLB_F8_SMALL_TILES=0: off)
static bool g_f8_mid_tiles = false;   // fp8: 80-row tiles on the 4-stage ring (experiment: SGLB_F8_SMALL_TILES=2 / set_fp8_native bit 2)
// Decode chain fusion level (sglb_set_chain_fusion).  1 (default): align + stage-1 operand preparation in ONE prologue
// launch.  2: also the int8 act quantisation inside the stage-2 producer and the top-k reduce inside the stage-2 kernel
// (3 launches per call) -- bit-identical, but measured SLOWER on B200 (gpurun r2n: Qwen3 int8 M = 128 step 0.130 ->
// 0.158 ms, bf16 0.204 -> 0.222 ms): under a saturated HBM every dependent global round trip costs 2-3 us, and the
// fence -> arrival atomic -> slot rows chain of a 4-9 us work item does not fit behind it.  Kept as an experiment.
static int g_chain_fusion = 1;
static std::vector<StageEvents> g_events;
static size_t g_events_used = 0;

struct StageTimer {
  cudaStream_t s;
  int idx = -1;
  StageTimer(int stage, cudaStream_t stream) : s(stream) {
    if (!g_profile) return;
    if (g_events_used == g_events.size()) {
      StageEvents e;
      cudaEventCreate(&e.a);
      cudaEventCreate(&e.b);
      g_events.push_back(e);
    }
    idx = (int)g_events_used++;
    g_events[idx].stage = stage;
    cudaEventRecord(g_events[idx].a, s);
  }
  ~StageTimer() {
    if (idx >= 0) cudaEventRecord(g_events[idx].b, s);
  }
};

int launch_shared_axpy(const void* y, const void* fused, float rsf, int64_t n, void* out,
                       cudaStream_t stream);
int launch_quant_rows_f32(const float* x, int64_t rows, int K, int8_t* q, float* scale,
                          cudaStream_t stream, const int32_t* n_valid = nullptr, int always_from = 0);
int launch_gather_rows(const int32_t* sorted_slot, int64_t S, int topk, int row_bytes, const void* x, const void* x2, void* y,
                       void* y2, cudaStream_t stream);

// ---- workspace carving -------------------------------------------------------------------------
struct MoeWorkspace {
  int32_t *expert_count, *expert_offset, *counters, *sorted_slot, *slot_pos, *chunk_list, *tile_list;
  int32_t* align_scratch;
  int32_t* zero_ids;          // [M] (shared expert / dense callers)
  __nv_bfloat16* act;         // [S, I]
  __nv_bfloat16* ic2;         // [S, H]
  int8_t* xq;                 // [M, H]
  float* xs;                  // [M]
  int8_t* actq;               // [S, I]
  float* acts;                // [S]
  int8_t *xq2, *actq2;        // fp8 tensor-native path: lo pieces (xq / actq hold the hi pieces)
  int8_t *xsorted, *xsorted2; // CTA-pair path: the stage-1 rows in expert-sorted order (bf16 / s8 / e4m3 hi, lo)
  // fused decode chain (M * topk <= kFuseMaxSlots): per-token valid slots, arrival counters, per-row max |act|
  int32_t *tok_valid, *red_cnt;
  uint32_t* row_amax;
  size_t bytes;
};

constexpr int kMinChunkRows = 8;
// Measured on B200 (Qwen3 / bench_moe.py shapes, r3i): the persistent TMA-fed tcgen05 kernel streams weights faster than
// the mma.sync kernel from 2 routed rows per expert upwards (M = 32 / 48: 110.6 vs 114.6 / 116.6 vs 122.5 us) and for a
// handful of slots (M = 1: 32.0 vs 37.2 us; M = 4: 103 vs 109 us bf16); in between the streaming kernel wins (M = 16:
// 88 vs 96 us).
constexpr double kUmmaMinRowsPerExpert = 2.0;
constexpr int64_t kUmmaTinySlots = 32;
constexpr double kPairMinRowsPerExpert = 128.0;
constexpr int kTileRows = kUmmaTileRows;
constexpr int kMinTileRows = 16;  // smallest tile_rows any path hands to moe_align (sizes tile_list)
// Decode chain fusion applies up to this many routed slots (the single-launch prologue's align blocks)
constexpr int64_t kFuseMaxSlots = (int64_t)kPrologueMaxAlignBlocks * 1024;
static int red_blocks_max(int H) { return ceil_div(H, 32) + 8; }  // counters per token: stream path, 32 columns per warp

static int64_t max_items(int64_t S, int E, int rows) {
  return (S < E ? S : E) + S / rows + 1;
}

static MoeWorkspace carve(void* base, int64_t M, int topk, int E, int H, int I, int fmt) {
  MoeWorkspace w;
  w.xq2 = nullptr;
  w.actq2 = nullptr;
  const int64_t S = M * topk + M;  // + M rows behind the routed ones: the shared expert riding along (align.cu)
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? (char*)base + off : nullptr;
    off += align_up(bytes ? bytes : 1, 256);
    return p;
  };
  w.expert_count = (int32_t*)take(sizeof(int32_t) * E);
  w.expert_offset = (int32_t*)take(sizeof(int32_t) * (E + 1));
  w.counters = (int32_t*)take(sizeof(int32_t) * 4);
  w.sorted_slot = (int32_t*)take(sizeof(int32_t) * S);
  w.slot_pos = (int32_t*)take(sizeof(int32_t) * S);
  w.chunk_list = (int32_t*)take(sizeof(int32_t) * 3 * max_items(S, E, kMinChunkRows));
  w.tile_list = (int32_t*)take(sizeof(int32_t) * 3 * (max_items(S, E, kMinTileRows) + M / kMinTileRows + 1));
  w.align_scratch = (int32_t*)take(sglb_moe_align_scratch_bytes(M, topk, E));
  w.zero_ids = (int32_t*)take(sizeof(int32_t) * M);
  // +kTileRows rows of slack: the tcgen05 path reads whole 128-row tiles past the last expert
  // int8 keeps the silu_and_mul output in f32 until it is re-quantised
  w.act = (__nv_bfloat16*)take((fmt == SGLB_FMT_INT8_W8A8 ? sizeof(float) : sizeof(__nv_bfloat16)) * (S + kTileRows) * I);
  w.ic2 = (__nv_bfloat16*)take(sizeof(__nv_bfloat16) * S * H);
  if (fmt == SGLB_FMT_INT8_W8A8) {
    w.xq = (int8_t*)take((size_t)M * H);
    w.xs = (float*)take(sizeof(float) * M);
    w.actq = (int8_t*)take((size_t)(S + kTileRows) * I);
    w.acts = (float*)take(sizeof(float) * (S + kTileRows));
  } else if (fmt == SGLB_FMT_FP8_W8A16 || fmt == SGLB_FMT_MXFP4) {
    w.xq = (int8_t*)take((size_t)M * H);
    w.xq2 = (int8_t*)take((size_t)M * H);
    w.xs = (float*)take(sizeof(float) * M);
    w.actq = (int8_t*)take((size_t)(S + kTileRows) * I);
    w.actq2 = (int8_t*)take((size_t)(S + kTileRows) * I);
    w.acts = (float*)take(sizeof(float) * (S + kTileRows));
  } else {
    w.xq = nullptr;
    w.xs = nullptr;
    w.actq = nullptr;
    w.acts = nullptr;
  }
  // the CTA-pair kernels take the stage-1 rows pre-gathered (one row per routed slot)
  const size_t eb = fmt == SGLB_FMT_BF16 ? 2 : 1;
  w.xsorted = (int8_t*)take((size_t)M * topk * H * eb);
  w.xsorted2 = fmt == SGLB_FMT_FP8_W8A16 ? (int8_t*)take((size_t)M * topk * H) : nullptr;
  const bool fusable = M * topk <= kFuseMaxSlots;
  w.tok_valid = (int32_t*)take(fusable ? sizeof(int32_t) * M : 0);
  w.red_cnt = (int32_t*)take(fusable ? sizeof(int32_t) * M * red_blocks_max(H) : 0);
  w.row_amax = (uint32_t*)take(fusable ? sizeof(uint32_t) * (S + kTileRows) : 0);
  w.bytes = off;
  return w;
}

// ---- plan for the streaming kernels ---------------------------------------------------------------
static StreamPlan make_plan(int64_t M, int topk, int E, int n_rows_stage, int stage) {
  StreamPlan p;
  const double S = (double)M * topk;
  double active = E * (1.0 - pow(1.0 - (double)(topk < E ? topk : E) / E, (double)M));
  if (E == 1) active = 1;
  if (active < 1) active = 1;
  const double avg = S / active;
  p.ng = avg <= 6.0 ? 1 : (avg <= 12.0 ? 2 : 4);
  const double chunks = active + S / (8.0 * p.ng);
  const int per_warp = stage == 1 ? 16 : 32;
  auto ctas = [&](int wn) { return chunks * ceil_div(n_rows_stage, per_warp * wn); };
  p.cfg = ctas(8) >= 2 * 148 ? 0 : (ctas(2) >= 2 * 148 ? 1 : 2);
  p.max_chunks = 0;  // filled by caller
  return p;
}

template <int STAGE>
static int dispatch_stream(int fmt, const StreamArgs& a, const StreamPlan& plan, cudaStream_t s) {
  int rc;
  switch (fmt) {
    case SGLB_FMT_BF16: rc = launch_stream_gemm<SGLB_FMT_BF16, STAGE>(a, plan, s); break;
    case SGLB_FMT_INT8_W8A8: rc = launch_stream_gemm<SGLB_FMT_INT8_W8A8, STAGE>(a, plan, s); break;
    case SGLB_FMT_FP8_W8A16: rc = launch_stream_gemm<SGLB_FMT_FP8_W8A16, STAGE>(a, plan, s); break;
    case SGLB_FMT_MXFP4: rc = launch_stream_gemm<SGLB_FMT_MXFP4, STAGE>(a, plan, s); break;
    default: set_error("unsupported weight format %d", fmt); return SGLB_ERR_INVALID;
  }
  if (rc == SGLB_OK) count_launch();
  return rc;
}

static int check_format(const char* who, int fmt, int H, int I, const void* s13, const void* s2,
                        int bn, int bk) {
  switch (fmt) {
    case SGLB_FMT_BF16:
      SGLB_CHECK(H % 8 == 0 && I % 8 == 0, "%s: bf16 needs H,I multiples of 8 (H=%d I=%d)", who, H, I);
      break;
    case SGLB_FMT_INT8_W8A8:
      SGLB_CHECK(H % 16 == 0 && I % 16 == 0, "%s: int8 needs H,I multiples of 16 (H=%d I=%d)", who, H, I);
      SGLB_CHECK(s13 && s2, "%s: int8 w8a8 needs w1_scale and w2_scale", who);
      break;
    case SGLB_FMT_FP8_W8A16:
      SGLB_CHECK(H % 16 == 0 && I % 16 == 0, "%s: fp8 needs H,I multiples of 16 (H=%d I=%d)", who, H, I);
      SGLB_CHECK(s13 && s2, "%s: fp8 w8a16 needs w1_scale and w2_scale", who);
      SGLB_CHECK(bn >= 8 && bn % 8 == 0 && bk >= 64 && bk % 64 == 0,
                 "%s: fp8 block_size [%d,%d] unsupported (block_n %% 8 == 0, block_k %% 64 == 0)", who, bn, bk);
      break;
    case SGLB_FMT_MXFP4:
      SGLB_CHECK(H % 32 == 0 && I % 32 == 0, "%s: mxfp4 needs H,I multiples of 32 (H=%d I=%d)", who, H, I);
      SGLB_CHECK(s13 && s2, "%s: mxfp4 needs e8m0 scales", who);
      break;
    default:
      set_error("%s: unsupported weight format %d", who, fmt);
      return SGLB_ERR_INVALID;
  }
  return SGLB_OK;
}

// Optional pieces of the two epilogues (gpt-oss expert bias + clamped swiglu; slot-addressed stage-2 output).
struct MlpExtra {
  const float* w13_bias = nullptr;  // [E, 2I] f32
  const float* w2_bias = nullptr;   // [E, H] f32
  int act = SGLB_ACT_SILU;
  float alpha = 1.f, limit = 0.f;
  int out_by_slot = 0;              // stage 2 writes row sorted_slot[p] (through out_map) of ic2_out
  const int32_t* out_map = nullptr;
  // the shared expert riding along as expert id E: M more rows behind the routed ones (align.cu)
  const void *sh_w13 = nullptr, *sh_w2 = nullptr, *sh_s13 = nullptr, *sh_s2 = nullptr;
  // fused decode chain: the prologue prepared the stage-1 operand; the stage-2 kernel finishes tokens into final_out
  bool fused = false;
  bool fused_reduce = false;  // bf16 / int8 on umma_gemm.cu (reducer warp) and the streaming kernels
  double rpe_hint = 0.0;      // expected routed rows per expert when the ids are mostly masked (expert parallel)
  void* final_out = nullptr;
  float shared_rsf = 1.f;
  bool any() const { return w13_bias || w2_bias || act != SGLB_ACT_SILU || out_by_slot || sh_w13; }
};

// One GEMM stage on the tcgen05 path over the tile list of an existing alignment.
static int run_umma_stage(int fmt, int stage, const StreamArgs& sa, int E, int64_t S, const MoeWorkspace& ws,
                          cudaStream_t stream, bool f8_native = false, bool pair = false,
                          const void* mx4_tiles = nullptr, int mx4_cap = 0, double rpe_hint = 0.0) {
  UmmaArgs u;
  memset(&u, 0, sizeof(u));
  u.W = sa.W;
  u.Wscale = sa.Wscale;
  u.X = sa.X;
  u.Xscale = sa.Xscale;
  u.tile_list = ws.tile_list;
  u.counters = ws.counters;
  u.sorted_slot = ws.sorted_slot;
  u.row_scale = sa.row_scale;
  u.out = sa.out;
  u.K = sa.K;
  u.N = sa.N;
  u.rows_per_expert = sa.rows_per_expert;
  u.E = E;
  u.topk = sa.topk;
  u.gather = sa.gather;
  // mx4_cap: row capacity of the tiles when it is not the default (mxfp4 kernel: 16..112; fp8 deep-ring variant: 32)
  const int tile_cap = mx4_cap > 0 ? mx4_cap : (pair ? umma2_tile_rows() : kTileRows);
  u.max_tiles = (int)max_items(S, E, tile_cap);
  u.block_n = sa.block_n;
  u.block_k = sa.block_k;
  u.scale_cols = sa.scale_cols;
  u.X2 = sa.X2;
  u.Wsf = mx4_tiles;
  u.tile_cap = mx4_cap;
  u.bias = sa.bias;
  u.ebias = sa.ebias;
  u.act = sa.act;
  u.act_alpha = sa.act_alpha;
  u.act_limit = sa.act_limit;
  u.out_by_slot = sa.out_by_slot;
  u.out_map = sa.out_map;
  u.Wsh = sa.Wsh;
  u.Wsh_scale = sa.Wsh_scale;
  u.shared_row0 = sa.shared_row0;
  u.x_rows = (int)(sa.gather ? S / sa.topk : S);  // token rows when gathered, else the sorted rows themselves
  static const int l2_hints = [] {
    // experiment, off: L2 eviction hints on the TMA loads of the CTA-pair fp8 kernel (weights evict-first, routed rows
    // evict-last).  Measured at DSV3 M = 8192 (r3g): w13 4251 -> 4180 us, w2 2060 -> 2090 us, whole step 8.01 -> 8.70 ms
    // -- the pinned lines cost the kernels around the GEMMs more than the GEMMs gain.
    const char* e = getenv("SGLB_L2_HINTS");
    return e ? atoi(e) : 0;
  }();
  u.l2_hints = l2_hints;
  u.red = sa.red;
  u.row_amax = sa.row_amax;
  u.Xraw = sa.Xraw;
  const int sh_tiles = sa.Wsh ? (int)ceil_div64(S / sa.topk, tile_cap) : 0;
  u.max_tiles += sh_tiles;
  if (!pair) {
    // Expected number of tiles when the call is weight-streaming bound (up to 64 rows per active expert: every item
    // then costs its weight bytes); the launchers size the persistent grid to whole waves of it (persistent_grid).
    // rpe = expected routed rows per expert: the caller's hint (an expert-parallel rank, whose ids are mostly masked)
    // or S / E; experts hit ~ Poisson.
    const double rpe = rpe_hint > 0.0 ? rpe_hint : (double)S / E;
    double active = E == 1 ? 1.0 : E * (1.0 - exp(-rpe));
    if (active < 1) active = 1;
    const double rows_active = rpe * E / active;
    if (rows_active <= 64.0) u.est_tiles = (int)(active * ceil(rows_active / tile_cap) + 0.5) + sh_tiles;
  }
  if (mx4_tiles) return launch_umma_mx4(stage, u, stream);
  if (f8_native) return pair ? launch_umma2_f8(stage, u, stream) : launch_umma_f8(stage, u, stream);
  if (pair) return launch_umma2_gemm(fmt, stage, u, stream);
  return launch_umma_gemm(fmt, stage, u, stream);
}

// Counters per token of the fused top-k reduce: one per output-column block a stage-2 CTA (tcgen05: 256 columns) or
// warp (streaming kernel: 32 columns) completes.
static int stage2_red_blocks(int path, int64_t M, int topk, int E, int H) {
  if (path != SGLB_PATH_STREAM) return ceil_div(H, 2 * 128);
  const StreamPlan p2 = make_plan(M, topk, E, H, 2);
  const int wn = p2.cfg == 0 ? 8 : (p2.cfg == 1 ? 2 : 1);
  return ceil_div(H, 32 * wn) * wn;
}

// Two grouped GEMMs over an existing alignment (path = SGLB_PATH_STREAM or SGLB_PATH_UMMA).
// `row_scale` = topk weights or NULL.
static int run_mlp(int path, int fmt, int64_t M, int topk, int E, int H, int I, const void* hidden,
                          const void* w13, const void* w2, const void* s13, const void* s2, int bn,
                          int bk, const float* row_scale, const MoeWorkspace& ws, void* ic2_out,
                          cudaStream_t stream, const void* s13_tiles = nullptr, const void* s2_tiles = nullptr,
                          int mx4_cap = 0, const MlpExtra& ex = MlpExtra()) {
  const int64_t S = M * topk;
  StreamArgs a;
  memset(&a, 0, sizeof(a));
  a.chunk_list = ws.chunk_list;
  a.counters = ws.counters;
  a.sorted_slot = ws.sorted_slot;
  a.topk = topk;
  a.block_n = bn;
  a.block_k = bk;

  StreamPlan plan = make_plan(M, topk, E, I, 1);
  plan.max_chunks = (int)max_items(S, E, 8 * plan.ng);
  SGLB_CHECK(path != SGLB_PATH_STREAM || plan.max_chunks <= 65535,
             "fused_experts(stream path): too many row chunks (%d)", plan.max_chunks);

  // stage 1
  a.W = w13;
  a.Wscale = s13;
  a.K = H;
  a.N = I;
  a.rows_per_expert = 2 * I;
  a.gather = 1;
  a.out = ws.act;
  a.row_scale = nullptr;
  a.scale_cols = fmt == SGLB_FMT_FP8_W8A16 ? ceil_div(H, bk) : H / 32;
  a.ebias = ex.w13_bias;
  a.Wsh = ex.sh_w13;
  a.Wsh_scale = ex.sh_s13;
  a.shared_row0 = ex.sh_w13 ? (int)S : 0;
  const int64_t act_rows = S + (ex.sh_w13 ? M : 0);  // rows of act / ic2 in use
  a.act = ex.act;
  a.act_alpha = ex.alpha;
  a.act_limit = ex.limit;
  int rc;
  // fp8 w8a16 on the fp8 tensor cores: activations split into two e4m3 pieces (umma_f8.cu)
  const bool pair = path == SGLB_PATH_UMMA2;
  const bool umma = path == SGLB_PATH_UMMA || pair;
  const bool mx4 = umma && !pair && fmt == SGLB_FMT_MXFP4 && s13_tiles && s2_tiles && mx4_cap > 0;
  const bool f8 = mx4 || (umma && fmt == SGLB_FMT_FP8_W8A16 && g_f8_native && umma_f8_supported(H, I, bn, bk));
  // fused decode chain: the single-launch prologue already prepared the stage-1 operand (run_prologue)
  const bool fused_int8 = ex.fused_reduce && fmt == SGLB_FMT_INT8_W8A8 && path == SGLB_PATH_UMMA;  // level 2
  if (fmt == SGLB_FMT_INT8_W8A8) {
    if (!ex.fused) {
      StageTimer t(ST_QUANT_X, stream);
      rc = sglb_quant_rows_int8(hidden, M, H, ws.xq, ws.xs, stream);
      if (rc) return rc;
    }
    a.X = ws.xq;
    a.Xscale = ws.xs;
    if (fused_int8) a.row_amax = ws.row_amax;  // the stage-1 epilogue leaves max |act| per row for stage 2
  } else if (f8) {
    if (!ex.fused) {
      StageTimer t(ST_QUANT_X, stream);
      rc = launch_split_rows_e4m3(hidden, M, H, ws.xq, ws.xq2, ws.xs, stream);
      if (rc) return rc;
    }
    a.X = ws.xq;
    a.X2 = ws.xq2;
    a.Xscale = ws.xs;
  } else {
    a.X = hidden;
  }
  if (pair) {  // routed rows into expert-sorted order: the pair kernels read them as plain TMA boxes
    StageTimer t(ST_QUANT_X, stream);
    rc = launch_gather_rows(ws.sorted_slot, S, topk, H * (fmt == SGLB_FMT_BF16 ? 2 : 1), a.X, a.X2, ws.xsorted,
                            a.X2 ? ws.xsorted2 : nullptr, stream);
    if (rc) return rc;
    a.X = ws.xsorted;
    if (a.X2) a.X2 = ws.xsorted2;
    a.gather = 0;
  }
  {
    StageTimer t(ST_GEMM1, stream);
    rc = umma ? run_umma_stage(fmt, 1, a, E, S, ws, stream, f8, pair, mx4 ? s13_tiles : nullptr, mx4_cap, ex.rpe_hint)
                                : dispatch_stream<1>(fmt, a, plan, stream);
  }
  if (rc) return rc;

  // stage 2
  StreamPlan plan2 = make_plan(M, topk, E, H, 2);
  plan2.ng = plan.ng;  // same chunk list
  plan2.max_chunks = plan.max_chunks;
  a.W = w2;
  a.Wscale = s2;
  a.K = I;
  a.N = H;
  a.rows_per_expert = H;
  a.gather = 0;
  a.out = ic2_out;
  a.row_scale = row_scale;
  a.scale_cols = fmt == SGLB_FMT_FP8_W8A16 ? ceil_div(I, bk) : I / 32;
  a.ebias = ex.w2_bias;
  a.Wsh = ex.sh_w2;
  a.Wsh_scale = ex.sh_s2;
  a.out_by_slot = ex.out_by_slot;
  a.out_map = ex.out_map;
  if (ex.fused_reduce) {
    a.red.cnt = ws.red_cnt;
    a.red.slot_pos = ws.slot_pos;
    a.red.tok_valid = ws.tok_valid;
    a.red.out = ex.final_out;
    a.red.M = (int)M;
    a.red.blocks = stage2_red_blocks(path, M, topk, E, H);
    a.red.need = topk + (ex.sh_w13 ? 1 : 0);
    a.red.shared_row0 = ex.sh_w13 ? (int)S : -1;
    a.red.rsf = ex.shared_rsf;
  }
  if (fused_int8) {
    // no quantisation launch: the gather warps of the stage-2 kernel quantise the f32 act rows on the way in
    a.X = nullptr;
    a.Xscale = nullptr;
    a.Xraw = (const float*)ws.act;
  } else if (fmt == SGLB_FMT_INT8_W8A8) {
    StageTimer t(ST_QUANT_ACT, stream);
    rc = launch_quant_rows_f32((const float*)ws.act, act_rows, I, ws.actq, ws.acts, stream, ws.counters, (int)S);
    if (rc) return rc;
    a.X = ws.actq;
    a.Xscale = ws.acts;
    a.row_amax = nullptr;
  } else if (f8) {
    StageTimer t(ST_QUANT_ACT, stream);
    rc = launch_split_rows_e4m3(ws.act, act_rows, I, ws.actq, ws.actq2, ws.acts, stream, ws.counters, (int)S);
    if (rc) return rc;
    a.X = ws.actq;
    a.X2 = ws.actq2;
    a.Xscale = ws.acts;
  } else {
    a.X = ws.act;
  }
  StageTimer t(ST_GEMM2, stream);
  return umma ? run_umma_stage(fmt, 2, a, E, S, ws, stream, f8, pair, mx4 ? s2_tiles : nullptr, mx4_cap, ex.rpe_hint)
                                : dispatch_stream<2>(fmt, a, plan2, stream);
}

int moe_align_impl(const int32_t* topk_ids, int64_t M, int topk, int E, int chunk_rows, int tile_rows,
                   const sglb_align_out* out, void* scratch, int shared_rows, cudaStream_t stream);

static int run_align(const int32_t* ids, int64_t M, int topk, int E, int chunk_rows,
                     const MoeWorkspace& ws, cudaStream_t stream, int tile_rows = kTileRows, int shared_rows = 0) {
  sglb_align_out ao;
  ao.expert_count = ws.expert_count;
  ao.expert_offset = ws.expert_offset;
  ao.sorted_slot = ws.sorted_slot;
  ao.slot_pos = ws.slot_pos;
  ao.chunk_list = ws.chunk_list;
  ao.tile_list = ws.tile_list;
  ao.counters = ws.counters;
  StageTimer t(ST_ALIGN, stream);
  return moe_align_impl(ids, M, topk, E, chunk_rows, tile_rows, &ao, ws.align_scratch, shared_rows, stream);
}

int moe_prologue_impl(const int32_t* topk_ids, int64_t M, int topk, int E, int chunk_rows, int tile_rows,
                      const sglb_align_out* out, int shared_rows, const PrologueRows& pr, cudaStream_t stream);

// Fused decode chain: align + stage-1 operand preparation (mode 1 int8 rows, 2 e4m3 hi / lo split) + the state of the
// fused top-k reduce, in ONE launch.
static int run_prologue(const int32_t* ids, const void* hidden, int64_t M, int topk, int E, int H, int mode, int red_blocks,  // red_blocks 0: no fused reduce
                        int chunk_rows, int tile_rows, int shared_rows, const MoeWorkspace& ws, cudaStream_t stream) {
  sglb_align_out ao;
  ao.expert_count = ws.expert_count;
  ao.expert_offset = ws.expert_offset;
  ao.sorted_slot = ws.sorted_slot;
  ao.slot_pos = ws.slot_pos;
  ao.chunk_list = ws.chunk_list;
  ao.tile_list = ws.tile_list;
  ao.counters = ws.counters;
  PrologueRows pr;
  memset(&pr, 0, sizeof(pr));
  pr.mode = mode;
  pr.x = (const __nv_bfloat16*)hidden;
  pr.M = (int)M;
  pr.K = H;
  pr.q = ws.xq;
  pr.q2 = ws.xq2;
  pr.xs = ws.xs;
  pr.tok_valid = red_blocks > 0 ? ws.tok_valid : nullptr;
  pr.red_cnt = red_blocks > 0 ? ws.red_cnt : nullptr;
  pr.red_blocks = red_blocks;
  pr.row_amax = red_blocks > 0 ? ws.row_amax : nullptr;  // level 2 only: maxed by the stage-1 epilogue (int8)
  pr.amax_rows = (int)(M * topk + M);
  StageTimer t(ST_ALIGN, stream);
  return moe_prologue_impl(ids, M, topk, E, chunk_rows, tile_rows, &ao, shared_rows, pr, stream);
}

}  // namespace sglb

using namespace sglb;

extern "C" int sglb_version(void) { return 100; }
extern "C" const char* sglb_last_error(void) { return g_err; }
extern "C" uint64_t sglb_launch_count(void) { return g_launches.load(); }

// Page-locks / releases a caller-owned HOST range for the host layer's staging (ops.py::_maybe_pin).  A failed
// registration (read-only or file-backed mapping, RLIMIT_MEMLOCK, ...) is reported and its error is cleared here, in
// this library's runtime, so it cannot surface later from an unrelated launch check.
extern "C" int sglb_host_register(void* host_ptr, size_t bytes) {
  SGLB_CHECK(host_ptr != nullptr && bytes > 0, "host_register: empty range");
  cudaError_t e = cudaHostRegister(host_ptr, bytes, cudaHostRegisterDefault);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    set_error("cudaHostRegister(%p, %zu) failed: %s", host_ptr, bytes, cudaGetErrorString(e));
    return SGLB_ERR_CUDA;
  }
  return SGLB_OK;
}
extern "C" int sglb_host_unregister(void* host_ptr) {
  cudaError_t e = cudaHostUnregister(host_ptr);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    set_error("cudaHostUnregister(%p) failed: %s", host_ptr, cudaGetErrorString(e));
    return SGLB_ERR_CUDA;
  }
  return SGLB_OK;
}

extern "C" void sglb_profile_enable(int on) {
  g_profile = on != 0;
  g_events_used = 0;
}
// Sums the recorded stage durations (ms) into ms_out[0..n) and the number of samples into
// calls_out[0..n); the caller must have synchronised the stream(s).  Resets the recording.
extern "C" int sglb_profile_read(float* ms_out, int* calls_out, int n) {
  for (int i = 0; i < n; ++i) {
    ms_out[i] = 0.f;
    calls_out[i] = 0;
  }
  for (size_t i = 0; i < g_events_used; ++i) {
    float ms = 0.f;
    SGLB_CUDA(cudaEventElapsedTime(&ms, g_events[i].a, g_events[i].b));
    if (g_events[i].stage < n) {
      ms_out[g_events[i].stage] += ms;
      calls_out[g_events[i].stage] += 1;
    }
  }
  g_events_used = 0;
  return SGLB_OK;
}

extern "C" void sglb_set_fp8_native(int on) {
  g_f8_native = (on & 1) != 0;
  g_f8_small_tiles = (on & 2) == 0;  // bit 1 set: keep 128-row tiles at decode too (A/B experiments, parity tests)
  g_f8_mid_tiles = (on & 4) != 0;    // bit 2 set: also the 80-row configuration
}
extern "C" void sglb_set_chain_fusion(int level) { g_chain_fusion = level < 0 ? 0 : (level > 2 ? 2 : level); }
extern "C" void sglb_set_mx4_native(int on) { g_mx4_native = on != 0; }
extern "C" size_t sglb_mx4_scale_tiles_bytes(int64_t E, int rows, int split, int K) {
  if (E <= 0 || rows <= 0 || K <= 0) return 0;
  return mx4_scale_tiles_bytes(E, rows, split, K);
}
extern "C" int sglb_mx4_tile_scales(const void* scales, void* tiles, int64_t E, int rows, int split, int K,
                                    sglb_stream_t stream) {
  SGLB_CHECK(E >= 0 && (E == 0 || (scales && tiles)), "mx4_tile_scales: null pointer");
  return launch_mx4_tile_scales(scales, tiles, E, rows, split, K, (cudaStream_t)stream);
}
extern "C" int sglb_split_rows_e4m3(const void* x_bf16, int64_t rows, int K, void* hi, void* lo, float* rexp,
                                    sglb_stream_t stream) {
  SGLB_CHECK(rows >= 0 && K > 0, "split_rows_e4m3: bad shape");
  if (rows == 0) return SGLB_OK;
  SGLB_CHECK(x_bf16 && hi && lo && rexp, "split_rows_e4m3: null pointer");
  return launch_split_rows_e4m3(x_bf16, rows, K, hi, lo, rexp, (cudaStream_t)stream);
}

extern "C" int64_t sglb_moe_align_max_items(int64_t M, int topk, int E, int rows) {
  if (rows <= 0) return 0;
  return max_items(M * topk, E, rows);
}

extern "C" size_t sglb_fused_experts_workspace_bytes(int64_t M, int topk, int E, int H, int I,
                                                     int fmt) {
  return carve(nullptr, M, topk, E, H, I, fmt).bytes;
}

namespace sglb {
int launch_topk_reduce(const void* ic2, const int32_t* slot_pos, int64_t M, int topk, int H, const void* shared_y,
                       float rsf, void* out, cudaStream_t stream);

// AUTO path rule of fused_experts (also what sglb_fused_experts_shared_rides consults)
static int auto_path(const sglb_moe_problem* p, bool mx4) {
  const double rows_per_expert = p->rows_per_expert_hint > 0.f ? (double)p->rows_per_expert_hint : (double)p->M * p->topk / p->E;
  // fp8 on the fp8 tensor cores streams weights faster than the mma.sync kernel at every batch size
  // measured (DSV3 M = 16: 82 % vs 59 % of the HBM roofline): no ALU work per weight byte
  const bool f8 = p->fmt == SGLB_FMT_FP8_W8A16 && g_f8_native && umma_f8_supported(p->H, p->I, p->block_n, p->block_k);
  const int path = ((f8 || mx4 || rows_per_expert >= kUmmaMinRowsPerExpert || p->M * p->topk <= kUmmaTinySlots) && umma_supported(p->fmt, p->M, p->topk, p->E, p->H, p->I))
                       ? SGLB_PATH_UMMA
                       : SGLB_PATH_STREAM;
  // Prefill: the CTA-pair kernels (M = 256 MMAs, routed rows pre-gathered into sorted order and fed by TMA) once
  // pair tiles fill up.  Measured (tools/pair_check.py, r2i; bit-identical to the single-CTA kernels): bf16 Qwen3
  // M = 4096 / 8192: 785 / 928 vs 732 / 829 TFLOP/s; fp8 DSV3 M = 8192: 806 vs 698; slower for int8 (744 vs 858
  // TOP/s) and for fp8 layers with little work per gathered row (Qwen3 I = 768: 562 vs 588, I = 384: 415 vs 468).
  // Not with expert biases / clamped swiglu / slot-addressed output (single-CTA epilogues only).
  const bool plain = !p->w13_bias && !p->w2_bias && p->act == SGLB_ACT_SILU && !p->slot_out;
  if (path == SGLB_PATH_UMMA && !mx4 && plain && rows_per_expert >= kPairMinRowsPerExpert &&
      (p->fmt == SGLB_FMT_BF16 || (f8 && (double)p->H * p->I >= 8.0e6)))
    return SGLB_PATH_UMMA2;
  return path;
}
}  // namespace sglb

extern "C" int sglb_fused_experts_shared_rides(const sglb_moe_problem* p) {
  if (p == nullptr || p->M <= 0 || p->E < 1 || p->topk < 1) return 0;
  if (p->slot_out != nullptr || p->act != SGLB_ACT_SILU) return 0;
  const bool f8 = p->fmt == SGLB_FMT_FP8_W8A16 && g_f8_native && umma_f8_supported(p->H, p->I, p->block_n, p->block_k);
  if (!(p->fmt == SGLB_FMT_BF16 || p->fmt == SGLB_FMT_INT8_W8A8 || f8)) return 0;
  if (!umma_supported(p->fmt, p->M, p->topk, p->E, p->H, p->I)) return 0;
  // where AUTO takes the CTA-pair kernels for the routed experts (prefill) the shared expert stays a call of its own
  const int path = p->path == SGLB_PATH_AUTO ? auto_path(p, false) : p->path;
  return path == SGLB_PATH_UMMA ? 1 : 0;
}

extern "C" int sglb_fused_experts(const sglb_moe_problem* p, sglb_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  SGLB_CHECK(p != nullptr, "fused_experts: null problem");
  SGLB_CHECK(p->M >= 0 && p->topk >= 1 && p->E >= 1 && p->H > 0 && p->I > 0,
             "fused_experts: bad shape M=%lld topk=%d E=%d H=%d I=%d", (long long)p->M, p->topk, p->E, p->H, p->I);
  int rc = check_format("fused_experts", p->fmt, p->H, p->I, p->w13_scale, p->w2_scale, p->block_n, p->block_k);
  if (rc) return rc;
  SGLB_CHECK(p->w13 && p->w2, "fused_experts: null weight pointer");
  if (p->M == 0) return SGLB_OK;  // empty batch: activation pointers may legitimately be null
  SGLB_CHECK(p->hidden && p->topk_weights && p->topk_ids && (p->out || p->slot_out), "fused_experts: null tensor pointer");
  SGLB_CHECK(p->act == SGLB_ACT_SILU || p->act == SGLB_ACT_SWIGLU_OAI, "fused_experts: unknown activation %d", p->act);
  SGLB_CHECK(p->slot_map == nullptr || p->slot_out != nullptr, "fused_experts: slot_map needs slot_out");
  MlpExtra ex;
  ex.w13_bias = p->w13_bias;
  ex.w2_bias = p->w2_bias;
  ex.act = p->act;
  ex.alpha = p->alpha;
  ex.limit = p->limit;
  ex.out_by_slot = p->slot_out != nullptr;
  ex.out_map = p->slot_map;
  ex.rpe_hint = p->rows_per_expert_hint > 0.f ? (double)p->rows_per_expert_hint : 0.0;
  const bool ride = p->shared_w13 != nullptr;
  if (ride) {
    SGLB_CHECK(p->shared_w2 != nullptr, "fused_experts: shared_w13 without shared_w2");
    SGLB_CHECK(p->fmt == SGLB_FMT_BF16 || (p->shared_w13_scale && p->shared_w2_scale), "fused_experts: the shared expert needs its scales");
    SGLB_CHECK(sglb_fused_experts_shared_rides(p),
               "fused_experts: the shared expert cannot ride along here (tcgen05 path with bf16 / int8 / tensor-native fp8 "
               "weights, plain silu, no slot_out); call sglb_shared_expert");
    ex.sh_w13 = p->shared_w13;
    ex.sh_w2 = p->shared_w2;
    ex.sh_s13 = p->shared_w13_scale;
    ex.sh_s2 = p->shared_w2_scale;
  }
  void* stage2_out = nullptr;  // set below: ws.ic2 (expert-sorted rows, reduced by K5) or the caller's slot rows
  const size_t need = sglb_fused_experts_workspace_bytes(p->M, p->topk, p->E, p->H, p->I, p->fmt);
  SGLB_CHECK(p->workspace && p->workspace_bytes >= need, "fused_experts: workspace too small (%zu < %zu)",
             p->workspace_bytes, need);
  SGLB_CHECK(((uintptr_t)p->workspace & 255) == 0, "fused_experts: workspace must be 256-byte aligned");
  MoeWorkspace ws = carve(p->workspace, p->M, p->topk, p->E, p->H, p->I, p->fmt);
  stage2_out = p->slot_out ? p->slot_out : (void*)ws.ic2;

  int path = p->path;
  // expected routed rows per expert decide the path; a caller whose ids are mostly masked (an expert-parallel
  // rank sees world x cap slots of which ~1/world are its own) passes its own estimate
  const double rows_per_expert = p->rows_per_expert_hint > 0.f ? (double)p->rows_per_expert_hint : (double)p->M * p->topk / p->E;
  // mxfp4 on the block-scaled tensor cores needs the re-tiled scales (sglb_mx4_tile_scales)
  const bool mx4 = p->fmt == SGLB_FMT_MXFP4 && g_mx4_native && p->w13_scale_tiles && p->w2_scale_tiles &&
                   umma_mx4_supported(p->H, p->I);
  const int mx4_cap = mx4 ? umma_mx4_tile_rows(rows_per_expert) : 0;
  if (path == SGLB_PATH_AUTO) {
    path = auto_path(p, mx4);
  }
  // Decode chain fusion (bit-identical to the separate launches): one prologue launch; at level 2 also the int8 act
  // quantisation inside the stage-2 producer and the top-k reduce inside the stage-2 kernel (not when the caller
  // reduces the slot rows itself: expert parallel).  Not for the CTA-pair kernels (prefill).
  const bool fuse = g_chain_fusion >= 1 && p->M * p->topk <= kFuseMaxSlots && p->H % 8 == 0 &&
                    (path == SGLB_PATH_UMMA || path == SGLB_PATH_STREAM);
  // The reduce moves into the stage-2 kernel where a warp off the critical path can take it: the reducer warp of
  // umma_gemm.cu (bf16 / int8) and the finishing warps of the streaming kernel.  The fp8 / mxfp4 tensor-native kernels
  // keep the separate launch (their epilogue warps promote partial sums all through an item).
  const bool fuse_reduce = fuse && g_chain_fusion >= 2 && !p->slot_out && (path == SGLB_PATH_STREAM || p->fmt == SGLB_FMT_BF16 || p->fmt == SGLB_FMT_INT8_W8A8);
  ex.fused = fuse;
  ex.fused_reduce = fuse_reduce;
  ex.final_out = p->out;
  ex.shared_rsf = p->shared_rsf;
  if (path == SGLB_PATH_UMMA || path == SGLB_PATH_UMMA2) {
    SGLB_CHECK(umma_supported(p->fmt, p->M, p->topk, p->E, p->H, p->I),
               "fused_experts: tcgen05 path does not support fmt=%d H=%d I=%d", p->fmt, p->H, p->I);
    SGLB_CHECK(path == SGLB_PATH_UMMA || p->fmt == SGLB_FMT_BF16 || p->fmt == SGLB_FMT_INT8_W8A8 ||
                   (p->fmt == SGLB_FMT_FP8_W8A16 && g_f8_native && umma_f8_supported(p->H, p->I, p->block_n, p->block_k)),
               "fused_experts: the CTA-pair tcgen05 path takes bf16, int8 and tensor-native fp8 weights only (fmt=%d)", p->fmt);
    SGLB_CHECK(path == SGLB_PATH_UMMA || !ex.any(),
               "fused_experts: the CTA-pair tcgen05 path has no expert bias / clamped swiglu / slot-addressed output");
    const bool use_mx4 = mx4 && path == SGLB_PATH_UMMA;
    // fp8 tensor-native kernel, decode: 32- or 80-row tiles run the deep-ring configurations (umma_f8.cu F8Cfg) -- when
    // an expert rarely gets more rows (Poisson(12) > 32: 1e-6; Poisson(64) > 80: 2 %; a longer expert is split and
    // streams twice) and a riding shared expert fits one tile
    const bool f8n = p->fmt == SGLB_FMT_FP8_W8A16 && g_f8_native && umma_f8_supported(p->H, p->I, p->block_n, p->block_k);
    int f8_cap = 0;
    static const int small_env = [] {
      const char* e = getenv("SGLB_F8_SMALL_TILES");  // 0 = 128-row tiles only, 1 (default) = 32-row at decode, 2 = also 80-row
      return e ? atoi(e) : 1;
    }();
    // Smaller tiles are a bet on balanced routing: an expert with more rows than the capacity is split and streams its
    // weights once per tile.  32-row tiles are therefore taken only up to 512 routed slots, where even the worst case
    // (every slot on one expert: 16 tiles of it) moves a fraction of what a balanced call streams; the 80-row
    // configuration gains 2 % on balanced routing (DSV3 M = 512: 0.931 -> 0.950) and can lose 60 % on a collapsed one,
    // so it stays an experiment switch.
    if (path == SGLB_PATH_UMMA && f8n && g_f8_small_tiles && small_env > 0) {
      if (rows_per_expert <= 12.0 && p->M * p->topk <= 512 && (!ride || p->M <= umma_f8_small_tile_rows()))
        f8_cap = umma_f8_small_tile_rows();
      else if ((small_env > 1 || g_f8_mid_tiles) && rows_per_expert <= 40.0 && (!ride || p->M <= umma_f8_mid_tile_rows()))
        f8_cap = umma_f8_mid_tile_rows();
    }
    const int cap_arg = use_mx4 ? mx4_cap : f8_cap;
    const int tile_rows = cap_arg > 0 ? cap_arg : (path == SGLB_PATH_UMMA2 ? umma2_tile_rows() : kTileRows);
    if (fuse) {
      const bool f8 = use_mx4 || (p->fmt == SGLB_FMT_FP8_W8A16 && g_f8_native && umma_f8_supported(p->H, p->I, p->block_n, p->block_k));
      const int mode = p->fmt == SGLB_FMT_INT8_W8A8 ? 1 : (f8 ? 2 : 0);
      rc = run_prologue(p->topk_ids, p->hidden, p->M, p->topk, p->E, p->H, mode,
                        fuse_reduce ? stage2_red_blocks(path, p->M, p->topk, p->E, p->H) : 0, 0, tile_rows,
                        ride ? (int)p->M : 0, ws, stream);
    } else {
      rc = run_align(p->topk_ids, p->M, p->topk, p->E, 0, ws, stream, tile_rows, ride ? (int)p->M : 0);
    }
    if (rc) return rc;
    rc = run_mlp(path, p->fmt, p->M, p->topk, p->E, p->H, p->I, p->hidden, p->w13, p->w2,
                 p->w13_scale, p->w2_scale, p->block_n, p->block_k, p->topk_weights, ws, stage2_out, stream,
                 use_mx4 ? p->w13_scale_tiles : nullptr, use_mx4 ? p->w2_scale_tiles : nullptr, cap_arg, ex);
    if (rc) return rc;
  } else {
    StreamPlan plan = make_plan(p->M, p->topk, p->E, p->I, 1);
    if (fuse)
      rc = run_prologue(p->topk_ids, p->hidden, p->M, p->topk, p->E, p->H, p->fmt == SGLB_FMT_INT8_W8A8 ? 1 : 0,
                        fuse_reduce ? stage2_red_blocks(path, p->M, p->topk, p->E, p->H) : 0, 8 * plan.ng, kTileRows, 0, ws,
                        stream);
    else
      rc = run_align(p->topk_ids, p->M, p->topk, p->E, 8 * plan.ng, ws, stream);
    if (rc) return rc;
    rc = run_mlp(SGLB_PATH_STREAM, p->fmt, p->M, p->topk, p->E, p->H, p->I, p->hidden, p->w13, p->w2,
                 p->w13_scale, p->w2_scale, p->block_n, p->block_k, p->topk_weights, ws, stage2_out, stream, nullptr,
                 nullptr, 0, ex);
    if (rc) return rc;
  }
  if (p->slot_out) return SGLB_OK;  // the caller reduces the slot rows (expert-parallel combine)
  if (fuse_reduce) return SGLB_OK;  // the stage-2 kernel finished the tokens itself
  StageTimer t(ST_REDUCE, stream);
  // riding shared expert: its rows sit behind the routed ones; the reduce applies shared_expert_cpu's epilogue
  const void* shared_y = ride ? (const void*)(ws.ic2 + (size_t)p->M * p->topk * p->H) : nullptr;
  return launch_topk_reduce(ws.ic2, ws.slot_pos, p->M, p->topk, p->H, shared_y, p->shared_rsf, p->out, stream);
}

extern "C" int sglb_shared_expert(const sglb_shared_problem* p, sglb_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  SGLB_CHECK(p != nullptr, "shared_expert: null problem");
  SGLB_CHECK(p->M >= 0 && p->H > 0 && p->I > 0, "shared_expert: bad shape");
  int rc = check_format("shared_expert", p->fmt, p->H, p->I, p->w1_scale, p->w2_scale, p->block_n, p->block_k);
  if (rc) return rc;
  SGLB_CHECK(p->w1 && p->w2, "shared_expert: null weight pointer");
  if (p->M == 0) return SGLB_OK;
  SGLB_CHECK(p->hidden && p->out, "shared_expert: null tensor pointer");
  const size_t need = sglb_fused_experts_workspace_bytes(p->M, 1, 1, p->H, p->I, p->fmt);
  SGLB_CHECK(p->workspace && p->workspace_bytes >= need, "shared_expert: workspace too small (%zu < %zu)",
             p->workspace_bytes, need);
  MoeWorkspace ws = carve(p->workspace, p->M, 1, 1, p->H, p->I, p->fmt);
  SGLB_CUDA(cudaMemsetAsync(ws.zero_ids, 0, sizeof(int32_t) * p->M, stream));
  // same path rule as fused_experts with E = 1 (every token is a routed row of the one expert)
  int path = p->path;
  if (path == SGLB_PATH_AUTO) {
    const bool f8 = p->fmt == SGLB_FMT_FP8_W8A16 && g_f8_native && umma_f8_supported(p->H, p->I, p->block_n, p->block_k);
    path = ((f8 || (double)p->M >= kUmmaMinRowsPerExpert) && umma_supported(p->fmt, p->M, 1, 1, p->H, p->I))
               ? SGLB_PATH_UMMA
               : SGLB_PATH_STREAM;
  }
  if (path == SGLB_PATH_UMMA) {
    SGLB_CHECK(umma_supported(p->fmt, p->M, 1, 1, p->H, p->I),
               "shared_expert: tcgen05 path does not support fmt=%d H=%d I=%d", p->fmt, p->H, p->I);
    rc = run_align(ws.zero_ids, p->M, 1, 1, 0, ws, stream);
  } else {
    StreamPlan plan = make_plan(p->M, 1, 1, p->I, 1);
    rc = run_align(ws.zero_ids, p->M, 1, 1, 8 * plan.ng, ws, stream);
  }
  if (rc) return rc;
  // without fused_out the MLP result goes straight to `out` (E = 1 keeps token order: sorted position = token)
  rc = run_mlp(path, p->fmt, p->M, 1, 1, p->H, p->I, p->hidden, p->w1, p->w2, p->w1_scale,
               p->w2_scale, p->block_n, p->block_k, nullptr, ws, p->fused_out ? (void*)ws.ic2 : p->out, stream);
  if (rc || !p->fused_out) return rc;
  return launch_shared_axpy(ws.ic2, p->fused_out, p->routed_scaling_factor, p->M * (int64_t)p->H, p->out, stream);
}

extern "C" int sglb_shared_axpy(const void* y, const void* fused, float routed_scaling_factor, int64_t n, void* out,
                                sglb_stream_t stream) {
  SGLB_CHECK(n >= 0 && (n == 0 || (y && fused && out)), "shared_axpy: bad arguments");
  return launch_shared_axpy(y, fused, routed_scaling_factor, n, out, (cudaStream_t)stream);
}

extern "C" int sglb_mxfp4_mm(const void* A, const void* Bq, const void* Bs, const float* bias, int64_t M,
                             int N, int K, void* out, void* workspace, size_t workspace_bytes,
                             sglb_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  SGLB_CHECK(Bq && Bs, "mxfp4_mm: null weight pointer");
  SGLB_CHECK(M >= 0 && N > 0 && K > 0 && K % 32 == 0, "mxfp4_mm: bad shape M=%lld N=%d K=%d (K %% 32 == 0)",
             (long long)M, N, K);
  if (M == 0) return SGLB_OK;
  SGLB_CHECK(A && out, "mxfp4_mm: null tensor pointer");
  const size_t need = sglb_fused_experts_workspace_bytes(M, 1, 1, K, N, SGLB_FMT_MXFP4);
  SGLB_CHECK(workspace && workspace_bytes >= need, "mxfp4_mm: workspace too small (%zu < %zu)", workspace_bytes, need);
  MoeWorkspace ws = carve(workspace, M, 1, 1, K, N, SGLB_FMT_MXFP4);
  SGLB_CUDA(cudaMemsetAsync(ws.zero_ids, 0, sizeof(int32_t) * M, stream));
  StreamPlan plan = make_plan(M, 1, 1, N, 2);
  plan.max_chunks = (int)max_items(M, 1, 8 * plan.ng);
  SGLB_CHECK(plan.max_chunks <= 65535, "mxfp4_mm: M too large for the streaming path");
  int rc = run_align(ws.zero_ids, M, 1, 1, 8 * plan.ng, ws, stream);
  if (rc) return rc;
  StreamArgs a;
  memset(&a, 0, sizeof(a));
  a.chunk_list = ws.chunk_list;
  a.counters = ws.counters;
  a.sorted_slot = ws.sorted_slot;
  a.topk = 1;
  a.W = Bq;
  a.Wscale = Bs;
  a.X = A;
  a.K = K;
  a.N = N;
  a.rows_per_expert = N;
  a.gather = 0;
  a.out = out;
  a.bias = bias;
  a.scale_cols = K / 32;
  return dispatch_stream<2>(SGLB_FMT_MXFP4, a, plan, stream);
}

// ---- dense quantised linears: stage 2 of the expert kernels with E = 1 ----------
// Workspace: the MoE carve with H = K and a token intermediate of 16 columns (act / ic2 are unused here).
extern "C" size_t sglb_linear_workspace_bytes(int64_t M, int N, int K, int fmt) {
  (void)N;
  if (M < 0 || K <= 0) return 0;
  return carve(nullptr, M, 1, 1, K, 16, fmt).bytes;
}

extern "C" int sglb_linear(const sglb_linear_problem* p, sglb_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  SGLB_CHECK(p != nullptr, "linear: null problem");
  const int fmt = p->fmt, N = p->N, K = p->K;
  const int64_t M = p->M;
  SGLB_CHECK(M >= 0 && N > 0 && K > 0, "linear: bad shape M=%lld N=%d K=%d", (long long)M, N, K);
  SGLB_CHECK(N % 8 == 0, "linear: N=%d must be a multiple of 8", N);
  switch (fmt) {
    case SGLB_FMT_BF16:
      SGLB_CHECK(K % 8 == 0, "linear: bf16 needs K %% 8 == 0 (K=%d)", K);
      break;
    case SGLB_FMT_INT8_W8A8:
      SGLB_CHECK(K % 16 == 0, "linear: int8 needs K %% 16 == 0 (K=%d)", K);
      SGLB_CHECK(p->w_scale, "linear: int8 w8a8 needs the per-channel weight scale");
      break;
    case SGLB_FMT_FP8_W8A16:
      SGLB_CHECK(K % 16 == 0, "linear: fp8 needs K %% 16 == 0 (K=%d)", K);
      SGLB_CHECK(p->w_scale && p->block_n > 0 && p->block_k > 0, "linear: fp8 w8a16 needs block scales and block_size");
      SGLB_CHECK(p->block_k % 64 == 0, "linear: fp8 block_k=%d must be a multiple of 64", p->block_k);
      break;
    case SGLB_FMT_MXFP4:
      SGLB_CHECK(K % 32 == 0, "linear: mxfp4 needs K %% 32 == 0 (K=%d)", K);
      SGLB_CHECK(p->w_scale, "linear: mxfp4 needs e8m0 scales");
      break;
    default:
      set_error("linear: unsupported weight format %d", fmt);
      return SGLB_ERR_INVALID;
  }
  SGLB_CHECK(p->a_scale == nullptr || fmt == SGLB_FMT_INT8_W8A8, "linear: a_scale is int8 w8a8 only");
  SGLB_CHECK(p->W, "linear: null weight pointer");
  if (M == 0) return SGLB_OK;
  SGLB_CHECK(p->A && p->out, "linear: null tensor pointer");
  const size_t need = sglb_linear_workspace_bytes(M, N, K, fmt);
  SGLB_CHECK(p->workspace && p->workspace_bytes >= need, "linear: workspace too small (%zu < %zu)",
             p->workspace_bytes, need);
  MoeWorkspace ws = carve(p->workspace, M, 1, 1, K, 16, fmt);
  SGLB_CUDA(cudaMemsetAsync(ws.zero_ids, 0, sizeof(int32_t) * M, stream));

  const bool f8 = fmt == SGLB_FMT_FP8_W8A16 && g_f8_native && umma_f8_supported(K, K, p->block_n, p->block_k);
  const bool umma_ok = umma_supported(fmt, M, 1, 1, K, K) != 0;
  int path = p->path;
  if (path == SGLB_PATH_AUTO)
    path = (umma_ok && (f8 || (double)M >= kUmmaMinRowsPerExpert)) ? SGLB_PATH_UMMA : SGLB_PATH_STREAM;
  SGLB_CHECK(path == SGLB_PATH_STREAM || path == SGLB_PATH_UMMA, "linear: path %d not available", path);
  SGLB_CHECK(path != SGLB_PATH_UMMA || umma_ok, "linear: tcgen05 path does not support fmt=%d K=%d", fmt, K);

  StreamArgs a;
  memset(&a, 0, sizeof(a));
  a.chunk_list = ws.chunk_list;
  a.counters = ws.counters;
  a.sorted_slot = ws.sorted_slot;
  a.topk = 1;
  a.block_n = p->block_n;
  a.block_k = p->block_k;
  a.W = p->W;
  a.Wscale = p->w_scale;
  a.K = K;
  a.N = N;
  a.rows_per_expert = N;
  a.out = p->out;
  a.bias = p->bias;
  a.scale_cols = fmt == SGLB_FMT_FP8_W8A16 ? ceil_div(K, p->block_k) : K / 32;
  a.X = p->A;
  int rc;
  if (fmt == SGLB_FMT_INT8_W8A8) {
    if (p->a_scale) {
      a.Xscale = p->a_scale;
    } else {
      StageTimer t(ST_QUANT_X, stream);
      rc = sglb_quant_rows_int8(p->A, M, K, ws.xq, ws.xs, stream);
      if (rc) return rc;
      a.X = ws.xq;
      a.Xscale = ws.xs;
    }
  } else if (path == SGLB_PATH_UMMA && f8) {
    StageTimer t(ST_QUANT_X, stream);
    rc = launch_split_rows_e4m3(p->A, M, K, ws.xq, ws.xq2, ws.xs, stream);
    if (rc) return rc;
    a.X = ws.xq;
    a.X2 = ws.xq2;
    a.Xscale = ws.xs;
  }
  if (path == SGLB_PATH_UMMA) {
    rc = run_align(ws.zero_ids, M, 1, 1, 0, ws, stream);
    if (rc) return rc;
    a.gather = 1;  // identity through sorted_slot (E = 1 keeps token order); rows are clamped to the tile
    StageTimer t(ST_GEMM2, stream);
    return run_umma_stage(fmt, 2, a, 1, M, ws, stream, f8);
  }
  StreamPlan plan = make_plan(M, 1, 1, N, 2);
  plan.max_chunks = (int)max_items(M, 1, 8 * plan.ng);
  SGLB_CHECK(plan.max_chunks <= 65535, "linear: M too large for the streaming path");
  rc = run_align(ws.zero_ids, M, 1, 1, 8 * plan.ng, ws, stream);
  if (rc) return rc;
  a.gather = 0;
  StageTimer t(ST_GEMM2, stream);
  return dispatch_stream<2>(fmt, a, plan, stream);
}
