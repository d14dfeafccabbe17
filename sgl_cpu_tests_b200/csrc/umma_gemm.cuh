// K3/K4 (prefill regime): interface of the tcgen05 / TMEM / TMA grouped GEMM (umma_gemm.cu).
#pragma once
#include "common.cuh"

namespace sglb {

struct UmmaArgs {
  const void* W;           // [E, rows_per_expert, K] row-major, K contiguous (bf16 or s8)
  const void* Wscale;      // int8: f32 [E, rows_per_expert]; fp8: f32 [E, rows/bn, K/bk]; mxfp4: u8 [E, rows, K/32]
  const void* X;           // routed rows, row stride K elements (bf16 or s8)
  const float* Xscale;     // int8: per X row
  const int32_t* tile_list;   // (expert, row_start, rows <= 128) from moe_align
  const int32_t* counters;    // counters[2] = number of tiles
  const int32_t* sorted_slot;
  const float* row_scale;  // stage 2: topk_weights flat [S] (indexed by slot) or nullptr
  void* out;               // stage 1: act [*, I] (bf16; f32 for int8); stage 2: ic2 [*, H] bf16
  int K;                   // reduction length in elements
  int N;                   // outputs per routed row: I (stage 1), H (stage 2)
  int rows_per_expert;     // 2I (stage 1), H (stage 2)
  int E;
  int topk;
  int gather;              // 1: X row = sorted_slot[p] / topk; 0: X row = p
  int max_tiles;           // host upper bound on counters[2]
  int block_n, block_k;    // fp8 scale block
  int scale_cols;          // fp8: ceil(K/bk); mxfp4: K/32
  const void* X2;          // fp8 / mxfp4 tensor-native paths: lo pieces of the routed rows (X = hi pieces, e4m3)
  const void* Wsf;         // mxfp4 tensor-native path: e8m0 scales re-tiled for tcgen05.cp (umma_mx4.cu)
  int tile_cap;            // row capacity of the tiles in tile_list when not 128: mxfp4 kernel 16..112, fp8 deep rings 32 / 80
  const float* bias;       // stage 2 optional f32 [N] added after row_scale (dense linears, E = 1)
  int est_tiles;           // umma_gemm.cu, decode regime: expected tile count (0 = unknown), sizes the persistent grid
  const float* ebias;      // per-expert f32 bias [E, rows_per_expert] or nullptr (gpt-oss w1_bias / w2_bias): added to
                           // the dequantised accumulator before the activation (stage 1) / before row_scale (stage 2)
  int act;                 // stage 1: SGLB_ACT_SILU or SGLB_ACT_SWIGLU_OAI (common.cuh glu_f)
  float act_alpha, act_limit;
  int out_by_slot;         // stage 2: write routed row p to row sorted_slot[p] (through out_map if set) instead of p
  const int32_t* out_map;
  // Shared expert riding along (umma_gemm.cu and umma_f8.cu only): tiles whose expert id is E (appended by moe_align,
  // rows [shared_row0, shared_row0 + M) of X / out) take their weights from Wsh (one expert of the same shape, through
  // a second tensor map) and their scales from Wsh_scale; they carry no per-expert bias and no topk weight.
  const void* Wsh;         // nullptr = none
  const void* Wsh_scale;
  int shared_row0;         // = M * topk when Wsh is set, else 0
  int x_rows;              // rows of X (CTA-pair kernels: bound of the TMA gather map)
  // ---- decode chain fusion (capi.cu: fewer launches per call, bit-identical results) ----
  FusedReduce red;         // stage 2: the epilogue finishes tokens itself (common.cuh); red.cnt == nullptr = off
  uint32_t* row_amax;      // int8: f32 bits of max |act| per act row.  Stage 1: the epilogue atomically maxes into it
                           // (zeroed by the prologue); stage 2 with Xraw set: the quantisation scale of each row
  int l2_hints;            // CTA-pair fp8 kernel: TMA loads carry L2 eviction hints (weights first, routed rows last)
  const float* Xraw;       // stage 2, int8 (umma_gemm.cu only): f32 act rows [*, K]; the gather warps quantise them into
                           // the operand tile with quant_rows_kernel's arithmetic (X / Xscale are unused)
};

constexpr int kUmmaTileRows = 128;  // routed rows per tile (tile_rows handed to moe_align)

// 1 if the tcgen05 path handles this problem (format and divisibility), else 0.
int umma_supported(int fmt, int64_t M, int topk, int E, int H, int I);
// stage 1: out = silu(X W13_gate^T) * (X W13_up^T);  stage 2: out = (X W2^T) * row_scale
int launch_umma_gemm(int fmt, int stage, const UmmaArgs& a, cudaStream_t stream);
// CTA-pair kernels (umma2_gemm.cu): tile_list must hold umma2_tile_rows()-row tiles; bf16 / int8
int umma2_tile_rows();
int launch_umma2_gemm(int fmt, int stage, const UmmaArgs& a, cudaStream_t stream);
// fp8 tensor-native kernel on a CTA pair (umma2_f8.cu): same operands as launch_umma_f8, 256-row tiles
int launch_umma2_f8(int stage, const UmmaArgs& a, cudaStream_t stream);
// fp8 w8a16 / mxfp4: weights dequantised to bf16 inside the kernel (umma_dq.cu)
int launch_umma_dq(int fmt, int stage, const UmmaArgs& a, cudaStream_t stream);
// fp8 w8a16 on the fp8 tensor cores (umma_f8.cu): X/X2 = hi/lo e4m3 pieces, Xscale = per-row 2^e factor
int umma_f8_supported(int H, int I, int block_n, int block_k);
int launch_umma_f8(int stage, const UmmaArgs& a, cudaStream_t stream);
int umma_f8_small_tile_rows();  // tile capacity of the deep-ring decode configuration (UmmaArgs.tile_cap)
int umma_f8_mid_tile_rows();    // 80-row tiles, 4-stage ring
// mxfp4 on the block-scaled fp8/fp4 tensor cores (umma_mx4.cu): weights e2m1 via TMA, e8m0 scales in TMEM
int umma_mx4_supported(int H, int I);
int umma_mx4_tile_rows(double rows_per_expert);
int launch_umma_mx4(int stage, const UmmaArgs& a, cudaStream_t stream);
// e8m0 scales [E, rows, K/32] -> 512-byte tcgen05.cp tiles [E, blocks(rows, split), ceil(K/128)]
size_t mx4_scale_tiles_bytes(int64_t E, int rows, int split, int K);
int launch_mx4_tile_scales(const void* src, void* dst, int64_t E, int rows, int split, int K, cudaStream_t stream);
int launch_split_rows_e4m3(const void* x_bf16, int64_t rows, int K, void* hi, void* lo, float* rexp,
                           cudaStream_t stream, const int32_t* n_valid = nullptr, int always_from = 0);

}  // namespace sglb
