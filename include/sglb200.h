/* sglb200.h -- C ABI of libsglb200.so, the sm_100a implementation of the quantized fused-MoE expert
 * path that mingfeima/sgl-cpu-tests drives through torch.ops.sgl_kernel.
 *
 * Plain pointers and sizes only: every pointer is a DEVICE pointer unless stated, every call is
 * asynchronous on the given CUDA stream (a cudaStream_t passed as void*), nothing here touches
 * torch.  A call returns SGLB_OK or a negative status; sglb_last_error() then holds a
 * thread-local message (the Python layer turns it into RuntimeError, the way the reference's
 * TORCH_CHECK surfaces).  The library keeps no global state apart from that message: the caller
 * owns all buffers including the workspace (one exception, stated there: sglb_moe_align_static).
 *
 * Each entry point cites the reference interface it stands behind (paths are relative to the
 * reference checkout, /root/reference in the build container).
 */
#ifndef SGLB200_H_
#define SGLB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* sglb_stream_t; /* cudaStream_t */

enum {
  SGLB_OK = 0,
  SGLB_ERR_INVALID = -1, /* bad argument / unsupported shape */
  SGLB_ERR_CUDA = -2     /* CUDA runtime error */
};

/* element types of caller tensors */
enum { SGLB_DT_BF16 = 0, SGLB_DT_F32 = 1, SGLB_DT_F16 = 2 };

/* weight formats; numbering follows CPUQuantMethod (test_moe.py:4,86) */
enum {
  SGLB_FMT_BF16 = 0,      /* UNQUANT   : w bf16                                         */
  SGLB_FMT_INT8_W8A8 = 1, /* INT8_W8A8 : w s8, f32 per-output-channel scale, dynamic    */
                          /*             per-token s8 activations (test_moe_int8.py)    */
  SGLB_FMT_FP8_W8A16 = 2, /* FP8_W8A16 : w e4m3, f32 scale per [block_n, block_k] block */
  SGLB_FMT_MXFP4 = 4      /* MXFP4     : w e2m1 nibbles (low nibble = even k), e8m0/32  */
};

/* which GEMM implementation sglb_fused_experts uses */
enum {
  SGLB_PATH_AUTO = 0,
  SGLB_PATH_STREAM = 1, /* mma.sync weight streaming (few rows per expert) */
  SGLB_PATH_UMMA = 2,   /* tcgen05, one CTA per tile */
  SGLB_PATH_UMMA2 = 3   /* tcgen05 cta_group::2, a CTA pair per 256 x 256 tile (bf16 / int8, prefill) */
};

int sglb_version(void);
const char* sglb_last_error(void);
/* number of kernels this library has launched in the calling process (bench.py reports it) */
uint64_t sglb_launch_count(void);

/* Page-locking of caller-owned HOST memory, for host layers that stage CPU tensors (the reference's scripts pass CPU
 * tensors: test_moe.py:96-99, bench_moe.py:52).  No counterpart in the reference.  host_ptr is a HOST pointer.  A
 * failure (read-only or file-backed mapping, locked-memory limit) returns SGLB_ERR_CUDA with a message and leaves no
 * pending CUDA error behind. */
int sglb_host_register(void* host_ptr, size_t bytes);
int sglb_host_unregister(void* host_ptr);

/* Optional per-stage timing with CUDA events on the launching stream (used by bench.py for the
 * roofline figure).  Stages: 0 align, 1 quant_x, 2 w13 GEMM, 3 quant_act, 4 w2 GEMM, 5 topk reduce.
 * sglb_profile_read sums ms and sample counts per stage since the last read; synchronise first. */
void sglb_profile_enable(int on);
int sglb_profile_read(float* ms_out, int* calls_out, int n);

/* fp8 w8a16 on the tcgen05 path: 1 (default) = weights stay e4m3 and run on the fp8 tensor cores with
 * the activations split into two e4m3 pieces (csrc/umma_f8.cu); 0 = weights dequantised to bf16 in
 * the kernel (csrc/umma_dq.cu).  Process-wide switch, used by the parity tests to cover both.  Bit 1 (on = 3)
 * additionally keeps 128-row tiles at decode instead of the 32-row deep-ring configuration (umma_f8.cu F8Cfg); bit 2
 * (on = 5) also takes the 80-row configuration (an experiment: it bets on balanced routing, capi.cu). */
void sglb_set_fp8_native(int on);

/* Decode chain fusion level, for calls of up to 16384 routed slots.  0 = every step its own launch (align_count,
 * align_scatter, activation preparation, w13, activation preparation, w2, topk_reduce).  1 (default) = align + the
 * stage-1 operand preparation (int8 rows / e4m3 hi-lo split of the hidden states) as ONE prologue launch.  2 =
 * additionally the int8 activation quantised inside the stage-2 producer and the top-k reduce performed inside the
 * stage-2 kernel (3 launches per call for bf16 / int8) -- measured slower on B200 and kept as an experiment
 * (DESIGN.md 4c).  Results are bit-identical at every level (tests/test_gpu_fused_chain.py). */
void sglb_set_chain_fusion(int level);

/* ---- K1 router ------------------------------------------------------------------------------
 * grouped_topk_cpu          (test_grouped_topk.py:61-69, test_moe.py:61-70)  : bias == NULL
 * biased_grouped_topk_cpu   (test_biased_grouped_topk.py:71-80)              : bias != NULL
 * gating [M,E] (gating_dtype), bias [E] (bias_dtype); out topk_weights [M,topk] f32,
 * topk_ids [M,topk] i32, picks in descending score order, ties to the lowest expert index.
 * The biased variant's sigmoid is the correctly rounded one, (float)(1/(1+exp(-(double)x))), so ids and
 * weights are bit-reproducible by the CPU oracle (csrc/router.cu header).
 * Serving arguments of the 9-argument form (test_moe.py:61-70; the reference always passes 0, None, None):
 *   num_fused_shared_experts n > 0: the last of the topk slots becomes the fused shared expert,
 *     id = E + (token % n), weight = (sum of routed weights) / routed_scaling_factor; renormalisation
 *     divides by the routed sum only.  routed_scaling_factor is used for that slot only.
 *   num_token_non_padded: device scalar (int32, or int64 when non_padded_is_i64) or NULL; rows at and past
 *     it get ids = -1 and weights = 0. */
int sglb_grouped_topk(const void* gating, int gating_dtype, const void* bias, int bias_dtype,
                      int64_t M, int E, int topk, int num_expert_group, int topk_group,
                      int renormalize, int num_fused_shared_experts, float routed_scaling_factor,
                      const void* num_token_non_padded, int non_padded_is_i64, float* topk_weights,
                      int32_t* topk_ids, sglb_stream_t stream);

/* ---- K2 moe_align -----------------------------------------------------------------------------
 * The sort/align pass inside fused_experts_cpu; layout documented in csrc/align.cu.
 * S = M*topk.  chunk_list / tile_list may be NULL. */
typedef struct sglb_align_out {
  int32_t* expert_count;  /* [E]   */
  int32_t* expert_offset; /* [E+1] */
  int32_t* sorted_slot;   /* [S]   */
  int32_t* slot_pos;      /* [S]   */
  int32_t* chunk_list;    /* [3 * sglb_moe_align_max_items(M,topk,E,chunk_rows)] */
  int32_t* tile_list;     /* [3 * sglb_moe_align_max_items(M,topk,E,tile_rows)]  */
  int32_t* counters;      /* [4] num_valid, num_chunks, num_tiles, num_active_experts */
} sglb_align_out;

size_t sglb_moe_align_scratch_bytes(int64_t M, int topk, int E);
int64_t sglb_moe_align_max_items(int64_t M, int topk, int E, int rows);
int sglb_moe_align(const int32_t* topk_ids, int64_t M, int topk, int E, int chunk_rows,
                   int tile_rows, const sglb_align_out* out, void* scratch, sglb_stream_t stream);

/* ---- K2 for callers that cannot allocate device memory ---------------------------------------------------
 * No counterpart in the reference (its align pass is internal to fused_experts_cpu and runs on host memory).
 * sglb_moe_align wants a caller-owned scratch buffer once a batch has more than 1024 routed slots.  A caller
 * that cannot supply one -- an engine that gave every free byte of HBM to its own pools at start-up, a plugin
 * that must not call cudaMalloc or its host's allocator, a step where an allocation's implicit device
 * synchronisation is unacceptable -- calls this variant instead: same inputs, same outputs, bit-identical
 * results, and the scratch lives in a 1 MiB array inside the library's device image (resident on every GPU
 * that loads the library until the process exits).  The call allocates nothing, on the device or through any
 * library.
 *   limit      ceil(M*topk/1024) * E * 4 bytes <= sglb_moe_align_static_max_bytes() (= 1 MiB; e.g. 131072
 *              tokens x top-8 over 256 experts); larger problems return SGLB_ERR_INVALID with a message.
 *   set-up     the first call on each device looks up the array's device address (cudaGetSymbolAddress, with
 *              that device current -- the one holding the buffers, as for every other entry point) and caches
 *              it; make that first call before starting any stream capture.
 *   sharing    every call on a device uses the same array and leaves its histograms behind.  Nothing needs
 *              clearing between calls (each call overwrites what it later reads), but two calls that can run
 *              at the same time on one GPU -- different streams, or different host threads -- race on it and
 *              give wrong indices: keep calls to this entry point ordered per device. */
size_t sglb_moe_align_static_max_bytes(void);
int sglb_moe_align_static(const int32_t* topk_ids, int64_t M, int topk, int E, int chunk_rows,
                          int tile_rows, const sglb_align_out* out, sglb_stream_t stream);

/* ---- K3-K5 fused_experts ------------------------------------------------------------------------
 * fused_experts_cpu (test_moe.py:79-92, test_moe_int8.py:129, test_moe_fp8_ext.py:118,
 * test_moe_offloading_cpu.py:103-117, bench_moe.py:65-130).
 * Weights are row-major ("is_vnni = False" layout): w13 [E, 2I, H] with rows 0..I-1 = gate and
 * I..2I-1 = up, w2 [E, H, I]; mxfp4 rows hold K/2 bytes.  The VNNI-packed layout produced by
 * convert_weight_packed is turned into this one by sglb_unpack_weight.
 *   out[m] = sum_j topk_weights[m,j] * W2[e_j] . silu_and_mul(W13[e_j] . x[m]),  e_j = topk_ids[m,j]
 * ids outside [0,E) contribute nothing.  out may alias hidden (the reference's inplace=True). */
typedef struct sglb_moe_problem {
  int64_t M;
  int32_t topk, E, H, I;
  int32_t fmt;  /* SGLB_FMT_* */
  int32_t path; /* SGLB_PATH_* */
  const void* hidden; /* [M,H] bf16 */
  const void* w13;
  const void* w2;
  const float* topk_weights; /* [M,topk] */
  const int32_t* topk_ids;   /* [M,topk] */
  const void* w13_scale; /* int8: f32 [E,2I]; fp8: f32 [E,ceil(2I/bn),ceil(H/bk)]; mxfp4: u8 [E,2I,H/32] */
  const void* w2_scale;  /* int8: f32 [E,H];  fp8: f32 [E,ceil(H/bn),ceil(I/bk)];  mxfp4: u8 [E,H,I/32]  */
  int32_t block_n, block_k; /* fp8 only */
  void* out;                /* [M,H] bf16 */
  void* workspace;
  size_t workspace_bytes; /* >= sglb_fused_experts_workspace_bytes(...) */
  /* mxfp4 only, optional (NULL = dequantising kernels): the e8m0 scales re-tiled by sglb_mx4_tile_scales
   * (split = I for w13, 0 for w2); with them the GEMMs run on the block-scaled fp4 tensor cores */
  const void* w13_scale_tiles;
  const void* w2_scale_tiles;
  /* 17-argument vintage (on-box vllm cpu_moe.py:308-326; gpt-oss): optional per-expert f32 biases added to the
   * dequantised GEMM results -- w13_bias [E,2I] before the activation, w2_bias [E,H] before the topk weight --
   * and the activation: act 0 = silu(gate)*up, act 1 = clamped swiglu g*sigmoid(alpha*g)*(u+1) with
   * g = min(gate,limit), u = clamp(up,-limit,limit) (limit <= 0: no clamp).  Gate = rows [0,I), up = rows [I,2I). */
  const float* w13_bias;
  const float* w2_bias;
  int32_t act;
  float alpha, limit;
  /* Slot-addressed output (expert-parallel return path): when slot_out != NULL stage 2 writes the bf16 row of
   * routed slot s = token*topk + j to row s of slot_out (row slot_map[s] when slot_map != NULL) and the top-k
   * reduce is skipped; rows of masked slots are not written; `out` may be NULL.  The caller reduces a token's
   * rows in slot order (sglb_topk_reduce's arithmetic), which makes n-GPU results identical to one GPU's. */
  void* slot_out;
  const int32_t* slot_map;
  /* > 0: expected routed rows per expert for the AUTO path rule (default M*topk/E) */
  float rows_per_expert_hint;
  /* The layer's shared expert riding along in the same launches (DeepSeek: fused_experts_cpu followed by
   * shared_expert_cpu, test_moe_fp8_ext.py:60-67,118-124).  With shared_w13 != NULL the call returns
   *   out = bf16(shared_mlp(x) + bf16(routed(x)) * shared_rsf)
   * bit-identical to sglb_fused_experts followed by sglb_shared_expert on its result: every token is one more
   * routed row of "expert E" whose weights are shared_w13 [2I,H] / shared_w2 [H,I] (same I, format and block
   * sizes as the routed experts; scales as for E = 1), so the shared expert's 3*H*I weight bytes stream inside
   * the routed GEMM launches instead of five small launches of their own.  Only where
   * sglb_fused_experts_shared_rides(p) returns 1 (tcgen05 path, bf16 / int8 / fp8 tensor-native, plain silu,
   * no slot_out); elsewhere call sglb_shared_expert. */
  const void* shared_w13;
  const void* shared_w2;
  const void* shared_w13_scale;
  const void* shared_w2_scale;
  float shared_rsf;
} sglb_moe_problem;

size_t sglb_fused_experts_workspace_bytes(int64_t M, int topk, int E, int H, int I, int fmt);
int sglb_fused_experts(const sglb_moe_problem* p, sglb_stream_t stream);
/* 1 if this problem (shape, format, path, options; the shared_* pointers themselves are not looked at) can carry
 * the shared expert inside its launches, else 0 */
int sglb_fused_experts_shared_rides(const sglb_moe_problem* p);

/* ---- K6 shared_expert -----------------------------------------------------------------------------
 * shared_expert_cpu (test_shared_experts.py:68,78; test_moe_fp8_ext.py:60-61; test_moe_fp8.py:87-88):
 *   out = silu_and_mul(x . W1^T) . W2^T + fused_out * routed_scaling_factor
 * w1 [2I,H], w2 [H,I]; same formats and scale layouts as above with E = 1. */
typedef struct sglb_shared_problem {
  int64_t M;
  int32_t H, I;
  int32_t fmt;
  int32_t path;
  const void* hidden; /* [M,H] bf16 */
  const void* w1;
  const void* w2;
  const void* w1_scale;
  const void* w2_scale;
  int32_t block_n, block_k;
  const void* fused_out; /* [M,H] bf16; NULL: out receives the MLP result alone (the caller applies
                          * sglb_shared_axpy later, e.g. after an expert-parallel combine) */
  float routed_scaling_factor;
  void* out; /* [M,H] bf16, may alias hidden */
  void* workspace;
  size_t workspace_bytes; /* >= sglb_fused_experts_workspace_bytes(M, 1, 1, H, I, fmt) */
} sglb_shared_problem;

int sglb_shared_expert(const sglb_shared_problem* p, sglb_stream_t stream);
/* the epilogue alone: out[i] = bf16(y[i] + fused[i] * routed_scaling_factor), n bf16 elements */
int sglb_shared_axpy(const void* y, const void* fused, float routed_scaling_factor, int64_t n, void* out,
                     sglb_stream_t stream);

/* ---- mxfp4 scale tiles for the block-scaled tensor cores ----------------------------------------------
 * e8m0 scales [E, rows, K/32] (test_mxfp4.py:69-88 layout) -> 512-byte tiles of 128 rows x 4 k-groups in
 * the order tcgen05.cp expects, [E, blocks, ceil(K/128), 512].  split > 0 tiles the rows [0, split) and
 * [split, rows) separately (gate and up halves of w13: split = I); padding entries hold 2^0.
 * Weights used with the tiles: K % 64 == 0, and when K % 128 != 0 the buffer must stay readable for 32
 * bytes past its end (the last row's final TMA box). */
size_t sglb_mx4_scale_tiles_bytes(int64_t E, int rows, int split, int K);
int sglb_mx4_tile_scales(const void* scales, void* tiles, int64_t E, int rows, int split, int K,
                         sglb_stream_t stream);
void sglb_set_mx4_native(int on); /* 0: keep mxfp4 on the dequantising kernels even when tiles are given */

/* ---- K9 mxfp4 linear ------------------------------------------------------------------------------
 * mxfp4_scaled_mm_cpu (test_mxfp4.py:167,196): out[M,N] bf16 = A[M,K] bf16 . dequant(Bq,Bs)^T + bias
 * Bq [N,K/2] u8 row-major, Bs [N,K/32] u8 row-major, bias f32 [N] or NULL.
 * workspace >= sglb_fused_experts_workspace_bytes(M, 1, 1, K, N, SGLB_FMT_MXFP4). */
int sglb_mxfp4_mm(const void* A, const void* Bq, const void* Bs, const float* bias, int64_t M, int N,
                  int K, void* out, void* workspace, size_t workspace_bytes, sglb_stream_t stream);

/* ---- K10 dense quantised linears ------------------------------------------------
 * One entry point behind the reference's dense ops; out[M,N] bf16 = A[M,K] . dequant(W[N,K])^T + bias:
 *   weight_packed_linear(mat1, mat2, bias?, is_vnni)                      test_gemm.py:22-25      fmt BF16
 *   int8_scaled_mm_cpu(Aq, Bq, As, Bs, bias?, out_dtype, is_vnni)         test_gemm_int8.py:69    fmt INT8, a_scale set
 *   int8_scaled_mm_with_quant(A, Bq, Bs, bias?, out_dtype, is_vnni)       test_gemm_int8.py:74    fmt INT8, a_scale NULL
 *   fp8_scaled_mm_cpu(A, Bq, Bs, block_size, bias?, out_dtype, is_vnni)   test_gemm_fp8.py:54-62  fmt FP8_W8A16
 * W is row-major [N,K] (the HBM twin registered by convert_weight_packed); w_scale: int8 f32 [N], fp8 f32
 * [ceil(N/bn), ceil(K/bk)], mxfp4 u8 [N,K/32].  A is bf16 [M,K], or int8 [M,K] with a_scale f32 [M]
 * (per_token_quant_int8_cpu's outputs).  bias f32 [N] or NULL, added in f32 before the one bf16 rounding.
 * Same kernels as fused_experts stage 2 with E = 1, topk = 1, no row weights. */
typedef struct sglb_linear_problem {
  int64_t M;
  int32_t N, K;
  int32_t fmt;
  int32_t path;         /* SGLB_PATH_AUTO / _STREAM / _UMMA */
  const void* A;
  const float* a_scale; /* int8 only; NULL = quantise A (bf16) per token inside the call */
  const void* W;
  const void* w_scale;
  int32_t block_n, block_k;
  const float* bias;
  void* out; /* [M,N] bf16 */
  void* workspace;
  size_t workspace_bytes; /* >= sglb_linear_workspace_bytes(M, N, K, fmt) */
} sglb_linear_problem;

size_t sglb_linear_workspace_bytes(int64_t M, int N, int K, int fmt);
int sglb_linear(const sglb_linear_problem* p, sglb_stream_t stream);

/* ---- K11 the ops either side of the MoE block ----------------------------------------
 * silu_and_mul_cpu       test_activation.py:10-28 (out, x), bench_silu_and_mul.py:31 (functional):
 *                        out[r, j] = T(silu(x[r, j])) * x[r, d + j], x [rows, 2d], out [rows, d]
 * rmsnorm_cpu            test_norm.py:43-48: out = T(x * rsqrt(mean(x^2) + eps)) * weight, rows of `hidden`
 * fused_add_rmsnorm_cpu  test_norm.py:50-63: x += residual in f32; residual = T(x); x = rmsnorm(x), in place
 * dtype = SGLB_DT_BF16 or SGLB_DT_F16 for every tensor of a call; rows are contiguous. */
int sglb_silu_and_mul(const void* x, void* out, int64_t rows, int d, int dtype, sglb_stream_t stream);
int sglb_rmsnorm(const void* x, const void* weight, float eps, void* out, int64_t rows, int hidden, int dtype,
                 sglb_stream_t stream);
int sglb_fused_add_rmsnorm(void* x, void* residual, const void* weight, float eps, int64_t rows, int hidden,
                           int dtype, sglb_stream_t stream);

/* ---- K7 prepack ---------------------------------------------------------------------------------
 * convert_weight_packed (test_moe.py:75-76,115-124; test_gemm.py:36-46; test_gemm_int8.py:80-93)
 * and convert_scale_packed (test_mxfp4.py:160,182-185) define a CPU (AMX/VNNI) tile layout.  These
 * kernels convert between the row-major layout the sm_100a kernels consume and that layout, so
 * tensors produced for / by the reference stay interchangeable:
 *   row-major [E, N, K] elements of elem_bytes  <->  [E, N/32, K/vnni, 32, vnni]
 *   bf16: elem_bytes 2, vnni 2 (pinned by test_gemm.py:38-44, test_moe.py:117-124)
 *   int8: elem_bytes 1, vnni 4, int8_comp 1 -- every 32-row block is followed by 32 int32
 *         compensation values 128 * sum_k w[n,k] (test_gemm_int8.py:84-93), so the packed buffer
 *         holds N*(K+4) bytes per expert
 *   fp8 / packed-fp4 bytes: elem_bytes 1, vnni 4, int8_comp 0 (layout not pinned by the reference)
 * to_packed != 0: row-major -> packed, else packed -> row-major (compensation is dropped). */
int sglb_repack_weight(const void* src, void* dst, int elem_bytes, int vnni, int64_t E, int64_t N,
                       int64_t K, int int8_comp, int to_packed, sglb_stream_t stream);
/* e8m0 scales [E, N, G] <-> [E, N/32, G, 32]  (G = K/32), test_mxfp4.py:185 */
int sglb_repack_scale(const void* src, void* dst, int64_t E, int64_t N, int64_t G, int to_packed,
                      sglb_stream_t stream);

/* ---- K8 expert-parallel dispatch / combine ---------------------------------------------------------
 * Nothing in the reference shards experts; its hook is that fused_experts_cpu ignores ids == -1
 * (test_moe_offloading_cpu.py:57-68), which is what a rank sees for experts it does not own.
 * Rank r owns experts [r*E/world, (r+1)*E/world).  These kernels are the device side of the exchange; the
 * collective itself (NCCL all-to-all) is issued by the host layer (ep.py).  Token rows travel once per
 * (token, owner rank); what comes back is one bf16 row per routed SLOT (the owner runs sglb_fused_experts with
 * slot_out / slot_map), and the source adds a token's rows in slot order in f32 with one rounding, so the
 * n-GPU result is bit-identical to the one-GPU result.
 *   plan:      dest_row[m*world + r] = row of token m in the segment sent to rank r (token order) or -1;
 *              send_count[r] = rows in that segment;  slot_index[m*topk + j] = index of slot (m, j) inside the
 *              slot segment returned by its owner (slots of one owner numbered in (token, j) order) or -1;
 *              send_slots[r] = slots owned by rank r
 *   pack:      send_x[row_seg(r) + dest_row] = hidden[m];  send_meta[row] = {topk local ids (-1 if not owned
 *              by r), topk weights as f32 bits};  send_rows = total rows of the send buffers
 *   ret_index: (owner) ret_index[i*topk + j] = position of received slot (row i, j) in the return buffer, the
 *              numbering the source computed in plan (-1 for masked slots); recv_ids rows are id_stride ints
 *              apart (2*topk for the packed meta rows); recv_count / recv_slots [world]: rows / slots received
 *              from each source (device arrays; NULL with a fixed capacity)
 *   combine:   out[m] = bf16( sum_j ret[slot_seg(owner(m,j)) + slot_index[m,j]] ), f32, slot order; with
 *              shared_y != NULL additionally out = bf16(shared_y[m] + that * routed_scaling_factor)
 *   row_seg(r) = r*cap, slot_seg(r) = r*cap*topk when cap > 0 (fixed-capacity exchange, padding rows read as
 *   all -1), else the exclusive scans of the row / slot counts (exact splits). */
int sglb_ep_plan(const int32_t* topk_ids, int64_t M, int topk, int E, int world, int32_t* dest_row,
                 int32_t* send_count, int32_t* slot_index, int32_t* send_slots, sglb_stream_t stream);
int sglb_ep_pack(const void* hidden, const int32_t* topk_ids, const float* topk_weights,
                 const int32_t* dest_row, const int32_t* send_count, int64_t M, int topk, int E,
                 int world, int H, int cap, int64_t send_rows, void* send_x, int32_t* send_meta,
                 sglb_stream_t stream);
int sglb_ep_ret_index(const int32_t* recv_ids, int id_stride, int64_t recv_rows, int topk, int world, int cap,
                      const int32_t* recv_count, const int32_t* recv_slots, int32_t* ret_index,
                      sglb_stream_t stream);
int sglb_ep_combine(const void* ret, const int32_t* topk_ids, const int32_t* slot_index,
                    const int32_t* send_slots, int64_t M, int topk, int E, int world, int H, int cap,
                    const void* shared_y, float routed_scaling_factor, void* out, sglb_stream_t stream);

/* ---- K8, decode regime: the same exchange over NVLink peer memory (csrc/ep_p2p.cu) ---------------------
 * Every rank allocates one region (sglb_p2p_alloc: cudaMalloc + CUDA IPC handle), the 64-byte handles are
 * exchanged by the host layer and opened with sglb_p2p_open; peer_bases[r] is rank r's region as mapped in
 * this process (the rank's own allocation for r == rank).  Calls are counted on the device (the region holds
 * the counter, advanced by dispatch), so the three kernels can sit in a CUDA graph; `parity` (0 / 1) selects
 * the buffer set and must alternate call by call, identically on every rank; `cap` = slots per source rank
 * (>= every rank's token count, equal on all ranks).
 *   dispatch:      stores rows / local ids / weights of this rank's M tokens (M may be 0) into every peer's
 *                  receive buffers (row rank*cap + m) and raises the per-source arrival flag on every peer
 *   wait_dispatch: returns (on the stream) once all `world` sources' rows of this call have arrived
 *   (the caller then runs sglb_fused_experts over the world*cap received rows with slot_out = the y buffer)
 *   combine:       raises the "rows ready" flag on every peer, waits for the owners of each token's experts
 *                  and adds the token's topk slot rows, read over NVLink, in slot order, f32 -> bf16
 *                  (+ the shared expert epilogue as in sglb_ep_combine) */
int sglb_p2p_alloc(size_t bytes, void** ptr, void* ipc_handle_64);
int sglb_p2p_open(const void* ipc_handle_64, void** peer_ptr);
int sglb_p2p_close(void* peer_ptr);
int sglb_p2p_free(void* ptr);
/* offsets[0..8) = recv_x[2], recv_ids[2], recv_weights[2], y[2] (index = parity); y holds world*cap*topk rows;
 * returns the region size */
size_t sglb_ep_p2p_region_bytes(int world, int cap, int H, int topk, size_t* offsets);
int sglb_ep_p2p_dispatch(const void* hidden, const int32_t* topk_ids, const float* topk_weights, int64_t M,
                         int topk, int E, int world, int rank, int H, int cap, void* const* peer_bases,
                         int parity, sglb_stream_t stream);
int sglb_ep_p2p_wait_dispatch(const void* local_base, int world, int cap, int H, int topk, sglb_stream_t stream);
int sglb_ep_p2p_combine(const int32_t* topk_ids, int64_t M, int topk, int E, int world, int rank, int H,
                        int cap, void* const* peer_bases, int parity, const void* shared_y,
                        float routed_scaling_factor, void* out, sglb_stream_t stream);

/* ---- building blocks exposed for tests and for the expert-parallel layer --------------------------
 * per-row dynamic int8 quantisation (test_moe_int8.py:23-31): q = round(x * 127/absmax), scale = absmax/127 */
int sglb_quant_rows_int8(const void* x_bf16, int64_t rows, int K, int8_t* q, float* scale,
                         sglb_stream_t stream);
/* bf16 rows -> two e4m3 pieces for the tensor-native fp8 path: x ~= (hi + lo) * rexp, rexp = 2^(e-7) with
 * e = floor(log2(rowmax)); exact for elements within 2^-9 of the row maximum (csrc/umma_f8.cu) */
int sglb_split_rows_e4m3(const void* x_bf16, int64_t rows, int K, void* hi_e4m3, void* lo_e4m3, float* rexp,
                         sglb_stream_t stream);
/* K5: out[m] = bf16( sum_j ic2[slot_pos[m*topk+j]] ) over valid slots, in slot order */
int sglb_topk_reduce(const void* ic2_bf16, const int32_t* slot_pos, int64_t M, int topk, int H,
                     void* out_bf16, sglb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* SGLB200_H_ */
