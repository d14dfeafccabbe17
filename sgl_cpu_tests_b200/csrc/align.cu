// K2: moe_align -- stable counting sort of the M*topk routed slots by expert.
//
// Stands where the reference kernel's internal moe_align_block_size pass stands (SURVEY.md 3b;
// not exposed by any op in /root/reference, so the layout below is this build's own contract,
// stated identically in oracle/moe_oracle.py::moe_align and held bit-exact by tests):
//
//   slot            = m*topk + j                    (flat index into topk_ids)
//   valid slot      = 0 <= topk_ids[slot] < E       (-1 and any other id contribute nothing,
//                                                    test_moe_offloading_cpu.py:57-68)
//   sorted_slot[p]  = valid slots ordered by (expert, slot) ascending; tail filled with S=M*topk
//   slot_pos[slot]  = p, or -1 for an invalid slot
//   expert_offset   = exclusive scan of per-expert counts, E+1 entries
//   chunk_list      = (expert, row_start, rows<=chunk_rows) work items for the streaming GEMM
//   tile_list       = (expert, row_start, rows<=tile_rows)  work items for the tcgen05 GEMM
//   counters        = {num_valid, num_chunks, num_tiles, num_active_experts}
//
// Shared expert riding along (capi.cu, sglb_moe_problem.shared_w13): `shared_rows` = M extra rows are appended
// BEHIND the routed ones, at sorted positions [S, S + M) with sorted_slot[S + m] = m * topk (so that the gather's
// slot / topk gives token m), as ceil(M / tile_rows) tiles of expert id E at the end of tile_list (counted in
// num_tiles).  Nothing else changes: slot_pos, expert_offset and the chunk list describe the routed slots only.
#include <mutex>

#include "common.cuh"
#include "rowops.cuh"

namespace sglb {

constexpr int kAlignThreads = 1024;

__device__ __forceinline__ int block_exclusive_scan(int v, int* warp_sums, int* total) {
  // 1024 threads.  Returns exclusive prefix of v over threadIdx.x; *total = sum.
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    int o = __shfl_up_sync(0xffffffffu, inc, off);
    if (lane >= off) inc += o;
  }
  __syncthreads();  // protect warp_sums reuse
  if (lane == 31) warp_sums[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int w = warp_sums[lane];
    int winc = w;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      int o = __shfl_up_sync(0xffffffffu, winc, off);
      if (lane >= off) winc += o;
    }
    warp_sums[lane] = winc - w;  // exclusive
    if (lane == 31) warp_sums[32] = winc;
  }
  __syncthreads();
  int res = warp_sums[warp] + inc - v;
  *total = warp_sums[32];
  return res;
}

// Per-block stable ranks.  On return: rank = number of earlier slots of this block with the same
// expert; warp_cnt[w*E + e] has been turned into the exclusive prefix over warps; block_total[e]
// (smem) = slots of expert e in this block.
__device__ __forceinline__ void block_rank(int e, bool valid, int E, uint16_t* warp_cnt,
                                           int* block_total, int& rank_in_block) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 32 * E; i += kAlignThreads) warp_cnt[i] = 0;
  __syncthreads();
  unsigned peers = __match_any_sync(0xffffffffu, valid ? e : -1 - lane);
  int rank_in_warp = __popc(peers & ((1u << lane) - 1u));
  if (valid && rank_in_warp == 0) warp_cnt[warp * E + e] = (uint16_t)__popc(peers);
  __syncthreads();
  for (int x = threadIdx.x; x < E; x += kAlignThreads) {
    int run = 0;
    for (int w = 0; w < 32; ++w) {
      int c = warp_cnt[w * E + x];
      warp_cnt[w * E + x] = (uint16_t)run;
      run += c;
    }
    block_total[x] = run;
  }
  __syncthreads();
  rank_in_block = valid ? (int)warp_cnt[warp * E + e] + rank_in_warp : 0;
}

__global__ void __launch_bounds__(kAlignThreads) align_count_kernel(const int32_t* __restrict__ ids,
                                                                   int S, int E,
                                                                   int32_t* __restrict__ block_hist) {
  extern __shared__ int smem_i[];
  int* hist = smem_i;
  pdl_wait();
  pdl_trigger();
  for (int x = threadIdx.x; x < E; x += kAlignThreads) hist[x] = 0;
  __syncthreads();
  int slot = blockIdx.x * kAlignThreads + threadIdx.x;
  if (slot < S) {
    int e = ids[slot];
    if (e >= 0 && e < E) atomicAdd(&hist[e], 1);
  }
  __syncthreads();
  for (int x = threadIdx.x; x < E; x += kAlignThreads) block_hist[(size_t)blockIdx.x * E + x] = hist[x];
}

// Body of one align block (blockIdx.x < nb).  block_hist: per-block expert histograms from align_count_kernel, or
// nullptr -- then a block counts the slots of the other blocks itself (the single-launch prologue, nb <= 8).
__device__ __forceinline__ void align_block(int* smem_i, const int32_t* __restrict__ ids, int S, int E, int nb,
                                            const int32_t* __restrict__ block_hist, int chunk_rows, int tile_rows,
                                            int32_t* __restrict__ expert_count, int32_t* __restrict__ expert_offset,
                                            int32_t* __restrict__ sorted_slot, int32_t* __restrict__ slot_pos,
                                            int32_t* __restrict__ chunk_list, int32_t* __restrict__ tile_list,
                                            int32_t* __restrict__ counters, int shared_rows, int topk) {
  int* block_total = smem_i;            // [E]
  int* base = block_total + E;          // [E]  expert_offset + slots in earlier blocks
  int* warp_sums = base + E;            // [33]
  uint16_t* warp_cnt = reinterpret_cast<uint16_t*>(warp_sums + 34);  // [32*E]
  int* all_tot = reinterpret_cast<int*>(warp_cnt + 32 * E);          // [E]  (self-counting only)
  int* all_before = all_tot + E;                                     // [E]
  const bool self_count = nb > 1 && block_hist == nullptr;
  if (self_count) {
    for (int x = threadIdx.x; x < E; x += kAlignThreads) {
      all_tot[x] = 0;
      all_before[x] = 0;
    }
    __syncthreads();
    for (int b = 0; b < nb; ++b) {
      const int sl = b * kAlignThreads + threadIdx.x;
      if (sl < S) {
        const int ee = ids[sl];
        if (ee >= 0 && ee < E) {
          atomicAdd(&all_tot[ee], 1);
          if (b < (int)blockIdx.x) atomicAdd(&all_before[ee], 1);
        }
      }
    }
    __syncthreads();
  }

  const int slot = blockIdx.x * kAlignThreads + threadIdx.x;
  int e = -1;
  if (slot < S) e = ids[slot];
  const bool valid = (e >= 0 && e < E);
  int rank;
  block_rank(e, valid, E, warp_cnt, block_total, rank);

  // totals over all blocks + slots of the same expert in earlier blocks
  int total_valid = 0;
  for (int x0 = 0; x0 < E; x0 += kAlignThreads) {  // E <= 1024 -> a single pass
    int x = x0 + threadIdx.x;
    int tot = 0, before = 0;
    if (x < E) {
      if (nb == 1) {
        tot = block_total[x];
      } else if (self_count) {
        tot = all_tot[x];
        before = all_before[x];
      } else {
        for (int b = 0; b < nb; ++b) {
          int c = block_hist[(size_t)b * E + x];
          if (b < (int)blockIdx.x) before += c;
          tot += c;
        }
      }
    }
    int off = block_exclusive_scan(tot, warp_sums, &total_valid);
    if (x < E) {
      base[x] = off + before;
      if (blockIdx.x == 0) {
        expert_count[x] = tot;
        expert_offset[x] = off;
      }
    }
    __syncthreads();

    if (blockIdx.x == 0) {
      if (threadIdx.x == 0) {
        expert_offset[E] = total_valid;
        counters[0] = total_valid;
      }
      // work lists
      int nact = 0;
      for (int pass = 0; pass < 2; ++pass) {
        const int rows = pass == 0 ? chunk_rows : tile_rows;
        int32_t* list = pass == 0 ? chunk_list : tile_list;
        int n = (x < E && rows > 0) ? (tot + rows - 1) / rows : 0;
        int ntot;
        int first = block_exclusive_scan(n, warp_sums, &ntot);
        if (list != nullptr) {
          for (int i = 0; i < n; ++i) {
            int r0 = i * rows;
            list[3 * (first + i) + 0] = x;
            list[3 * (first + i) + 1] = off + r0;
            list[3 * (first + i) + 2] = min(rows, tot - r0);
          }
        }
        int nsh = 0;  // the shared expert's tiles: expert id E, rows [S, S + shared_rows)
        if (pass == 1 && list != nullptr && shared_rows > 0 && rows > 0) {
          nsh = (shared_rows + rows - 1) / rows;
          for (int i = threadIdx.x; i < nsh; i += kAlignThreads) {
            list[3 * (ntot + i) + 0] = E;
            list[3 * (ntot + i) + 1] = S + i * rows;
            list[3 * (ntot + i) + 2] = min(rows, shared_rows - i * rows);
          }
        }
        if (threadIdx.x == 0) counters[1 + pass] = (list != nullptr) ? ntot + nsh : 0;
        __syncthreads();
      }
      int act = (x < E && tot > 0) ? 1 : 0;
      block_exclusive_scan(act, warp_sums, &nact);
      if (threadIdx.x == 0) counters[3] = nact;
    }
  }
  __syncthreads();

  if (slot < S) {
    if (valid) {
      int p = base[e] + rank;
      sorted_slot[p] = slot;
      slot_pos[slot] = p;
    } else {
      slot_pos[slot] = -1;
    }
  }
  // sentinel tail
  for (int p = total_valid + blockIdx.x * kAlignThreads + threadIdx.x; p < S;
       p += nb * kAlignThreads)
    sorted_slot[p] = S;
  for (int m = blockIdx.x * kAlignThreads + threadIdx.x; m < shared_rows; m += nb * kAlignThreads)
    sorted_slot[S + m] = m * topk;
}

__global__ void __launch_bounds__(kAlignThreads)
    align_scatter_kernel(const int32_t* __restrict__ ids, int S, int E, int nb,
                         const int32_t* __restrict__ block_hist, int chunk_rows, int tile_rows,
                         int32_t* __restrict__ expert_count, int32_t* __restrict__ expert_offset,
                         int32_t* __restrict__ sorted_slot, int32_t* __restrict__ slot_pos,
                         int32_t* __restrict__ chunk_list, int32_t* __restrict__ tile_list,
                         int32_t* __restrict__ counters, int shared_rows, int topk) {
  extern __shared__ int smem_i[];
  pdl_wait();
  pdl_trigger();
  align_block(smem_i, ids, S, E, nb, block_hist, chunk_rows, tile_rows, expert_count, expert_offset, sorted_slot, slot_pos,
              chunk_list, tile_list, counters, shared_rows, topk);
}

// ---- prologue of the fused decode chain: ONE launch in front of the two GEMMs --------------------------------
// Blocks [0, nb): the align pass above (each block counts the other blocks' slots itself, nb <= kPrologueMaxAlignBlocks).
// Blocks [nb, nb + ceil(M / 4)): four token rows per block, 256 threads each:
//   * activation operand of stage 1: per-row int8 quantisation (mode 1, quant_row_int8) or the e4m3 hi / lo split
//     (mode 2, split_row_e4m3) -- the device functions the stand-alone kernels run, so the bytes are identical;
//   * state of the fused top-k reduce: tok_valid[m], red_cnt[m, :] = topk - valid (masked slots never arrive);
//   * zeroes row_amax (int8: maxed by the stage-1 epilogue).
__global__ void __launch_bounds__(kAlignThreads)
    moe_prologue_kernel(const int32_t* __restrict__ ids, int S, int E, int nb, int chunk_rows, int tile_rows,
                        int32_t* __restrict__ expert_count, int32_t* __restrict__ expert_offset,
                        int32_t* __restrict__ sorted_slot, int32_t* __restrict__ slot_pos, int32_t* __restrict__ chunk_list,
                        int32_t* __restrict__ tile_list, int32_t* __restrict__ counters, int shared_rows, int topk,
                        const PrologueRows pr) {
  extern __shared__ int smem_i[];
  pdl_wait();
  pdl_trigger();
  if ((int)blockIdx.x < nb) {
    align_block(smem_i, ids, S, E, nb, nullptr, chunk_rows, tile_rows, expert_count, expert_offset, sorted_slot, slot_pos,
                chunk_list, tile_list, counters, shared_rows, topk);
    return;
  }
  const int rb = blockIdx.x - nb, nrb = gridDim.x - nb;
  const int grp = threadIdx.x >> 8, t = threadIdx.x & 255;
  if (pr.row_amax != nullptr)
    for (int i = rb * kAlignThreads + threadIdx.x; i < pr.amax_rows; i += nrb * kAlignThreads) pr.row_amax[i] = 0u;
  const int m = rb * 4 + grp;
  if (m >= pr.M) return;  // whole 256-thread group
  int valid = 0;
  for (int j = 0; j < topk; ++j) {
    const int e = ids[(size_t)m * topk + j];
    valid += (e >= 0 && e < E) ? 1 : 0;
  }
  if (pr.red_cnt != nullptr) {
    if (t == 0) pr.tok_valid[m] = valid;
    for (int b = t; b < pr.red_blocks; b += 256) pr.red_cnt[(size_t)m * pr.red_blocks + b] = topk - valid;
  }
  // a token no routed slot points at is never gathered (an expert-parallel rank: most received rows); with a riding
  // shared expert every token row is an operand
  if (valid == 0 && shared_rows == 0) return;
  float* red_s = reinterpret_cast<float*>(smem_i) + grp * 8;
  if (pr.mode == 1)
    quant_row_int8<__nv_bfloat16>(pr.x + (size_t)m * pr.K, pr.K, (int8_t*)pr.q + (size_t)m * pr.K, pr.xs + m, t, red_s,
                                  1 + grp);
  else if (pr.mode == 2)
    split_row_e4m3(pr.x + (size_t)m * pr.K, pr.K, (uint8_t*)pr.q + (size_t)m * pr.K, (uint8_t*)pr.q2 + (size_t)m * pr.K,
                   pr.xs + m, t, red_s, 1 + grp);
}

int moe_align_impl(const int32_t* topk_ids, int64_t M, int topk, int E, int chunk_rows, int tile_rows,
                   const sglb_align_out* out, void* scratch, int shared_rows, cudaStream_t stream) {
  SGLB_CHECK(out != nullptr, "moe_align: null output struct");
  SGLB_CHECK(E >= 1 && E <= 1024, "moe_align: E=%d must be in [1,1024]", E);
  SGLB_CHECK(M >= 0 && topk >= 1, "moe_align: bad shape");
  SGLB_CHECK(M * (int64_t)topk < (1ll << 30), "moe_align: M*topk too large");
  const int S = (int)(M * topk);
  SGLB_CHECK(out->expert_count && out->expert_offset && out->sorted_slot && out->slot_pos &&
                 out->counters,
             "moe_align: null output pointer");
  const int nb = S > 0 ? ceil_div(S, kAlignThreads) : 1;
  if (nb > 1) {
    SGLB_CHECK(scratch != nullptr, "moe_align: scratch required for M*topk > %d", kAlignThreads);
    SGLB_CUDA(launch_pdl(align_count_kernel, dim3(nb), dim3(kAlignThreads), E * sizeof(int), stream,
                         topk_ids, S, E, (int32_t*)scratch));
  }
  size_t smem = (size_t)(2 * E + 34) * sizeof(int) + (size_t)32 * E * sizeof(uint16_t);
  if (smem > 48 * 1024) {
    SGLB_CUDA(cudaFuncSetAttribute(align_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)smem));
  }
  SGLB_CUDA(launch_pdl(align_scatter_kernel, dim3(nb), dim3(kAlignThreads), smem, stream, topk_ids, S, E,
                       nb, (const int32_t*)scratch, chunk_rows, tile_rows, out->expert_count,
                       out->expert_offset, out->sorted_slot, out->slot_pos, out->chunk_list,
                       out->tile_list, out->counters, shared_rows, topk));
  count_launch(nb > 1 ? 2 : 1);
  return SGLB_OK;
}

// One launch: align + stage-1 operand preparation + fused-reduce state (see moe_prologue_kernel).
int moe_prologue_impl(const int32_t* topk_ids, int64_t M, int topk, int E, int chunk_rows, int tile_rows,
                      const sglb_align_out* out, int shared_rows, const PrologueRows& pr, cudaStream_t stream) {
  SGLB_CHECK(E >= 1 && E <= 1024, "moe_prologue: E=%d must be in [1,1024]", E);
  const int S = (int)(M * topk);
  const int nb = S > 0 ? ceil_div(S, kAlignThreads) : 1;
  SGLB_CHECK(nb <= kPrologueMaxAlignBlocks, "moe_prologue: M*topk=%d too large for the single-launch prologue", S);
  const size_t smem = (size_t)(4 * E + 34) * sizeof(int) + (size_t)32 * E * sizeof(uint16_t);
  SGLB_CHECK(smem <= 96 * 1024, "moe_prologue: E=%d needs too much shared memory", E);
  static bool attr_set = false;
  if (!attr_set) {
    SGLB_CUDA(cudaFuncSetAttribute(moe_prologue_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    attr_set = true;
  }
  const int nrb = (int)ceil_div64(M, 4);
  SGLB_CUDA(launch_pdl(moe_prologue_kernel, dim3(nb + nrb), dim3(kAlignThreads), smem, stream, topk_ids, S, E, nb,
                       chunk_rows, tile_rows, out->expert_count, out->expert_offset, out->sorted_slot, out->slot_pos,
                       out->chunk_list, out->tile_list, out->counters, shared_rows, topk, pr));
  count_launch();
  return SGLB_OK;
}

}  // namespace sglb

using namespace sglb;

extern "C" size_t sglb_moe_align_scratch_bytes(int64_t M, int topk, int E) {
  int64_t S = M * topk;
  int64_t nb = ceil_div64(S > 0 ? S : 1, kAlignThreads);
  return (size_t)(nb * E) * sizeof(int32_t);
}

extern "C" int sglb_moe_align(const int32_t* topk_ids, int64_t M, int topk, int E, int chunk_rows,
                              int tile_rows, const sglb_align_out* out, void* scratch,
                              sglb_stream_t stream_) {
  return moe_align_impl(topk_ids, M, topk, E, chunk_rows, tile_rows, out, scratch, 0, (cudaStream_t)stream_);
}

// ---- moe_align for callers that cannot allocate device memory -------------------------------------------------
// sglb_moe_align needs a caller-provided scratch buffer (the per-block expert histograms that align_count_kernel
// hands to align_scatter_kernel) as soon as a batch has more than 1024 routed slots.  Some callers have nowhere
// to get one: a serving loop that handed all free HBM to its KV-cache pool at start-up, a plugin that is
// forbidden from touching the host's caching allocator, code that runs where cudaMalloc would implicitly
// synchronise the device in the middle of a latency-critical step.  For them the same two kernels run over a
// buffer that is part of this library's device image: the array below.  It costs kAlignStaticBytes of HBM on
// every GPU that loads libsglb200.so, from load until the process exits, whether or not the entry point is used.
//
// In host code the NAME g_align_static_scratch is only a handle for the runtime -- it is not a device address,
// and a kernel handed `g_align_static_scratch` from the host would fault (or, on a machine where the GPU can reach
// host memory, quietly scribble over the wrong bytes).  The device address is looked up once per device with
// cudaGetSymbolAddress, with that device current, and cached.
constexpr size_t kAlignStaticBytes = (size_t)1 << 20;  // 262144 histogram entries: blocks(M*topk / 1024) * E
constexpr int kAlignStaticMaxDevices = 64;
static __device__ __align__(256) int32_t g_align_static_scratch[kAlignStaticBytes / sizeof(int32_t)];

namespace {
std::mutex g_align_static_mu;
int32_t* g_align_static_ptr[kAlignStaticMaxDevices] = {};  // device address of the array, per device ordinal

// Device address of g_align_static_scratch on the CURRENT device (first use on a device does the look-up).
int align_static_scratch(int32_t** ptr) {
  int dev = -1;
  SGLB_CUDA(cudaGetDevice(&dev));
  SGLB_CHECK(dev >= 0 && dev < kAlignStaticMaxDevices, "moe_align_static: device ordinal %d not supported (max %d)", dev,
             kAlignStaticMaxDevices - 1);
  std::lock_guard<std::mutex> lock(g_align_static_mu);
  if (g_align_static_ptr[dev] == nullptr) {
    void* p = nullptr;
    SGLB_CUDA(cudaGetSymbolAddress(&p, g_align_static_scratch));
    SGLB_CHECK(p != nullptr, "moe_align_static: cudaGetSymbolAddress returned a null address on device %d", dev);
    g_align_static_ptr[dev] = (int32_t*)p;
  }
  *ptr = g_align_static_ptr[dev];
  return SGLB_OK;
}
}  // namespace

extern "C" size_t sglb_moe_align_static_max_bytes(void) { return kAlignStaticBytes; }

extern "C" int sglb_moe_align_static(const int32_t* topk_ids, int64_t M, int topk, int E, int chunk_rows,
                                     int tile_rows, const sglb_align_out* out, sglb_stream_t stream_) {
  SGLB_CHECK(M >= 0 && topk >= 1 && E >= 1, "moe_align_static: bad shape M=%lld topk=%d E=%d", (long long)M, topk, E);
  SGLB_CHECK(M * (int64_t)topk < (1ll << 30), "moe_align_static: M*topk too large");
  const size_t need = sglb_moe_align_scratch_bytes(M, topk, E);
  SGLB_CHECK(need <= kAlignStaticBytes,
             "moe_align_static: M=%lld topk=%d E=%d needs %zu bytes of scratch, the built-in buffer holds %zu "
             "(ceil(M*topk/1024)*E*4 must not exceed it); use sglb_moe_align with a caller-owned scratch buffer",
             (long long)M, topk, E, need, kAlignStaticBytes);
  int32_t* scratch = nullptr;
  if (int rc = align_static_scratch(&scratch)) return rc;
  // Nothing in the buffer has to start at zero: align_count_kernel stores every histogram entry of the
  // ceil(M*topk/1024) blocks this call uses before align_scatter_kernel reads them, and entries past those
  // (leftovers of an earlier, larger call) are never read.
  return moe_align_impl(topk_ids, M, topk, E, chunk_rows, tile_rows, out, scratch, 0, (cudaStream_t)stream_);
}
