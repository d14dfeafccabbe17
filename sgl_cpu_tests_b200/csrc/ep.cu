// K8: expert-parallel dispatch / combine -- the device side of the exchange step around
// fused_experts when the E experts are sharded over `world` GPUs.  Nothing in the
// reference does this; the hook it offers is that fused_experts_cpu ignores ids == -1
// (test_moe_offloading_cpu.py:57-68), which is exactly what a rank sees for non-local experts.
//
// Rank r owns experts [r*E_local, (r+1)*E_local).  A token row is sent ONCE to every rank that owns at
// least one of its topk experts, together with its topk ids rewritten to that rank's local numbering
// (-1 for experts that live elsewhere) and its topk weights.  The owner runs the ordinary local
// fused_experts over the received rows with SLOT-ADDRESSED output (sglb_moe_problem.slot_out / slot_map):
// one bf16 row per routed slot, exactly the row a single GPU would hold in its ic2 buffer.  The rows
// travel back and the source adds a token's rows in slot order j = 0..topk-1 in f32 with one rounding --
// the arithmetic of sglb_topk_reduce -- so the n-GPU result is bit-identical to the one-GPU result
// (SURVEY.md 8e "bit-exactness under EP").
//
//   plan       dest_row[m, r]   index of token m inside the row segment sent to rank r (token order), -1 if
//                               rank r owns none of its experts;  send_count[r] = rows in that segment
//              slot_index[m, j] index of slot (m, j) inside the slot segment that comes back from its owner:
//                               slots of the same owner are numbered in (token, j) order;  -1 for ids
//                               outside [0, E);  send_slots[r] = slots owned by rank r
//   pack       send_x   [row_seg(r) + dest_row, H]      bf16
//              send_meta[row_seg(r) + dest_row, 2*topk] i32: local ids | topk weights (f32 bits)
//   ret_index  (owner) ret_index[i, j] = slot_seg(s) + running count of valid slots among the rows received
//              from source s, in row order -- the same numbering the source computed, so no index travels;
//              -1 for masked slots;  this table is the slot_map of the local fused_experts call, whose
//              stage-2 epilogue therefore writes straight into the return buffer
//   combine    out[m] = bf16( sum_j f32(ret[slot_seg(owner_j) + slot_index[m, j]]) ), j ascending, valid only
//              (+ the shared expert: out = bf16(f32(y_shared[m]) + f32(that bf16 sum) * rsf), the arithmetic
//              of shared_expert_cpu's epilogue, test_shared_experts.py:34-40)
//   row_seg(r)  = r*cap       when cap > 0 (fixed-capacity exchange: no host sync; rows past send_count[r]
//   slot_seg(r) = r*cap*topk   carry ids = -1), else the exclusive scans of the row / slot counts.
#include "common.cuh"

namespace sglb {

constexpr int kPlanThreads = 1024;

// Block-wide exclusive scan of one int per thread (kPlanThreads threads); returns the exclusive prefix and
// the block total through `total`.  `smem` holds 33 ints.
__device__ __forceinline__ int block_exclusive_scan(int v, int* smem, int& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, inc, off);
    if (lane >= off) inc += o;
  }
  if (lane == 31) smem[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const int w = smem[lane];
    int winc = w;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, winc, off);
      if (lane >= off) winc += o;
    }
    smem[lane] = winc - w;
    if (lane == 31) smem[32] = winc;
  }
  __syncthreads();
  const int excl = smem[warp] + inc - v;
  total = smem[32];
  __syncthreads();
  return excl;
}

// one block per destination rank r
__global__ void __launch_bounds__(kPlanThreads)
    ep_plan_kernel(const int32_t* __restrict__ ids, int M, int topk, int E_local, int world,
                   int32_t* __restrict__ dest_row, int32_t* __restrict__ send_count,
                   int32_t* __restrict__ slot_index, int32_t* __restrict__ send_slots) {
  __shared__ int scan_a[33], scan_b[33];
  const int r = blockIdx.x;
  const int lo = r * E_local, hi = lo + E_local;
  int rows_before = 0, slots_before = 0;
  for (int base = 0; base < M; base += kPlanThreads) {
    const int m = base + threadIdx.x;
    int cnt = 0;
    if (m < M) {
      for (int j = 0; j < topk; ++j) {
        const int e = ids[(size_t)m * topk + j];
        cnt += (e >= lo && e < hi);
      }
    }
    int rows_total, slots_total;
    const int row_excl = block_exclusive_scan(cnt > 0, scan_a, rows_total);
    const int slot_excl = block_exclusive_scan(cnt, scan_b, slots_total);
    if (m < M) {
      dest_row[(size_t)m * world + r] = cnt > 0 ? rows_before + row_excl : -1;
      int k = slots_before + slot_excl;
      for (int j = 0; j < topk; ++j) {
        const int e = ids[(size_t)m * topk + j];
        if (e >= lo && e < hi) slot_index[(size_t)m * topk + j] = k++;
        else if (r == 0 && (e < 0 || e >= E_local * world)) slot_index[(size_t)m * topk + j] = -1;
      }
    }
    rows_before += rows_total;
    slots_before += slots_total;
  }
  if (threadIdx.x == 0) {
    send_count[r] = rows_before;
    send_slots[r] = slots_before;
  }
}

__device__ __forceinline__ int prefix_of(const int32_t* count, int r) {
  int s = 0;
  for (int q = 0; q < r; ++q) s += count[q];
  return s;
}

// grid (M, world); 128 threads copy one row with 16-B vectors.
__global__ void __launch_bounds__(128)
    ep_pack_kernel(const __nv_bfloat16* __restrict__ x, const int32_t* __restrict__ ids,
                   const float* __restrict__ w, const int32_t* __restrict__ dest_row,
                   const int32_t* __restrict__ send_count, int topk, int E_local, int world, int H,
                   int cap, __nv_bfloat16* __restrict__ send_x, int32_t* __restrict__ send_meta) {
  const int m = blockIdx.x, r = blockIdx.y;
  const int d = dest_row[(size_t)m * world + r];
  if (d < 0) return;
  const size_t row = (size_t)(cap > 0 ? r * cap : prefix_of(send_count, r)) + d;
  const __nv_bfloat16* src = x + (size_t)m * H;
  __nv_bfloat16* dst = send_x + row * H;
  if (H % 8 == 0) {
    for (int i = threadIdx.x; i < H / 8; i += 128)
      reinterpret_cast<int4*>(dst)[i] = reinterpret_cast<const int4*>(src)[i];
  } else {
    for (int i = threadIdx.x; i < H; i += 128) dst[i] = src[i];
  }
  const int lo = r * E_local, hi = lo + E_local;
  for (int j = threadIdx.x; j < topk; j += 128) {
    const int e = ids[(size_t)m * topk + j];
    const bool mine = e >= lo && e < hi;
    send_meta[row * 2 * topk + j] = mine ? e - lo : -1;
    send_meta[row * 2 * topk + topk + j] = mine ? __float_as_int(w[(size_t)m * topk + j]) : 0;
  }
}

// owner side, one block per source rank s: numbers the valid slots of the rows received from s
__global__ void __launch_bounds__(kPlanThreads)
    ep_ret_index_kernel(const int32_t* __restrict__ recv_ids, int id_stride, int topk, int world, int cap,
                        const int32_t* __restrict__ recv_count, const int32_t* __restrict__ recv_slots,
                        int32_t* __restrict__ ret_index) {
  __shared__ int scan_a[33];
  const int s = blockIdx.x;
  const int row0 = cap > 0 ? s * cap : prefix_of(recv_count, s);
  const int nrows = cap > 0 ? cap : recv_count[s];
  int before = cap > 0 ? s * cap * topk : prefix_of(recv_slots, s);
  for (int base = 0; base < nrows; base += kPlanThreads) {
    const int i = base + threadIdx.x;
    int cnt = 0;
    if (i < nrows)
      for (int j = 0; j < topk; ++j) cnt += recv_ids[(size_t)(row0 + i) * id_stride + j] >= 0;
    int total;
    const int excl = block_exclusive_scan(cnt, scan_a, total);
    if (i < nrows) {
      int k = before + excl;
      for (int j = 0; j < topk; ++j)
        ret_index[(size_t)(row0 + i) * topk + j] = recv_ids[(size_t)(row0 + i) * id_stride + j] >= 0 ? k++ : -1;
    }
    before += total;
  }
}

// grid (M, ceil(H/8/256))
__global__ void __launch_bounds__(256)
    ep_combine_kernel(const __nv_bfloat16* __restrict__ ret, const int32_t* __restrict__ ids,
                      const int32_t* __restrict__ slot_index, const int32_t* __restrict__ send_slots, int topk,
                      int E_local, int world, int H, int cap, const __nv_bfloat16* __restrict__ shared_y, float rsf,
                      __nv_bfloat16* __restrict__ out) {
  const int64_t m = blockIdx.x;
  const int h0 = (blockIdx.y * 256 + threadIdx.x) * 8;
  if (h0 >= H) return;
  const bool full = (H % 8 == 0);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int j = 0; j < topk; ++j) {
    const int e = ids[m * topk + j];
    if (e < 0 || e >= E_local * world) continue;
    const int r = e / E_local;
    const int seg = cap > 0 ? r * cap * topk : prefix_of(send_slots, r);
    const __nv_bfloat16* src = ret + ((size_t)seg + slot_index[m * topk + j]) * H + h0;
    if (full) {
      const int4 v = ldg_stream16(src);
      const uint32_t* wv = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        acc[2 * i] += bf16_bits_to_float(wv[i] & 0xffffu);
        acc[2 * i + 1] += bf16_bits_to_float(wv[i] >> 16);
      }
    } else {
      for (int i = 0; i < 8 && h0 + i < H; ++i) acc[i] += __bfloat162float(src[i]);
    }
  }
  __nv_bfloat16* dst = out + (size_t)m * H + h0;
  if (shared_y != nullptr) {
    const __nv_bfloat16* y = shared_y + (size_t)m * H + h0;
    for (int i = 0; i < 8 && h0 + i < H; ++i) acc[i] = fmaf(bf16_round(acc[i]), rsf, __bfloat162float(y[i]));
  }
  if (full) {
    uint32_t o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) o[i] = pack_bf16x2(acc[2 * i], acc[2 * i + 1]);
    *reinterpret_cast<int4*>(dst) = make_int4(o[0], o[1], o[2], o[3]);
  } else {
    for (int i = 0; i < 8 && h0 + i < H; ++i) dst[i] = __float2bfloat16_rn(acc[i]);
  }
}

}  // namespace sglb

using namespace sglb;

extern "C" int sglb_ep_plan(const int32_t* topk_ids, int64_t M, int topk, int E, int world,
                            int32_t* dest_row, int32_t* send_count, int32_t* slot_index, int32_t* send_slots,
                            sglb_stream_t stream) {
  SGLB_CHECK(world >= 1 && world <= 1024 && E >= world && E % world == 0,
             "ep_plan: the %d experts must split evenly over %d ranks", E, world);
  SGLB_CHECK(M >= 0 && M < (1ll << 31) / (topk > 0 ? topk : 1) && topk >= 1, "ep_plan: bad shape");
  SGLB_CHECK(send_count && send_slots && (M == 0 || (topk_ids && dest_row && slot_index)), "ep_plan: null pointer");
  ep_plan_kernel<<<world, kPlanThreads, 0, (cudaStream_t)stream>>>(topk_ids, (int)M, topk, E / world, world, dest_row,
                                                                   send_count, slot_index, send_slots);
  SGLB_LAUNCH_CHECK();
  count_launch();
  return SGLB_OK;
}

extern "C" int sglb_ep_pack(const void* hidden, const int32_t* topk_ids, const float* topk_weights,
                            const int32_t* dest_row, const int32_t* send_count, int64_t M, int topk,
                            int E, int world, int H, int cap, int64_t send_rows, void* send_x,
                            int32_t* send_meta, sglb_stream_t stream) {
  SGLB_CHECK(world >= 1 && E % world == 0 && H > 0 && topk >= 1 && M >= 0, "ep_pack: bad shape");
  SGLB_CHECK(cap == 0 || (cap >= M && send_rows == (int64_t)cap * world),
             "ep_pack: fixed capacity %d must hold all %lld tokens and send_rows must be cap*world", cap,
             (long long)M);
  if (send_rows > 0) {
    SGLB_CHECK(send_meta, "ep_pack: null pointer");
    // rows nobody writes (fixed-capacity padding) must read as "all experts masked"
    SGLB_CUDA(cudaMemsetAsync(send_meta, 0xff, sizeof(int32_t) * 2 * topk * send_rows, (cudaStream_t)stream));
  }
  if (M == 0) return SGLB_OK;
  SGLB_CHECK(hidden && topk_ids && topk_weights && dest_row && send_count && send_x, "ep_pack: null pointer");
  dim3 grid((unsigned)M, (unsigned)world);
  ep_pack_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)hidden, topk_ids, topk_weights, dest_row, send_count, topk, E / world, world,
      H, cap, (__nv_bfloat16*)send_x, send_meta);
  SGLB_LAUNCH_CHECK();
  count_launch();
  return SGLB_OK;
}

extern "C" int sglb_ep_ret_index(const int32_t* recv_ids, int id_stride, int64_t recv_rows, int topk, int world, int cap,
                                 const int32_t* recv_count, const int32_t* recv_slots, int32_t* ret_index,
                                 sglb_stream_t stream) {
  SGLB_CHECK(world >= 1 && world <= 1024 && topk >= 1 && id_stride >= topk && recv_rows >= 0, "ep_ret_index: bad shape");
  SGLB_CHECK(cap > 0 ? recv_rows == (int64_t)cap * world : (recv_count && recv_slots),
             "ep_ret_index: fixed capacity needs recv_rows == cap*world, exact splits need the two count arrays");
  if (recv_rows == 0) return SGLB_OK;
  SGLB_CHECK(recv_ids && ret_index, "ep_ret_index: null pointer");
  ep_ret_index_kernel<<<world, kPlanThreads, 0, (cudaStream_t)stream>>>(recv_ids, id_stride, topk, world, cap, recv_count,
                                                                        recv_slots, ret_index);
  SGLB_LAUNCH_CHECK();
  count_launch();
  return SGLB_OK;
}

extern "C" int sglb_ep_combine(const void* ret, const int32_t* topk_ids, const int32_t* slot_index,
                               const int32_t* send_slots, int64_t M, int topk, int E, int world, int H, int cap,
                               const void* shared_y, float routed_scaling_factor, void* out, sglb_stream_t stream) {
  SGLB_CHECK(world >= 1 && E % world == 0 && H > 0 && M >= 0 && topk >= 1, "ep_combine: bad shape");
  if (M == 0) return SGLB_OK;
  SGLB_CHECK(topk_ids && slot_index && send_slots && out, "ep_combine: null pointer");
  dim3 grid((unsigned)M, (unsigned)ceil_div(ceil_div(H, 8), 256));
  ep_combine_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)ret, topk_ids, slot_index, send_slots,
                                                            topk, E / world, world, H, cap,
                                                            (const __nv_bfloat16*)shared_y, routed_scaling_factor,
                                                            (__nv_bfloat16*)out);
  SGLB_LAUNCH_CHECK();
  count_launch();
  return SGLB_OK;
}
