// K8: expert-parallel dispatch / combine -- the device side of the exchange step around
// fused_experts when the E experts are sharded over `world` GPUs.  Nothing in the
// reference does this; the hook it offers is that fused_experts_cpu ignores ids == -1
// (test_moe_offloading_cpu.py:57-68), which is exactly what a rank sees for non-local experts.
//
// Rank r owns experts [r*E_local, (r+1)*E_local).  A token is sent ONCE to every rank that owns at
// least one of its topk experts, together with its topk ids rewritten to that rank's local numbering
// (-1 for experts that live elsewhere) and its topk weights.  The receiving rank runs the ordinary
// local fused_experts over the received rows and returns one partial row per received row; the
// source sums the partial rows of a token in rank order 0..world-1 in f32 (deterministic).
//
//   plan     dest_row[m, r] = index of token m inside the segment sent to rank r (token order), -1 if
//            rank r owns none of its experts; send_count[r] = rows in that segment
//   pack     send_x   [seg_start(r) + dest_row, H]      bf16
//            send_meta[seg_start(r) + dest_row, 2*topk] i32: local ids | topk weights (f32 bits)
//   combine  out[m] = bf16( sum_r ret[seg_start(r) + dest_row[m, r]] )
//   seg_start(r) = r*cap when cap > 0 (fixed-capacity exchange: no host sync, rows past send_count[r]
//   carry ids = -1), else the exclusive scan of send_count (exact split sizes).
#include "common.cuh"

namespace sglb {

constexpr int kPlanThreads = 1024;

__global__ void __launch_bounds__(kPlanThreads)
    ep_plan_kernel(const int32_t* __restrict__ ids, int M, int topk, int E_local,
                   int world, int32_t* __restrict__ dest_row, int32_t* __restrict__ send_count) {
  __shared__ int warp_sums[32];
  __shared__ int chunk_total;
  const int r = blockIdx.x;
  const int lo = r * E_local, hi = lo + E_local;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int running = 0;
  for (int base = 0; base < M; base += kPlanThreads) {
    const int m = base + threadIdx.x;
    bool flag = false;
    if (m < M) {
      for (int j = 0; j < topk; ++j) {
        const int e = ids[(size_t)m * topk + j];
        flag |= (e >= lo && e < hi);
      }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, flag);
    const int in_warp = __popc(bal & ((1u << lane) - 1u));
    if (lane == 0) warp_sums[warp] = __popc(bal);
    __syncthreads();
    if (warp == 0) {
      int v = warp_sums[lane], inc = v;
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        int o = __shfl_up_sync(0xffffffffu, inc, off);
        if (lane >= off) inc += o;
      }
      warp_sums[lane] = inc - v;
      if (lane == 31) chunk_total = inc;
    }
    __syncthreads();
    if (m < M) dest_row[(size_t)m * world + r] = flag ? running + warp_sums[warp] + in_warp : -1;
    running += chunk_total;
    __syncthreads();
  }
  if (threadIdx.x == 0) send_count[r] = running;
}

__device__ __forceinline__ int seg_start(const int32_t* send_count, int r, int cap) {
  if (cap > 0) return r * cap;
  int s = 0;
  for (int q = 0; q < r; ++q) s += send_count[q];
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
  const size_t row = (size_t)seg_start(send_count, r, cap) + d;
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

// grid (M, ceil(H/8/256))
__global__ void __launch_bounds__(256)
    ep_combine_kernel(const __nv_bfloat16* __restrict__ ret, const int32_t* __restrict__ dest_row,
                      const int32_t* __restrict__ send_count, int world, int H, int cap,
                      __nv_bfloat16* __restrict__ out) {
  const int64_t m = blockIdx.x;
  const int h0 = (blockIdx.y * 256 + threadIdx.x) * 8;
  if (h0 >= H) return;
  const bool full = (H % 8 == 0);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  int start = 0;
  for (int r = 0; r < world; ++r) {
    const int d = dest_row[m * world + r];
    const int seg = cap > 0 ? r * cap : start;
    start += send_count[r];
    if (d < 0) continue;
    const __nv_bfloat16* src = ret + ((size_t)seg + d) * H + h0;
    if (full) {
      int4 v = ldg_stream16(src);
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
                            int32_t* dest_row, int32_t* send_count, sglb_stream_t stream) {
  SGLB_CHECK(world >= 1 && world <= 1024 && E >= world && E % world == 0,
             "ep_plan: the %d experts must split evenly over %d ranks", E, world);
  SGLB_CHECK(M >= 0 && M < (1ll << 31) && topk >= 1, "ep_plan: bad shape");
  SGLB_CHECK(send_count && (M == 0 || (topk_ids && dest_row)), "ep_plan: null pointer");
  ep_plan_kernel<<<world, kPlanThreads, 0, (cudaStream_t)stream>>>(topk_ids, (int)M, topk, E / world,
                                                                   world, dest_row, send_count);
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

extern "C" int sglb_ep_combine(const void* ret, const int32_t* dest_row, const int32_t* send_count,
                               int64_t M, int world, int H, int cap, void* out, sglb_stream_t stream) {
  SGLB_CHECK(world >= 1 && H > 0 && M >= 0, "ep_combine: bad shape");
  if (M == 0) return SGLB_OK;
  SGLB_CHECK(ret && dest_row && send_count && out, "ep_combine: null pointer");
  dim3 grid((unsigned)M, (unsigned)ceil_div(ceil_div(H, 8), 256));
  ep_combine_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)ret, dest_row, send_count,
                                                            world, H, cap, (__nv_bfloat16*)out);
  SGLB_LAUNCH_CHECK();
  count_launch();
  return SGLB_OK;
}
