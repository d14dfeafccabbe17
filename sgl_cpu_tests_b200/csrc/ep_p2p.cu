// K8, decode regime: expert-parallel dispatch / combine over NVLink peer memory, no collective library
// on the data path.  At decode batch sizes the NCCL exchange of ep.py (all-to-alls of a few hundred KB) costs
// more than the experts themselves (2 GPUs, Qwen3 int8, 128 tokens per rank: 250 us per step against 120 us
// on one GPU): the exchange is pure latency.  Here every rank owns one cudaMalloc'd region that all peers map
// through CUDA IPC, and two kernels do the exchange:
//
//   dispatch  block (m, r) stores token m's row straight into rank r's receive buffer (slot [src, m]:
//             fixed capacity, no counts to agree on) together with its topk ids in r's local numbering
//             (-1 where r does not own the expert, and for the unused slots m >= M); the last block to
//             finish raises flag[src] = seq on every peer (system-scope release)
//   experts   the ordinary local fused_experts over the world*cap received rows, after a one-block
//             kernel has seen every source's flag (system-scope acquire), with SLOT-ADDRESSED output: the
//             stage-2 epilogue writes the bf16 row of slot (row i, j) to y[i*topk + j] -- exactly the row a
//             single GPU holds in its ic2 buffer for that (token, expert) pair
//   combine   raises the "y ready" flag on every peer, waits for the peers that own experts of its
//             token, then reads the token's topk rows over NVLink (L1-bypassing loads) and adds them in
//             slot order j = 0..topk-1 in f32 with one rounding -- sglb_topk_reduce's arithmetic, so the
//             result is bit-identical to the single-GPU fused_experts (and to the NCCL path of ep.cu);
//             optionally fuses shared_expert_cpu's epilogue, out = bf16(y_shared + bf16(sum) * rsf)
//
// Call counting lives on the device: every rank's region holds `seq`, advanced by the dispatch kernel, and
// the three kernels read it from there, so a captured CUDA graph can be replayed (no per-call kernel
// argument).  Buffers are double-buffered on `par`, a kernel argument: callers alternate par = 0, 1, 0, ...
// (two captured graphs).  Reuse is safe without further handshakes: a rank can only start dispatch(i+2)
// after its combine(i+1), which waited for every peer's experts(i+1) to finish, i.e. after every reader of
// the buffers of call i is done (same argument for `y`).
#include <stdlib.h>

#include "common.cuh"

namespace sglb {

namespace {

constexpr int kMaxWorld = 16;

struct P2PLayout {
  size_t x[2], ids[2], tw[2], y[2], flags, counter, seq, total;
};

__host__ __device__ inline P2PLayout p2p_layout(int world, int cap, int H, int topk) {
  P2PLayout L;
  size_t off = 0;
  const size_t rows = (size_t)world * cap;
  auto take = [&](size_t bytes) {
    const size_t o = off;
    off += (bytes + 255) / 256 * 256;
    return o;
  };
  for (int p = 0; p < 2; ++p) {
    L.x[p] = take(rows * H * 2);
    L.ids[p] = take(rows * topk * 4);
    L.tw[p] = take(rows * topk * 4);
    L.y[p] = take(rows * topk * H * 2);  // one row per routed slot
  }
  L.flags = take(sizeof(uint32_t) * 2 * kMaxWorld);
  L.counter = take(256);
  L.seq = take(256);
  L.total = off;
  return L;
}

struct PeerPtrs {
  char* base[kMaxWorld];
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ int4 ld_peer16(const void* p) {  // bypasses L1: the line changes between calls
  int4 v;
  asm volatile("ld.relaxed.sys.global.v4.s32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// Spins until *flag has reached seq (wrap-safe).  A peer that never arrives (a crashed rank) ends in a trap
// after `timeout_ns` rather than in a hung GPU; 0 waits for ever (SGLB_EP_TIMEOUT_S, default 60 s: long
// enough for a peer held up by a prefill batch, a GC pause or checkpoint I/O).
__device__ __forceinline__ void wait_flag(const uint32_t* flag, uint32_t seq, uint64_t timeout_ns) {
  const uint64_t t0 = globaltimer_ns();
  while ((int32_t)(ld_acquire_sys(flag) - seq) < 0) {
    __nanosleep(64);
    if (timeout_ns != 0 && globaltimer_ns() - t0 > timeout_ns) __trap();
  }
}

// grid (cap, world), 128 threads
__global__ void __launch_bounds__(128)
    ep_p2p_dispatch_kernel(const __nv_bfloat16* __restrict__ x, const int32_t* __restrict__ ids,
                           const float* __restrict__ tw, int M, int topk, int E_local, int world, int rank, int H,
                           int cap, PeerPtrs peers, int par) {
  const int m = blockIdx.x, r = blockIdx.y;
  const P2PLayout L = p2p_layout(world, cap, H, topk);
  // this call's number: the local counter + 1 (the last block stores it back, after every block has read it)
  const uint32_t seq = *reinterpret_cast<const volatile uint32_t*>(peers.base[rank] + L.seq) + 1u;
  const size_t slot = (size_t)rank * cap + m;
  char* pb = peers.base[r];
  int32_t* pid = reinterpret_cast<int32_t*>(pb + L.ids[par]) + slot * topk;
  float* ptw = reinterpret_cast<float*>(pb + L.tw[par]) + slot * topk;
  const int lo = r * E_local, hi = lo + E_local;
  __shared__ int any_mine;
  if (threadIdx.x == 0) any_mine = 0;
  __syncthreads();
  for (int j = threadIdx.x; j < topk; j += 128) {
    int e = -1;
    float w = 0.f;
    if (m < M) {
      e = ids[(size_t)m * topk + j];
      w = tw[(size_t)m * topk + j];
    }
    const bool mine = e >= lo && e < hi;
    pid[j] = mine ? e - lo : -1;
    ptw[j] = mine ? w : 0.f;
    if (mine) any_mine = 1;
  }
  __syncthreads();
  if (any_mine) {
    const int4* src = reinterpret_cast<const int4*>(x + (size_t)m * H);
    int4* dst = reinterpret_cast<int4*>(pb + L.x[par] + slot * H * 2);
#pragma unroll 4  // four 16-byte loads in flight per thread before the first peer store
    for (int i = threadIdx.x; i < H / 8; i += 128) dst[i] = src[i];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* counter = reinterpret_cast<uint32_t*>(peers.base[rank] + L.counter);
    const uint32_t total = gridDim.x * gridDim.y;
    if (atomicAdd(counter, 1u) == total - 1) {
      *counter = 0;
      *reinterpret_cast<volatile uint32_t*>(peers.base[rank] + L.seq) = seq;
      __threadfence_system();
      for (int q = 0; q < world; ++q)
        st_release_sys(reinterpret_cast<uint32_t*>(peers.base[q] + L.flags) + rank, seq);
    }
  }
}

// one block: every source's rows of call `seq` have landed in this rank's receive buffers
__global__ void ep_p2p_wait_kernel(const char* base, size_t flags_off, size_t seq_off, int world, uint64_t timeout_ns) {
  const uint32_t seq = *reinterpret_cast<const volatile uint32_t*>(base + seq_off);
  if ((int)threadIdx.x < world)
    wait_flag(reinterpret_cast<const uint32_t*>(base + flags_off) + threadIdx.x, seq, timeout_ns);
}

// grid (max(M, 1), ceil(H/8/256)), 256 threads
__global__ void __launch_bounds__(256)
    ep_p2p_combine_kernel(const int32_t* __restrict__ ids, int M, int topk, int E_local, int world, int rank, int H,
                          int cap, PeerPtrs peers, int par, uint64_t timeout_ns,
                          const __nv_bfloat16* __restrict__ shared_y, float rsf, __nv_bfloat16* __restrict__ out) {
  const int64_t m = blockIdx.x;
  const P2PLayout L = p2p_layout(world, cap, H, topk);
  const uint32_t seq = *reinterpret_cast<const volatile uint32_t*>(peers.base[rank] + L.seq);
  // this rank's slot rows are complete (previous kernels of the stream): block (0, 0) tells every peer
  if (blockIdx.x == 0 && blockIdx.y == 0 && (int)threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(peers.base[threadIdx.x] + L.flags) + kMaxWorld + rank, seq);
  }
  if (m >= M) return;  // a rank without tokens only announces its rows
  __shared__ uint32_t need;  // bit r: rank r owns at least one expert of token m
  if (threadIdx.x == 0) need = 0;
  __syncthreads();
  for (int j = threadIdx.x; j < topk; j += 256) {
    const int e = ids[m * topk + j];
    if (e >= 0 && e < E_local * world) atomicOr(&need, 1u << (e / E_local));
  }
  __syncthreads();
  const uint32_t nd = need;
  if ((int)threadIdx.x < world && ((nd >> threadIdx.x) & 1u))
    wait_flag(reinterpret_cast<const uint32_t*>(peers.base[rank] + L.flags) + kMaxWorld + threadIdx.x, seq, timeout_ns);
  __syncthreads();
  const int h0 = (blockIdx.y * 256 + threadIdx.x) * 8;
  if (h0 >= H) return;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  // slot order: the single-GPU reduction order.  Eight slots per pass: their peer loads are all in flight before the
  // first addition (one NVLink round trip per pass instead of one per slot).
  for (int j0 = 0; j0 < topk; j0 += 8) {
    int e[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      e[i] = j0 + i < topk ? ids[m * topk + j0 + i] : -1;
      if (e[i] >= E_local * world) e[i] = -1;
    }
    int4 v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
      v[i] = e[i] >= 0 ? ld_peer16(peers.base[e[i] / E_local] + L.y[par] +
                                   ((((size_t)rank * cap + m) * topk + j0 + i) * H + h0) * 2)
                       : make_int4(0, 0, 0, 0);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (e[i] < 0) continue;
      const uint32_t* wv = reinterpret_cast<const uint32_t*>(&v[i]);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        acc[2 * q] += bf16_bits_to_float(wv[q] & 0xffffu);
        acc[2 * q + 1] += bf16_bits_to_float(wv[q] >> 16);
      }
    }
  }
  if (shared_y != nullptr) {  // shared_expert_cpu's epilogue on the rounded sum (misc.cu shared_axpy_kernel)
    const int4 yv = *reinterpret_cast<const int4*>(shared_y + (size_t)m * H + h0);
    const uint32_t* yw = reinterpret_cast<const uint32_t*>(&yv);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      acc[2 * i] = fmaf(bf16_round(acc[2 * i]), rsf, bf16_bits_to_float(yw[i] & 0xffffu));
      acc[2 * i + 1] = fmaf(bf16_round(acc[2 * i + 1]), rsf, bf16_bits_to_float(yw[i] >> 16));
    }
  }
  uint32_t o[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) o[i] = pack_bf16x2(acc[2 * i], acc[2 * i + 1]);
  *reinterpret_cast<int4*>(out + (size_t)m * H + h0) = make_int4(o[0], o[1], o[2], o[3]);
}

uint64_t timeout_ns() {
  static const uint64_t v = [] {
    const char* e = getenv("SGLB_EP_TIMEOUT_S");
    const double s = e ? atof(e) : 60.0;
    return s <= 0 ? (uint64_t)0 : (uint64_t)(s * 1e9);
  }();
  return v;
}

int fill_peers(void* const* peer_bases, int world, PeerPtrs* p) {
  SGLB_CHECK(world >= 1 && world <= kMaxWorld, "ep_p2p: world size %d not in [1, %d]", world, kMaxWorld);
  SGLB_CHECK(peer_bases != nullptr, "ep_p2p: null peer table");
  for (int r = 0; r < kMaxWorld; ++r) p->base[r] = r < world ? (char*)peer_bases[r] : nullptr;
  for (int r = 0; r < world; ++r) SGLB_CHECK(p->base[r] != nullptr, "ep_p2p: peer %d is not mapped", r);
  return SGLB_OK;
}

}  // namespace
}  // namespace sglb

using namespace sglb;

// ---- peer-mapped regions (CUDA IPC) ------------------------------------------------------------------
extern "C" int sglb_p2p_alloc(size_t bytes, void** ptr, void* ipc_handle_64) {
  SGLB_CHECK(ptr && ipc_handle_64 && bytes > 0, "p2p_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  SGLB_CUDA(cudaMalloc(&p, bytes));
  SGLB_CUDA(cudaMemset(p, 0, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    set_error("p2p_alloc: cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    return SGLB_ERR_CUDA;
  }
  memcpy(ipc_handle_64, &h, sizeof(h));
  *ptr = p;
  return SGLB_OK;
}
extern "C" int sglb_p2p_open(const void* ipc_handle_64, void** peer_ptr) {
  SGLB_CHECK(ipc_handle_64 && peer_ptr, "p2p_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, ipc_handle_64, sizeof(h));
  SGLB_CUDA(cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return SGLB_OK;
}
extern "C" int sglb_p2p_close(void* peer_ptr) {
  if (peer_ptr) SGLB_CUDA(cudaIpcCloseMemHandle(peer_ptr));
  return SGLB_OK;
}
extern "C" int sglb_p2p_free(void* ptr) {
  if (ptr) SGLB_CUDA(cudaFree(ptr));
  return SGLB_OK;
}

// offsets[0..8): recv_x[0], recv_x[1], recv_ids[0], recv_ids[1], recv_tw[0], recv_tw[1], y[0], y[1]; returns bytes
extern "C" size_t sglb_ep_p2p_region_bytes(int world, int cap, int H, int topk, size_t* offsets) {
  if (world < 1 || world > kMaxWorld || cap < 1 || H < 1 || topk < 1) return 0;
  const P2PLayout L = p2p_layout(world, cap, H, topk);
  if (offsets) {
    for (int p = 0; p < 2; ++p) {
      offsets[p] = L.x[p];
      offsets[2 + p] = L.ids[p];
      offsets[4 + p] = L.tw[p];
      offsets[6 + p] = L.y[p];
    }
  }
  return L.total;
}

extern "C" int sglb_ep_p2p_dispatch(const void* hidden, const int32_t* topk_ids, const float* topk_weights, int64_t M,
                                    int topk, int E, int world, int rank, int H, int cap, void* const* peer_bases,
                                    int parity, sglb_stream_t stream) {
  SGLB_CHECK(E % world == 0 && rank >= 0 && rank < world && H % 8 == 0 && topk >= 1, "ep_p2p_dispatch: bad shape");
  SGLB_CHECK(M >= 0 && M <= cap, "ep_p2p_dispatch: %lld tokens exceed the capacity %d", (long long)M, cap);
  SGLB_CHECK(M == 0 || (hidden && topk_ids && topk_weights), "ep_p2p_dispatch: null pointer");
  PeerPtrs peers;
  int rc = fill_peers(peer_bases, world, &peers);
  if (rc) return rc;
  ep_p2p_dispatch_kernel<<<dim3((unsigned)cap, (unsigned)world), 128, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)hidden, topk_ids, topk_weights, (int)M, topk, E / world, world, rank, H, cap, peers, parity & 1);
  SGLB_LAUNCH_CHECK();
  count_launch();
  return SGLB_OK;
}

extern "C" int sglb_ep_p2p_wait_dispatch(const void* local_base, int world, int cap, int H, int topk,
                                         sglb_stream_t stream) {
  SGLB_CHECK(local_base && world >= 1 && world <= kMaxWorld, "ep_p2p_wait_dispatch: bad arguments");
  const P2PLayout L = p2p_layout(world, cap, H, topk);
  ep_p2p_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>((const char*)local_base, L.flags, L.seq, world, timeout_ns());
  SGLB_LAUNCH_CHECK();
  count_launch();
  return SGLB_OK;
}

extern "C" int sglb_ep_p2p_combine(const int32_t* topk_ids, int64_t M, int topk, int E, int world, int rank, int H,
                                   int cap, void* const* peer_bases, int parity, const void* shared_y,
                                   float routed_scaling_factor, void* out, sglb_stream_t stream) {
  SGLB_CHECK(E % world == 0 && rank >= 0 && rank < world && H % 8 == 0 && topk >= 1 && M >= 0 && M <= cap,
             "ep_p2p_combine: bad shape");
  PeerPtrs peers;
  int rc = fill_peers(peer_bases, world, &peers);
  if (rc) return rc;
  // every rank runs this kernel, tokens or not: it tells the peers that this rank's slot rows are ready
  SGLB_CHECK(M == 0 || (topk_ids && out), "ep_p2p_combine: null pointer");
  const unsigned gx = (unsigned)(M > 0 ? M : 1);
  ep_p2p_combine_kernel<<<dim3(gx, (unsigned)ceil_div(ceil_div(H, 8), 256)), 256, 0, (cudaStream_t)stream>>>(
      topk_ids, (int)M, topk, E / world, world, rank, H, cap, peers, parity & 1, timeout_ns(),
      (const __nv_bfloat16*)shared_y, routed_scaling_factor, (__nv_bfloat16*)out);
  SGLB_LAUNCH_CHECK();
  count_launch();
  return SGLB_OK;
}
