// Helpers shared by the CTA-pair (cta_group::2) tcgen05 kernels: cluster addressing, barriers that
// cross the pair, the pair MMA / commit wrappers and the pair work-item decode.
#pragma once
#include "umma_common.cuh"

namespace sglb {
namespace ummak {

constexpr int kPairRows = 256;  // routed rows per pair tile

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// shared::cluster address of `addr` (a shared::cta address of this CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// Remote arrive on a barrier of another CTA of the pair (TMEM hand-back by the epilogue warps: ordered by
// tcgen05.wait::ld + tcgen05.fence).  NOT sufficient for the relay of the peer's cp.async rows -- see
// mbar_arrive_cluster_release below; the text that follows records why the plain form was tried there.
// Deliberately without .release.cluster /
// .acquire.cluster on the waiting side: ptxas turns cluster-scope acquire into CCTL.IVALL (an L1
// invalidate) and release into a full membar, which serialised every pipeline stage (~1.1 us per
// stage, ncu r1f).  Nothing crosses the pair through the generic proxy here: the data the barriers
// guard is read by the tensor core (async proxy, after fence.proxy.async in the writer's CTA) or
// lives in TMEM (ordered by tcgen05.fence), the same discipline CUTLASS' 2-SM kernels use.
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// Same arrive with cluster-scope release semantics: everything the arriving thread has observed (the peer's
// cp.async rows, through the CTA-scope barrier it waited on) is ordered before the arrive for the whole
// cluster; the leader's wait stays a plain (CTA-scope) wait.  Used by the relay warp.  Measured
// (tools/dbg_pair_determinism.py, Qwen3 bf16 / int8, 256 rows per expert): with the plain arrive 1-20 rows per
// 4096 differed from run to run; with this one 12 of 12 runs are bit-identical to the single-CTA kernel -- but
// the membar serialises the relay at ~1 us per stage (bf16 M = 8192: 965 -> 440 TFLOP/s), so the pair kernels are
// explicit-only until the peer's rows arrive through the async proxy (TMA gather4 with the completion counted on
// the leader's barrier).
__device__ __forceinline__ void mbar_arrive_cluster_release(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load whose completion bytes are counted on a barrier of the pair's leader CTA
__device__ __forceinline__ void tma_load_3d_pair(uint32_t dst, const CUtensorMap* map, uint32_t leader_bar,
                                                 int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, "
      "%4, %5}], [%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(leader_bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
// 2-D TMA tile of routed rows (rows [r0, r0 + box rows) x 128 bytes at byte column c0 of a [rows, K] array, 128-B
// swizzle), bytes counted on a barrier of the pair's LEADER CTA like the weights': the routed rows of both CTAs reach
// shared memory through the async proxy and complete on the barrier the MMA issuer waits on -- no generic-proxy
// writes, no relay warp, no cluster-scope fence (umma2_gemm.cu header).
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, uint32_t leader_bar, int c0, int r0) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(leader_bar), "r"(c0), "r"(r0)
      : "memory");
}
// The same two loads with an L2 eviction-priority hint (createpolicy-style constants, as CUTLASS's CacheHintSm90):
// weights are read once per call (evict first), the routed rows of a tile are re-read by every weight block of the
// tile (evict last) -- ncu r2p: 6.5 GB of those re-reads missed L2 in the fp8 prefill kernel.
constexpr uint64_t kL2EvictFirst = 0x12F0000000000000ull, kL2EvictLast = 0x14F0000000000000ull;
__device__ __forceinline__ void tma_load_3d_pair_hint(uint32_t dst, const CUtensorMap* map, uint32_t leader_bar,
                                                      int c0, int c1, int c2, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, "
      "%4, %5}], [%2], %6;" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(leader_bar), "r"(c0), "r"(c1), "r"(c2), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair_hint(uint32_t dst, const CUtensorMap* map, uint32_t leader_bar, int c0,
                                                      int r0, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(
          dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(leader_bar), "r"(c0), "r"(r0), "l"(policy)
      : "memory");
}
template <int FMT>
__device__ __forceinline__ void umma_pair(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  if constexpr (FMT == SGLB_FMT_INT8_W8A8) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// arrives (once all MMAs issued so far have completed) on the barrier at this offset in both CTAs
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
      "h"((uint16_t)3)
      : "memory");
}
// instruction descriptor for the pair MMA: M = 256, N = n
template <int FMT>
__device__ __forceinline__ uint32_t instr_desc_pair(int n) {
  const uint32_t cfmt = FMT == SGLB_FMT_INT8_W8A8 ? 2u : 1u;
  return (cfmt << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ float sel8(const float (&v)[8], int i) {
  float r = v[0];
#pragma unroll
  for (int k = 1; k < 8; ++k) r = i == k ? v[k] : r;
  return r;
}

struct PairItem {
  int e, row_start, rows, nb;
};
__device__ __forceinline__ PairItem get_pair_item(const UmmaArgs& a, int w, int n_blocks) {
  PairItem it;
  const int t = w / n_blocks;
  it.nb = w - t * n_blocks;
  it.e = a.tile_list[3 * t + 0];
  it.row_start = a.tile_list[3 * t + 1];
  it.rows = a.tile_list[3 * t + 2];
  return it;
}


}  // namespace ummak
}  // namespace sglb
