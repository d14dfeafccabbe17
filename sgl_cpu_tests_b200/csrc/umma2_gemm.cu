// K3/K4 (prefill regime), CTA-pair version: grouped GEMM with tcgen05.mma.cta_group::2.
//
// Why: ncu on the single-CTA kernel (profiles/README.md, r1e) shows the MMA issuer blocked on the
// tensor pipe's operand fetch, not on math: with a 128 x 128 tile both operands come from shared
// memory, 8 KB per 64-cycle MMA = 128 B/clk, which is all the shared-memory bandwidth an SM has --
// and the same memory is absorbing the TMA and cp.async writes that refill the ring.  A CTA pair
// issuing one M = 256 MMA halves that: each SM feeds its own 128 weight rows (A) and only its half
// of the routed rows (B); the hardware shares the B halves between the two tensor cores.
//
//   pair tile   = up to 256 routed rows of one expert  x  256 weight rows
//   CTA r       = weight rows half r (A, 128 rows)  +  routed rows half r (B, n_eff / 2 rows)
//   D (per CTA) = its 128 weight rows x all n_eff routed rows, f32/s32 in TMEM, double-buffered
//
// Stage 1 keeps silu_and_mul fused by giving every CTA 64 gate rows and the matching 64 up rows
// (two 64-row TMA boxes into one 128-row operand tile); the epilogue warps that hold `up` hand
// their values to the warps that hold `gate` through shared memory.
//
// Routed rows come through the TMA unit as well: the caller hands both stages rows that are ALREADY in
// expert-sorted order (stage 1: capi.cu gathers the token rows once per call, gather_rows_kernel; stage 2: act is
// produced in that order), so each CTA's half of a tile is one 2-D TMA box whose bytes are counted on the LEADER's
// full barrier like the weights'.  History: the first version gathered with cp.async and relayed "my half is in
// place" to the leader with a remote mbarrier arrive; rows written through the generic proxy in one CTA and read by
// a tensor core driven from the other need a cluster-scope release on that arrive (without it 1-20 rows per 4096
// differed from run to run; with it the membar cost ~1 us per stage).  TMA tile::gather4 (four arbitrary rows per
// instruction, tools/gather4_probe.cu) removes the proxy problem but 32-64 instructions of 4 x 128 B per stage
// run at half the speed of the single-CTA kernel (tools/pair_check.py, r2i).  One box per stage does not.
//
// Barriers (s = smem stage, b = TMEM buffer; "L" lives in the leader CTA only):
//   full[s]  L   weights and routed rows of both CTAs landed (TMA cta_group::2 complete_tx on the leader's barrier)
//   empty[s]     tcgen05.commit multicast to both CTAs: stage may be refilled
//   tfull[b]     tcgen05.commit multicast: accumulators complete
//   tempty[b] L  one arrive per epilogue warp of both CTAs: buffer drained
#include "umma_pair.cuh"

namespace sglb {

using namespace ummak;

namespace {

constexpr int kStages2 = 5;
constexpr int kStageBytes2 = kWTileBytes + kXTileBytes;       // 32 KB: A half (128 rows) + B half (<= 128 rows)
constexpr int kFirstEpiWarp2 = 4;                             // warp 0: TMA (weights + routed rows), 1: MMA issuer, 2-3 idle
constexpr int kThreads2 = (kFirstEpiWarp2 + kEpiWarps) * 32;  // 448
constexpr int kXchgBytes = 4 * 2 * 2048;                      // 4 warp pairs x 2 buffers x (16 tokens x 32 lanes f32)
constexpr int kSmemBytes2 = kStages2 * kStageBytes2 + kEpiWarps * kEpiBufBytes + kXchgBytes + 1024 + 256;

template <int FMT, int STAGE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads2, 1)
    umma2_moe_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
                     const UmmaArgs a) {
  extern __shared__ uint8_t smem_raw[];
  constexpr bool kInt8 = FMT == SGLB_FMT_INT8_W8A8;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t epi_base = smem_base + kStages2 * kStageBytes2;
  const uint32_t xchg_base = epi_base + kEpiWarps * kEpiBufBytes;
  const uint32_t bar_base = xchg_base + kXchgBytes;
  auto w_tile = [&](int s) { return smem_base + s * kStageBytes2; };
  auto x_tile = [&](int s) { return smem_base + s * kStageBytes2 + kWTileBytes; };
  auto full_bar = [&](int s) { return bar_base + 8 * s; };
  auto empty_bar = [&](int s) { return bar_base + 8 * (2 * kStages2 + s); };
  auto tfull_bar = [&](int b) { return bar_base + 8 * (3 * kStages2 + b); };
  auto tempty_bar = [&](int b) { return bar_base + 8 * (3 * kStages2 + 2 + b); };
  const uint32_t tmem_slot = bar_base + 8 * (3 * kStages2 + 4);
  static_assert(8 * (3 * kStages2 + 4) + 4 <= 256, "barrier region overflow");
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();  // 0 = leader (issues the MMAs), 1 = peer
  constexpr int EB = kInt8 ? 1 : 2;
  const int k_bytes = a.K * EB;
  const int num_kb = k_bytes / kRowBytes;
  const int n_blocks = num_weight_blocks<STAGE>(a);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_x)) : "memory");
    for (int s = 0; s < kStages2; ++s) {
      mbar_init(full_bar(s), 1);                 // the leader's expect_tx arrive; everything else is TMA bytes
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tfull_bar(b), 1);
      mbar_init(tempty_bar(b), 2 * kEpiWarps);   // one arrive per epilogue warp of both CTAs
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                 "n"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();  // barriers initialised and TMEM allocated in both CTAs
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  pdl_wait();
  pdl_trigger();
  const int total = a.counters[2] * n_blocks;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

  if (warp == 0) {
    // ================= TMA producer: this CTA's 128 weight rows =================
    uint32_t it = 0;
    for (int w = cluster_id; w < total; w += num_clusters) {
      const PairItem wi = get_pair_item(a, w, n_blocks);
      const int n_eff = (wi.rows + 15) & ~15;
      for (int kb = 0; kb < num_kb; ++kb, ++it) {
        const int s = it % kStages2;
        mbar_wait(empty_bar(s), ((it / kStages2) & 1) ^ 1);
        if (elect_one()) {
          const uint32_t lbar = mapa(full_bar(s), 0);
          // both CTAs' weight halves + both CTAs' routed-row boxes (whole 128-row boxes: rows past the tile's half
          // belong to the next tile or are zero-filled past the array, and are never multiplied: N = n_eff)
          if (rank == 0) mbar_arrive_expect_tx(full_bar(s), 2 * kWTileBytes + 2 * kXTileBytes);
          tma_load_2d_pair(x_tile(s), &tmap_x, lbar, kb * kRowBytes, wi.row_start + (int)rank * (n_eff >> 1));
          if constexpr (STAGE == 1) {  // 64 gate rows + the matching 64 up rows
            const int g0 = wi.nb * kWSub + (int)rank * 64;
            tma_load_3d_pair(w_tile(s), &tmap_w, lbar, kb * kRowBytes, g0, wi.e);
            tma_load_3d_pair(w_tile(s) + kWTileBytes / 2, &tmap_w, lbar, kb * kRowBytes, a.N + g0, wi.e);
          } else {
            tma_load_3d_pair(w_tile(s), &tmap_w, lbar, kb * kRowBytes, wi.nb * 2 * kWSub + (int)rank * kWSub, wi.e);
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    if (rank == 0) {
      // ================= MMA issuer (leader CTA) =================
      uint32_t it = 0, li = 0;
      for (int w = cluster_id; w < total; w += num_clusters, ++li) {
        const PairItem wi = get_pair_item(a, w, n_blocks);
        const int n_eff = (wi.rows + 15) & ~15;
        const uint32_t idesc = instr_desc_pair<FMT>(n_eff);
        const int buf = li & 1;
        mbar_wait(tempty_bar(buf), ((li >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t acc = tmem_base + buf * kAccCols;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % kStages2;
          mbar_wait(full_bar(s), (it / kStages2) & 1);
          tc_fence_after();
          if (elect_one()) {
#pragma unroll
            for (int j = 0; j < kRowBytes / 32; ++j) {
              const uint64_t ad = smem_desc_sw128(w_tile(s) + j * 32);
              const uint64_t bd = smem_desc_sw128(x_tile(s) + j * 32);
              umma_pair<FMT>(acc, ad, bd, idesc, (kb > 0 || j > 0) ? 1u : 0u);
            }
            umma_commit_pair(empty_bar(s));
            if (kb == num_kb - 1) umma_commit_pair(tfull_bar(buf));
          }
          __syncwarp();
        }
      }
    }
  } else if (warp < kFirstEpiWarp2) {
    // idle
  } else {
    // ================= epilogue =================
    // warp (q, par): TMEM lanes [32q, 32q+32) of this CTA, 16-column (= 16-token) chunks of parity par.
    const int ew = warp - kFirstEpiWarp2;
    const int q = warp & 3, par = ew >> 2;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    constexpr int OB = (kInt8 && STAGE == 1) ? 4 : 2;  // output element bytes (f32 act for int8 stage 1)
    constexpr int ROWB = 32 * OB, SEGS = ROWB / 16, TPI = 32 / SEGS, NRD = 16 / TPI;
    uint8_t* ebuf = smem_raw + (epi_base + ew * kEpiBufBytes - smem_u32(smem_raw));
    const int rd_tok = lane / SEGS, rd_seg = lane % SEGS;
    // stage 1: lanes 0..63 hold gate, 64..127 hold up of the same 64 outputs; warp (q, par) with q >= 2
    // passes its up values to warp (q - 2, par) through xbuf (two buffers, one named barrier per chunk)
    const bool is_up = STAGE == 1 && q >= 2;
    const int pair_idx = (q & 1) + 2 * par;
    float* xbuf = reinterpret_cast<float*>(smem_raw + (xchg_base + pair_idx * 4096 - smem_u32(smem_raw)));
    const uint32_t tempty_l0 = mapa(tempty_bar(0), 0), tempty_l1 = mapa(tempty_bar(1), 0);
    uint32_t li = 0;
    // Exchange-buffer parity runs over ALL chunks this warp pair processes, not per item: with a per-item counter a
    // ragged tile with an odd number of chunks is followed by a chunk in the SAME buffer, and the `up` warp -- which
    // has nothing left to do for the item once it has passed the last barrier, and whose next accumulators are
    // usually ready -- overwrote values the `gate` warp was still reading (the run-to-run differences that round 1
    // blamed on the relay alone; tools/pair_check.py).
    uint32_t ci = 0;
    for (int w = cluster_id; w < total; w += num_clusters, ++li) {
      const PairItem wi = get_pair_item(a, w, n_blocks);
      const int n_eff = (wi.rows + 15) & ~15;
      const int buf = li & 1;
      // per-token scalars: lane l keeps tokens l + 32 i
      float xs_r[8], rs_r[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        xs_r[i] = 1.f;
        rs_r[i] = 1.f;
        const int p = wi.row_start + min(lane + 32 * i, wi.rows - 1);
        if (32 * i < n_eff) {
          if constexpr (STAGE == 1) {
            if (kInt8) xs_r[i] = a.Xscale[a.sorted_slot[p] / a.topk];
          } else {
            if (kInt8) xs_r[i] = a.Xscale[p];
            if (a.row_scale != nullptr) rs_r[i] = a.row_scale[a.sorted_slot[p]];
          }
        }
      }
      // output column of this lane and its weight scale
      int nbase;
      if constexpr (STAGE == 1) nbase = wi.nb * kWSub + (int)rank * 64 + (q & 1) * 32;
      else nbase = wi.nb * 2 * kWSub + (int)rank * kWSub + q * 32;
      const int n = nbase + lane;
      float wsc = 1.f;
      if (kInt8 && n < a.N) {
        const float* ws = (const float*)a.Wscale + (size_t)wi.e * a.rows_per_expert;
        wsc = (STAGE == 1 && is_up) ? ws[a.N + n] : ws[n];
      }
      mbar_wait(tfull_bar(buf), (li >> 1) & 1);
      tc_fence_after();
      const uint32_t acc = tmem_base + lane_off + buf * kAccCols;
#pragma unroll 1
      for (int c0 = 16 * par; c0 < n_eff; c0 += 32, ++ci) {  // n_eff is uniform over the CTA
        const float xsel = sel8(xs_r, c0 >> 5), rsel = sel8(rs_r, c0 >> 5);
        uint32_t v[16];
        tmem_ld16(acc + c0, v);
        tmem_ld_wait(v);
        float val[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          if constexpr (kInt8) {
            const float xs = __shfl_sync(0xffffffffu, xsel, 16 * par + j);
            val[j] = xs * (float)(int)v[j] * wsc;
          } else {
            val[j] = __uint_as_float(v[j]);
          }
        }
        if constexpr (STAGE == 1) {
          float* xb = xbuf + (ci & 1) * 512;
          if (is_up) {
#pragma unroll
            for (int j = 0; j < 16; ++j) xb[j * 32 + lane] = val[j];
            named_bar_sync(1 + pair_idx, 64);
            continue;
          }
          named_bar_sync(1 + pair_idx, 64);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float r = silu_f(val[j]) * xb[j * 32 + lane];
            if constexpr (kInt8) *reinterpret_cast<float*>(ebuf + j * ROWB + lane * 4) = r;
            else *reinterpret_cast<__nv_bfloat16*>(ebuf + j * ROWB + lane * 2) = __float2bfloat16_rn(r);
          }
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float rs = __shfl_sync(0xffffffffu, rsel, 16 * par + j);
            *reinterpret_cast<__nv_bfloat16*>(ebuf + j * ROWB + lane * 2) = __float2bfloat16_rn(val[j] * rs);
          }
        }
        __syncwarp();
        const int nn = nbase + rd_seg * (16 / OB);
#pragma unroll
        for (int i = 0; i < NRD; ++i) {
          const int tok = rd_tok + TPI * i;
          const int4 piece = *reinterpret_cast<const int4*>(ebuf + tok * ROWB + rd_seg * 16);
          if (c0 + tok < wi.rows && nn < a.N)
            *reinterpret_cast<int4*>((char*)a.out + ((size_t)(wi.row_start + c0 + tok) * a.N + nn) * OB) = piece;
        }
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(buf ? tempty_l1 : tempty_l0);
    }
  }

  tc_fence_before();
  cluster_sync_all();  // the peer's smem, barriers and TMEM stay valid until both CTAs are done
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols)
                 : "memory");
  }
}

template <int FMT, int STAGE>
int launch_pair(const UmmaArgs& a, cudaStream_t stream) {
  constexpr int EB = FMT == SGLB_FMT_INT8_W8A8 ? 1 : 2;
  CUtensorMap map;
  int rc = weight_tensor_map(a.W, (uint64_t)a.K * EB, (uint64_t)a.rows_per_expert, (uint64_t)a.E, kRowBytes,
                             STAGE == 1 ? 64 : kWSub, 1, &map);
  if (rc) return rc;
  CUtensorMap map_x;
  SGLB_CHECK(a.gather == 0, "CTA-pair tcgen05 path: routed rows must be in expert-sorted order (gather = 0)");
  rc = rows_tensor_map(a.X, (uint64_t)a.K * EB, (uint64_t)a.x_rows, kUmmaTileRows, &map_x);
  if (rc) return rc;
  static bool attr_set = false;  // per instantiation
  if (!attr_set) {
    SGLB_CUDA(cudaFuncSetAttribute(umma2_moe_kernel<FMT, STAGE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   kSmemBytes2));
    attr_set = true;
  }
  const int n_blocks = STAGE == 1 ? ceil_div(a.N, kWSub) : ceil_div(a.N, 2 * kWSub);
  const int64_t max_items = (int64_t)a.max_tiles * n_blocks;
  const int max_clusters = sm_count() / 2;
  const int clusters = (int)(max_items < max_clusters ? (max_items > 0 ? max_items : 1) : max_clusters);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * clusters);
  cfg.blockDim = dim3(kThreads2);
  cfg.dynamicSmemBytes = kSmemBytes2;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  SGLB_CUDA(cudaLaunchKernelEx(&cfg, umma2_moe_kernel<FMT, STAGE>, map, map_x, a));
  return SGLB_OK;
}

}  // namespace

int umma2_tile_rows() { return kPairRows; }

int launch_umma2_gemm(int fmt, int stage, const UmmaArgs& a, cudaStream_t stream) {
  int rc;
  if (fmt == SGLB_FMT_BF16)
    rc = stage == 1 ? launch_pair<SGLB_FMT_BF16, 1>(a, stream) : launch_pair<SGLB_FMT_BF16, 2>(a, stream);
  else if (fmt == SGLB_FMT_INT8_W8A8)
    rc = stage == 1 ? launch_pair<SGLB_FMT_INT8_W8A8, 1>(a, stream) : launch_pair<SGLB_FMT_INT8_W8A8, 2>(a, stream);
  else {
    set_error("CTA-pair tcgen05 path: unsupported weight format %d", fmt);
    return SGLB_ERR_INVALID;
  }
  if (rc == SGLB_OK) count_launch();
  return rc;
}

}  // namespace sglb
