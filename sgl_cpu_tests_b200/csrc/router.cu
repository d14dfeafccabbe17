// K1: grouped top-k router, one warp per token.
//
// Replaces the reference ops grouped_topk_cpu (test_grouped_topk.py:61-69, oracle :20-39) and
// biased_grouped_topk_cpu (test_biased_grouped_topk.py:71-80, oracle :20-47).
//
// Tie-break rule (the reference leaves it open): among equal scores the lowest
// index wins; picks are emitted in descending score order.  oracle/moe_oracle.py states the
// same rule and the tests hold this kernel bit-exact on ids to it.
//
// Sigmoid contract (biased variant): sigmoid_f64() below -- evaluated in f64 by a FIXED SEQUENCE OF IEEE-754
// OPERATIONS (Cody-Waite reduction, degree-13 Taylor polynomial in Horner form, separate multiplies and adds, one
// division) and rounded once to f32.  Every step is a correctly rounded basic operation, so the CPU oracle
// (oracle/moe_oracle.py:sigmoid_f64, the same sequence with torch f64 tensor ops) reproduces it bit for bit: ids,
// their order AND the weights of the biased router are bit-exact.  Two things that do not work: the reference's own
// expression, torch's f32 sigmoid, is a ~1-ulp approximation (it differs from the correctly rounded value on ~30 % of
// inputs) that no second implementation can reproduce; and exp() of two math libraries agrees only to an ulp of
// f64, which is enough to flip the f32 rounding where sigmoid(x) = 0.5 + x/4 - ... sits on a half-ulp of f32
// (bf16 x = 9.894e-6 did, on the B200 box).
//
// Serving arguments of the 9-argument call (test_moe.py:61-70; semantics from sglang's python router, not
// pinned by the reference): num_fused_shared_experts n > 0 makes the LAST of the topk slots the fused shared
// expert: id = E + (token % n), weight = (sum of the routed weights) / routed_scaling_factor, and the
// renormalisation divides by the routed sum only; num_token_non_padded (device scalar) masks rows at and
// past it: ids = -1, weights = 0.  routed_scaling_factor is used for that slot only -- the op never
// scales the routed weights (sglang asserts apply_routed_scaling_factor_on_output == False on this path;
// shared_expert_cpu applies the factor).
#include "common.cuh"

namespace sglb {

constexpr int kMaxPerLane = 32;  // E <= 1024

struct ValIdx {
  float v;
  int i;
};

__device__ __forceinline__ bool better(float v, int i, float bv, int bi) {
  // NaNs never win; equal values -> lower index.
  return (v > bv) || (v == bv && i < bi);
}

__device__ __forceinline__ ValIdx warp_argmax(float v, int i) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    float ov = __shfl_xor_sync(0xffffffffu, v, off);
    int oi = __shfl_xor_sync(0xffffffffu, i, off);
    if (better(ov, oi, v, i)) {
      v = ov;
      i = oi;
    }
  }
  return {v, i};
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, off));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}

// 1 / (1 + exp(-x)) in f64, every operation an explicitly rounded IEEE-754 one (no contraction into fma), then one
// rounding to f32.  exp(t) = 2^k * P(r), k = rint(t / ln 2), r = t - k ln2_hi - k ln2_lo, P = sum_{n<=13} r^n / n!
// (|r| <= 0.347: truncation 4e-18).  t is clamped to [-50, 120]: below, 1 + e = 1 exactly; above, the quotient is
// far under the smallest f32 subnormal.  oracle/moe_oracle.py:sigmoid_f64 is the same sequence.
__device__ __forceinline__ float sigmoid_f64(float x) {
  double t = -(double)x;
  if (!(t == t)) return x;  // NaN stays NaN (never wins a comparison in the selection below)
  t = t < -50.0 ? -50.0 : (t > 120.0 ? 120.0 : t);
  const double k = rint(__dmul_rn(t, 0x1.71547652b82fep+0));
  const double r = __dsub_rn(__dsub_rn(t, __dmul_rn(k, 0x1.62e42fee00000p-1)), __dmul_rn(k, 0x1.a39ef35793c76p-33));
  const double c[14] = {0x1.0000000000000p+0,  0x1.0000000000000p+0,  0x1.0000000000000p-1,  0x1.5555555555555p-3,
                        0x1.5555555555555p-5,  0x1.1111111111111p-7,  0x1.6c16c16c16c17p-10, 0x1.a01a01a01a01ap-13,
                        0x1.a01a01a01a01ap-16, 0x1.71de3a556c734p-19, 0x1.27e4fb7789f5cp-22, 0x1.ae64567f544e4p-26,
                        0x1.1eed8eff8d898p-29, 0x1.6124613a86d09p-33};
  double p = c[13];
#pragma unroll
  for (int n = 12; n >= 0; --n) p = __dadd_rn(__dmul_rn(p, r), c[n]);
  const double e = __dmul_rn(p, __longlong_as_double(((long long)k + 1023ll) << 52));
  return (float)__ddiv_rn(1.0, __dadd_rn(1.0, e));
}

// PER = ceil(E/32) rounded up to a compiled size.  Lane l owns experts l, l+32, ...
template <int PER, typename TG, typename TB, bool BIASED>
__global__ void __launch_bounds__(256) router_kernel(const TG* __restrict__ gating,
                                                    const TB* __restrict__ bias, int64_t M, int E,
                                                    int topk, int G, int topk_group, int renorm,
                                                    int n_shared, float routed_scale,
                                                    const void* __restrict__ non_padded, int non_padded_i64,
                                                    float* __restrict__ out_w,
                                                    int32_t* __restrict__ out_id, int wpt) {
  // wpt = warps per token.  The f64 sigmoid is ~120 f64 instructions and the f64 pipe issues one warp instruction in
  // ~35 clk per scheduler (measured: one warp with 8 sigmoids per lane = 18 us, whatever M), so at decode sizes a
  // token's experts are spread over wpt warps of the block (each lane evaluates PER / wpt sigmoids, values meet in
  // shared memory) and warp 0 of the token runs the selection on exactly the values the one-warp form computes.
  extern __shared__ float sig_sm[];  // [tokens per block][E], wpt > 1 only
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool multi = BIASED && wpt > 1;
  const int tpb = multi ? (int)(blockDim.x >> 5) / wpt : (int)(blockDim.x >> 5);
  const int slot = multi ? warp / wpt : warp, sub = multi ? warp % wpt : 0;
  const int64_t tok = (int64_t)blockIdx.x * tpb + slot;
  bool padded = false;
  if (tok < M && non_padded != nullptr) {
    const int64_t n_valid = non_padded_i64 ? *reinterpret_cast<const int64_t*>(non_padded)
                                           : (int64_t) * reinterpret_cast<const int32_t*>(non_padded);
    padded = tok >= n_valid;
  }
  const float NEG = -INFINITY;
  float s[PER];  // weight source: softmax prob or sigmoid
  float c[PER];  // choice score
  if (tok < M && !padded) {
    const TG* row = gating + tok * E;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
      int e = lane + 32 * k;
      s[k] = (e < E) ? load_as_float<TG>(row + e) : NEG;
    }
  }
  if (multi) {
    if (tok < M && !padded) {
      const int per_w = PER / wpt;  // host guarantees PER % wpt == 0
#pragma unroll
      for (int k = 0; k < PER; ++k) {
        const int e = lane + 32 * k;
        if (k / per_w == sub && e < E) sig_sm[slot * E + e] = sigmoid_f64(s[k]);
      }
    }
    __syncthreads();
    if (sub != 0) return;
  }
  if (tok >= M) return;
  if (padded) {  // padded row: every slot masked
    if (lane < topk) {
      out_w[tok * topk + lane] = 0.f;
      out_id[tok * topk + lane] = -1;
    }
    return;
  }
  if (!BIASED) {
    float mx = NEG;
#pragma unroll
    for (int k = 0; k < PER; ++k) mx = fmaxf(mx, s[k]);
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
      int e = lane + 32 * k;
      s[k] = (e < E) ? expf(s[k] - mx) : 0.f;
      sum += s[k];
    }
    sum = warp_sum(sum);
#pragma unroll
    for (int k = 0; k < PER; ++k) {
      s[k] = s[k] / sum;
      c[k] = s[k];
    }
  } else {
#pragma unroll
    for (int k = 0; k < PER; ++k) {
      int e = lane + 32 * k;
      if (e < E) {
        s[k] = multi ? sig_sm[slot * E + e] : sigmoid_f64(s[k]);  // bit-reproducible on the CPU: see the file header
        c[k] = s[k] + load_as_float<TB>(bias + e);
      } else {
        s[k] = 0.f;
        c[k] = NEG;
      }
    }
  }

  // ---- group scores: lane g ends up holding the score of group g (G <= 32) ----
  const int gs = E / G;
  // group of each expert this lane owns, computed once: an integer division by a run-time value is ~25 dependent
  // instructions, and the kernel is one long dependency chain per warp (ncu r2x: 970 warp instructions at 16 clk each)
  int grp[PER];
#pragma unroll
  for (int k = 0; k < PER; ++k) grp[k] = (lane + 32 * k) / gs;
  uint32_t group_mask = 0xffffffffu;
  if (G > 1 && topk_group < G) {
    float my_group_score = NEG;
#pragma unroll 4  // the butterflies of different groups are independent: let them overlap
    for (int g = 0; g < G; ++g) {
      float m1 = NEG, m2 = NEG;  // top-2 of the group (biased) / max (softmax)
#pragma unroll
      for (int k = 0; k < PER; ++k) {
        int e = lane + 32 * k;
        if (e < E && grp[k] == g) {
          float v = c[k];
          if (v > m1) {
            m2 = m1;
            m1 = v;
          } else if (v > m2) {
            m2 = v;
          }
        }
      }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
        float o1 = __shfl_xor_sync(0xffffffffu, m1, off);
        float o2 = __shfl_xor_sync(0xffffffffu, m2, off);
        float n1 = fmaxf(m1, o1);
        float n2 = fmaxf(fminf(m1, o1), fmaxf(m2, o2));
        m1 = n1;
        m2 = n2;
      }
      float score = (BIASED && gs >= 2) ? (m1 + m2) : m1;
      if (lane == g) my_group_score = score;
    }
    group_mask = 0u;
    float v = (lane < G) ? my_group_score : NEG;
    int idx = (lane < G) ? lane : 0x7fffffff;
    for (int r = 0; r < topk_group; ++r) {
      ValIdx b = warp_argmax(v, idx);
      group_mask |= (1u << b.i);
      if (lane == b.i) {
        v = NEG;
        idx = 0x7fffffff;  // taken: can no longer win a tie between -inf entries
      }
    }
  }

  // ---- mask: softmax variant fills 0.0 (test_grouped_topk.py:33), biased fills -inf (:41-43) ----
  bool taken[PER];
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    int e = lane + 32 * k;
    taken[k] = !(e < E);
    if (e < E && !((group_mask >> grp[k]) & 1u)) c[k] = BIASED ? NEG : 0.0f;
  }

  // ---- top-k by repeated warp argmax ----
  float wsum = 0.f;
  float my_w = 0.f;
  int my_id = 0;
  const int routed = n_shared > 0 ? topk - 1 : topk;
  for (int r = 0; r < routed; ++r) {
    float bv = NEG;
    int bi = 0x7fffffff;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
      int e = lane + 32 * k;
      if (!taken[k] && (bi == 0x7fffffff || better(c[k], e, bv, bi))) {
        bv = c[k];
        bi = e;
      }
    }
    ValIdx b = warp_argmax(bv, bi);
    // weight of the winner lives in its owner lane
    float w = 0.f;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
      int e = lane + 32 * k;
      if (e == b.i) {
        taken[k] = true;
        w = BIASED ? s[k] : c[k];
      }
    }
    w = __shfl_sync(0xffffffffu, w, b.i & 31);
    wsum += w;
    if (lane == r) {
      my_w = w;
      my_id = b.i;
    }
  }
  if (lane < topk) {
    float w = my_w;
    if (n_shared > 0 && lane == topk - 1) {
      w = wsum / routed_scale;
      my_id = E + (int)(tok % n_shared);
    }
    if (renorm) w = w / wsum;
    out_w[tok * topk + lane] = w;
    out_id[tok * topk + lane] = my_id;
  }
}

template <typename TG, typename TB, bool BIASED>
static int launch_router(const void* gating, const void* bias, int64_t M, int E, int topk, int G,
                         int topk_group, int renorm, int n_shared, float routed_scale,
                         const void* non_padded, int non_padded_i64, float* out_w,
                         int32_t* out_id, cudaStream_t stream) {
  const int per = ceil_div(E, 32);
  const int per_c = per <= 1 ? 1 : (per <= 2 ? 2 : (per <= 4 ? 4 : (per <= 8 ? 8 : (per <= 16 ? 16 : 32))));  // compiled PER
  // biased router at decode sizes: up to 8 warps per token (see the kernel), as long as ~2048 warps are not exceeded
  int wpt = 1;
  if (BIASED) {
    wpt = 8;
    while (wpt > 1 && (M * wpt > 2048 || per_c % wpt != 0)) wpt /= 2;
  }
  const int warps = wpt > 1 ? 8 : 4;
  const int tpb = warps / wpt;
  dim3 grid((unsigned)ceil_div64(M, tpb)), block(warps * 32);
  const size_t smem = wpt > 1 ? sizeof(float) * tpb * E : 0;
#define SGLB_ROUTER_CASE(P)                                                                      \
  router_kernel<P, TG, TB, BIASED><<<grid, block, smem, stream>>>(                               \
      (const TG*)gating, (const TB*)bias, M, E, topk, G, topk_group, renorm, n_shared, routed_scale, \
      non_padded, non_padded_i64, out_w, out_id, wpt)
  if (per <= 1) SGLB_ROUTER_CASE(1);
  else if (per <= 2) SGLB_ROUTER_CASE(2);
  else if (per <= 4) SGLB_ROUTER_CASE(4);
  else if (per <= 8) SGLB_ROUTER_CASE(8);
  else if (per <= 16) SGLB_ROUTER_CASE(16);
  else SGLB_ROUTER_CASE(32);
#undef SGLB_ROUTER_CASE
  SGLB_LAUNCH_CHECK();
  count_launch();
  return SGLB_OK;
}

template <typename TG, bool BIASED>
static int dispatch_bias(int bias_dtype, const void* gating, const void* bias, int64_t M, int E,
                         int topk, int G, int topk_group, int renorm, int n_shared, float routed_scale,
                         const void* non_padded, int non_padded_i64, float* out_w, int32_t* out_id,
                         cudaStream_t stream) {
  switch (bias_dtype) {
    case SGLB_DT_BF16:
      return launch_router<TG, __nv_bfloat16, BIASED>(gating, bias, M, E, topk, G, topk_group, renorm, n_shared,
                                                      routed_scale, non_padded, non_padded_i64, out_w, out_id, stream);
    case SGLB_DT_F32:
      return launch_router<TG, float, BIASED>(gating, bias, M, E, topk, G, topk_group, renorm, n_shared,
                                              routed_scale, non_padded, non_padded_i64, out_w, out_id, stream);
    case SGLB_DT_F16:
      return launch_router<TG, __half, BIASED>(gating, bias, M, E, topk, G, topk_group, renorm, n_shared,
                                               routed_scale, non_padded, non_padded_i64, out_w, out_id, stream);
  }
  set_error("router: unsupported bias dtype %d", bias_dtype);
  return SGLB_ERR_INVALID;
}

}  // namespace sglb

using namespace sglb;

extern "C" int sglb_grouped_topk(const void* gating, int gating_dtype, const void* bias,
                                 int bias_dtype, int64_t M, int E, int topk, int num_expert_group,
                                 int topk_group, int renormalize, int num_fused_shared_experts,
                                 float routed_scaling_factor, const void* num_token_non_padded,
                                 int non_padded_is_i64, float* topk_weights, int32_t* topk_ids,
                                 sglb_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  SGLB_CHECK(M >= 0 && E > 0, "router: bad shape M=%lld E=%d", (long long)M, E);
  SGLB_CHECK(E <= 32 * kMaxPerLane, "router: E=%d exceeds %d", E, 32 * kMaxPerLane);
  const int n_shared = num_fused_shared_experts;
  SGLB_CHECK(n_shared >= 0, "router: num_fused_shared_experts=%d must be >= 0", n_shared);
  SGLB_CHECK(topk >= 1 && topk <= 32 && topk - (n_shared > 0) <= E && topk - (n_shared > 0) >= 1,
             "router: topk=%d must be in [1, min(32,E)] (one more slot with fused shared experts)", topk);
  SGLB_CHECK(n_shared == 0 || routed_scaling_factor != 0.f, "router: routed_scaling_factor must not be 0");
  if (num_expert_group <= 0) num_expert_group = 1;
  if (topk_group <= 0) topk_group = num_expert_group;
  SGLB_CHECK(num_expert_group <= 32 && E % num_expert_group == 0,
             "router: num_expert_group=%d must divide E=%d and be <= 32", num_expert_group, E);
  SGLB_CHECK(topk_group <= num_expert_group, "router: topk_group=%d > num_expert_group=%d",
             topk_group, num_expert_group);
  SGLB_CHECK(gating && topk_weights && topk_ids, "router: null pointer");
  if (M == 0) return SGLB_OK;
  const bool biased = bias != nullptr;
#define SGLB_R(TG)                                                                                 \
  (biased ? dispatch_bias<TG, true>(bias_dtype, gating, bias, M, E, topk, num_expert_group,        \
                                    topk_group, renormalize, n_shared, routed_scaling_factor,      \
                                    num_token_non_padded, non_padded_is_i64, topk_weights,         \
                                    topk_ids, stream)                                              \
          : launch_router<TG, float, false>(gating, nullptr, M, E, topk, num_expert_group,         \
                                            topk_group, renormalize, n_shared,                     \
                                            routed_scaling_factor, num_token_non_padded,           \
                                            non_padded_is_i64, topk_weights, topk_ids, stream))
  switch (gating_dtype) {
    case SGLB_DT_BF16: return SGLB_R(__nv_bfloat16);
    case SGLB_DT_F32: return SGLB_R(float);
    case SGLB_DT_F16: return SGLB_R(__half);
  }
#undef SGLB_R
  set_error("router: unsupported gating dtype %d", gating_dtype);
  return SGLB_ERR_INVALID;
}
