// Microbenchmark: dense tcgen05.mma throughput per data kind on sm_100a, the denominators of the tensor-bound
// roofline figures for int8 / fp8 / block-scaled mxf8f6f4 work (SURVEY.md 6 and BASELINE.md 2: "must be measured
// before any TC fraction is claimed for those formats"; MEASURED_PEAKS.json only carries bf16).
//
// One persistent CTA per SM issues back-to-back M128 x N256 MMAs whose operands already sit in shared memory
// (128-B-swizzled K-major tiles of random, finite values; no loads in the loop), alternating between two TMEM
// accumulators, for ~2 s per kind so the figure is the SUSTAINED one under the board power cap -- the same
// regime as MEASURED_PEAKS.json's bf16_tflops_sustained, which this tool re-measures as a cross-check.
//
// Build + run (on the GPU box):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/mma_peak tools/mma_peak.cu && gpurun_out/mma_peak
// Prints one JSON object (profiles/peaks_r2.json).
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc_sw128(uint32_t addr) {
  return (uint64_t)((addr >> 4) & 0x3FFFu) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}

enum { KIND_BF16 = 0, KIND_I8 = 1, KIND_F8 = 2, KIND_MXF8 = 3 };

template <int KIND>
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc, uint32_t sfa, uint32_t sfb) {
  if constexpr (KIND == KIND_BF16)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                 "l"(a), "l"(b), "r"(idesc), "r"(acc)
                 : "memory");
  else if constexpr (KIND == KIND_I8)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                 "l"(a), "l"(b), "r"(idesc), "r"(acc)
                 : "memory");
  else if constexpr (KIND == KIND_F8)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                 "l"(a), "l"(b), "r"(idesc), "r"(acc)
                 : "memory");
  else
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::mxf8f6f4.block_scale [%0], %1, %2, %3, [%5], [%6], p;\n\t}" ::"r"(d),
        "l"(a), "l"(b), "r"(idesc), "r"(acc), "r"(sfa), "r"(sfb)
        : "memory");
}

template <int KIND>
__device__ __forceinline__ uint32_t idesc(int n) {
  const uint32_t mn = ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  if (KIND == KIND_BF16) return (1u << 4) | (1u << 7) | (1u << 10) | mn;   // D f32, A/B bf16
  if (KIND == KIND_I8) return (2u << 4) | (1u << 7) | (1u << 10) | mn;     // D s32, A/B s8
  if (KIND == KIND_F8) return (1u << 4) | mn;                             // D f32, A/B e4m3
  return (0u << 7) | (0u << 10) | mn | (1u << 23);                        // block scaled: A/B e4m3, e8m0 scales
}

constexpr int kABytes = 128 * 128, kBBytes = 256 * 128;

template <int KIND>
__global__ void __launch_bounds__(128, 1) peak_kernel(long long iters, const uint8_t* __restrict__ seed) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) uint64_t bar;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* tiles = smem_raw + (base - smem_u32(smem_raw));
  for (int i = threadIdx.x; i < (kABytes + kBBytes) / 16; i += blockDim.x)
    reinterpret_cast<int4*>(tiles)[i] = reinterpret_cast<const int4*>(seed)[i];
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tm = tmem_slot;
  if (KIND == KIND_MXF8) {  // e8m0 scale columns: 2^0 everywhere (garbage could be NaN scales and change the power draw)
    const uint32_t one = 0x7F7F7F7Fu;
    for (int c = 256; c < 384; ++c)
      asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(tm + ((uint32_t)(warp * 32) << 16) + c), "r"(one) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  if (warp == 0) {
    const uint32_t id = idesc<KIND>(256);
    // two 256-column accumulators fill TMEM; the block-scaled kind needs columns for its scales, so it keeps one
    // accumulator at column 0 (consecutive k-steps of a real GEMM accumulate into one tile as well) and the scales
    // at columns 256..
    const uint32_t idn = id;
    for (long long it = 0; it < iters; ++it) {
      if (elect_one()) {
        const uint32_t acc = tm + (KIND == KIND_MXF8 ? 0u : (uint32_t)(it & 1) * 256u);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          mma<KIND>(acc, desc_sw128(base + j * 32), desc_sw128(base + kABytes + j * 32), idn, j > 0 ? 1u : 0u, tm + 256, tm + 320);
      }
      __syncwarp();
    }
    if (elect_one())
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    __syncwarp();
    uint32_t ok = 0;
    while (!ok)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(ok)
                   : "r"(smem_u32(&bar))
                   : "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
  }
}

template <int KIND>
double run(const char* name, int sms, const uint8_t* seed, double seconds, double* burst) {
  const int smem = kABytes + kBBytes + 1024;
  cudaFuncSetAttribute(peak_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int n = 256;
  const int kelems = KIND == KIND_BF16 ? 16 : 32;                       // per 32-byte k step
  const double ops_per_iter = 4.0 * 2.0 * 128.0 * n * kelems;           // 4 k-steps of one 128-byte stage
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  auto timed = [&](long long iters) {
    cudaEventRecord(a);
    peak_kernel<KIND><<<sms, 128, smem>>>(iters, seed);
    cudaEventRecord(b);
    cudaError_t e = cudaEventSynchronize(b);
    if (e != cudaSuccess) {
      fprintf(stderr, "%s: %s\n", name, cudaGetErrorString(e));
      exit(1);
    }
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    return (double)ms * 1e-3;
  };
  timed(2000);
  const long long probe = 200000;
  const double t_probe = timed(probe);                                   // ~10-50 ms: the burst figure
  *burst = ops_per_iter * probe * sms / t_probe / 1e12;
  const long long iters = (long long)(probe * seconds / t_probe);
  const double t = timed(iters);
  return ops_per_iter * iters * sms / t / 1e12;
}

int main(int argc, char** argv) {
  const double seconds = argc > 1 ? atof(argv[1]) : 2.0;
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount;
  const int nbytes = kABytes + kBBytes;
  uint8_t* h = (uint8_t*)malloc(nbytes);
  uint32_t s = 12345;
  for (int i = 0; i < nbytes; ++i) {
    s = s * 1664525u + 1013904223u;
    h[i] = (uint8_t)(s >> 24) & 0xBF;       // finite e4m3 / bf16 patterns (top exponent bit clear), arbitrary s8
  }
  uint8_t* d;
  cudaMalloc(&d, nbytes);
  cudaMemcpy(d, h, nbytes, cudaMemcpyHostToDevice);
  double b0, b1, b2, b3;
  const double bf16 = run<KIND_BF16>("bf16", sms, d, seconds, &b0);
  const double i8 = run<KIND_I8>("int8", sms, d, seconds, &b1);
  const double f8 = run<KIND_F8>("fp8", sms, d, seconds, &b2);
  const double mx = run<KIND_MXF8>("mxf8f6f4", sms, d, seconds, &b3);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"how\": \"tools/mma_peak.cu: one CTA per SM, back-to-back tcgen05.mma cta_group::1 "
         "M128 x N256 from shared memory, %.1f s per kind (sustained, power-capped) after a ~%d ms burst probe\", "
         "\"bf16_tflops_sustained\": %.1f, \"bf16_tflops_burst\": %.1f, "
         "\"int8_tops_sustained\": %.1f, \"int8_tops_burst\": %.1f, "
         "\"fp8_tops_sustained\": %.1f, \"fp8_tops_burst\": %.1f, "
         "\"mxf8f6f4_tops_sustained\": %.1f, \"mxf8f6f4_tops_burst\": %.1f}\n",
         p.name, sms, seconds, 30, bf16, b0, i8, b1, f8, b2, mx, b3);
  return 0;
}
