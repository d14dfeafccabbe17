// Microbenchmark: TMEM read (tcgen05.ld) throughput per SM on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/tmem_ld_bw tools/tmem_ld_bw.cu
// Prints bytes/clk/SM for 4 and 8 reader warps and for x16 / x64 / x128 shapes.  Used to check the
// "drain = 2x math" bound stated in DESIGN.md 4a for the fp8 block-scaled kernel.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void ld16(uint32_t a, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(a));
}
__device__ __forceinline__ void ld32(uint32_t a, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(a));
}
__device__ __forceinline__ void ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// MODE 0: x16 then wait each; 1: x32 then wait each; 2: 4 x x32 in flight then one wait
template <int MODE>
__global__ void __launch_bounds__(256) k(int iters, long long* clk, uint32_t* sink) {
  __shared__ uint32_t tmem_base;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(
        (uint32_t)__cvta_generic_to_shared(&tmem_base)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 256;
  uint32_t acc = 0;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int c = 0; c < 256; c += 16) {
        uint32_t v[16];
        ld16(base + c, v);
        ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) acc ^= v[i];
      }
    } else if (MODE == 1) {
#pragma unroll
      for (int c = 0; c < 256; c += 32) {
        uint32_t v[32];
        ld32(base + c, v);
        ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) acc ^= v[i];
      }
    } else {
#pragma unroll
      for (int c = 0; c < 256; c += 64) {
        uint32_t v0[32], v1[32];
        ld32(base + c, v0);
        ld32(base + c + 32, v1);
        ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) acc ^= v0[i] ^ v1[i];
      }
    }
  }
  __syncthreads();
  long long t1 = clock64();
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base));
}

template <int MODE>
void run(int warps, const char* name) {
  long long* clk;
  uint32_t* sink;
  cudaMalloc(&clk, 148 * 8);
  cudaMalloc(&sink, 148 * 256 * 4);
  const int iters = 2000;
  k<MODE><<<148, warps * 32>>>(10, clk, sink);
  k<MODE><<<148, warps * 32>>>(iters, clk, sink);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
  // each warp reads 32 lanes x 256 columns x 4 B per iteration
  double bytes = (double)warps * 32 * 256 * 4 * iters;
  printf("%-28s warps=%d  %s  clk=%lld  bytes/clk/SM=%.1f\n", name, warps, cudaGetErrorString(e), h[0],
         bytes / (double)h[0]);
  cudaFree(clk);
  cudaFree(sink);
}

int main() {
  run<0>(4, "x16 + wait");
  run<0>(8, "x16 + wait");
  run<1>(4, "x32 + wait");
  run<1>(8, "x32 + wait");
  run<2>(4, "2 x x32 in flight");
  run<2>(8, "2 x x32 in flight");
  return 0;
}
