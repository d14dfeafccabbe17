// Probe: TMA tile::gather4 (cp.async.bulk.tensor.2d ... tile::gather4) into a 128-B-swizzled K-major tile.
// Question answered: which box shape the descriptor needs and whether 32 gather4 pieces of 4 rows x 128 B written at
// 512-B steps form exactly the [128 rows x 128 B] SWIZZLE_128B tile the UMMA descriptors of umma_common.cuh read
// (16-B chunk c of row r at r * 128 + ((c ^ (r & 7)) << 4)).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o gpurun_out/gather4_probe tools/gather4_probe.cu && gpurun_out/gather4_probe
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128) probe(const __grid_constant__ CUtensorMap map, const int* __restrict__ idx, int kb,
                                             uint8_t* __restrict__ out, int rows) {
  extern __shared__ uint8_t raw[];
  __shared__ __align__(8) uint64_t bar;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* tile = raw + (base - smem_u32(raw));
  for (int i = threadIdx.x; i < 128 * 128; i += blockDim.x) tile[i] = 0xEE;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0)
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(rows * 128) : "memory");
  __syncthreads();
  if (threadIdx.x < rows / 4) {
    const int g = threadIdx.x;
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, "
        "%7}], [%2];" ::"r"(base + g * 512),
        "l"(reinterpret_cast<uint64_t>(&map)), "r"(smem_u32(&bar)), "r"(kb * 128), "r"(idx[4 * g]), "r"(idx[4 * g + 1]),
        "r"(idx[4 * g + 2]), "r"(idx[4 * g + 3])
        : "memory");
  }
  uint32_t ok = 0;
  while (!ok)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok)
                 : "r"(smem_u32(&bar))
                 : "memory");
  __syncthreads();
  for (int i = threadIdx.x; i < 128 * 128; i += blockDim.x) out[i] = tile[i];
}

int main() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaFree(0);
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) return 2;
  EncodeTiledFn enc = (EncodeTiledFn)fn;
  const int R = 1000, KB = 1024;  // rows x bytes per row
  uint8_t* h = (uint8_t*)malloc((size_t)R * KB);
  for (int r = 0; r < R; ++r)
    for (int c = 0; c < KB; ++c) h[(size_t)r * KB + c] = (uint8_t)((r * 131 + c * 7 + (c >> 4) * 13) & 0xff);
  uint8_t *d, *dout;
  cudaMalloc(&d, (size_t)R * KB);
  cudaMemcpy(d, h, (size_t)R * KB, cudaMemcpyHostToDevice);
  cudaMalloc(&dout, 128 * 128);
  int hidx[128];
  uint32_t s = 99;
  for (int i = 0; i < 128; ++i) {
    s = s * 1664525u + 1013904223u;
    hidx[i] = (s >> 8) % R;
  }
  int* didx;
  cudaMalloc(&didx, sizeof(hidx));
  cudaMemcpy(didx, hidx, sizeof(hidx), cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 128 + 1024);
  for (int boxrows : {1, 4}) {
    for (int rows : {128, 16}) {
      CUtensorMap m;
      const cuuint64_t gdim[2] = {(cuuint64_t)KB, (cuuint64_t)R};
      const cuuint64_t gstride[1] = {(cuuint64_t)KB};
      const cuuint32_t box[2] = {128, (cuuint32_t)boxrows};
      const cuuint32_t estr[2] = {1, 1};
      CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) {
        printf("box rows %d: encode failed (%d)\n", boxrows, (int)r);
        continue;
      }
      const int kb = 3;
      cudaMemset(dout, 0, 128 * 128);
      probe<<<1, 128, 128 * 128 + 1024>>>(m, didx, kb, dout, rows);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) {
        printf("box rows %d, %d rows: kernel failed: %s\n", boxrows, rows, cudaGetErrorString(e));
        return 1;  // sticky
      }
      uint8_t o[128 * 128];
      cudaMemcpy(o, dout, sizeof(o), cudaMemcpyDeviceToHost);
      int bad = 0, untouched_ok = 1;
      for (int rr = 0; rr < 128; ++rr)
        for (int c = 0; c < 8; ++c)
          for (int b = 0; b < 16; ++b) {
            const uint8_t got = o[rr * 128 + ((c ^ (rr & 7)) << 4) + b];
            if (rr < rows) {
              const uint8_t want = h[(size_t)hidx[rr] * KB + kb * 128 + c * 16 + b];
              bad += got != want;
            } else if (got != 0xEE) {
              untouched_ok = 0;
            }
          }
      printf("box rows %d, %d gathered rows: %s (%d bytes differ), rows beyond %s\n", boxrows, rows, bad ? "MISMATCH" : "LAYOUT OK",
             bad, untouched_ok ? "untouched" : "OVERWRITTEN");
    }
  }
  return 0;
}
