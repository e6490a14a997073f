// tcgen05.mma issue-rate probe: cycles per MMA for different shared-memory layouts / shapes (data is garbage).
//   tc_rate <N> <layout: 0 interleave(lbo=16R,sbo=128) | 1 interleave(lbo=128,sbo=nkc*128) | 2 swizzle128B> <b_mn_major> <grid>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include "../../pytorch-vae_b200/csrc/tc.cuh"
using namespace vk::tc;

__global__ void __launch_bounds__(128, 1) rate(int N, int layout, int bmn, int nmma, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tslot;
  for (int i = threadIdx.x; i < 196608 / 4; i += 128) ((uint32_t*)smem)[i] = 0x3c003c00u + i;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
  if (threadIdx.x < 32) tmem_alloc(smem_u32(&tslot), 512);
  fence_proxy_async();
  fence_before_sync(); __syncthreads(); fence_after_sync();
  const uint32_t tm = tslot;
  if (threadIdx.x == 0) {
    const uint32_t sa = (smem_u32(smem) + 1023) & ~1023u, sb = sa + 65536;
    const uint32_t idesc = idesc_bf16(128, N, 0, bmn);
    const int nkc = 16;   // pretend K = 128
    long long t0 = clock64();
    for (int i = 0; i < nmma; ++i) {
      const int k = i & 7;
      uint64_t ad, bd;
      if (layout == 0) {
        ad = smem_desc(sa + k * 2 * 128 * 16, 128 * 16, 128);
        bd = bmn ? smem_desc(sb + k * 256, 128, N * 16) : smem_desc(sb + k * 2 * N * 16, N * 16, 128);
      } else if (layout == 1) {
        ad = smem_desc(sa + k * 256, 128, nkc * 128);
        bd = smem_desc(sb + k * 256, 128, nkc * 128);
      } else {
        ad = (uint64_t)(((sa + k * 32) >> 4) & 0x3FFF) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
        bd = (uint64_t)(((sb + k * 32) >> 4) & 0x3FFF) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
      }
      mma_bf16(tm, ad, bd, idesc, i > 0);
    }
    mma_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
  fence_before_sync(); __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 512);
}

int main(int argc, char** argv) {
  const int N = argc > 1 ? atoi(argv[1]) : 128, layout = argc > 2 ? atoi(argv[2]) : 0, bmn = argc > 3 ? atoi(argv[3]) : 0;
  const int grid = argc > 4 ? atoi(argv[4]) : 1, nmma = 4096;
  long long* d; cudaMalloc(&d, grid * 8);
  cudaFuncSetAttribute(rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int rep = 0; rep < 2; ++rep) rate<<<grid, 128, 200 * 1024>>>(N, layout, bmn, nmma, d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[256]; cudaMemcpy(h, d, grid * 8, cudaMemcpyDeviceToHost);
  long long mx = 0; for (int i = 0; i < grid; ++i) if (h[i] > mx) mx = h[i];
  printf("N=%3d layout=%d b_mn=%d grid=%3d: %s  %.1f cycles/MMA (floor %.0f)\n", N, layout, bmn, grid, cudaGetErrorString(e),
         (double)mx / nmma, 128.0 * N / 256);
  return 0;
}
