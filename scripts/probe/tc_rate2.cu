// tcgen05.mma issue-rate probe 2: is a small-N MMA bound by (a) the dependent accumulate chain, (b) the issuing
// thread, (c) shared-memory operand fetch?
//   tc_rate2 <N> <nacc: accumulators rotated> <issuers: 1|2 warps issuing> <sameA: 1 = A operand address fixed>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include "../../pytorch-vae_b200/csrc/tc.cuh"
using namespace vk::tc;

__global__ void __launch_bounds__(128, 1) rate(int N, int nacc, int issuers, int sameA, int nmma, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar[2];
  __shared__ uint32_t tslot;
  for (int i = threadIdx.x; i < 196608 / 4; i += 128) ((uint32_t*)smem)[i] = 0x3c003c00u + i;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar[0]), 1); mbar_init(smem_u32(&bar[1]), 1); mbar_fence_init(); }
  if (threadIdx.x < 32) tmem_alloc(smem_u32(&tslot), 512);
  fence_proxy_async();
  fence_before_sync(); __syncthreads(); fence_after_sync();
  const uint32_t tm = tslot;
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0 && w < issuers) {
    const uint32_t sa = ((smem_u32(smem) + 1023) & ~1023u) + w * 32768, sb = sa + 65536;
    const uint32_t idesc = idesc_bf16(128, N, 0, 0);
    const uint64_t ad0 = smem_desc(sa, 128 * 16, 128), bd0 = smem_desc(sb, N * 16, 128);
    const uint64_t astep = sameA ? 0 : (2 * 128 * 16) >> 4, bstep = (2 * N * 16) >> 4;
    const uint32_t tbase = tm + w * 256;
    long long t0 = clock64();
    for (int i = 0; i < nmma; i += 8) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t d = tbase + ((i / 8 * 8 + k) % nacc) * N;
        mma_bf16(d, ad0 + k * astep, bd0 + k * bstep, idesc, 1);
      }
    }
    mma_commit(smem_u32(&bar[w]));
    mbar_wait(smem_u32(&bar[w]), 0);
    long long t1 = clock64();
    out[blockIdx.x * 2 + w] = t1 - t0;
  }
  fence_before_sync(); __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 512);
}

int main(int argc, char** argv) {
  const int N = argc > 1 ? atoi(argv[1]) : 128, nacc = argc > 2 ? atoi(argv[2]) : 1, issuers = argc > 3 ? atoi(argv[3]) : 1;
  const int sameA = argc > 4 ? atoi(argv[4]) : 0, nmma = 4096, grid = 1;
  if (nacc * N > 256) { printf("nacc*N must be <= 256\n"); return 1; }
  long long* d; cudaMalloc(&d, 16); cudaMemset(d, 0, 16);
  cudaFuncSetAttribute(rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int rep = 0; rep < 2; ++rep) rate<<<grid, 128, 200 * 1024>>>(N, nacc, issuers, sameA, nmma, d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  const long long mx = h[0] > h[1] ? h[0] : h[1];
  printf("N=%3d nacc=%d issuers=%d sameA=%d: %s  %.1f cycles per MMA per issuer, %.1f per MMA overall (floor %.0f)\n", N, nacc,
         issuers, sameA, cudaGetErrorString(e), (double)mx / nmma, (double)mx / (nmma * issuers), 128.0 * N / 256);
  return 0;
}
