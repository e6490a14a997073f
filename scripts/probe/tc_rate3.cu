// tcgen05.mma issue-rate probe 3: the MMA warp stays converged, one elected lane issues.
//   tc_rate3 <N> <mode: 0 predicate inside asm | 1 C++ if (leader)> <b_mn: 1 = B operand MN-major, 64-row stage as in the MMD gradient GEMM>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include "../../pytorch-vae_b200/csrc/tc.cuh"
using namespace vk::tc;

template <int MODE>
__global__ void __launch_bounds__(128, 1) rate(int N, int bmn, int nmma, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tslot;
  for (int i = threadIdx.x; i < 196608 / 4; i += 128) ((uint32_t*)smem)[i] = 0x3c003c00u + i;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
  if (threadIdx.x < 32) tmem_alloc(smem_u32(&tslot), 512);
  fence_proxy_async();
  fence_before_sync(); __syncthreads(); fence_after_sync();
  const uint32_t tm = tslot;
  if (threadIdx.x < 32) {
    const uint32_t leader = elect_one();
    const uint32_t sa = (smem_u32(smem) + 1023) & ~1023u, sb = sa + 65536;
    const uint32_t idesc = idesc_bf16(128, N, 0, bmn);
    const uint64_t ad0 = smem_desc(sa, 128 * 16, 128), bd0 = bmn ? smem_desc(sb, 128, 64 * 16) : smem_desc(sb, N * 16, 128);
    const uint64_t astep = (2 * 128 * 16) >> 4, bstep = bmn ? (256 >> 4) : ((2 * N * 16) >> 4);
    long long t0 = clock64();
    for (int i = 0; i < nmma; i += 8) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int kk = bmn ? (k & 3) : k;
        if (MODE == 0) mma_bf16_if(leader, tm, ad0 + kk * astep, bd0 + kk * bstep, idesc, 1);
        else if (leader) mma_bf16(tm, ad0 + kk * astep, bd0 + kk * bstep, idesc, 1);
      }
    }
    mma_commit_if(leader, smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    long long t1 = clock64();
    if (leader) out[blockIdx.x] = t1 - t0;
  }
  fence_before_sync(); __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 512);
}

int main(int argc, char** argv) {
  const int N = argc > 1 ? atoi(argv[1]) : 128, mode = argc > 2 ? atoi(argv[2]) : 0, bmn = argc > 3 ? atoi(argv[3]) : 0, nmma = 4096;
  long long* d; cudaMalloc(&d, 16); cudaMemset(d, 0, 16);
  cudaFuncSetAttribute(rate<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(rate<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int rep = 0; rep < 2; ++rep) {
    if (mode == 0) rate<0><<<1, 128, 200 * 1024>>>(N, bmn, nmma, d); else rate<1><<<1, 128, 200 * 1024>>>(N, bmn, nmma, d);
  }
  cudaError_t e = cudaDeviceSynchronize();
  long long h = 0; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  printf("N=%3d mode=%d b_mn=%d: %s  %.1f cycles per MMA (floor %.0f)\n", N, mode, bmn, cudaGetErrorString(e), (double)h / nmma, 128.0 * N / 256);
  return 0;
}
