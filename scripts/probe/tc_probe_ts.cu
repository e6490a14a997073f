// Bring-up probe: tcgen05.mma with the A operand in TENSOR MEMORY (written by tcgen05.st), B in shared memory.
//   C[128,N] = A[128,K] * B[N,K]^T ; also times the MMA stream.   tc_probe_ts <N> <b_mn>
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "../../pytorch-vae_b200/csrc/tc.cuh"
using namespace vk::tc;

constexpr int K = 128;

__global__ void __launch_bounds__(128, 1) probe(const uint16_t* A, const uint8_t* Bimg, float* C, int N, int bmn, int reps,
                                                long long* cyc) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ uint32_t tmem_base;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bars[0]), 1); mbar_init(smem_u32(&bars[1]), 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base), 512);
  fence_before_sync(); __syncthreads(); fence_after_sync();
  const uint32_t tm = tmem_base, tmA = tm + 256;
  const int row = warp * 32 + lane;
  // A row -> TMEM columns [256, 256 + K/2): column j = elements 2j (low half) and 2j+1 (high half)
  for (int c0 = 0; c0 < K / 2; c0 += 16) {
    uint32_t r[16];
    for (int j = 0; j < 16; ++j) r[j] = (uint32_t)A[row * K + 2 * (c0 + j)] | ((uint32_t)A[row * K + 2 * (c0 + j) + 1] << 16);
    tmem_st16(tmA + ((uint32_t)(warp * 32) << 16) + c0, r);
  }
  tmem_st_wait();
  if (threadIdx.x == 0) {
    mbar_expect_tx(smem_u32(&bars[0]), N * K * 2);
    bulk_g2s(smem_u32(smem), Bimg, N * K * 2, smem_u32(&bars[0]));
  }
  mbar_wait(smem_u32(&bars[0]), 0);
  fence_before_sync(); __syncthreads(); fence_after_sync();
  if (warp == 0) {
    const uint32_t leader = elect_one();
    const uint32_t idesc = idesc_bf16(128, N, 0, bmn);
    // K-major B: [K/8][N][8] ; MN-major B (B given as [K rows][N] -> image [N/8][K][8])
    const uint64_t bd0 = bmn ? smem_desc(smem_u32(smem), 128, K * 16) : smem_desc(smem_u32(smem), N * 16, 128);
    const uint64_t bstep = bmn ? (256 >> 4) : ((2 * N * 16) >> 4);
    const long long t0 = clock64();
    for (int rep = 0; rep < reps; ++rep) {
#pragma unroll
      for (int k = 0; k < K / 16; ++k) mma_bf16_ts_if(leader, tm, tmA + k * 8, bd0 + k * bstep, idesc, (rep | k) != 0);
    }
    mma_commit_if(leader, smem_u32(&bars[1]));
    mbar_wait(smem_u32(&bars[1]), 0);
    if (leader) *cyc = clock64() - t0;
  }
  mbar_wait(smem_u32(&bars[1]), 0);
  fence_after_sync();
  for (int c0 = 0; c0 < N; c0 += 32) {
    uint32_t r[32];
    tmem_ld32(tm + ((uint32_t)(warp * 32) << 16) + c0, r);
    tmem_ld_wait();
    for (int i = 0; i < 32; ++i) C[(size_t)row * N + c0 + i] = __uint_as_float(r[i]);
  }
  fence_before_sync(); __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 512);
}

static uint16_t f2bf(float x) { uint32_t u; memcpy(&u, &x, 4); u += 0x7FFF + ((u >> 16) & 1); return (uint16_t)(u >> 16); }
static float bf2f(uint16_t h) { uint32_t u = (uint32_t)h << 16; float f; memcpy(&f, &u, 4); return f; }

int main(int argc, char** argv) {
  const int N = argc > 1 ? atoi(argv[1]) : 128, bmn = argc > 2 ? atoi(argv[2]) : 0;
  srand(2);
  std::vector<uint16_t> a(128 * K), bimg((size_t)N * K);
  std::vector<float> af(128 * K), bf((size_t)N * K);      // b[j][k]
  for (size_t i = 0; i < a.size(); ++i) { a[i] = f2bf((float)rand() / RAND_MAX * 2 - 1); af[i] = bf2f(a[i]); }
  for (size_t i = 0; i < bf.size(); ++i) bf[i] = bf2f(f2bf((float)rand() / RAND_MAX * 2 - 1));
  if (!bmn) {   // K-major image of B[N,K]: chunk c over k, row j
    for (int c = 0; c < K / 8; ++c) for (int j = 0; j < N; ++j) for (int e = 0; e < 8; ++e)
      bimg[((size_t)c * N + j) * 8 + e] = f2bf(bf[(size_t)j * K + c * 8 + e]);
  } else {      // MN-major: matrix Z[K rows][N] with Z[k][j] = b[j][k]; image chunk c over j (N), row k
    for (int c = 0; c < N / 8; ++c) for (int k = 0; k < K; ++k) for (int e = 0; e < 8; ++e)
      bimg[((size_t)c * K + k) * 8 + e] = f2bf(bf[(size_t)(c * 8 + e) * K + k]);
  }
  uint16_t* dA; uint8_t* dB; float* dC; long long* dcyc;
  cudaMalloc(&dA, a.size() * 2); cudaMalloc(&dB, bimg.size() * 2); cudaMalloc(&dC, 128 * N * 4); cudaMalloc(&dcyc, 8);
  cudaMemcpy(dA, a.data(), a.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, bimg.data(), bimg.size() * 2, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  for (int reps : {1, 512}) {
    cudaMemset(dC, 0xFF, 128 * N * 4);
    probe<<<1, 128, 100 * 1024>>>(dA, dB, dC, N, bmn, reps, dcyc);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<float> c(128 * N);
    long long cyc = 0;
    cudaMemcpy(c.data(), dC, c.size() * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost);
    double mx = 0, mr = 0;
    for (int i = 0; i < 128; ++i) for (int j = 0; j < N; ++j) {
      double s = 0; for (int k = 0; k < K; ++k) s += (double)af[i * K + k] * bf[(size_t)j * K + k];
      s *= reps;
      const double d = fabs((double)c[i * N + j] - s); if (!(d <= mx)) mx = d; if (fabs(s) > mr) mr = fabs(s);
    }
    printf("TS N=%d b_mn=%d reps=%d: %s maxabs=%.3e (ref max %.3e) %s  %.1f cycles/MMA\n", N, bmn, reps, cudaGetErrorString(e), mx, mr,
           mx < 1e-3 * mr ? "OK" : "BAD", (double)cyc / (reps * (K / 16)));
    if (e != cudaSuccess) break;
  }
  return 0;
}
