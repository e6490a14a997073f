// Stand-alone bring-up probe for the tcgen05 path (not part of the library): one CTA computes
// C[128,N] = A * B with bf16 operands staged by 1-D bulk TMA, for several descriptor hypotheses.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tc_probe tc_probe.cu && timeout 60 ./tc_probe
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include "../../pytorch-vae_b200/csrc/tc.cuh"
using namespace vk::tc;

struct P {
  uint32_t a_bytes, b_bytes, a_lbo, a_sbo, a_kadv, b_lbo, b_sbo, b_kadv, idesc, nk, N;
};

__global__ void __launch_bounds__(128, 1) probe(const uint8_t* A, const uint8_t* B, float* C, P p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ uint32_t tmem_base;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* sa = smem;
  uint8_t* sb = smem + ((p.a_bytes + 1023) & ~1023u);
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bars[0]), 1);
    mbar_init(smem_u32(&bars[1]), 1);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base), 256);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tm = tmem_base;
  if (threadIdx.x == 0) {
    mbar_expect_tx(smem_u32(&bars[0]), p.a_bytes + p.b_bytes);
    bulk_g2s(smem_u32(sa), A, p.a_bytes, smem_u32(&bars[0]));
    bulk_g2s(smem_u32(sb), B, p.b_bytes, smem_u32(&bars[0]));
    mbar_wait(smem_u32(&bars[0]), 0);
    fence_after_sync();
    for (uint32_t k = 0; k < p.nk; ++k) {
      const uint64_t ad = smem_desc(smem_u32(sa) + k * p.a_kadv, p.a_lbo, p.a_sbo);
      const uint64_t bd = smem_desc(smem_u32(sb) + k * p.b_kadv, p.b_lbo, p.b_sbo);
      mma_bf16(tm, ad, bd, p.idesc, k > 0);
    }
    mma_commit(smem_u32(&bars[1]));
  }
  __syncwarp();
  mbar_wait(smem_u32(&bars[1]), 0);
  fence_after_sync();
  const int row = warp * 32 + lane;
  for (uint32_t c0 = 0; c0 < p.N; c0 += 32) {
    uint32_t r[32];
    tmem_ld32(tm + ((uint32_t)(warp * 32) << 16) + c0, r);
    tmem_ld_wait();
    for (int i = 0; i < 32; ++i) C[(size_t)row * p.N + c0 + i] = __uint_as_float(r[i]);
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 256);
}

static uint16_t f2bf(float x) { uint32_t u; memcpy(&u, &x, 4); u += 0x7FFF + ((u >> 16) & 1); return (uint16_t)(u >> 16); }
static float bf2f(uint16_t h) { uint32_t u = (uint32_t)h << 16; float f; memcpy(&f, &u, 4); return f; }

// image of a row-major [rows, cols] matrix: chunk c (8 cols), row r -> 16 bytes at (c*rows + r)*16
static std::vector<uint16_t> image(const std::vector<float>& m, int rows, int cols) {
  std::vector<uint16_t> im((size_t)rows * cols);
  for (int c = 0; c < cols / 8; ++c)
    for (int r = 0; r < rows; ++r)
      for (int e = 0; e < 8; ++e) im[((size_t)c * rows + r) * 8 + e] = f2bf(m[(size_t)r * cols + c * 8 + e]);
  return im;
}

static double run(const char* name, const std::vector<uint16_t>& ia, const std::vector<uint16_t>& ib, P p,
                  const std::vector<double>& ref) {
  uint8_t *dA, *dB; float* dC;
  cudaMalloc(&dA, ia.size() * 2); cudaMalloc(&dB, ib.size() * 2); cudaMalloc(&dC, 128 * p.N * 4);
  cudaMemcpy(dA, ia.data(), ia.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, ib.data(), ib.size() * 2, cudaMemcpyHostToDevice);
  cudaMemset(dC, 0xFF, 128 * p.N * 4);
  p.a_bytes = ia.size() * 2; p.b_bytes = ib.size() * 2;
  size_t sm = ((p.a_bytes + 1023) & ~1023u) + p.b_bytes + 1024;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
  probe<<<1, 128, sm>>>(dA, dB, dC, p);
  cudaError_t e = cudaDeviceSynchronize();
  std::vector<float> c(128 * p.N);
  cudaMemcpy(c.data(), dC, c.size() * 4, cudaMemcpyDeviceToHost);
  double mx = 0, mr = 0;
  for (size_t i = 0; i < c.size(); ++i) { double d = fabs((double)c[i] - ref[i]); if (!(d <= mx)) mx = d; if (fabs(ref[i]) > mr) mr = fabs(ref[i]); }
  printf("%-40s err=%s maxabs=%.3e (ref max %.3e) %s\n", name, cudaGetErrorString(e), mx, mr, mx < 1e-3 * mr ? "OK" : "BAD");
  fflush(stdout);
  cudaFree(dA); cudaFree(dB); cudaFree(dC);
  return mx;
}

int main(int argc, char** argv) {
  srand(1);
  const int which = argc > 1 ? atoi(argv[1]) : -1; int id = 0;
#define RUN(...) do { if (which < 0 || which == id) run(__VA_ARGS__); ++id; } while (0)
  auto rnd = [](int n) { std::vector<float> v(n); for (auto& x : v) x = bf2f(f2bf((float)rand() / RAND_MAX * 2 - 1)); return v; };
  // ---- case A: C[128,N] = A[128,K] * B[N,K]^T, both K-major ----
  for (int N : {128, 64, 256}) {
    const int K = 128;
    auto a = rnd(128 * K), b = rnd(N * K);
    std::vector<double> ref(128 * N);
    for (int i = 0; i < 128; ++i) for (int j = 0; j < N; ++j) { double s = 0; for (int k = 0; k < K; ++k) s += (double)a[i * K + k] * b[j * K + k]; ref[i * N + j] = s; }
    auto ia = image(a, 128, K), ib = image(b, N, K);
    P p{}; p.N = N; p.nk = K / 16; p.idesc = idesc_bf16(128, N, 0, 0);
    p.a_lbo = 128 * 16; p.a_sbo = 128; p.a_kadv = 2 * 128 * 16;
    p.b_lbo = N * 16; p.b_sbo = 128; p.b_kadv = 2 * N * 16;
    char nm[64]; snprintf(nm, 64, "NT K-major N=%d lbo=Kchunk sbo=128", N); RUN(nm, ia, ib, p, ref);
    if (N == 128) {
      P q = p; q.a_lbo = 128; q.a_sbo = 128 * 16; q.b_lbo = 128; q.b_sbo = N * 16;
      RUN("NT K-major N=128 SWAPPED lbo/sbo", ia, ib, q, ref);
    }
  }
  // ---- case B: C[128,Nd] = A[128,J] * Z[J,Nd]; A K-major (K=J), Z read MN-major (N = d contiguous) ----
  for (int J : {64, 128}) {
    const int Nd = 128;
    auto a = rnd(128 * J), z = rnd(J * Nd);
    std::vector<double> ref(128 * Nd);
    for (int i = 0; i < 128; ++i) for (int d = 0; d < Nd; ++d) { double s = 0; for (int j = 0; j < J; ++j) s += (double)a[i * J + j] * z[j * Nd + d]; ref[i * Nd + d] = s; }
    auto ia = image(a, 128, J), iz = image(z, J, Nd);   // iz: chunk over d, row j
    P p{}; p.N = Nd; p.nk = J / 16; p.idesc = idesc_bf16(128, Nd, 0, 1);
    p.a_lbo = 128 * 16; p.a_sbo = 128; p.a_kadv = 2 * 128 * 16;
    p.b_sbo = J * 16; p.b_lbo = 128; p.b_kadv = 16 * 16;      // H1: sbo = MN-chunk stride, lbo = K 8-group stride
    char nm[64]; snprintf(nm, 64, "NN B MN-major J=%d H1 (sbo=MNchunk)", J); RUN(nm, ia, iz, p, ref);
    P q = p; q.b_lbo = J * 16; q.b_sbo = 128;                  // H2: swapped
    snprintf(nm, 64, "NN B MN-major J=%d H2 (lbo=MNchunk)", J); RUN(nm, ia, iz, q, ref);
  }
  return 0;
}
