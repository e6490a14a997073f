// K4b: the kernel matrix on its own -- WAE_MMD / InfoVAE compute_kernel, compute_rbf, compute_inv_mult_quad as public
// methods (reference: models/wae_mmd.py:125-191, models/info_vae.py:150-201).  The hot-path caller compute_mmd never
// needs the [N, N] matrix (mmd.cu sweeps it on tensor cores); these entry points exist so the three methods keep their
// reference signatures and results when a user calls them directly.
//
// Inputs are the reference's [N1, N2, D] operands given as base pointer + three element strides, so the expand() views
// compute_kernel builds (stride 0 on the broadcast axis, wae_mmd.py:140-141) are read without materialising anything.
//   rbf  k_ij = exp(-(mean_d (x1 - x2)^2) / sigma),   sigma = 2 D z_var                       (wae_mmd.py:164-167)
//   imq  k_ij = C / (eps + C + sum_d (x1 - x2)^2),    C = 2 D z_var;  result = sum k - sum diag k   (:185-189)
// SIMT, one warp per pair with lanes over the latent dim; fp32 like the reference, sums of the imq scalar in double
// with a fixed-order two-stage reduction.
#include "common.cuh"

namespace vk {
namespace {

struct PairView { const float* p; int64_t si, sj, sd; };

__device__ __forceinline__ float pair_sq(const PairView& a, const PairView& b, int64_t i, int64_t j, int D, int lane) {
  const float* pa = a.p + i * a.si + j * a.sj;
  const float* pb = b.p + i * b.si + j * b.sj;
  float s = 0.f;
  for (int d = lane; d < D; d += 32) {
    const float t = __ldg(pa + d * a.sd) - __ldg(pb + d * b.sd);
    s = fmaf(t, t, s);
  }
  return warp_sum(s);
}

// kernel value from the squared distance, literal association of the reference
__device__ __forceinline__ float kval(float s, int type, float Df, float sigma, float c, float epc) {
  return type == 0 ? expf(-((s / Df) / sigma)) : c / (epc + s);
}

constexpr int kPairThreads = 256;

__global__ void __launch_bounds__(kPairThreads)
pair_fwd_kernel(PairView a, PairView b, int64_t N1, int64_t N2, int D, int type, float sigma, float c, float epc,
                float* __restrict__ kmat, float* __restrict__ imq_out, unsigned int* ticket, double* partials) {
  __shared__ double red[2 * kPairThreads / 32];
  const int lane = threadIdx.x & 31;
  const int64_t wpb = kPairThreads / 32;
  const int64_t pairs = N1 * N2;
  double tot = 0.0, dia = 0.0;
  for (int64_t w = (int64_t)blockIdx.x * wpb + (threadIdx.x >> 5); w < pairs; w += (int64_t)gridDim.x * wpb) {
    const int64_t i = w / N2, j = w - i * N2;
    const float k = kval(pair_sq(a, b, i, j, D, lane), type, (float)D, sigma, c, epc);
    if (lane == 0) {
      kmat[w] = k;
      tot += (double)k;
      if (i == j) dia += (double)k;
    }
  }
  if (!imq_out) return;
  double v[2] = {tot, dia};
  block_sum<2>(v, red);
  if (threadIdx.x == 0) { partials[2 * blockIdx.x] = v[0]; partials[2 * blockIdx.x + 1] = v[1]; }
  if (last_block_ticket(ticket, gridDim.x)) {
    double t[2] = {0.0, 0.0};
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) { t[0] += partials[2 * i]; t[1] += partials[2 * i + 1]; }
    __syncthreads();
    block_sum<2>(t, red);
    if (threadIdx.x == 0) imq_out[0] = (float)t[0] - (float)t[1];           // kernel.sum() - kernel.diag().sum()
  }
}

// d k_ij / d s_ij times the upstream gradient:  rbf  g_ij k (-1 / (D sigma));  imq  g (1 - delta_ij) (-k^2 / C)
__device__ __forceinline__ float pair_coef(float k, float g, bool diag, int type, float Df, float sigma, float c) {
  if (type == 0) return g * k * (-1.f / (Df * sigma));
  return diag ? 0.f : g * (-(k * k) / c);
}

// Row form (x1 [N1, D] against x2 [N2, D], the compute_kernel case): out_r = sign * sum_c 2 coef(r, c) (xa_r - xb_c).
// The same kernel serves d/dx1 (r = i, c = j) and d/dx2 (r = j, c = i, sign = -1 and (xa, xb) = (x2, x1) swapped,
// which turns (x2_j - x1_i) into the needed -(x1_i - x2_j)).
__global__ void __launch_bounds__(kPairThreads)
pair_bwd_rows_kernel(const float* __restrict__ xa, int64_t a_sr, int64_t a_sd, const float* __restrict__ xb, int64_t b_sr,
                     int64_t b_sd, int64_t R, int64_t Cn, int D, const float* __restrict__ kmat, int64_t k_sr, int64_t k_sc,
                     const float* __restrict__ g, int64_t g_sr, int64_t g_sc, int type, float sigma, float c,
                     float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * (kPairThreads / 32) + (threadIdx.x >> 5);
  if (r >= R) return;
  for (int d0 = 0; d0 < D; d0 += 128) {                       // 4 latent dims per lane per sweep
    float xr[4], acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int u = 0; u < 4; ++u) { const int d = d0 + lane + 32 * u; xr[u] = d < D ? __ldg(xa + r * a_sr + d * a_sd) : 0.f; }
    for (int64_t cc = 0; cc < Cn; ++cc) {
      const float co = 2.f * pair_coef(__ldg(kmat + r * k_sr + cc * k_sc), __ldg(g + r * g_sr + cc * g_sc), r == cc, type,
                                       (float)D, sigma, c);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int d = d0 + lane + 32 * u;
        if (d < D) acc[u] = fmaf(co, xr[u] - __ldg(xb + cc * b_sr + d * b_sd), acc[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) { const int d = d0 + lane + 32 * u; if (d < D) out[r * D + d] = acc[u]; }
  }
}

// Dense form (general [N1, N2, D] operands): g1[i, j, :] = 2 coef_ij (x1_ij - x2_ij), g2 = -g1.
__global__ void __launch_bounds__(kPairThreads)
pair_bwd_dense_kernel(PairView a, PairView b, int64_t N1, int64_t N2, int D, const float* __restrict__ kmat,
                      const float* __restrict__ g, int64_t g_sr, int64_t g_sc, int type, float sigma, float c,
                      float* __restrict__ g1, float* __restrict__ g2) {
  const int lane = threadIdx.x & 31;
  const int64_t wpb = kPairThreads / 32, pairs = N1 * N2;
  for (int64_t w = (int64_t)blockIdx.x * wpb + (threadIdx.x >> 5); w < pairs; w += (int64_t)gridDim.x * wpb) {
    const int64_t i = w / N2, j = w - i * N2;
    const float co = 2.f * pair_coef(__ldg(kmat + w), __ldg(g + i * g_sr + j * g_sc), i == j, type, (float)D, sigma, c);
    const float* pa = a.p + i * a.si + j * a.sj;
    const float* pb = b.p + i * b.si + j * b.sj;
    for (int d = lane; d < D; d += 32) {
      const float v = co * (__ldg(pa + d * a.sd) - __ldg(pb + d * b.sd));
      if (g1) g1[w * D + d] = v;
      if (g2) g2[w * D + d] = -v;
    }
  }
}

inline int pair_consts(int kernel_type, int64_t D, float z_var, float eps, float& sigma, float& c, float& epc) {
  VK_REQUIRE(kernel_type == VK_KERNEL_RBF || kernel_type == VK_KERNEL_IMQ, VK_ERR_BAD_ARG, "pair_kernel: kernel_type %d",
             kernel_type);
  const double C = 2.0 * (double)D * (double)z_var;           // Python float arithmetic in the reference
  sigma = (float)C;
  c = (float)C;
  epc = (float)((double)eps + C);                             // eps + C is added in double before it meets the tensor
  return VK_OK;
}

}  // namespace
}  // namespace vk

using namespace vk;

extern "C" int vk_pair_kernel_fwd(const float* x1, const int64_t* s1, const float* x2, const int64_t* s2, int64_t N1,
                                  int64_t N2, int64_t D, int kernel_type, float z_var, float eps, float* kmat,
                                  float* imq_out, void* workspace, int64_t workspace_bytes, void* stream) {
  VK_REQUIRE(x1 && x2 && s1 && s2 && kmat && workspace, VK_ERR_BAD_ARG, "pair_kernel_fwd: null pointer");
  VK_REQUIRE(N1 >= 1 && N2 >= 1 && D >= 1 && D < (1 << 30), VK_ERR_BAD_ARG, "pair_kernel_fwd: empty shape %lldx%lldx%lld",
             (long long)N1, (long long)N2, (long long)D);
  VK_REQUIRE(workspace_bytes >= kWsBytes, VK_ERR_WORKSPACE, "pair_kernel_fwd: workspace too small");
  VK_REQUIRE((kernel_type == VK_KERNEL_IMQ) == (imq_out != nullptr), VK_ERR_BAD_ARG,
             "pair_kernel_fwd: imq_out must be given for the imq kernel and only for it");
  float sigma, c, epc;
  int rc = pair_consts(kernel_type, D, z_var, eps, sigma, c, epc);
  if (rc != VK_OK) return rc;
  const int64_t pairs = N1 * N2, wpb = kPairThreads / 32;
  int64_t blocks = (pairs + wpb - 1) / wpb;
  if (blocks > kMaxReduceBlocks) blocks = kMaxReduceBlocks;
  const PairView a{x1, s1[0], s1[1], s1[2]}, b{x2, s2[0], s2[1], s2[2]};
  pair_fwd_kernel<<<(int)blocks, kPairThreads, 0, (cudaStream_t)stream>>>(
      a, b, N1, N2, (int)D, kernel_type, sigma, c, epc, kmat, imq_out, (unsigned int*)workspace,
      (double*)((uint8_t*)workspace + kWsHeader));
  return check_launch("pair_kernel_fwd");
}

extern "C" int vk_pair_kernel_bwd_rows(const float* x1, int64_t x1_sr, int64_t x1_sd, const float* x2, int64_t x2_sr,
                                       int64_t x2_sd, int64_t N1, int64_t N2, int64_t D, int kernel_type, float z_var,
                                       float eps, const float* kmat, const float* g, int64_t g_sr, int64_t g_sc,
                                       float* g_x1, float* g_x2, void* stream) {
  VK_REQUIRE(x1 && x2 && kmat && g, VK_ERR_BAD_ARG, "pair_kernel_bwd_rows: null pointer");
  VK_REQUIRE(N1 >= 1 && N2 >= 1 && D >= 1 && D < (1 << 30), VK_ERR_BAD_ARG, "pair_kernel_bwd_rows: empty shape");
  float sigma, c, epc;
  int rc = pair_consts(kernel_type, D, z_var, eps, sigma, c, epc);
  if (rc != VK_OK) return rc;
  const int wpb = kPairThreads / 32;
  cudaStream_t st = (cudaStream_t)stream;
  if (g_x1)
    pair_bwd_rows_kernel<<<(unsigned)((N1 + wpb - 1) / wpb), kPairThreads, 0, st>>>(
        x1, x1_sr, x1_sd, x2, x2_sr, x2_sd, N1, N2, (int)D, kmat, N2, 1, g, g_sr, g_sc, kernel_type, sigma, c, g_x1);
  if (g_x2)
    pair_bwd_rows_kernel<<<(unsigned)((N2 + wpb - 1) / wpb), kPairThreads, 0, st>>>(
        x2, x2_sr, x2_sd, x1, x1_sr, x1_sd, N2, N1, (int)D, kmat, 1, N2, g, g_sc, g_sr, kernel_type, sigma, c, g_x2);
  return check_launch("pair_kernel_bwd_rows");
}

extern "C" int vk_pair_kernel_bwd_dense(const float* x1, const int64_t* s1, const float* x2, const int64_t* s2, int64_t N1,
                                        int64_t N2, int64_t D, int kernel_type, float z_var, float eps, const float* kmat,
                                        const float* g, int64_t g_sr, int64_t g_sc, float* g_x1, float* g_x2,
                                        void* stream) {
  VK_REQUIRE(x1 && x2 && s1 && s2 && kmat && g, VK_ERR_BAD_ARG, "pair_kernel_bwd_dense: null pointer");
  VK_REQUIRE(N1 >= 1 && N2 >= 1 && D >= 1 && D < (1 << 30), VK_ERR_BAD_ARG, "pair_kernel_bwd_dense: empty shape");
  float sigma, c, epc;
  int rc = pair_consts(kernel_type, D, z_var, eps, sigma, c, epc);
  if (rc != VK_OK) return rc;
  const int64_t pairs = N1 * N2, wpb = kPairThreads / 32;
  int64_t blocks = (pairs + wpb - 1) / wpb;
  if (blocks > sm_count() * 64) blocks = sm_count() * 64;
  const PairView a{x1, s1[0], s1[1], s1[2]}, b{x2, s2[0], s2[1], s2[2]};
  pair_bwd_dense_kernel<<<(int)blocks, kPairThreads, 0, (cudaStream_t)stream>>>(a, b, N1, N2, (int)D, kmat, g, g_sr, g_sc,
                                                                                  kernel_type, sigma, c, g_x1, g_x2);
  return check_launch("pair_kernel_bwd_dense");
}
