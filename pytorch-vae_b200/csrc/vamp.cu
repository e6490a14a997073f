// K9: the VampPrior term of models/vampvae.py:124-160.
//
//   log q rows   lq_i  = sum_d -0.5 (lv_id + (z_id - mu_id)^2) / exp(lv_id)                       (vampvae.py:136-138)
//   prior rows   t_ik  = sum_d -0.5 (plv_kd + (z_id - pmu_kd)^2) / exp(plv_kd) - log K            (vampvae.py:153-155)
//                lse_i = logsumexp_k t_ik                                                         (vampvae.py:158)
//
// (pmu, plv) are the encodings of the K pseudo-inputs.  The reference materialises the [B, K, D] broadcast four times
// forward (sub, pow, add, div) and again in autograd; here the [B, K] table t is the only intermediate, 4 B K bytes.
// Forward: one CTA per sample, one warp per component at a time, lanes over D, z row staged in shared memory.
// Backward: softmax weights w_ik = exp(t_ik - lse_i) recomputed from the saved table; g_z by one CTA per sample,
// (g_pmu, g_plv) by one CTA per component with the sample loop in a fixed order -- no atomics, bit-reproducible.
#include "common.cuh"

namespace vk {
namespace {

constexpr int kVpThreads = 256;

__global__ void __launch_bounds__(kVpThreads)
vamp_prior_fwd_kernel(const float* __restrict__ z, const float* __restrict__ pmu, const float* __restrict__ plv, int64_t B,
                      int K, int D, float log_k, float* __restrict__ t_out, float* __restrict__ lse_out) {
  extern __shared__ float sm[];
  float* z_s = sm;            // [D]
  float* t_s = sm + D;        // [K]
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int64_t i = blockIdx.x; i < B; i += gridDim.x) {
    for (int d = threadIdx.x; d < D; d += kVpThreads) z_s[d] = z[i * D + d];
    __syncthreads();
    for (int k = w; k < K; k += kVpThreads / 32) {
      const float* m = pmu + (size_t)k * D;
      const float* v = plv + (size_t)k * D;
      float acc = 0.f;
      for (int d = lane; d < D; d += 32) {
        const float diff = z_s[d] - m[d];
        acc += (-0.5f * (v[d] + diff * diff)) / expf(v[d]);
      }
      acc = warp_sum(acc);
      if (lane == 0) {
        const float t = acc - log_k;
        t_s[k] = t;
        t_out[i * K + k] = t;
      }
    }
    __syncthreads();
    if (w == 0) {
      float mx = -INFINITY;
      for (int k = lane; k < K; k += 32) mx = fmaxf(mx, t_s[k]);
      mx = warp_max(mx);
      float s = 0.f;
      for (int k = lane; k < K; k += 32) s += expf(t_s[k] - mx);
      s = warp_sum(s);
      if (lane == 0) lse_out[i] = mx + logf(s);
    }
    __syncthreads();
  }
}

// g_z[i, d] = g_i sum_k w_ik (-(z_id - pmu_kd) exp(-plv_kd))
__global__ void __launch_bounds__(kVpThreads)
vamp_prior_bwd_z_kernel(const float* __restrict__ z, const float* __restrict__ pmu, const float* __restrict__ plv,
                        const float* __restrict__ t, const float* __restrict__ lse, const float* __restrict__ g_rows,
                        int64_t B, int K, int D, float* __restrict__ g_z) {
  extern __shared__ float sm[];
  float* w_s = sm;            // [K]
  for (int64_t i = blockIdx.x; i < B; i += gridDim.x) {
    const float l = lse[i];
    for (int k = threadIdx.x; k < K; k += kVpThreads) w_s[k] = expf(t[i * K + k] - l);
    __syncthreads();
    const float g = g_rows[i];
    for (int d = threadIdx.x; d < D; d += kVpThreads) {
      const float zi = z[i * D + d];
      float acc = 0.f;
      for (int k = 0; k < K; ++k) {
        const float diff = zi - pmu[(size_t)k * D + d];
        acc = fmaf(w_s[k], -diff * expf(-plv[(size_t)k * D + d]), acc);
      }
      g_z[i * D + d] = g * acc;
    }
    __syncthreads();
  }
}

// g_pmu[k, d] = sum_i g_i w_ik (z_id - pmu_kd) e,   g_plv[k, d] = sum_i g_i w_ik 0.5 e (plv_kd + diff^2 - 1),  e = exp(-plv_kd)
// blockIdx.x = component, blockIdx.y = slice of D; the sample loop runs in index order (coefficients staged per 256 samples).
__global__ void __launch_bounds__(kVpThreads)
vamp_prior_bwd_p_kernel(const float* __restrict__ z, const float* __restrict__ pmu, const float* __restrict__ plv,
                        const float* __restrict__ t, const float* __restrict__ lse, const float* __restrict__ g_rows,
                        int64_t B, int K, int D, float* __restrict__ g_pmu, float* __restrict__ g_plv) {
  __shared__ float c_s[kVpThreads];
  const int k = blockIdx.x;
  const int d = blockIdx.y * kVpThreads + threadIdx.x;
  const bool live = d < D;
  const float m = live ? pmu[(size_t)k * D + d] : 0.f;
  const float v = live ? plv[(size_t)k * D + d] : 0.f;
  const float e = expf(-v);
  float am = 0.f, av = 0.f;
  for (int64_t i0 = 0; i0 < B; i0 += kVpThreads) {
    const int64_t i = i0 + threadIdx.x;
    c_s[threadIdx.x] = i < B ? g_rows[i] * expf(t[i * K + k] - lse[i]) : 0.f;
    __syncthreads();
    const int n = (int)((B - i0) < kVpThreads ? (B - i0) : kVpThreads);
    if (live) {
      for (int j = 0; j < n; ++j) {
        const float diff = z[(i0 + j) * D + d] - m;
        const float c = c_s[j];
        am = fmaf(c, diff * e, am);
        av = fmaf(c, 0.5f * e * (v + diff * diff - 1.f), av);
      }
    }
    __syncthreads();
  }
  if (live) {
    g_pmu[(size_t)k * D + d] = am;
    g_plv[(size_t)k * D + d] = av;
  }
}

// rows[i] = sum_d -0.5 (lv + (z - mu)^2) / exp(lv): one warp per row
__global__ void __launch_bounds__(kVpThreads)
gauss_logq_rows_fwd_kernel(const float* __restrict__ z, const float* __restrict__ mu, const float* __restrict__ lv, int64_t B,
                           int D, float* __restrict__ rows) {
  const int lane = threadIdx.x & 31;
  const int64_t wpb = kVpThreads / 32;
  for (int64_t i = (int64_t)blockIdx.x * wpb + (threadIdx.x >> 5); i < B; i += (int64_t)gridDim.x * wpb) {
    float acc = 0.f;
    for (int d = lane; d < D; d += 32) {
      const int64_t e = i * D + d;
      const float diff = z[e] - mu[e];
      acc += (-0.5f * (lv[e] + diff * diff)) / expf(lv[e]);
    }
    acc = warp_sum(acc);
    if (lane == 0) rows[i] = acc;
  }
}

__global__ void __launch_bounds__(kVpThreads)
gauss_logq_rows_bwd_kernel(const float* __restrict__ z, const float* __restrict__ mu, const float* __restrict__ lv,
                           const float* __restrict__ g_rows, int64_t B, int D, float* __restrict__ g_z,
                           float* __restrict__ g_mu, float* __restrict__ g_lv) {
  const int64_t n = B * D;
  for (int64_t e = (int64_t)blockIdx.x * kVpThreads + threadIdx.x; e < n; e += (int64_t)gridDim.x * kVpThreads) {
    const float g = g_rows[e / D];
    const float diff = z[e] - mu[e];
    const float v = lv[e];
    const float ex = expf(-v);
    const float gz = -g * diff * ex;
    g_z[e] = gz;
    g_mu[e] = -gz;
    g_lv[e] = g * 0.5f * ex * (v + diff * diff - 1.f);
  }
}

inline unsigned capped(int64_t blocks) {
  const int64_t cap = (int64_t)sm_count() * 8;
  return (unsigned)(blocks < cap ? (blocks < 1 ? 1 : blocks) : cap);
}

inline bool vamp_shape_ok(int64_t B, int64_t K, int64_t D) {
  return B >= 1 && K >= 1 && D >= 1 && K <= 8192 && D <= 8192;      // (D + K) floats of shared memory per CTA
}

}  // namespace
}  // namespace vk

using namespace vk;

extern "C" int vk_vamp_prior_fwd(const float* z, const float* pmu, const float* plv, int64_t B, int64_t K, int64_t D,
                                 float* t_out, float* lse_out, void* stream) {
  VK_REQUIRE(z && pmu && plv && t_out && lse_out, VK_ERR_BAD_ARG, "vamp_prior_fwd: null pointer");
  VK_REQUIRE(vamp_shape_ok(B, K, D), VK_ERR_BAD_ARG, "vamp_prior_fwd: bad shape B=%lld K=%lld D=%lld", (long long)B,
             (long long)K, (long long)D);
  const float log_k = logf((float)K);                       // torch.log(torch.tensor(K).float()), vampvae.py:155
  const size_t smem = (size_t)(D + K) * sizeof(float);
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(vamp_prior_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  vamp_prior_fwd_kernel<<<capped(B), kVpThreads, smem, (cudaStream_t)stream>>>(z, pmu, plv, B, (int)K, (int)D, log_k, t_out,
                                                                              lse_out);
  return check_launch("vamp_prior_fwd");
}

extern "C" int vk_vamp_prior_bwd(const float* z, const float* pmu, const float* plv, const float* t, const float* lse,
                                 const float* g_rows, int64_t B, int64_t K, int64_t D, float* g_z, float* g_pmu,
                                 float* g_plv, void* stream) {
  VK_REQUIRE(z && pmu && plv && t && lse && g_rows && g_z && g_pmu && g_plv, VK_ERR_BAD_ARG, "vamp_prior_bwd: null pointer");
  VK_REQUIRE(vamp_shape_ok(B, K, D), VK_ERR_BAD_ARG, "vamp_prior_bwd: bad shape B=%lld K=%lld D=%lld", (long long)B,
             (long long)K, (long long)D);
  cudaStream_t s = (cudaStream_t)stream;
  vamp_prior_bwd_z_kernel<<<capped(B), kVpThreads, (size_t)K * sizeof(float), s>>>(z, pmu, plv, t, lse, g_rows, B, (int)K,
                                                                                  (int)D, g_z);
  const int rc = check_launch("vamp_prior_bwd (z)");
  if (rc != VK_OK) return rc;
  const dim3 grid((unsigned)K, (unsigned)((D + kVpThreads - 1) / kVpThreads));
  vamp_prior_bwd_p_kernel<<<grid, kVpThreads, 0, s>>>(z, pmu, plv, t, lse, g_rows, B, (int)K, (int)D, g_pmu, g_plv);
  return check_launch("vamp_prior_bwd (prior)");
}

extern "C" int vk_gauss_logq_rows_fwd(const float* z, const float* mu, const float* log_var, int64_t B, int64_t D,
                                      float* rows_out, void* stream) {
  VK_REQUIRE(z && mu && log_var && rows_out, VK_ERR_BAD_ARG, "gauss_logq_rows_fwd: null pointer");
  VK_REQUIRE(B >= 1 && D >= 1 && D < (1 << 24), VK_ERR_BAD_ARG, "gauss_logq_rows_fwd: bad shape B=%lld D=%lld", (long long)B,
             (long long)D);
  const int64_t wpb = kVpThreads / 32;
  gauss_logq_rows_fwd_kernel<<<capped((B + wpb - 1) / wpb), kVpThreads, 0, (cudaStream_t)stream>>>(z, mu, log_var, B, (int)D,
                                                                                                  rows_out);
  return check_launch("gauss_logq_rows_fwd");
}

extern "C" int vk_gauss_logq_rows_bwd(const float* z, const float* mu, const float* log_var, const float* g_rows, int64_t B,
                                      int64_t D, float* g_z, float* g_mu, float* g_lv, void* stream) {
  VK_REQUIRE(z && mu && log_var && g_rows && g_z && g_mu && g_lv, VK_ERR_BAD_ARG, "gauss_logq_rows_bwd: null pointer");
  VK_REQUIRE(B >= 1 && D >= 1 && D < (1 << 24), VK_ERR_BAD_ARG, "gauss_logq_rows_bwd: bad shape B=%lld D=%lld", (long long)B,
             (long long)D);
  gauss_logq_rows_bwd_kernel<<<capped((B * D + kVpThreads - 1) / kVpThreads), kVpThreads, 0, (cudaStream_t)stream>>>(
      z, mu, log_var, g_rows, B, (int)D, g_z, g_mu, g_lv);
  return check_launch("gauss_logq_rows_bwd");
}
