// K8: the Gumbel-softmax reparameterisation and the categorical KL term of
// models/cat_vae.py:118-131, 133-173 (the same trick is joint_vae.py:142-167).
//
//   reparameterize   u ~ U(0,1);  g = -log(-log(u + eps) + eps);  s = softmax((q + g) / temp, dim = -1)
//   KL rows          p = softmax(q, -1);  kl_b = sum_{l,c} p (log(p + eps) - log(1/C + eps))      (per sample b)
//
// One warp per categorical row (C classes, 40 in the shipped config): lanes over classes, warp-shuffle max / sum, noise
// from the contract Philox stream keyed by the element index (one uniform per element, 23-bit, never 0 or 1), so the
// result does not depend on the launch geometry or on how a batch is sharded.  The reference runs 9 launches forward
// for the sample and 8 for the KL; here one each, and one each backward.
#include "common.cuh"
#include "philox.cuh"

namespace vk {
namespace {

constexpr int kGsThreads = 256;

__device__ __forceinline__ float philox_uniform1(uint64_t seed, uint64_t offset, uint64_t e) {
  const uint4 w = philox_quad(seed, offset, e >> 2);
  const int c = (int)(e & 3);
  return u23(c == 0 ? w.x : c == 1 ? w.y : c == 2 ? w.z : w.w);
}

// s[r, :] = softmax((q[r, :] + gumbel) / temp); u_out (optional) receives the uniforms for oracle injection
__global__ void __launch_bounds__(kGsThreads)
gumbel_softmax_fwd_kernel(const float* __restrict__ q, int64_t R, int C, float temp, float eps, uint64_t seed,
                          uint64_t offset, const float* __restrict__ u_in, float* __restrict__ s, float* __restrict__ u_out) {
  const int lane = threadIdx.x & 31;
  const int64_t wpb = kGsThreads / 32;
  for (int64_t r = (int64_t)blockIdx.x * wpb + (threadIdx.x >> 5); r < R; r += (int64_t)gridDim.x * wpb) {
    const float* qr = q + r * C;
    float mx = -INFINITY;
    for (int c = lane; c < C; c += 32) {
      const int64_t e = r * C + c;
      const float u = u_in ? u_in[e] : philox_uniform1(seed, offset, (uint64_t)e);
      if (u_out) u_out[e] = u;
      const float g = -logf(-logf(u + eps) + eps);
      const float y = (qr[c] + g) / temp;
      s[e] = y;                                       // staged in the output row, normalised below
      mx = fmaxf(mx, y);
    }
    mx = warp_max(mx);
    float sum = 0.f;
    for (int c = lane; c < C; c += 32) {
      const float ex = expf(s[r * C + c] - mx);
      s[r * C + c] = ex;
      sum += ex;
    }
    sum = warp_sum(sum);
    for (int c = lane; c < C; c += 32) s[r * C + c] = s[r * C + c] / sum;
  }
}

// g_q = s (g_s - sum_c g_s s) / temp
__global__ void __launch_bounds__(kGsThreads)
gumbel_softmax_bwd_kernel(const float* __restrict__ s, const float* __restrict__ g_s, int64_t R, int C, float temp,
                          float* __restrict__ g_q) {
  const int lane = threadIdx.x & 31;
  const int64_t wpb = kGsThreads / 32;
  for (int64_t r = (int64_t)blockIdx.x * wpb + (threadIdx.x >> 5); r < R; r += (int64_t)gridDim.x * wpb) {
    float dot = 0.f;
    for (int c = lane; c < C; c += 32) dot = fmaf(g_s[r * C + c], s[r * C + c], dot);
    dot = warp_sum(dot);
    for (int c = lane; c < C; c += 32) g_q[r * C + c] = s[r * C + c] * (g_s[r * C + c] - dot) / temp;
  }
}

// rows_out[b] = sum over the L categorical rows of sample b of sum_c p (log(p + eps) - k),  k = log(1/C + eps)
// One CTA per sample (grid-stride), one warp per categorical row at a time, block reduce in double.
template <bool BWD>
__global__ void __launch_bounds__(kGsThreads)
cat_kl_kernel(const float* __restrict__ q, int64_t B, int L, int C, float eps, float k, const float* __restrict__ g_rows,
              float* __restrict__ rows_out, float* __restrict__ g_q) {
  __shared__ double red[kGsThreads / 32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
    double acc = 0.0;
    const float gb = BWD ? g_rows[b] : 0.f;
    for (int l = w; l < L; l += kGsThreads / 32) {
      const float* qr = q + ((size_t)b * L + l) * C;
      float mx = -INFINITY;
      for (int c = lane; c < C; c += 32) mx = fmaxf(mx, qr[c]);
      mx = warp_max(mx);
      float sum = 0.f;
      for (int c = lane; c < C; c += 32) sum += expf(qr[c] - mx);
      sum = warp_sum(sum);
      float part = 0.f, pd = 0.f;
      for (int c = lane; c < C; c += 32) {
        const float p = expf(qr[c] - mx) / sum;
        const float lg = logf(p + eps);
        part += p * lg - p * k;                                   // h1 - h2, cat_vae.py:164-168
        if (BWD) pd = fmaf(p, lg - k + p / (p + eps), pd);        // sum_j p_j d f_j / d p_j
      }
      if (BWD) {
        pd = warp_sum(pd);
        float* gr = g_q + ((size_t)b * L + l) * C;
        for (int c = lane; c < C; c += 32) {
          const float p = expf(qr[c] - mx) / sum;
          const float d = logf(p + eps) - k + p / (p + eps);
          gr[c] = gb * p * (d - pd);
        }
      } else {
        acc += (double)warp_sum(part);
      }
    }
    if (!BWD) {
      if (lane == 0) red[w] = acc;
      __syncthreads();
      if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < kGsThreads / 32; ++i) t += red[i];
        rows_out[b] = (float)t;
      }
      __syncthreads();
    }
  }
}

inline unsigned row_grid(int64_t rows_of_warps) {
  const int64_t wpb = kGsThreads / 32, cap = (int64_t)sm_count() * 8;
  const int64_t b = (rows_of_warps + wpb - 1) / wpb;
  return (unsigned)(b < cap ? b : cap);
}

}  // namespace
}  // namespace vk

using namespace vk;

extern "C" int vk_gumbel_softmax_fwd(const float* q, int64_t R, int64_t C, float temp, float eps, uint64_t seed,
                                     uint64_t offset, const float* u_in, float* s_out, float* u_out, void* stream) {
  VK_REQUIRE(q && s_out, VK_ERR_BAD_ARG, "gumbel_softmax_fwd: null pointer");
  VK_REQUIRE(R >= 1 && C >= 1 && C < (1 << 20), VK_ERR_BAD_ARG, "gumbel_softmax_fwd: empty shape R=%lld C=%lld",
             (long long)R, (long long)C);
  VK_REQUIRE(temp > 0.f, VK_ERR_BAD_ARG, "gumbel_softmax_fwd: temperature %g", (double)temp);
  gumbel_softmax_fwd_kernel<<<row_grid(R), kGsThreads, 0, (cudaStream_t)stream>>>(q, R, (int)C, temp, eps, seed, offset, u_in,
                                                                                 s_out, u_out);
  return check_launch("gumbel_softmax_fwd");
}

extern "C" int vk_gumbel_softmax_bwd(const float* s, const float* g_s, int64_t R, int64_t C, float temp, float* g_q,
                                     void* stream) {
  VK_REQUIRE(s && g_s && g_q, VK_ERR_BAD_ARG, "gumbel_softmax_bwd: null pointer");
  VK_REQUIRE(R >= 1 && C >= 1 && C < (1 << 20) && temp > 0.f, VK_ERR_BAD_ARG, "gumbel_softmax_bwd: bad shape / temperature");
  gumbel_softmax_bwd_kernel<<<row_grid(R), kGsThreads, 0, (cudaStream_t)stream>>>(s, g_s, R, (int)C, temp, g_q);
  return check_launch("gumbel_softmax_bwd");
}

extern "C" int vk_cat_kl_rows_fwd(const float* q, int64_t B, int64_t L, int64_t C, float eps, float* rows_out, void* stream) {
  VK_REQUIRE(q && rows_out, VK_ERR_BAD_ARG, "cat_kl_rows_fwd: null pointer");
  VK_REQUIRE(B >= 1 && L >= 1 && C >= 1 && L < (1 << 24) && C < (1 << 20), VK_ERR_BAD_ARG, "cat_kl_rows_fwd: bad shape");
  const float k = (float)log(1.0 / (double)C + (double)eps);              // np.log(1. / categorical_dim + eps): double
  const int64_t cap = (int64_t)sm_count() * 8;
  cat_kl_kernel<false><<<(unsigned)(B < cap ? B : cap), kGsThreads, 0, (cudaStream_t)stream>>>(q, B, (int)L, (int)C, eps, k,
                                                                                             nullptr, rows_out, nullptr);
  return check_launch("cat_kl_rows_fwd");
}

extern "C" int vk_cat_kl_rows_bwd(const float* q, const float* g_rows, int64_t B, int64_t L, int64_t C, float eps, float* g_q,
                                  void* stream) {
  VK_REQUIRE(q && g_rows && g_q, VK_ERR_BAD_ARG, "cat_kl_rows_bwd: null pointer");
  VK_REQUIRE(B >= 1 && L >= 1 && C >= 1 && L < (1 << 24) && C < (1 << 20), VK_ERR_BAD_ARG, "cat_kl_rows_bwd: bad shape");
  const float k = (float)log(1.0 / (double)C + (double)eps);
  const int64_t cap = (int64_t)sm_count() * 8;
  cat_kl_kernel<true><<<(unsigned)(B < cap ? B : cap), kGsThreads, 0, (cudaStream_t)stream>>>(q, B, (int)L, (int)C, eps, k,
                                                                                            g_rows, nullptr, g_q);
  return check_launch("cat_kl_rows_bwd");
}
