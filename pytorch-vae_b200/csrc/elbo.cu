// K2: streaming reconstruction term + KL combine, forward and backward.
// Replaces the ~13 fwd / ~25 bwd ATen launches of models/vanilla_vae.py:134-146,
// beta_vae.py:129-152, logcosh_vae.py:125-155 and the recon terms of
// betatc_vae.py:163, swae.py:121-122, wae_mmd.py:118, vq_vae.py:205.
// HBM-bound: 2 x 128-bit streaming loads per 4 pixels forward (8 B/pixel),
// 2 loads + 1 store backward (12 B/pixel).  Deterministic two-stage reduction
// (per-block partials in a workspace, last block combines in fixed order).
#include "common.cuh"

namespace vk {

constexpr float kLog2f = 0.69314718055994530942f;  // torch.log(torch.tensor(2.0)) in fp32

template <int MODE>
__device__ __forceinline__ void recon_accum(float r, float x, float alpha, float neg2alpha, float& s0, float& s1) {
  const float t = r - x;
  if (MODE == VK_RECON_BCE_MEAN) {
    // F.binary_cross_entropy (ATen binary_cross_entropy kernel): both logs clamped at -100, log(1 - r) as log1p(-r)
    s0 += (x - 1.0f) * fmaxf(log1pf(-r), -100.0f) - x * fmaxf(logf(r), -100.0f);
  } else if (MODE == VK_RECON_LOGCOSH) {
    // literal (unstabilised) form of logcosh_vae.py:146-148 so overflow behaviour matches
    s0 += (alpha * t + logf(1.0f + expf(neg2alpha * t))) - kLog2f;
  } else {
    s0 += t * t;
    if (MODE == VK_RECON_MSE_L1_MEAN) s1 += fabsf(t);
  }
}

struct ElboParams {
  int recon_mode, combine_mode;
  float alpha, w, C;
  int64_t numel, rows;
};

__device__ __forceinline__ void elbo_finalize(const ElboParams& p, double s0, double s1, double skl, bool has_kl,
                                              float* out) {
  float recon;
  const double n = (double)p.numel;
  switch (p.recon_mode) {
    case VK_RECON_MSE_MEAN: recon = (float)(s0 / n); break;
    case VK_RECON_MSE_SUM: recon = (float)s0; break;
    case VK_RECON_BCE_MEAN: recon = (float)(s0 / n); break;
    case VK_RECON_BCE_SUM: recon = (float)s0; break;
    case VK_RECON_LOGCOSH: recon = (1.0f / p.alpha) * (float)(s0 / n); break;
    default: recon = (float)(s0 / n) + (float)(s1 / n); break;
  }
  const float kld = has_kl ? (float)(skl / (double)p.rows) : 0.0f;
  float loss = recon;
  if (has_kl && p.combine_mode == VK_COMBINE_LINEAR) loss = recon + p.w * kld;
  if (has_kl && p.combine_mode == VK_COMBINE_ABS) loss = recon + p.w * fabsf(kld - p.C);
  if (p.combine_mode == VK_COMBINE_SCALED) loss = p.C * recon + p.w * kld;      // kld = 0 without kl_rows
  out[0] = loss;
  out[1] = recon;
  out[2] = kld;
  out[3] = -kld;
}

template <int MODE>
__global__ void __launch_bounds__(256)
elbo_fwd_kernel(const float* __restrict__ recons, const float* __restrict__ input,
                const float* __restrict__ kl_rows, ElboParams p, float* __restrict__ out,
                unsigned int* __restrict__ ticket, double* __restrict__ partials) {
  __shared__ double red[3 * 8];
  const float neg2alpha = -2.0f * p.alpha;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, b0 = 0.f, b1 = 0.f, b2 = 0.f, b3 = 0.f;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  const bool vec = ((reinterpret_cast<uintptr_t>(recons) | reinterpret_cast<uintptr_t>(input)) & 15) == 0;
  const int64_t nvec = vec ? (p.numel >> 2) : 0;
  const float4* r4 = reinterpret_cast<const float4*>(recons);
  const float4* x4 = reinterpret_cast<const float4*>(input);
  int64_t i = tid;
  // 4 independent 128-bit load pairs in flight per thread
  for (; i + 3 * nthreads < nvec; i += 4 * nthreads) {
    float4 r[4], x[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) { r[u] = ldg_stream(r4 + i + u * nthreads); x[u] = ldg_stream(x4 + i + u * nthreads); }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      recon_accum<MODE>(r[u].x, x[u].x, p.alpha, neg2alpha, a0, b0);
      recon_accum<MODE>(r[u].y, x[u].y, p.alpha, neg2alpha, a1, b1);
      recon_accum<MODE>(r[u].z, x[u].z, p.alpha, neg2alpha, a2, b2);
      recon_accum<MODE>(r[u].w, x[u].w, p.alpha, neg2alpha, a3, b3);
    }
  }
  for (; i < nvec; i += nthreads) {
    const float4 r = ldg_stream(r4 + i), x = ldg_stream(x4 + i);
    recon_accum<MODE>(r.x, x.x, p.alpha, neg2alpha, a0, b0);
    recon_accum<MODE>(r.y, x.y, p.alpha, neg2alpha, a1, b1);
    recon_accum<MODE>(r.z, x.z, p.alpha, neg2alpha, a2, b2);
    recon_accum<MODE>(r.w, x.w, p.alpha, neg2alpha, a3, b3);
  }
  for (int64_t j = 4 * nvec + tid; j < p.numel; j += nthreads)  // tail / unaligned path
    recon_accum<MODE>(recons[j], input[j], p.alpha, neg2alpha, a0, b0);
  double v[3];
  v[0] = ((double)a0 + (double)a1) + ((double)a2 + (double)a3);
  v[1] = ((double)b0 + (double)b1) + ((double)b2 + (double)b3);
  v[2] = 0.0;
  if (kl_rows)
    for (int64_t j = tid; j < p.rows; j += nthreads) v[2] += (double)kl_rows[j];
  block_sum<3>(v, red);
  if (threadIdx.x == 0) {
    partials[3 * blockIdx.x + 0] = v[0];
    partials[3 * blockIdx.x + 1] = v[1];
    partials[3 * blockIdx.x + 2] = v[2];
  }
  if (last_block_ticket(ticket, gridDim.x)) {
    double t[3] = {0.0, 0.0, 0.0};
    for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) {
      t[0] += partials[3 * b + 0];
      t[1] += partials[3 * b + 1];
      t[2] += partials[3 * b + 2];
    }
    __syncthreads();
    block_sum<3>(t, red);
    if (threadIdx.x == 0) elbo_finalize(p, t[0], t[1], t[2], kl_rows != nullptr, out);
  }
}

__device__ __forceinline__ float sgnf(float v) { return (float)((v > 0.0f) - (v < 0.0f)); }

template <int MODE>
__device__ __forceinline__ float recon_grad(float r, float x, float c, float alpha, float neg2alpha) {
  const float t = r - x;
  if (MODE == VK_RECON_BCE_MEAN || MODE == VK_RECON_BCE_SUM) {
    return c * (t / fmaxf((1.0f - r) * r, 1e-12f));      // binary_cross_entropy_backward: (r - x) / max((1 - r) r, 1e-12)
  } else if (MODE == VK_RECON_LOGCOSH) {
    const float e = expf(neg2alpha * t);
    return c * (alpha + neg2alpha * (e / (1.0f + e)));   // c = g / (alpha * numel)
  } else if (MODE == VK_RECON_MSE_L1_MEAN) {
    return c * (2.0f * t + sgnf(t));                     // c = g / numel
  } else {
    return c * (2.0f * t);                               // c = g / numel (mean) or g (sum)
  }
}

template <int MODE>
__global__ void __launch_bounds__(256)
elbo_bwd_kernel(const float* __restrict__ recons, const float* __restrict__ input, ElboParams p,
                const float* __restrict__ out, const float* __restrict__ g_loss, const float* __restrict__ g_recon,
                const float* __restrict__ g_kld, const float* __restrict__ g_negkld,
                float* __restrict__ g_recons, float* __restrict__ g_kl_rows) {
  const float gl = g_loss ? *g_loss : 0.0f;
  const float neg2alpha = -2.0f * p.alpha;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  if (g_kl_rows) {
    float dk = (g_kld ? *g_kld : 0.0f) - (g_negkld ? *g_negkld : 0.0f);
    if (p.combine_mode == VK_COMBINE_LINEAR || p.combine_mode == VK_COMBINE_SCALED) dk += gl * p.w;
    if (p.combine_mode == VK_COMBINE_ABS) dk += gl * p.w * sgnf(out[2] - p.C);
    dk /= (float)p.rows;
    for (int64_t j = tid; j < p.rows; j += nthreads) g_kl_rows[j] = dk;
  }
  if (!g_recons) return;
  float c = (p.combine_mode == VK_COMBINE_SCALED ? gl * p.C : gl) + (g_recon ? *g_recon : 0.0f);
  if (MODE == VK_RECON_LOGCOSH) c = c / (p.alpha * (float)p.numel);
  else if (MODE != VK_RECON_MSE_SUM && MODE != VK_RECON_BCE_SUM) c = c / (float)p.numel;
  const bool vec = ((reinterpret_cast<uintptr_t>(recons) | reinterpret_cast<uintptr_t>(input) |
                     reinterpret_cast<uintptr_t>(g_recons)) & 15) == 0;
  const int64_t nvec = vec ? (p.numel >> 2) : 0;
  const float4* r4 = reinterpret_cast<const float4*>(recons);
  const float4* x4 = reinterpret_cast<const float4*>(input);
  float4* g4 = reinterpret_cast<float4*>(g_recons);
  int64_t i = tid;
  for (; i + 3 * nthreads < nvec; i += 4 * nthreads) {
    float4 r[4], x[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) { r[u] = ldg_stream(r4 + i + u * nthreads); x[u] = ldg_stream(x4 + i + u * nthreads); }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float4 g;
      g.x = recon_grad<MODE>(r[u].x, x[u].x, c, p.alpha, neg2alpha);
      g.y = recon_grad<MODE>(r[u].y, x[u].y, c, p.alpha, neg2alpha);
      g.z = recon_grad<MODE>(r[u].z, x[u].z, c, p.alpha, neg2alpha);
      g.w = recon_grad<MODE>(r[u].w, x[u].w, c, p.alpha, neg2alpha);
      stg_stream(g4 + i + u * nthreads, g);
    }
  }
  for (; i < nvec; i += nthreads) {
    const float4 r = ldg_stream(r4 + i), x = ldg_stream(x4 + i);
    float4 g;
    g.x = recon_grad<MODE>(r.x, x.x, c, p.alpha, neg2alpha);
    g.y = recon_grad<MODE>(r.y, x.y, c, p.alpha, neg2alpha);
    g.z = recon_grad<MODE>(r.z, x.z, c, p.alpha, neg2alpha);
    g.w = recon_grad<MODE>(r.w, x.w, c, p.alpha, neg2alpha);
    stg_stream(g4 + i, g);
  }
  for (int64_t j = 4 * nvec + tid; j < p.numel; j += nthreads)
    g_recons[j] = recon_grad<MODE>(recons[j], input[j], c, p.alpha, neg2alpha);
}

static int stream_grid(int64_t numel) {
  // 4 float4 per thread per trip; cap at 8 resident CTAs per SM (persistent grid-stride)
  int64_t want = (numel / 4 + 256 * 4 - 1) / (256 * 4);
  const int64_t cap = (int64_t)sm_count() * 8;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  return (int)want;
}

}  // namespace vk

using namespace vk;

extern "C" int vk_elbo_fwd(const float* recons, const float* input, int64_t numel, const float* kl_rows,
                           int64_t rows, int recon_mode, float alpha, int combine_mode, float w, float C,
                           float* out, void* workspace, int64_t workspace_bytes, void* stream) {
  VK_REQUIRE(recons && input && out && workspace, VK_ERR_BAD_ARG, "elbo_fwd: null pointer");
  VK_REQUIRE(numel > 0 && rows >= 0, VK_ERR_BAD_ARG, "elbo_fwd: bad sizes");
  VK_REQUIRE(recon_mode >= 0 && recon_mode <= 5 && combine_mode >= 0 && combine_mode <= 3, VK_ERR_BAD_ARG,
             "elbo_fwd: unknown mode");
  VK_REQUIRE(workspace_bytes >= kWsBytes, VK_ERR_WORKSPACE, "elbo_fwd: workspace needs %lld bytes", (long long)kWsBytes);
  VK_REQUIRE(!kl_rows || rows > 0, VK_ERR_BAD_ARG, "elbo_fwd: kl_rows given but rows == 0");
  ElboParams p{recon_mode, combine_mode, alpha, w, C, numel, rows};
  const int grid = stream_grid(numel);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(workspace);
  double* partials = reinterpret_cast<double*>(reinterpret_cast<char*>(workspace) + kWsHeader);
  cudaStream_t s = (cudaStream_t)stream;
  switch (recon_mode) {
    case VK_RECON_MSE_MEAN:
    case VK_RECON_MSE_SUM:
      elbo_fwd_kernel<VK_RECON_MSE_MEAN><<<grid, 256, 0, s>>>(recons, input, kl_rows, p, out, ticket, partials); break;
    case VK_RECON_LOGCOSH:
      elbo_fwd_kernel<VK_RECON_LOGCOSH><<<grid, 256, 0, s>>>(recons, input, kl_rows, p, out, ticket, partials); break;
    case VK_RECON_BCE_MEAN:
    case VK_RECON_BCE_SUM:
      elbo_fwd_kernel<VK_RECON_BCE_MEAN><<<grid, 256, 0, s>>>(recons, input, kl_rows, p, out, ticket, partials); break;
    default:
      elbo_fwd_kernel<VK_RECON_MSE_L1_MEAN><<<grid, 256, 0, s>>>(recons, input, kl_rows, p, out, ticket, partials); break;
  }
  return check_launch("elbo_fwd");
}

extern "C" int vk_elbo_bwd(const float* recons, const float* input, int64_t numel, int64_t rows, int recon_mode,
                           float alpha, int combine_mode, float w, float C, const float* out, const float* g_loss,
                           const float* g_recon, const float* g_kld, const float* g_negkld, float* g_recons,
                           float* g_kl_rows, void* stream) {
  VK_REQUIRE(recons && input && out, VK_ERR_BAD_ARG, "elbo_bwd: null pointer");
  VK_REQUIRE(numel > 0 && rows >= 0, VK_ERR_BAD_ARG, "elbo_bwd: bad sizes");
  VK_REQUIRE(recon_mode >= 0 && recon_mode <= 5 && combine_mode >= 0 && combine_mode <= 3, VK_ERR_BAD_ARG,
             "elbo_bwd: unknown mode");
  if (!g_recons && !g_kl_rows) return VK_OK;
  ElboParams p{recon_mode, combine_mode, alpha, w, C, numel, rows};
  const int grid = g_recons ? stream_grid(numel) : (int)((rows + 255) / 256 > 1024 ? 1024 : (rows + 255) / 256);
  cudaStream_t s = (cudaStream_t)stream;
  switch (recon_mode) {
    case VK_RECON_MSE_MEAN:
      elbo_bwd_kernel<VK_RECON_MSE_MEAN><<<grid, 256, 0, s>>>(recons, input, p, out, g_loss, g_recon, g_kld, g_negkld, g_recons, g_kl_rows); break;
    case VK_RECON_MSE_SUM:
      elbo_bwd_kernel<VK_RECON_MSE_SUM><<<grid, 256, 0, s>>>(recons, input, p, out, g_loss, g_recon, g_kld, g_negkld, g_recons, g_kl_rows); break;
    case VK_RECON_LOGCOSH:
      elbo_bwd_kernel<VK_RECON_LOGCOSH><<<grid, 256, 0, s>>>(recons, input, p, out, g_loss, g_recon, g_kld, g_negkld, g_recons, g_kl_rows); break;
    case VK_RECON_BCE_MEAN:
      elbo_bwd_kernel<VK_RECON_BCE_MEAN><<<grid, 256, 0, s>>>(recons, input, p, out, g_loss, g_recon, g_kld, g_negkld, g_recons, g_kl_rows); break;
    case VK_RECON_BCE_SUM:
      elbo_bwd_kernel<VK_RECON_BCE_SUM><<<grid, 256, 0, s>>>(recons, input, p, out, g_loss, g_recon, g_kld, g_negkld, g_recons, g_kl_rows); break;
    default:
      elbo_bwd_kernel<VK_RECON_MSE_L1_MEAN><<<grid, 256, 0, s>>>(recons, input, p, out, g_loss, g_recon, g_kld, g_negkld, g_recons, g_kl_rows); break;
  }
  return check_launch("elbo_bwd");
}

// ---- per-image MSE (the trainer's difficulty-sampling side output) ------------------------------------------------
// experiment.py:58-62 recomputes F.mse_loss(recons, input, reduction='none').mean(dim=[1,2,3]) after the loss: a
// [B,C,H,W] temporary, a reduction and one .item() per image.  Here: one streaming pass, one CTA per image (grid-stride
// over images), fp32 sums per thread combined in double, out[b] = mean over the image's P pixels.
namespace vk {
namespace {
__global__ void __launch_bounds__(256)
per_image_mse_kernel(const float* __restrict__ recons, const float* __restrict__ input, int64_t B, int64_t P,
                     float* __restrict__ out) {
  __shared__ double red[8];
  for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
    const float* r = recons + b * P;
    const float* x = input + b * P;
    float s = 0.f;
    const bool vec = (P & 3) == 0 && ((reinterpret_cast<uintptr_t>(r) | reinterpret_cast<uintptr_t>(x)) & 15) == 0;
    if (vec) {
      const float4* r4 = reinterpret_cast<const float4*>(r);
      const float4* x4 = reinterpret_cast<const float4*>(x);
      for (int64_t i = threadIdx.x; i < P / 4; i += blockDim.x) {
        const float4 a = ldg_stream(r4 + i), c = ldg_stream(x4 + i);
        const float d0 = a.x - c.x, d1 = a.y - c.y, d2 = a.z - c.z, d3 = a.w - c.w;
        s += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
      }
    } else {
      for (int64_t i = threadIdx.x; i < P; i += blockDim.x) { const float d = r[i] - x[i]; s += d * d; }
    }
    double v[1] = {(double)s};
    block_sum<1>(v, red);
    if (threadIdx.x == 0) out[b] = (float)(v[0] / (double)P);
    __syncthreads();
  }
}
}  // namespace
}  // namespace vk

extern "C" int vk_per_image_mse(const float* recons, const float* input, int64_t B, int64_t P, float* out, void* stream) {
  VK_REQUIRE(recons && input && out, VK_ERR_BAD_ARG, "per_image_mse: null pointer");
  VK_REQUIRE(B >= 1 && P >= 1, VK_ERR_BAD_ARG, "per_image_mse: empty shape B=%lld P=%lld", (long long)B, (long long)P);
  const int64_t cap = (int64_t)sm_count() * 8;
  per_image_mse_kernel<<<(unsigned)(B < cap ? B : cap), 256, 0, (cudaStream_t)stream>>>(recons, input, B, P, out);
  return check_launch("per_image_mse");
}
