// K3: IWAE / MIWAE importance-weighted bound, forward and backward.
// Replaces models/iwae.py:145-158 and models/miwae.py:148-162: the reference
// materialises input.repeat(K) (1 GB at B=4096,K=5), then makes >= 4 passes over
// the [B,K,P] tensors.  Here one CTA owns one batch item: the input image is
// staged once in shared memory and every one of its G*K reconstructions is
// streamed past it exactly once (forward: read-only; backward: read + write
// g_recons).  softmax over K and the closed-form backward run in
// registers of a few threads; there is no [B,K,P] temporary at all.
#include "common.cuh"

namespace vk {

struct IwParams {
  int64_t B, G, K, P;
  float M_N;
};

// sum over one reconstruction of (r - x)^2, x from shared memory (or global when it does not fit)
template <bool XS>
__device__ __forceinline__ float sq_diff_partial(const float* __restrict__ r, const float* __restrict__ xg,
                                                 const float* xs, int64_t P, bool vec) {
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  const int t = threadIdx.x, nt = blockDim.x;
  const int64_t nvec = vec ? (P >> 2) : 0;
  const float4* r4 = reinterpret_cast<const float4*>(r);
  const float4* x4 = reinterpret_cast<const float4*>(XS ? xs : xg);
  int64_t i = t;
  for (; i + 3 * nt < nvec; i += 4 * nt) {
    float4 rv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) rv[u] = ldg_stream(r4 + i + u * nt);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const float4 x = x4[i + u * nt];
      const float d0 = rv[u].x - x.x, d1 = rv[u].y - x.y, d2 = rv[u].z - x.z, d3 = rv[u].w - x.w;
      a0 += d0 * d0; a1 += d1 * d1; a2 += d2 * d2; a3 += d3 * d3;
    }
  }
  for (; i < nvec; i += nt) {
    const float4 rv = ldg_stream(r4 + i);
    const float4 x = x4[i];
    const float d0 = rv.x - x.x, d1 = rv.y - x.y, d2 = rv.z - x.z, d3 = rv.w - x.w;
    a0 += d0 * d0; a1 += d1 * d1; a2 += d2 * d2; a3 += d3 * d3;
  }
  const float* xsrc = XS ? xs : xg;
  for (int64_t j = 4 * nvec + t; j < P; j += nt) {
    const float d = r[j] - xsrc[j];
    a0 += d * d;
  }
  return (a0 + a1) + (a2 + a3);
}

template <bool XS>
__device__ __forceinline__ void stage_input(const float* __restrict__ xg, float* xs, int64_t P, bool vec) {
  if (!XS) return;
  const int t = threadIdx.x, nt = blockDim.x;
  const int64_t nvec = vec ? (P >> 2) : 0;
  for (int64_t i = t; i < nvec; i += nt)
    reinterpret_cast<float4*>(xs)[i] = ldg_stream(reinterpret_cast<const float4*>(xg) + i);
  for (int64_t j = 4 * nvec + t; j < P; j += nt) xs[j] = xg[j];
  __syncthreads();
}

// softmax statistics of one (b,g) row held by one thread
__device__ __forceinline__ void row_stats(const float* lp, const float* kl, int64_t K, float M_N, float& mx,
                                          float& denom, float& S) {
  mx = -INFINITY;
  for (int64_t k = 0; k < K; ++k) mx = fmaxf(mx, lp[k] + M_N * kl[k]);
  denom = 0.f;
  float num = 0.f;
  for (int64_t k = 0; k < K; ++k) {
    const float lw = lp[k] + M_N * kl[k];
    const float e = expf(lw - mx);
    denom += e;
    num += e * lw;
  }
  S = num / denom;
}

// dynamic smem: [P floats input (if XS)] [GK * nwarps floats warp partials] [GK floats lp]
template <bool XS>
__global__ void __launch_bounds__(256)
iw_bound_fwd_kernel(const float* __restrict__ recons, const float* __restrict__ input, const float* __restrict__ kl,
                    IwParams p, float* __restrict__ lp_out, float* __restrict__ out,
                    unsigned int* __restrict__ ticket, double* __restrict__ partials) {
  extern __shared__ __align__(16) float smem[];
  __shared__ double red[3 * 8];
  const int64_t b = blockIdx.x;
  const int64_t GK = p.G * p.K;
  const int nw = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t Ppad = XS ? ((p.P + 3) & ~(int64_t)3) : 0;
  float* xs = smem;
  float* wpart = smem + Ppad;
  float* lps = wpart + GK * nw;
  const float* xg = input + b * p.P;
  const float* rb = recons + b * GK * p.P;
  const bool vec = (p.P % 4 == 0) &&
                   (((reinterpret_cast<uintptr_t>(recons) | reinterpret_cast<uintptr_t>(input)) & 15) == 0);
  stage_input<XS>(xg, xs, p.P, vec);
  for (int64_t s = 0; s < GK; ++s) {
    float v = sq_diff_partial<XS>(rb + s * p.P, xg, xs, p.P, vec);
    v = warp_sum(v);
    if (lane == 0) wpart[s * nw + wid] = v;
  }
  __syncthreads();
  const float invP = 1.0f / (float)p.P;
  for (int64_t s = threadIdx.x; s < GK; s += blockDim.x) {
    float t = 0.f;
    for (int w = 0; w < nw; ++w) t += wpart[s * nw + w];
    const float lp = t * invP;
    lps[s] = lp;
    lp_out[b * GK + s] = lp;
  }
  __syncthreads();
  double v[3] = {0.0, 0.0, 0.0};
  for (int64_t g = threadIdx.x; g < p.G; g += blockDim.x) {
    float mx, denom, S;
    row_stats(lps + g * p.K, kl + (b * p.G + g) * p.K, p.K, p.M_N, mx, denom, S);
    v[0] += (double)S;
  }
  for (int64_t s = threadIdx.x; s < GK; s += blockDim.x) {
    v[1] += (double)lps[s];
    v[2] += (double)kl[b * GK + s];
  }
  block_sum<3>(v, red);
  if (threadIdx.x == 0) {
    partials[3 * b + 0] = v[0];
    partials[3 * b + 1] = v[1];
    partials[3 * b + 2] = v[2];
  }
  if (last_block_ticket(ticket, gridDim.x)) {
    double t[3] = {0.0, 0.0, 0.0};
    for (int64_t i = threadIdx.x; i < p.B; i += blockDim.x) {
      t[0] += partials[3 * i + 0];
      t[1] += partials[3 * i + 1];
      t[2] += partials[3 * i + 2];
    }
    __syncthreads();
    block_sum<3>(t, red);
    if (threadIdx.x == 0) {
      const double rows = (double)(p.B * p.G), all = (double)(p.B * GK);
      out[0] = (float)(t[0] / rows);
      out[1] = (float)(t[1] / all);
      out[2] = (float)(t[2] / all);
      out[3] = -(float)(t[2] / all);
    }
  }
}

// grid = (B, splits): block (b, y) handles pixels [y*chunk, (y+1)*chunk) of every reconstruction of item b.
// dynamic smem: [chunk floats input (if XS)] [GK floats coefficient]
template <bool XS>
__global__ void __launch_bounds__(256)
iw_bound_bwd_kernel(const float* __restrict__ recons, const float* __restrict__ input, const float* __restrict__ lp,
                    const float* __restrict__ kl, IwParams p, int64_t chunk, const float* __restrict__ g_loss,
                    const float* __restrict__ g_recon, const float* __restrict__ g_kld,
                    const float* __restrict__ g_negkld, float* __restrict__ g_recons, float* __restrict__ g_kl) {
  extern __shared__ __align__(16) float smem[];
  const int64_t b = blockIdx.x;
  const int64_t GK = p.G * p.K;
  const int64_t p0 = (int64_t)blockIdx.y * chunk;
  const int64_t plen = (p.P - p0 < chunk) ? (p.P - p0) : chunk;
  const int64_t cpad = XS ? ((chunk + 3) & ~(int64_t)3) : 0;
  float* xs = smem;
  float* coef = smem + cpad;
  const float gl = g_loss ? *g_loss : 0.f;
  const float gr = g_recon ? *g_recon : 0.f;
  const float gk = (g_kld ? *g_kld : 0.f) - (g_negkld ? *g_negkld : 0.f);
  const float inv_rows = 1.0f / (float)(p.B * p.G), inv_all = 1.0f / (float)(p.B * GK);
  for (int64_t g = threadIdx.x; g < p.G; g += blockDim.x) {
    const float* lpr = lp + (b * p.G + g) * p.K;
    const float* klr = kl + (b * p.G + g) * p.K;
    float mx, denom, S;
    row_stats(lpr, klr, p.K, p.M_N, mx, denom, S);
    for (int64_t k = 0; k < p.K; ++k) {
      const float lw = lpr[k] + p.M_N * klr[k];
      const float w = expf(lw - mx) / denom;
      const float c = w * (1.0f + lw - S) * inv_rows;          // dL/dlw  
      coef[g * p.K + k] = (gl * c + gr * inv_all) * (2.0f / (float)p.P);
      if (g_kl && blockIdx.y == 0) g_kl[(b * p.G + g) * p.K + k] = gl * c * p.M_N + gk * inv_all;
    }
  }
  if (!g_recons || plen <= 0) return;
  const float* xg = input + b * p.P + p0;
  const bool vec = (p.P % 4 == 0) && (chunk % 4 == 0) &&
                   (((reinterpret_cast<uintptr_t>(recons) | reinterpret_cast<uintptr_t>(input) |
                      reinterpret_cast<uintptr_t>(g_recons)) & 15) == 0);
  if (XS) stage_input<true>(xg, xs, plen, vec); else __syncthreads();
  const int t = threadIdx.x, nt = blockDim.x;
  const int64_t nvec = vec ? (plen >> 2) : 0;
  const float4* x4 = reinterpret_cast<const float4*>(XS ? xs : xg);
  const float* xsrc = XS ? xs : xg;
  for (int64_t s = 0; s < GK; ++s) {
    const float c = coef[s];
    const float* r = recons + (b * GK + s) * p.P + p0;
    float* go = g_recons + (b * GK + s) * p.P + p0;
    const float4* r4 = reinterpret_cast<const float4*>(r);
    float4* g4 = reinterpret_cast<float4*>(go);
    int64_t i = t;
    for (; i + 3 * nt < nvec; i += 4 * nt) {
      float4 rv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) rv[u] = ldg_stream(r4 + i + u * nt);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float4 x = x4[i + u * nt];
        stg_stream(g4 + i + u * nt, make_float4(c * (rv[u].x - x.x), c * (rv[u].y - x.y), c * (rv[u].z - x.z),
                                                c * (rv[u].w - x.w)));
      }
    }
    for (; i < nvec; i += nt) {
      const float4 rv = ldg_stream(r4 + i);
      const float4 x = x4[i];
      stg_stream(g4 + i, make_float4(c * (rv.x - x.x), c * (rv.y - x.y), c * (rv.z - x.z), c * (rv.w - x.w)));
    }
    for (int64_t j = 4 * nvec + t; j < plen; j += nt) go[j] = c * (r[j] - xsrc[j]);
  }
}

constexpr int64_t kIwMaxSmemInput = 160 * 1024;  // bytes of input image staged per CTA

}  // namespace vk

using namespace vk;

extern "C" int vk_iw_bound_fwd(const float* recons, const float* input, const float* kl, int64_t B, int64_t G,
                               int64_t K, int64_t P, float M_N, float* lp_out, float* out, void* workspace,
                               int64_t workspace_bytes, void* stream) {
  VK_REQUIRE(recons && input && kl && lp_out && out && workspace, VK_ERR_BAD_ARG, "iw_bound_fwd: null pointer");
  VK_REQUIRE(B > 0 && G > 0 && K > 0 && P > 0, VK_ERR_BAD_ARG, "iw_bound_fwd: bad sizes");
  VK_REQUIRE(B < (1ll << 31), VK_ERR_UNSUPPORTED, "iw_bound_fwd: B too large");
  const int64_t need_ws = kWsHeader + B * 3 * (int64_t)sizeof(double);
  VK_REQUIRE(workspace_bytes >= need_ws, VK_ERR_WORKSPACE, "iw_bound_fwd: workspace needs %lld bytes", (long long)need_ws);
  const int64_t GK = G * K;
  IwParams p{B, G, K, P, M_N};
  const bool xs = P * 4 <= kIwMaxSmemInput;
  const int64_t smem = ((xs ? ((P + 3) & ~(int64_t)3) : 0) + GK * 8 + GK) * 4;
  VK_REQUIRE(smem <= 220 * 1024, VK_ERR_UNSUPPORTED, "iw_bound_fwd: G*K=%lld too large for shared memory", (long long)GK);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(workspace);
  double* partials = reinterpret_cast<double*>(reinterpret_cast<char*>(workspace) + kWsHeader);
  cudaStream_t s = (cudaStream_t)stream;
  if (xs) {
    cudaFuncSetAttribute(iw_bound_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    iw_bound_fwd_kernel<true><<<(unsigned)B, 256, smem, s>>>(recons, input, kl, p, lp_out, out, ticket, partials);
  } else {
    cudaFuncSetAttribute(iw_bound_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    iw_bound_fwd_kernel<false><<<(unsigned)B, 256, smem, s>>>(recons, input, kl, p, lp_out, out, ticket, partials);
  }
  return check_launch("iw_bound_fwd");
}

extern "C" int vk_iw_bound_bwd(const float* recons, const float* input, const float* lp, const float* kl, int64_t B,
                               int64_t G, int64_t K, int64_t P, float M_N, const float* g_loss, const float* g_recon,
                               const float* g_kld, const float* g_negkld, float* g_recons, float* g_kl, void* stream) {
  VK_REQUIRE(recons && input && lp && kl, VK_ERR_BAD_ARG, "iw_bound_bwd: null pointer");
  VK_REQUIRE(B > 0 && G > 0 && K > 0 && P > 0, VK_ERR_BAD_ARG, "iw_bound_bwd: bad sizes");
  VK_REQUIRE(B < (1ll << 31), VK_ERR_UNSUPPORTED, "iw_bound_bwd: B too large");
  if (!g_recons && !g_kl) return VK_OK;
  const int64_t GK = G * K;
  IwParams p{B, G, K, P, M_N};
  // split P over blockIdx.y when the batch alone cannot fill the machine
  int64_t splits = 1;
  const int64_t want_ctas = (int64_t)sm_count() * 4;
  if (g_recons && B < want_ctas) {
    splits = (want_ctas + B - 1) / B;
    const int64_t max_splits = (P + 4095) / 4096;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
  }
  int64_t chunk = (P + splits - 1) / splits;
  chunk = (chunk + 3) & ~(int64_t)3;
  splits = (P + chunk - 1) / chunk;
  VK_REQUIRE(splits <= 65535, VK_ERR_UNSUPPORTED, "iw_bound_bwd: P too large");
  const bool xs = chunk * 4 <= kIwMaxSmemInput;
  const int64_t smem = ((xs ? chunk : 0) + GK) * 4;
  VK_REQUIRE(smem <= 220 * 1024, VK_ERR_UNSUPPORTED, "iw_bound_bwd: G*K=%lld too large for shared memory", (long long)GK);
  dim3 grid((unsigned)B, (unsigned)splits);
  cudaStream_t s = (cudaStream_t)stream;
  if (xs) {
    cudaFuncSetAttribute(iw_bound_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    iw_bound_bwd_kernel<true><<<grid, 256, smem, s>>>(recons, input, lp, kl, p, chunk, g_loss, g_recon, g_kld, g_negkld, g_recons, g_kl);
  } else {
    cudaFuncSetAttribute(iw_bound_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    iw_bound_bwd_kernel<false><<<grid, 256, smem, s>>>(recons, input, lp, kl, p, chunk, g_loss, g_recon, g_kld, g_negkld, g_recons, g_kl);
  }
  return check_launch("iw_bound_bwd");
}
