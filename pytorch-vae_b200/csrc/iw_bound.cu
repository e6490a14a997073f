// K3: IWAE / MIWAE importance-weighted bound, forward and backward.
// Replaces models/iwae.py:145-158 and models/miwae.py:148-162: the reference
// materialises input.repeat(K) (1 GB at B=4096,K=5), then makes >= 4 passes over
// the [B,K,P] tensors.  Here one CTA owns one batch item: the input image is
// held in registers column by column and every one of its G*K reconstructions is
// streamed past it exactly once (forward: read-only; backward: read + write
// g_recons).  softmax over K and the closed-form backward run in
// registers of a few threads; there is no [B,K,P] temporary at all.
#include "common.cuh"

namespace vk {

struct IwParams {
  int64_t B, G, K, P;
  float M_N;
};

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

constexpr int kSC = 8;   // reconstructions processed together per pixel column (independent loads in flight)

// Column-major traversal: a thread owns pixel columns i = t, t+nt, ...; for each column the input pixel
// quad is loaded ONCE into registers and the same column of up to kSC reconstructions is streamed past it,
// so kSC independent 128-bit loads are in flight per thread and nothing is staged in shared memory.
// dynamic smem: [GK * nwarps floats warp partials] [GK floats lp]
__global__ void __launch_bounds__(256, 3)
iw_bound_fwd_kernel(const float* __restrict__ recons, const float* __restrict__ input, const float* __restrict__ kl,
                    IwParams p, float* __restrict__ lp_out, float* __restrict__ out,
                    unsigned int* __restrict__ ticket, double* __restrict__ partials) {
  extern __shared__ __align__(16) float smem[];
  __shared__ double red[3 * 8];
  const int64_t b = blockIdx.x;
  const int64_t GK = p.G * p.K;
  const int nw = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t = threadIdx.x, nt = blockDim.x;
  float* wpart = smem;
  float* lps = wpart + GK * nw;
  const float* xg = input + b * p.P;
  const float* rb = recons + b * GK * p.P;
  const bool vec = (p.P % 4 == 0) &&
                   (((reinterpret_cast<uintptr_t>(recons) | reinterpret_cast<uintptr_t>(input)) & 15) == 0);
  const int64_t nvec = vec ? (p.P >> 2) : 0;
  const float4* x4 = reinterpret_cast<const float4*>(xg);
  for (int64_t s0 = 0; s0 < GK; s0 += kSC) {
    const int ns = (int)((GK - s0 < kSC) ? (GK - s0) : kSC);
    float acc[kSC];
#pragma unroll
    for (int j = 0; j < kSC; ++j) acc[j] = 0.f;
    if (ns == kSC) {
      for (int64_t i = t; i < nvec; i += nt) {
        const float4 x = __ldg(x4 + i);
        float4 rv[kSC];
#pragma unroll
        for (int j = 0; j < kSC; ++j) rv[j] = ldg_stream(reinterpret_cast<const float4*>(rb + (s0 + j) * p.P) + i);
#pragma unroll
        for (int j = 0; j < kSC; ++j) {
          const float d0 = rv[j].x - x.x, d1 = rv[j].y - x.y, d2 = rv[j].z - x.z, d3 = rv[j].w - x.w;
          acc[j] += (d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3);
        }
      }
    } else {
      for (int64_t i = t; i < nvec; i += nt) {
        const float4 x = __ldg(x4 + i);
        float4 rv[kSC];
#pragma unroll
        for (int j = 0; j < kSC; ++j)
          if (j < ns) rv[j] = ldg_stream(reinterpret_cast<const float4*>(rb + (s0 + j) * p.P) + i);
#pragma unroll
        for (int j = 0; j < kSC; ++j)
          if (j < ns) {
            const float d0 = rv[j].x - x.x, d1 = rv[j].y - x.y, d2 = rv[j].z - x.z, d3 = rv[j].w - x.w;
            acc[j] += (d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3);
          }
      }
    }
    for (int64_t i = 4 * nvec + t; i < p.P; i += nt) {     // scalar tail / unaligned path
      const float x = xg[i];
#pragma unroll
      for (int j = 0; j < kSC; ++j)
        if (j < ns) { const float d = rb[(s0 + j) * p.P + i] - x; acc[j] += d * d; }
    }
#pragma unroll
    for (int j = 0; j < kSC; ++j) {
      const float v = warp_sum(acc[j]);
      if (lane == 0 && j < ns) wpart[(s0 + j) * nw + wid] = v;
    }
  }
  __syncthreads();
  const float invP = 1.0f / (float)p.P;
  for (int64_t s = threadIdx.x; s < GK; s += blockDim.x) {
    float tt = 0.f;
    for (int w = 0; w < nw; ++w) tt += wpart[s * nw + w];
    const float lp = tt * invP;
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
    double tt[3] = {0.0, 0.0, 0.0};
    for (int64_t i = threadIdx.x; i < p.B; i += blockDim.x) {
      tt[0] += partials[3 * i + 0];
      tt[1] += partials[3 * i + 1];
      tt[2] += partials[3 * i + 2];
    }
    __syncthreads();
    block_sum<3>(tt, red);
    if (threadIdx.x == 0) {
      const double rows = (double)(p.B * p.G), all = (double)(p.B * GK);
      out[0] = (float)(tt[0] / rows);
      out[1] = (float)(tt[1] / all);
      out[2] = (float)(tt[2] / all);
      out[3] = -(float)(tt[2] / all);
    }
  }
}

// grid = (B, splits): block (b, y) handles pixels [y*chunk, (y+1)*chunk) of every reconstruction of item b.
// dynamic smem: [GK floats coefficient]
__global__ void __launch_bounds__(256, 3)
iw_bound_bwd_kernel(const float* __restrict__ recons, const float* __restrict__ input, const float* __restrict__ lp,
                    const float* __restrict__ kl, IwParams p, int64_t chunk, const float* __restrict__ g_loss,
                    const float* __restrict__ g_recon, const float* __restrict__ g_kld,
                    const float* __restrict__ g_negkld, float* __restrict__ g_recons, float* __restrict__ g_kl) {
  extern __shared__ __align__(16) float smem[];
  const int64_t b = blockIdx.x;
  const int64_t GK = p.G * p.K;
  const int64_t p0 = (int64_t)blockIdx.y * chunk;
  const int64_t plen = (p.P - p0 < chunk) ? (p.P - p0) : chunk;
  float* coef = smem;
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
  __syncthreads();
  if (!g_recons || plen <= 0) return;
  const float* xg = input + b * p.P + p0;
  const bool vec = (p.P % 4 == 0) && (chunk % 4 == 0) &&
                   (((reinterpret_cast<uintptr_t>(recons) | reinterpret_cast<uintptr_t>(input) |
                      reinterpret_cast<uintptr_t>(g_recons)) & 15) == 0);
  const int t = threadIdx.x, nt = blockDim.x;
  const int64_t nvec = vec ? (plen >> 2) : 0;
  const float4* x4 = reinterpret_cast<const float4*>(xg);
  const float* rb = recons + b * GK * p.P + p0;
  float* gb = g_recons + b * GK * p.P + p0;
  for (int64_t s0 = 0; s0 < GK; s0 += kSC) {
    const int ns = (int)((GK - s0 < kSC) ? (GK - s0) : kSC);
    float c[kSC];
#pragma unroll
    for (int j = 0; j < kSC; ++j) c[j] = (j < ns) ? coef[s0 + j] : 0.f;
    for (int64_t i = t; i < nvec; i += nt) {
      const float4 x = __ldg(x4 + i);
      float4 rv[kSC];
#pragma unroll
      for (int j = 0; j < kSC; ++j)
        if (j < ns) rv[j] = ldg_stream(reinterpret_cast<const float4*>(rb + (s0 + j) * p.P) + i);
#pragma unroll
      for (int j = 0; j < kSC; ++j)
        if (j < ns)
          stg_stream(reinterpret_cast<float4*>(gb + (s0 + j) * p.P) + i,
                     make_float4(c[j] * (rv[j].x - x.x), c[j] * (rv[j].y - x.y), c[j] * (rv[j].z - x.z),
                                 c[j] * (rv[j].w - x.w)));
    }
    for (int64_t i = 4 * nvec + t; i < plen; i += nt) {
      const float x = xg[i];
#pragma unroll
      for (int j = 0; j < kSC; ++j)
        if (j < ns) gb[(s0 + j) * p.P + i] = c[j] * (rb[(s0 + j) * p.P + i] - x);
    }
  }
}

constexpr int64_t kIwMaxSmem = 200 * 1024;

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
  const int64_t smem = (GK * 8 + GK) * 4;
  VK_REQUIRE(smem <= kIwMaxSmem, VK_ERR_UNSUPPORTED, "iw_bound_fwd: G*K=%lld too large for shared memory", (long long)GK);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(workspace);
  double* partials = reinterpret_cast<double*>(reinterpret_cast<char*>(workspace) + kWsHeader);
  cudaStream_t s = (cudaStream_t)stream;
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(iw_bound_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  iw_bound_fwd_kernel<<<(unsigned)B, 256, smem, s>>>(recons, input, kl, p, lp_out, out, ticket, partials);
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
  const int64_t smem = GK * 4;
  VK_REQUIRE(smem <= kIwMaxSmem, VK_ERR_UNSUPPORTED, "iw_bound_bwd: G*K=%lld too large for shared memory", (long long)GK);
  dim3 grid((unsigned)B, (unsigned)splits);
  cudaStream_t s = (cudaStream_t)stream;
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(iw_bound_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  iw_bound_bwd_kernel<<<grid, 256, smem, s>>>(recons, input, lp, kl, p, chunk, g_loss, g_recon, g_kld, g_negkld, g_recons, g_kl);
  return check_launch("iw_bound_bwd");
}

// ---- allocation-free variant of the forward --------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  vk_iw_bound_fwd takes its scratch (the
// ticket header + 3 doubles per batch item) from a caller-provided workspace; this entry point keeps that scratch in a
// file-scope device array instead, so the caller passes inputs and outputs only.  The array is 256 + 24 * 16384 = 393 472
// bytes and every GPU that loads the library holds it until the process exits.
//
// In HOST code `g_iw_static_scratch` is only a stand-in for the symbol, not a device address (a kernel handed it would
// fault, or silently use host memory where the GPU can reach it): the address is resolved with cudaGetSymbolAddress once
// per device, with that device current, and cached.
// Every call on a device shares the one array: calls that overlap on the GPU (two streams, two host threads) race on it.
namespace vk {
namespace {
constexpr int64_t kIwStaticMaxB = VK_IW_STATIC_MAX_BATCH;
constexpr int kIwStaticMaxDevices = 64;
static __device__ __align__(256) unsigned char g_iw_static_scratch[kWsHeader + kIwStaticMaxB * 3 * sizeof(double)];

// device address of the array on the CURRENT device (resolved on first use per device; no allocation involved)
int iw_static_scratch(unsigned char** out) {
  static void* cached[kIwStaticMaxDevices] = {nullptr};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(VK_ERR_CUDA, "iw_bound_fwd_static: cudaGetDevice: %s", cudaGetErrorString(e));
  VK_REQUIRE(dev >= 0 && dev < kIwStaticMaxDevices, VK_ERR_UNSUPPORTED, "iw_bound_fwd_static: device index %d out of range", dev);
  void* p = __atomic_load_n(&cached[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    e = cudaGetSymbolAddress(&p, g_iw_static_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return fail(VK_ERR_CUDA, "iw_bound_fwd_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
    }
    __atomic_store_n(&cached[dev], p, __ATOMIC_RELEASE);      // the same value whichever thread resolves it first
  }
  *out = static_cast<unsigned char*>(p);
  return VK_OK;
}
}  // namespace
}  // namespace vk

extern "C" int vk_iw_bound_fwd_static(const float* recons, const float* input, const float* kl, int64_t B, int64_t G,
                                      int64_t K, int64_t P, float M_N, float* lp_out, float* out, void* stream) {
  VK_REQUIRE(recons && input && kl && lp_out && out, VK_ERR_BAD_ARG, "iw_bound_fwd_static: null pointer");
  VK_REQUIRE(B > 0 && G > 0 && K > 0 && P > 0, VK_ERR_BAD_ARG, "iw_bound_fwd_static: bad sizes");
  VK_REQUIRE(B <= kIwStaticMaxB, VK_ERR_UNSUPPORTED,
             "iw_bound_fwd_static: batch %lld exceeds the static scratch maximum of %lld (use vk_iw_bound_fwd with a workspace)",
             (long long)B, (long long)kIwStaticMaxB);
  const int64_t GK = G * K;
  IwParams p{B, G, K, P, M_N};
  const int64_t smem = (GK * 8 + GK) * 4;
  VK_REQUIRE(smem <= kIwMaxSmem, VK_ERR_UNSUPPORTED, "iw_bound_fwd_static: G*K=%lld too large for shared memory", (long long)GK);
  unsigned char* scratch = nullptr;
  int rc = iw_static_scratch(&scratch);
  if (rc != VK_OK) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  // the array keeps the last call's contents: the ticket header must start at zero (the kernel leaves it zeroed, but an
  // aborted launch would not); the partials are fully rewritten for the B blocks of this call before they are read
  cudaError_t e = cudaMemsetAsync(scratch, 0, kWsHeader, s);
  if (e != cudaSuccess) return fail(VK_ERR_CUDA, "iw_bound_fwd_static: cudaMemsetAsync: %s", cudaGetErrorString(e));
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(iw_bound_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  iw_bound_fwd_kernel<<<(unsigned)B, 256, smem, s>>>(recons, input, kl, p, lp_out, out,
                                                     reinterpret_cast<unsigned int*>(scratch),
                                                     reinterpret_cast<double*>(scratch + kWsHeader));
  return check_launch("iw_bound_fwd_static");
}
