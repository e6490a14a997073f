// K6: SWAE sliced-Wasserstein distance, compute_swd (reference: models/swae.py:151-179), value and d/dz in one call.
//
//   project   latT[s][n] = z[n,:] . theta_s,  priT[s][n] = prior_z[n,:] . theta_s   (fp32 FFMA GEMM, stored
//             projection-major so every projection is one contiguous sort segment; the reference's
//             z.matmul(proj) is [N,S] and is transposed before the sort, swae.py:176)
//   sort      segmented LSD radix sort (seg_sort.cuh): latT rows with their sample indices, priT rows key-only
//   loss      w = sort(lat) - sort(pri);  sum w^p  (deterministic two-stage reduction) and, through the sort
//             permutation, d(sum w^p)/d lat scattered back to sample order
//   grad      dz[n,:] = sum_s dlat[s][n] * theta_s                                  (fp32 FFMA GEMM)
//
// The projection count S may be a shard of the model's projections (multi-GPU: sharded by projection, the sums
// and the [N,D] gradients of the shards add up).
#include "common.cuh"
#include "seg_sort.cuh"
#include "gemm3x.cuh"

namespace vk {
namespace {

// ---- fp32 GEMM  C(m,n) = sum_k A(m,k) B(k,n)  with element strides, 128 x 64 x 16 tiles, 8 x 4 per thread -------
constexpr int BM = 128, BN = 64, BK = 16, GT = 256;

struct GemmArgs {
  const float* A; const float* B; float* C;
  int64_t M, N, K;
  int64_t a_sm, a_sk, b_sk, b_sn, c_sm, c_sn;
};

// A_KC: A is contiguous along k (row-major [M,K]); else contiguous along m.  B_KC: B contiguous along k; else along n.
// C_MC: C contiguous along m (transposed store); else along n.
template <bool A_KC, bool B_KC, bool C_MC>
__global__ void __launch_bounds__(GT)
sgemm_kernel(GemmArgs g) {
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int mi = C_MC ? tid % 16 : tid / 16, ni = C_MC ? tid / 16 : tid % 16;
  const int64_t m0 = (int64_t)blockIdx.x * BM, n0 = (int64_t)blockIdx.y * BN;
  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int64_t k0 = 0; k0 < g.K; k0 += BK) {
#pragma unroll
    for (int i = 0; i < BM * BK / GT; ++i) {
      const int e = tid + i * GT;
      const int k = A_KC ? e % BK : e / BM, m = A_KC ? e / BK : e % BM;
      const bool ok = (m0 + m < g.M) && (k0 + k < g.K);
      As[k][m] = ok ? __ldg(g.A + (m0 + m) * g.a_sm + (k0 + k) * g.a_sk) : 0.f;
    }
#pragma unroll
    for (int i = 0; i < BN * BK / GT; ++i) {
      const int e = tid + i * GT;
      const int k = B_KC ? e % BK : e / BN, n = B_KC ? e / BK : e % BN;
      const bool ok = (n0 + n < g.N) && (k0 + k < g.K);
      Bs[k][n] = ok ? __ldg(g.B + (k0 + k) * g.b_sk + (n0 + n) * g.b_sn) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[k][mi * 8]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[k][mi * 8 + 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[k][ni * 4]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float b[4] = {b0.x, b0.y, b0.z, b0.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  const int64_t mb = m0 + mi * 8, nb = n0 + ni * 4;
  if (C_MC) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (nb + j >= g.N) continue;
      float* c = g.C + (nb + j) * g.c_sn + mb;
      if (mb + 8 <= g.M && ((reinterpret_cast<uintptr_t>(c) & 15) == 0)) {
        *reinterpret_cast<float4*>(c) = make_float4(acc[0][j], acc[1][j], acc[2][j], acc[3][j]);
        *reinterpret_cast<float4*>(c + 4) = make_float4(acc[4][j], acc[5][j], acc[6][j], acc[7][j]);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) if (mb + i < g.M) c[i] = acc[i][j];
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (mb + i >= g.M) continue;
      float* c = g.C + (mb + i) * g.c_sm + nb;
      if (nb + 4 <= g.N && ((reinterpret_cast<uintptr_t>(c) & 15) == 0)) {
        *reinterpret_cast<float4*>(c) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) if (nb + j < g.N) c[j] = acc[i][j];
      }
    }
  }
}

// (the loss and its gradient through the permutation are fused into the last sort pass: segsort::LossFuse)

// out[0] = reg * (float)(sum / count)   (reg_weight * w_dist.mean(), swae.py:179)
__global__ void mean_scale_kernel(const double* sum, double count, float reg, float* out) {
  out[0] = reg * (float)(sum[0] / count);
}

struct SwdWs {
  size_t partials, keyZ0, keyZ1, idx0, idx1, keyP0, keyP1, counts, bimg, lossp, total;
};
constexpr size_t kBimgBytes = 3u * 256u * 256u * 2u;      // packed theta images of the tcgen05 GEMMs (upper bound)
inline SwdWs swd_ws(int64_t N, int64_t S) {
  SwdWs w;
  const size_t seg = ((size_t)S * N * 4 + 255) & ~(size_t)255;
  size_t o = kWsHeader;
  w.partials = o; o += (size_t)kMaxReduceBlocks * 8 * sizeof(double);
  o = (o + 255) & ~(size_t)255;
  w.keyZ0 = o; o += seg;
  w.keyZ1 = o; o += seg;
  w.idx0 = o; o += seg;
  w.idx1 = o; o += seg;
  w.keyP0 = o; o += seg;
  w.keyP1 = o; o += seg;
  w.counts = o; o += segsort::counts_bytes(S, N);
  o = (o + 255) & ~(size_t)255;
  w.bimg = o; o += kBimgBytes;
  w.lossp = o; o += (size_t)S * segsort::chunks_for(N) * sizeof(double);     // per-chunk partials of the fused loss
  w.total = o;
  return w;
}


// Sorts + loss + d/dlat for S projection-major segments of N keys each.  The prior goes first; the last pass of the
// latent sort (key + source index) then consumes the sorted prior and produces the loss partials and d/dlat in sample
// order (segsort::LossFuse) instead of storing its result: d/dlat lands in `lat` (the fused pass reads keyZ1 / idx1
// and would have written lat / idx0).  `pri` holds the sorted prior projections afterwards.
inline int swd_segments(float* lat, float* pri, uint32_t* keyZ1, uint32_t* idx0, uint32_t* idx1, uint32_t* keyP1,
                        uint32_t* counts, double* lossp, int64_t N, int64_t S, float p, double* sum_out, bool want_grad,
                        cudaStream_t st) {
  int rc = segsort::sort<false>((uint32_t*)pri, keyP1, nullptr, nullptr, S, N, counts, st);
  if (rc != VK_OK) return rc;
  const segsort::LossFuse fuse{pri, want_grad ? lat : nullptr, lossp, p};
  rc = segsort::sort<true>((uint32_t*)lat, keyZ1, idx0, idx1, S, N, counts, st, &fuse);
  if (rc != VK_OK) return rc;
  segsort::sum_partials_kernel<<<1, 1024, 0, st>>>(lossp, S * (int64_t)segsort::chunks_for(N), sum_out);
  return check_launch("swd loss sum");
}

struct SegWs {
  size_t keyZ1, idx0, idx1, keyP1, counts, lossp, total;
};
inline SegWs seg_ws(int64_t N, int64_t S) {
  SegWs w;
  const size_t seg = ((size_t)S * N * 4 + 255) & ~(size_t)255;
  size_t o = kWsHeader;
  w.keyZ1 = o; o += seg;
  w.idx0 = o; o += seg;
  w.idx1 = o; o += seg;
  w.keyP1 = o; o += seg;
  w.counts = o; o += segsort::counts_bytes(S, N);
  o = (o + 255) & ~(size_t)255;
  w.lossp = o; o += (size_t)S * segsort::chunks_for(N) * sizeof(double);
  w.total = o;
  return w;
}

}  // namespace
}  // namespace vk

using namespace vk;

extern "C" int64_t vk_swd_workspace_bytes(int64_t N, int64_t S) {
  if (N < 1 || S < 1) return -1;
  return (int64_t)swd_ws(N, S).total;
}

extern "C" int vk_swd_fwd(const float* z, const float* prior, const float* proj, int64_t N, int64_t D, int64_t S, float p,
                          double* sum_out, float* grad_unscaled, void* workspace, int64_t workspace_bytes, void* stream) {
  VK_REQUIRE(z && prior && proj && sum_out && workspace, VK_ERR_BAD_ARG, "swd_fwd: null pointer");
  VK_REQUIRE(N >= 1 && D >= 1 && S >= 1, VK_ERR_BAD_ARG, "swd_fwd: empty shape N=%lld D=%lld S=%lld",
             (long long)N, (long long)D, (long long)S);
  VK_REQUIRE(N < ((int64_t)1 << 32) && S <= 65535, VK_ERR_UNSUPPORTED, "swd_fwd: N=%lld or S=%lld too large",
             (long long)N, (long long)S);
  const SwdWs w = swd_ws(N, S);
  VK_REQUIRE(workspace_bytes >= (int64_t)w.total, VK_ERR_WORKSPACE, "swd_fwd: workspace %lld < %lld",
             (long long)workspace_bytes, (long long)w.total);
  cudaStream_t st = (cudaStream_t)stream;
  uint8_t* ws = (uint8_t*)workspace;
  float* latT = (float*)(ws + w.keyZ0);
  float* priT = (float*)(ws + w.keyP0);
  // 1. projections, stored [S][N]
  const dim3 pg((unsigned)((N + BM - 1) / BM), (unsigned)((S + BN - 1) / BN));
  int rc;
  if (g3::fits((int)S, (int)D)) {
    // tcgen05 path: latT(n, s) = sum_d z(n, d) theta(s, d), six bf16 passes of a three-way split (fp32-grade keys)
    rc = g3::launch(z, D, 1, proj, D, 1, latT, 1, N, N, (int)S, (int)D, ws + w.bimg, st, "swd project (tcgen05)");
    if (rc != VK_OK) return rc;
    rc = g3::launch(prior, D, 1, proj, D, 1, priT, 1, N, N, (int)S, (int)D, ws + w.bimg, st, "swd project (tcgen05)");
  } else {
    GemmArgs ga{z, proj, latT, N, S, D, D, 1, 1, D, 1, N};
    sgemm_kernel<true, true, true><<<pg, GT, 0, st>>>(ga);
    GemmArgs gp{prior, proj, priT, N, S, D, D, 1, 1, D, 1, N};
    sgemm_kernel<true, true, true><<<pg, GT, 0, st>>>(gp);
    rc = check_launch("swd project");
  }
  if (rc != VK_OK) return rc;
  // 2.+3. sorts, loss and d/dlat (lands in keyZ0, sample order)
  rc = swd_segments(latT, priT, (uint32_t*)(ws + w.keyZ1), (uint32_t*)(ws + w.idx0), (uint32_t*)(ws + w.idx1),
                    (uint32_t*)(ws + w.keyP1), (uint32_t*)(ws + w.counts), (double*)(ws + w.lossp), N, S, p, sum_out,
                    grad_unscaled != nullptr, st);
  if (rc != VK_OK) return rc;
  float* g_lat = grad_unscaled ? latT : nullptr;
  // 4. dz = g_lat^T theta
  if (grad_unscaled) {
    if (g3::fits((int)D, (int)S)) {
      // dz(n, d) = sum_s g(s, n) theta(s, d):  A(n, k = s) = g_lat[s * N + n],  B(j = d, k = s) = theta[s * D + d]
      rc = g3::launch(g_lat, 1, N, proj, 1, D, grad_unscaled, D, 1, N, (int)D, (int)S, ws + w.bimg, st,
                      "swd grad gemm (tcgen05)");
    } else {
      const dim3 gg((unsigned)((N + BM - 1) / BM), (unsigned)((D + BN - 1) / BN));
      GemmArgs gb{g_lat, proj, grad_unscaled, N, D, S, 1, N, D, 1, D, 1};
      sgemm_kernel<false, false, false><<<gg, GT, 0, st>>>(gb);
      rc = check_launch("swd grad gemm");
    }
  }
  return rc;
}

extern "C" int64_t vk_swd_segments_workspace_bytes(int64_t N, int64_t S) {
  if (N < 1 || S < 1) return -1;
  return (int64_t)seg_ws(N, S).total;
}

extern "C" int vk_swd_segments_fwd(float* lat, float* pri, int64_t N, int64_t S, float p, double* sum_out, int want_grad,
                                   void* workspace, int64_t workspace_bytes, void* stream) {
  VK_REQUIRE(lat && pri && sum_out && workspace, VK_ERR_BAD_ARG, "swd_segments_fwd: null pointer");
  VK_REQUIRE(N >= 1 && S >= 1, VK_ERR_BAD_ARG, "swd_segments_fwd: empty shape N=%lld S=%lld", (long long)N, (long long)S);
  VK_REQUIRE(N < ((int64_t)1 << 32) && S <= 65535, VK_ERR_UNSUPPORTED, "swd_segments_fwd: N=%lld or S=%lld too large",
             (long long)N, (long long)S);
  const SegWs w = seg_ws(N, S);
  VK_REQUIRE(workspace_bytes >= (int64_t)w.total, VK_ERR_WORKSPACE, "swd_segments_fwd: workspace %lld < %lld",
             (long long)workspace_bytes, (long long)w.total);
  uint8_t* ws = (uint8_t*)workspace;
  return swd_segments(lat, pri, (uint32_t*)(ws + w.keyZ1), (uint32_t*)(ws + w.idx0), (uint32_t*)(ws + w.idx1),
                      (uint32_t*)(ws + w.keyP1), (uint32_t*)(ws + w.counts), (double*)(ws + w.lossp), N, S, p, sum_out,
                      want_grad != 0, (cudaStream_t)stream);
}

extern "C" int64_t vk_gemm_nt_workspace_bytes(void) { return (int64_t)(kWsHeader + kBimgBytes); }

extern "C" int vk_gemm_nt(const float* A, int64_t a_sn, int64_t a_sk, const float* B, int64_t b_sj, int64_t b_sk, float* C,
                          int64_t c_sn, int64_t c_sj, int64_t N, int64_t J, int64_t K, void* workspace,
                          int64_t workspace_bytes, void* stream) {
  VK_REQUIRE(A && B && C && workspace, VK_ERR_BAD_ARG, "gemm_nt: null pointer");
  VK_REQUIRE(N >= 1 && J >= 1 && K >= 1, VK_ERR_BAD_ARG, "gemm_nt: empty shape N=%lld J=%lld K=%lld", (long long)N,
             (long long)J, (long long)K);
  VK_REQUIRE(J <= 256 && K <= 256 && g3::fits((int)J, (int)K), VK_ERR_UNSUPPORTED,
             "gemm_nt: J=%lld, K=%lld do not fit the tensor-memory / shared-memory budget of the tcgen05 GEMM", (long long)J,
             (long long)K);
  VK_REQUIRE(workspace_bytes >= vk_gemm_nt_workspace_bytes(), VK_ERR_WORKSPACE, "gemm_nt: workspace too small");
  return g3::launch(A, a_sn, a_sk, B, b_sj, b_sk, C, c_sn, c_sj, N, (int)J, (int)K, (uint8_t*)workspace + kWsHeader,
                    (cudaStream_t)stream, "gemm_nt");
}

extern "C" int vk_gemm_nt_tab(const float* const* a_cols, const float* A, int64_t a_sn, int64_t a_sk, const float* B,
                              int64_t b_sj, int64_t b_sk, float* const* c_cols, float* C, int64_t c_sn, int64_t c_sj,
                              int64_t N, int64_t J, int64_t K, void* workspace, int64_t workspace_bytes, void* stream) {
  VK_REQUIRE((A || a_cols) && B && (C || c_cols) && workspace, VK_ERR_BAD_ARG, "gemm_nt_tab: null pointer");
  VK_REQUIRE(N >= 1 && J >= 1 && K >= 1, VK_ERR_BAD_ARG, "gemm_nt_tab: empty shape N=%lld J=%lld K=%lld", (long long)N,
             (long long)J, (long long)K);
  VK_REQUIRE(J <= 256 && K <= 256 && g3::fits((int)J, (int)K), VK_ERR_UNSUPPORTED,
             "gemm_nt_tab: J=%lld, K=%lld do not fit the tensor-memory / shared-memory budget of the tcgen05 GEMM",
             (long long)J, (long long)K);
  VK_REQUIRE(workspace_bytes >= vk_gemm_nt_workspace_bytes(), VK_ERR_WORKSPACE, "gemm_nt_tab: workspace too small");
  return g3::launch(A, a_sn, a_sk, B, b_sj, b_sk, C, c_sn, c_sj, N, (int)J, (int)K, (uint8_t*)workspace + kWsHeader,
                    (cudaStream_t)stream, "gemm_nt_tab", a_cols, c_cols);
}

extern "C" int vk_mean_scale(const double* sum, double count, float scale, float* out, void* stream) {
  VK_REQUIRE(sum && out && count > 0, VK_ERR_BAD_ARG, "mean_scale: bad argument");
  mean_scale_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(sum, count, scale, out);
  return check_launch("mean_scale");
}

// Stand-alone segmented sort entry point (tests, and callers that only need the sort): keys [segments, n] fp32 are
// sorted ascending in place; perm_out [segments, n] (uint32, may be NULL) receives the source index of every
// sorted element.  scratch must hold vk_seg_sort_workspace_bytes(segments, n).
extern "C" int64_t vk_seg_sort_workspace_bytes(int64_t segments, int64_t n) {
  if (segments < 1 || n < 1) return -1;
  const size_t seg = ((size_t)segments * n * 4 + 255) & ~(size_t)255;
  return (int64_t)(kWsHeader + 2 * seg + segsort::counts_bytes(segments, n) + 256);
}
extern "C" int vk_seg_sort(float* keys, uint32_t* perm_out, int64_t segments, int64_t n, void* workspace,
                           int64_t workspace_bytes, void* stream) {
  VK_REQUIRE(keys && workspace && segments >= 1 && n >= 1, VK_ERR_BAD_ARG, "seg_sort: bad argument");
  VK_REQUIRE(n < ((int64_t)1 << 32) && segments <= 65535, VK_ERR_UNSUPPORTED, "seg_sort: shape too large");
  VK_REQUIRE(workspace_bytes >= vk_seg_sort_workspace_bytes(segments, n), VK_ERR_WORKSPACE, "seg_sort: workspace too small");
  const size_t seg = ((size_t)segments * n * 4 + 255) & ~(size_t)255;
  uint8_t* ws = (uint8_t*)workspace + kWsHeader;          // the ticket header of a shared workspace stays zero
  uint32_t* key1 = (uint32_t*)ws;
  uint32_t* val1 = (uint32_t*)(ws + seg);
  uint32_t* counts = (uint32_t*)(ws + 2 * seg);
  if (perm_out) return segsort::sort<true>((uint32_t*)keys, key1, perm_out, val1, segments, n, counts, (cudaStream_t)stream);
  return segsort::sort<false>((uint32_t*)keys, key1, nullptr, nullptr, segments, n, counts, (cudaStream_t)stream);
}
