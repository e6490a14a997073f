// K5: Beta-TC-VAE minibatch-stratified total-correlation terms
// (reference: models/betatc_vae.py:132-142 log_density_gaussian, :165-211 loss_function).
//
// For every pair (i, j) and latent dim d the reference materialises  ld_ijd = log N(z_id; mu_jd, exp(lv_jd))  as a
// [B,B,D] tensor, adds the [B,B] log importance weights to EVERY latent dim (so they count D times in log_q_z),
// and takes  log_q_z = LSE_j sum_d,  log_prod_q_z = sum_d LSE_j.  Every one of the B^2 D densities needs its own
// exp for the per-dimension LSE, so the path is bound by the MUFU/FP32 pipes, not by a contraction: the sum over
// d (the "Gram" half) rides along at one FADD per density and no tensor-core formulation saves work.
//
// Kernels (anchor = one warp, partners = lanes x 4, latent dims in chunks of 32 staged in shared memory):
//   tc_prep        h = 0.5 exp(-lv), c0 = -0.5 (log 2pi + lv)
//   tc_fwd         anchor row i: online LSE over j of A_ij = sum_d ld + D lw_ij and, per d, of ld + lw_ij;
//                  row terms log q(z|x), log p(z); deterministic two-stage reduction -> {mi, tc, kld, weighted sum}
//   tc_bwd<ROWS>   anchor row i  -> d/dz_i ;  tc_bwd<COLS> anchor column j -> d/dmu_j, d/dlog_var_j
//                  (softmax weights p_ij, r_ijd recomputed from the saved log_q_z[i] and LSE_j[i,d])
#include <stdlib.h>
#include "common.cuh"

namespace vk {
namespace {

constexpr int kW = 8;             // anchors (warps) per CTA
constexpr int kJT = 4;            // partners per lane
constexpr int kPT = 32 * kJT;     // partners per tile: 128
constexpr int kDC = 32;           // latent dims per chunk
constexpr int kLD = kDC + 1;      // padded row of a tile
constexpr float kLog2Pi = 1.8378770664093453f;
constexpr float kLog2e = 1.4426950408889634f;

struct TcParams {
  int B, D;
  int r0, r1;                      // anchor rows / columns this call covers (a rank's row block; [0, B) on one GPU)
  float lw_ds, lw_strat, lw_bm1;   // log importance weights: column 0, column 1 (and [B-2,0]), everything else
  float alpha, beta, gamma_anneal; // loss = alpha*mi + (beta*tc + anneal*gamma*kld)
};

__device__ __forceinline__ float log_w(const TcParams& p, int i, int j) {
  if (j == 0) return i == p.B - 2 ? p.lw_strat : p.lw_ds;      // betatc_vae.py:180,182
  if (j == 1) return p.lw_strat;                               // betatc_vae.py:181
  return p.lw_bm1;                                             // betatc_vae.py:179
}
__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void tc_prep_kernel(const float* __restrict__ lv, int64_t n, float* __restrict__ h, float* __restrict__ c0) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = lv[i];
    h[i] = 0.5f * expf(-v);
    c0[i] = -0.5f * (kLog2Pi + v);
  }
}

// cooperative load of a [kPT partners x kDC dims] tile of a row-major [B, D] array into smem [kPT][kLD]
__device__ __forceinline__ void load_tile(float* dst, const float* __restrict__ src, int B, int D, int p0, int d0, float fill) {
  for (int e = threadIdx.x; e < kPT * kDC; e += blockDim.x) {
    const int r = e / kDC, c = e % kDC;
    const int row = p0 + r, col = d0 + c;
    dst[r * kLD + c] = (row < B && col < D) ? __ldg(src + (size_t)row * D + col) : fill;
  }
}

// ---- forward ------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32 * kW)
tc_fwd_kernel(const float* __restrict__ Z, const float* __restrict__ MU, const float* __restrict__ LV,
              const float* __restrict__ H, const float* __restrict__ C0, TcParams p, float* __restrict__ lqz_out,
              float* __restrict__ lse_out, double* __restrict__ partials, unsigned int* ticket, float* __restrict__ out) {
  extern __shared__ float sm[];
  float* Tmu = sm;                        // [kPT][kLD]
  float* Th = Tmu + kPT * kLD;
  float* Tc = Th + kPT * kLD;
  float* Zs = Tc + kPT * kLD;             // [kW][Dpad]
  const int Dpad = (p.D + kDC - 1) / kDC * kDC;
  float* Ms = Zs + kW * Dpad;             // running max  [kW][Dpad]
  float* Ss = Ms + kW * Dpad;             // running sum  [kW][Dpad]
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = p.r0 + blockIdx.x * kW + w;      // anchor rows [r0, r1): a rank's row block, [0, B) on one GPU
  const bool alive = i < p.r1;
  for (int d = lane; d < Dpad; d += 32) {
    Zs[w * Dpad + d] = (alive && d < p.D) ? Z[(size_t)i * p.D + d] : 0.f;
    Ms[w * Dpad + d] = -INFINITY;
    Ss[w * Dpad + d] = 0.f;
  }
  float mrow = -INFINITY, srow = 0.f;
  for (int j0 = 0; j0 < p.B; j0 += kPT) {
    float A[kJT], lw[kJT];
#pragma unroll
    for (int k = 0; k < kJT; ++k) {
      const int j = j0 + lane + 32 * k;
      A[k] = 0.f;
      lw[k] = j < p.B ? log_w(p, i, j) : -INFINITY;
    }
    for (int d0 = 0; d0 < p.D; d0 += kDC) {
      __syncthreads();
      load_tile(Tmu, MU, p.B, p.D, j0, d0, 0.f);
      load_tile(Th, H, p.B, p.D, j0, d0, 0.f);
      load_tile(Tc, C0, p.B, p.D, j0, d0, 0.f);
      __syncthreads();
      const int dn = p.D - d0 < kDC ? p.D - d0 : kDC;
      for (int c = 0; c < dn; ++c) {
        const float zd = Zs[w * Dpad + d0 + c];
        float E[kJT], lmax = -INFINITY;
#pragma unroll
        for (int k = 0; k < kJT; ++k) {
          const int r = (lane + 32 * k) * kLD + c;
          const float df = zd - Tmu[r];
          const float ld = fmaf(-df * df, Th[r], Tc[r]);          // -0.5(log 2pi + lv) - 0.5 (z-mu)^2 exp(-lv)
          A[k] += ld;
          E[k] = ld + lw[k];
          lmax = fmaxf(lmax, E[k]);
        }
        const float wmax = warp_max(lmax);
        const float Mo = Ms[w * Dpad + d0 + c];
        const float Mn = fmaxf(Mo, wmax);
        float e = 0.f;
#pragma unroll
        for (int k = 0; k < kJT; ++k) e += ex2f((E[k] - Mn) * kLog2e);
        e = warp_sum(e);
        if (lane == 0) {
          Ss[w * Dpad + d0 + c] = Ss[w * Dpad + d0 + c] * ex2f((Mo - Mn) * kLog2e) + e;
          Ms[w * Dpad + d0 + c] = Mn;
        }
      }
    }
    // A_ij complete for this partner tile: online LSE over j of sum_d ld + D * lw   (betatc_vae.py:185-187)
    float lmax = -INFINITY;
#pragma unroll
    for (int k = 0; k < kJT; ++k) { A[k] += (float)p.D * lw[k]; lmax = fmaxf(lmax, A[k]); }
    const float wmax = warp_max(lmax);
    const float mn = fmaxf(mrow, wmax);
    float e = 0.f;
#pragma unroll
    for (int k = 0; k < kJT; ++k) e += ex2f((A[k] - mn) * kLog2e);
    e = warp_sum(e);
    srow = srow * ex2f((mrow - mn) * kLog2e) + e;
    mrow = mn;
  }
  __syncwarp();
  // row terms
  double t_qzx = 0.0, t_qz = 0.0, t_pq = 0.0, t_pz = 0.0;
  if (alive) {
    const float lqz = mrow + logf(srow);
    float lpq = 0.f, lqzx = 0.f, lpz = 0.f;
    for (int d = lane; d < p.D; d += 32) {
      const float l = Ms[w * Dpad + d] + logf(Ss[w * Dpad + d]);
      lse_out[(size_t)i * p.D + d] = l;
      lpq += l;
      const float z = Zs[w * Dpad + d];
      const float df = z - MU[(size_t)i * p.D + d];
      lqzx += fmaf(-df * df, H[(size_t)i * p.D + d], C0[(size_t)i * p.D + d]);
      lpz += -0.5f * kLog2Pi - 0.5f * z * z;
    }
    lpq = warp_sum(lpq); lqzx = warp_sum(lqzx); lpz = warp_sum(lpz);
    if (lane == 0) { lqz_out[i] = lqz; lqz_out[p.B + i] = 0.f; }   // [B, 2B): fp32 residual (fast path only)
    t_qzx = lqzx; t_qz = lqz; t_pq = lpq; t_pz = lpz;
  }
  __shared__ double red[kW * 4];
  if (lane == 0) { red[w * 4 + 0] = t_qzx; red[w * 4 + 1] = t_qz; red[w * 4 + 2] = t_pq; red[w * 4 + 3] = t_pz; }
  __syncthreads();
  if (threadIdx.x < 4) {
    double t = 0.0;
    for (int k = 0; k < kW; ++k) t += red[k * 4 + threadIdx.x];
    partials[(size_t)blockIdx.x * 4 + threadIdx.x] = t;
  }
  if (last_block_ticket(ticket, gridDim.x)) {
    if (threadIdx.x < 4) {
      double t = 0.0;
      for (unsigned int b = 0; b < gridDim.x; ++b) t += ((volatile double*)partials)[(size_t)b * 4 + threadIdx.x];
      red[threadIdx.x] = t / (double)p.B;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      const float qzx = (float)red[0], qz = (float)red[1], pq = (float)red[2], pz = (float)red[3];
      const float mi = qzx - qz, tc = qz - pq, kld = pq - pz;                 // betatc_vae.py:190-192 (means)
      out[0] = p.alpha * mi + (p.beta * tc + p.gamma_anneal * kld);           // the latent part of betatc_vae.py:202-205
      out[1] = kld;
      out[2] = tc;
      out[3] = mi;
    }
  }
}

// ---- backward -----------------------------------------------------------------------------------------------------
// L = (1/B) sum_i [ alpha lqzx_i + (beta - alpha) lqz_i + (ga - beta) lpq_i - ga lpz_i ],  ga = anneal * gamma
//   d/d ld_ijd = [ (beta - alpha) p_ij + (ga - beta) r_ijd ] / B,   p = softmax_j(A), r = softmax_j(ld + lw) per d
template <bool ROWS>
__global__ void __launch_bounds__(32 * kW)
tc_bwd_kernel(const float* __restrict__ Z, const float* __restrict__ MU, const float* __restrict__ H,
              const float* __restrict__ C0, const float* __restrict__ lqz, const float* __restrict__ lse, TcParams p,
              const float* __restrict__ g_loss, float* __restrict__ g_z, float* __restrict__ g_mu, float* __restrict__ g_lv) {
  extern __shared__ float sm[];
  // ROWS: partner tiles hold (mu, h, c0) of columns j ; COLS: partner tiles hold (z, lse) of rows i
  float* T0 = sm;
  float* T1 = T0 + kPT * kLD;
  float* T2 = T1 + kPT * kLD;
  const int Dpad = (p.D + kDC - 1) / kDC * kDC;
  float* A0 = T2 + kPT * kLD;             // anchor vectors [kW][Dpad] x 3
  float* A1 = A0 + kW * Dpad;
  float* A2 = A1 + kW * Dpad;
  float* G0 = A2 + kW * Dpad;             // gradient accumulators [kW][Dpad] x 2
  float* G1 = G0 + kW * Dpad;
  float* Tq = G1 + kW * Dpad;             // COLS: lqz of the partner rows [kPT]
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int a = p.r0 + blockIdx.x * kW + w;      // anchor index (row i or column j) inside [r0, r1)
  const bool alive = a < p.r1;
  const float g = g_loss ? *g_loss : 1.f;
  const float invB = 1.f / (float)p.B;
  const float cq = (p.beta - p.alpha) * invB * g, cp = (p.gamma_anneal - p.beta) * invB * g;
  for (int d = lane; d < Dpad; d += 32) {
    const bool ok = alive && d < p.D;
    const size_t o = (size_t)a * p.D + d;
    if (ROWS) { A0[w * Dpad + d] = ok ? Z[o] : 0.f; A1[w * Dpad + d] = ok ? lse[o] : 0.f; A2[w * Dpad + d] = 0.f; }
    else { A0[w * Dpad + d] = ok ? MU[o] : 0.f; A1[w * Dpad + d] = ok ? H[o] : 0.f; A2[w * Dpad + d] = ok ? C0[o] : 0.f; }
    G0[w * Dpad + d] = 0.f;
    G1[w * Dpad + d] = 0.f;
  }
  const float lqz_a = (ROWS && alive) ? lqz[a] : 0.f;
  for (int p0 = 0; p0 < p.B; p0 += kPT) {
    float A[kJT], lw[kJT], pij[kJT];
#pragma unroll
    for (int k = 0; k < kJT; ++k) {
      const int b = p0 + lane + 32 * k;
      A[k] = 0.f;
      lw[k] = (b < p.B && alive) ? (ROWS ? log_w(p, a, b) : log_w(p, b, a)) : -INFINITY;
    }
    if (!ROWS) {
      __syncthreads();
      for (int e = threadIdx.x; e < kPT; e += blockDim.x) Tq[e] = p0 + e < p.B ? lqz[p0 + e] : 0.f;
    }
    // pass 1: A_ij = sum_d ld_ijd
    for (int d0 = 0; d0 < p.D; d0 += kDC) {
      __syncthreads();
      if (ROWS) { load_tile(T0, MU, p.B, p.D, p0, d0, 0.f); load_tile(T1, H, p.B, p.D, p0, d0, 0.f); load_tile(T2, C0, p.B, p.D, p0, d0, 0.f); }
      else { load_tile(T0, Z, p.B, p.D, p0, d0, 0.f); }
      __syncthreads();
      const int dn = p.D - d0 < kDC ? p.D - d0 : kDC;
      for (int c = 0; c < dn; ++c) {
        const int ad = w * Dpad + d0 + c;
#pragma unroll
        for (int k = 0; k < kJT; ++k) {
          const int r = (lane + 32 * k) * kLD + c;
          const float df = ROWS ? A0[ad] - T0[r] : T0[r] - A0[ad];
          A[k] += ROWS ? fmaf(-df * df, T1[r], T2[r]) : fmaf(-df * df, A1[ad], A2[ad]);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < kJT; ++k) {
      const float q = ROWS ? lqz_a : Tq[lane + 32 * k];
      pij[k] = ex2f((A[k] + (float)p.D * lw[k] - q) * kLog2e);              // softmax weight of (i, j) in log_q_z[i]
    }
    // pass 2: per-dimension weights and gradient contributions
    for (int d0 = 0; d0 < p.D; d0 += kDC) {
      __syncthreads();
      if (ROWS) { load_tile(T0, MU, p.B, p.D, p0, d0, 0.f); load_tile(T1, H, p.B, p.D, p0, d0, 0.f); load_tile(T2, C0, p.B, p.D, p0, d0, 0.f); }
      else { load_tile(T0, Z, p.B, p.D, p0, d0, 0.f); load_tile(T1, lse, p.B, p.D, p0, d0, 0.f); }
      __syncthreads();
      const int dn = p.D - d0 < kDC ? p.D - d0 : kDC;
      for (int c = 0; c < dn; ++c) {
        const int ad = w * Dpad + d0 + c;
        float s0 = 0.f, s1 = 0.f;
#pragma unroll
        for (int k = 0; k < kJT; ++k) {
          const int r = (lane + 32 * k) * kLD + c;
          float df, h, ld, l;
          if (ROWS) { df = A0[ad] - T0[r]; h = T1[r]; ld = fmaf(-df * df, h, T2[r]); l = A1[ad]; }
          else { df = T0[r] - A0[ad]; h = A1[ad]; ld = fmaf(-df * df, h, A2[ad]); l = T1[r]; }
          const float rr = ex2f((ld + lw[k] - l) * kLog2e);                 // softmax weight inside LSE_j[i, d]
          const float cf = cq * pij[k] + cp * rr;                          // dL / d ld_ijd  (0 for masked partners)
          const float t = 2.f * h * df;                                    // (z - mu) exp(-lv)
          if (ROWS) { s0 -= cf * t; }
          else { s0 += cf * t; s1 += cf * (df * df * h - 0.5f); }
        }
        s0 = warp_sum(s0);
        if (!ROWS) s1 = warp_sum(s1);
        if (lane == 0) { G0[ad] += s0; if (!ROWS) G1[ad] += s1; }
      }
    }
  }
  __syncwarp();
  if (alive) {
    const float ca = p.alpha * invB * g, cz = p.gamma_anneal * invB * g;
    for (int d = lane; d < p.D; d += 32) {
      const size_t o = (size_t)a * p.D + d;
      const float z = Z[o], df = z - MU[o], h = H[o];
      const float t = 2.f * h * df;
      if (ROWS) {
        // diagonal terms: alpha/B * d log q(z_i|x_i)/dz  and  -ga/B * d log p(z_i)/dz = +ga/B * z
        g_z[o] = G0[w * Dpad + d] - ca * t + cz * z;
      } else {
        g_mu[o] = G0[w * Dpad + d] + ca * t;
        g_lv[o] = G1[w * Dpad + d] + ca * (df * df * h - 0.5f);
      }
    }
  }
}

inline size_t tc_smem_fwd(int D) { const int Dpad = (D + kDC - 1) / kDC * kDC; return (size_t)(3 * kPT * kLD + 3 * kW * Dpad) * 4; }
inline size_t tc_smem_bwd(int D) { const int Dpad = (D + kDC - 1) / kDC * kDC; return (size_t)(3 * kPT * kLD + 5 * kW * Dpad + kPT) * 4; }

int make_params(int64_t B, int64_t D, float M_N, float alpha, float beta, float gamma_anneal, TcParams& p) {
  VK_REQUIRE(B >= 2 && D >= 1, VK_ERR_BAD_ARG, "betatc: needs batch >= 2 and latent dim >= 1 (got %lld, %lld)", (long long)B, (long long)D);
  VK_REQUIRE(B <= (1 << 20) && D <= 1024, VK_ERR_UNSUPPORTED, "betatc: B=%lld or D=%lld too large", (long long)B, (long long)D);
  p.B = (int)B; p.D = (int)D;
  p.r0 = 0; p.r1 = (int)B;
  // betatc_vae.py:177-183: the table is fp32; its log is taken in fp32
  const double ds = (1.0 / (double)M_N) * (double)B;
  const double strat = (ds - (double)B + 1.0) / (ds * (double)(B - 1));
  p.lw_ds = logf((float)(1.0 / ds));
  p.lw_strat = logf((float)strat);
  p.lw_bm1 = logf((float)(1.0 / (double)(B - 1)));
  p.alpha = alpha; p.beta = beta; p.gamma_anneal = gamma_anneal;
  return VK_OK;
}

#include "betatc_tiled.cuh"

}  // namespace
}  // namespace vk

using namespace vk;

extern "C" int64_t vk_tc_workspace_bytes(int64_t B, int64_t D) {
  if (B < 2 || D < 1) return -1;
  const int64_t blocks = (B + kW - 1) / kW;
  // header | per-CTA partials | H, C0 [B, D] | Cb [B] doubles | partner-split partial sums of the backward (3 x [B, D])
  return (int64_t)kWsHeader + blocks * 4 * 8 + 256 + 2 * B * D * 4 + 256 + B * 8 + 256 + 3 * B * D * 4 + 256 + 16 * B;
}

static int tc_fwd_impl(const float* z, const float* mu, const float* log_var, int64_t B, int64_t D, int64_t row_begin,
                       int64_t row_end, float M_N, float alpha, float beta, float gamma_anneal, float* lqz_out,
                       float* lse_out, float* out, void* workspace, int64_t workspace_bytes, void* stream) {
  VK_REQUIRE(z && mu && log_var && lqz_out && lse_out && out && workspace, VK_ERR_BAD_ARG, "tc_fwd: null pointer");
  TcParams p;
  int rc = make_params(B, D, M_N, alpha, beta, gamma_anneal, p);
  if (rc != VK_OK) return rc;
  VK_REQUIRE(0 <= row_begin && row_begin < row_end && row_end <= B, VK_ERR_BAD_ARG, "tc_fwd: bad row range [%lld, %lld) of %lld",
             (long long)row_begin, (long long)row_end, (long long)B);
  p.r0 = (int)row_begin; p.r1 = (int)row_end;
  VK_REQUIRE(workspace_bytes >= vk_tc_workspace_bytes(B, D), VK_ERR_WORKSPACE, "tc_fwd: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  uint8_t* ws = (uint8_t*)workspace;
  const int blocks = (int)((B + kW - 1) / kW);
  double* partials = (double*)(ws + kWsHeader);
  float* H = (float*)(ws + kWsHeader + (((size_t)blocks * 32 + 255) & ~(size_t)255));
  float* C0 = H + B * D;
  const int prep_blocks = (int)((B * D + 255) / 256 < (int64_t)sm_count() * 8 ? (B * D + 255) / 256 : (int64_t)sm_count() * 8);
  if (D <= 128) {
    // lanes-own-dims path (betatc_tiled.cuh)
    // D <= 32: one dim per lane; <= 64: two; <= 128: four.  (Two warps per anchor group with two dims per lane each --
    // NS = 2 -- doubles the warps but the per-step exchange barrier costs more than it hides: 5.1 -> 7.4 ms, measured.)
    const int cfg = D <= 32 ? 0 : (D <= 64 ? 1 : 2);
    const int DP = cfg == 0 ? 32 : (cfg == 1 ? 64 : 128), NS = 1;
    const int per_cta = tiled::kTW / NS * tiled::kR;
    const int tb = (int)((p.r1 - p.r0 + per_cta - 1) / per_cta);
    const size_t tsm = tiled::smem_bytes(DP);
    double* Cb = (double*)(((uintptr_t)(C0 + B * D) + 255) & ~(uintptr_t)255);
    tiled::prep_kernel<<<prep_blocks, 256, 0, st>>>(log_var, p, H, C0);
    tiled::cbar_kernel<<<(int)((B * 32 + 255) / 256), 256, 0, st>>>(C0, p, Cb);
    auto kern = cfg == 0 ? tiled::fwd_kernel<1, 1> : (cfg == 1 ? tiled::fwd_kernel<2, 1> : tiled::fwd_kernel<4, 1>);
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm) != cudaSuccess)
      return check_launch("tc_fwd (shared memory attribute)");
    // row block of a multi-GPU shard: split the partner range over gridDim.y (see tc_bwd_impl) and merge the LSE states
    const int64_t rows = p.r1 - p.r0;
    int nsplit = rows == B ? 1 : (int)(B / rows < 8 ? B / rows : 8);
    if (nsplit < 1 || getenv("VAEKIT_TC_NO_FWD_SPLIT")) nsplit = 1;
    double2* part_r = (double2*)(((uintptr_t)(Cb + B) + 255) & ~(uintptr_t)255);
    float2* part_ms = (float2*)(((uintptr_t)(part_r + (size_t)nsplit * rows) + 255) & ~(uintptr_t)255);
    kern<<<dim3(tb, nsplit), tiled::kThreads, tsm, st>>>(z, mu, log_var, H, C0, Cb, p, lqz_out, lse_out, partials,
                                                          (unsigned int*)ws, out, part_ms, part_r);
    if (nsplit > 1) {
      const unsigned mb = (unsigned)((rows + 7) / 8 < blocks ? (rows + 7) / 8 : blocks);      // <= blocks: partials fit
      tiled::fwd_merge_kernel<<<mb, 256, 0, st>>>(z, mu, log_var, p, nsplit, part_ms, part_r, lqz_out, lse_out, partials,
                                                  (unsigned int*)ws, out);
    }
    return check_launch("tc_fwd");
  }
  tc_prep_kernel<<<prep_blocks, 256, 0, st>>>(log_var, B * D, H, C0);
  const size_t sm = tc_smem_fwd((int)D);
  if (cudaFuncSetAttribute(tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm) != cudaSuccess)
    return check_launch("tc_fwd (shared memory attribute)");
  // latent dims > 128: first-generation kernels (one anchor per warp, partners over lanes), anchors of the row range only
  const int rblocks = (int)((p.r1 - p.r0 + kW - 1) / kW);
  tc_fwd_kernel<<<rblocks, 32 * kW, sm, st>>>(z, mu, log_var, H, C0, p, lqz_out, lse_out, partials, (unsigned int*)ws, out);
  return check_launch("tc_fwd");
}

extern "C" int vk_tc_fwd(const float* z, const float* mu, const float* log_var, int64_t B, int64_t D, float M_N,
                         float alpha, float beta, float gamma_anneal, float* lqz_out, float* lse_out, float* out,
                         void* workspace, int64_t workspace_bytes, void* stream) {
  return tc_fwd_impl(z, mu, log_var, B, D, 0, B, M_N, alpha, beta, gamma_anneal, lqz_out, lse_out, out, workspace,
                     workspace_bytes, stream);
}

extern "C" int vk_tc_fwd_rows(const float* z, const float* mu, const float* log_var, int64_t B, int64_t D, int64_t row_begin,
                              int64_t row_end, float M_N, float alpha, float beta, float gamma_anneal, float* lqz_out,
                              float* lse_out, float* out, void* workspace, int64_t workspace_bytes, void* stream) {
  return tc_fwd_impl(z, mu, log_var, B, D, row_begin, row_end, M_N, alpha, beta, gamma_anneal, lqz_out, lse_out, out,
                     workspace, workspace_bytes, stream);
}

static int tc_bwd_impl(const float* z, const float* mu, const float* log_var, const float* lqz, const float* lse, int64_t B,
                       int64_t D, int64_t row_begin, int64_t row_end, float M_N, float alpha, float beta, float gamma_anneal,
                       const float* g_loss, float* g_z, float* g_mu, float* g_lv, void* workspace, int64_t workspace_bytes,
                       void* stream) {
  VK_REQUIRE(z && mu && log_var && lqz && lse && g_z && g_mu && g_lv && workspace, VK_ERR_BAD_ARG, "tc_bwd: null pointer");
  TcParams p;
  int rc = make_params(B, D, M_N, alpha, beta, gamma_anneal, p);
  if (rc != VK_OK) return rc;
  VK_REQUIRE(0 <= row_begin && row_begin < row_end && row_end <= B, VK_ERR_BAD_ARG, "tc_bwd: bad row range [%lld, %lld) of %lld",
             (long long)row_begin, (long long)row_end, (long long)B);
  p.r0 = (int)row_begin; p.r1 = (int)row_end;
  VK_REQUIRE(workspace_bytes >= vk_tc_workspace_bytes(B, D), VK_ERR_WORKSPACE, "tc_bwd: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  uint8_t* ws = (uint8_t*)workspace;
  const int blocks = (int)((B + kW - 1) / kW);
  float* H = (float*)(ws + kWsHeader + (((size_t)blocks * 32 + 255) & ~(size_t)255));
  float* C0 = H + B * D;
  const int prep_blocks = (int)((B * D + 255) / 256 < (int64_t)sm_count() * 8 ? (B * D + 255) / 256 : (int64_t)sm_count() * 8);
  if (D <= 128) {
    const int cfg = D <= 32 ? 0 : (D <= 64 ? 1 : 2);
    const int DP = cfg == 0 ? 32 : (cfg == 1 ? 64 : 128), NS = 1;
    const int per_cta = tiled::kTW / NS * tiled::kR;
    const int tb = (int)((p.r1 - p.r0 + per_cta - 1) / per_cta);
    const size_t tsm = tiled::smem_bytes(DP);
    double* Cb = (double*)(((uintptr_t)(C0 + B * D) + 255) & ~(uintptr_t)255);
    tiled::prep_kernel<<<prep_blocks, 256, 0, st>>>(log_var, p, H, C0);
    tiled::cbar_kernel<<<(int)((B * 32 + 255) / 256), 256, 0, st>>>(C0, p, Cb);
    auto kr = cfg == 0 ? tiled::bwd_rows_kernel<1, 1> : (cfg == 1 ? tiled::bwd_rows_kernel<2, 1> : tiled::bwd_rows_kernel<4, 1>);
    auto kc = cfg == 0 ? tiled::bwd_cols_kernel<1, 1> : (cfg == 1 ? tiled::bwd_cols_kernel<2, 1> : tiled::bwd_cols_kernel<4, 1>);
    if (cudaFuncSetAttribute(kr, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm) != cudaSuccess ||
        cudaFuncSetAttribute(kc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm) != cudaSuccess)
      return check_launch("tc_bwd (shared memory attribute)");
    // A row block of a multi-GPU shard has too few anchor groups to fill the GPU (every CTA sweeps all B partners, so
    // fewer CTAs are not faster): split the partner range over gridDim.y and merge in a fixed order.
    const int64_t rows = p.r1 - p.r0;
    int nsplit = rows == B ? 1 : (int)(B / rows < 8 ? B / rows : 8);
    if (nsplit < 1) nsplit = 1;
    float* part = (float*)(((uintptr_t)(Cb + B) + 255) & ~(uintptr_t)255);
    float* part_mu = part + (size_t)nsplit * rows * D;
    float* part_lv = part_mu + (size_t)nsplit * rows * D;
    kr<<<dim3(tb, nsplit), tiled::kThreads, tsm, st>>>(z, mu, log_var, H, C0, Cb, lqz, lse, p, g_loss, g_z, part);
    kc<<<dim3(tb, nsplit), tiled::kThreads, tsm, st>>>(z, mu, log_var, H, C0, Cb, lqz, lse, p, g_loss, g_mu, g_lv, part_mu, part_lv);
    if (nsplit > 1) {
      const int64_t n = rows * D;
      tiled::bwd_merge_kernel<<<(unsigned)((n + 255) / 256 < (int64_t)sm_count() * 8 ? (n + 255) / 256 : (int64_t)sm_count() * 8), 256, 0, st>>>(
          z, mu, log_var, p, g_loss, nsplit, part, part_mu, part_lv, g_z, g_mu, g_lv);
    }
    return check_launch("tc_bwd");
  }
  tc_prep_kernel<<<prep_blocks, 256, 0, st>>>(log_var, B * D, H, C0);
  const size_t sm = tc_smem_bwd((int)D);
  if (cudaFuncSetAttribute(tc_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm) != cudaSuccess ||
      cudaFuncSetAttribute(tc_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm) != cudaSuccess)
    return check_launch("tc_bwd (shared memory attribute)");
  const int rblocks = (int)((p.r1 - p.r0 + kW - 1) / kW);
  tc_bwd_kernel<true><<<rblocks, 32 * kW, sm, st>>>(z, mu, H, C0, lqz, lse, p, g_loss, g_z, g_mu, g_lv);
  tc_bwd_kernel<false><<<rblocks, 32 * kW, sm, st>>>(z, mu, H, C0, lqz, lse, p, g_loss, g_z, g_mu, g_lv);
  return check_launch("tc_bwd");
}

extern "C" int vk_tc_bwd(const float* z, const float* mu, const float* log_var, const float* lqz, const float* lse,
                         int64_t B, int64_t D, float M_N, float alpha, float beta, float gamma_anneal,
                         const float* g_loss, float* g_z, float* g_mu, float* g_lv, void* workspace,
                         int64_t workspace_bytes, void* stream) {
  return tc_bwd_impl(z, mu, log_var, lqz, lse, B, D, 0, B, M_N, alpha, beta, gamma_anneal, g_loss, g_z, g_mu, g_lv, workspace,
                     workspace_bytes, stream);
}

extern "C" int vk_tc_bwd_rows(const float* z, const float* mu, const float* log_var, const float* lqz, const float* lse,
                              int64_t B, int64_t D, int64_t row_begin, int64_t row_end, float M_N, float alpha, float beta,
                              float gamma_anneal, const float* g_loss, float* g_z, float* g_mu, float* g_lv,
                              void* workspace, int64_t workspace_bytes, void* stream) {
  return tc_bwd_impl(z, mu, log_var, lqz, lse, B, D, row_begin, row_end, M_N, alpha, beta, gamma_anneal, g_loss, g_z, g_mu,
                     g_lv, workspace, workspace_bytes, stream);
}
