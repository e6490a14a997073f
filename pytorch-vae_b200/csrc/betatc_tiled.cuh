// K5 fast path (latent dim <= 128): lanes own latent dims, anchors are register-blocked.
//
// The first version of these sweeps gave one anchor row to a warp and spread the PARTNERS over the lanes, so every
// (anchor, dim) sum needed a warp reduction per 128 partners and three shared-memory loads per density.  Here a warp
// owns kR anchors and lane l owns dims l, l+32, ...: the per-(anchor, dim) sums (the per-dimension logsumexp of
// betatc_vae.py:188, the gradient rows) never leave the lane, a staged partner value is used by kR anchors, and only
// the per-pair sum over dims  A_ij = sum_d ld_ijd  crosses lanes -- sixteen pairs at a time through a transposing
// butterfly (one shuffle per pair).  Everything is carried in log2 units (operands pre-scaled by log2 e) so each
// exponential is one FADD + ex2.approx; the log importance weight of column j (betatc_vae.py:179-182) is folded into
// the column's constant term, the single (B-2, 0) exception is a compile-time variant of the step taken only by the
// warp and step that contain it.
#pragma once
#include <type_traits>

namespace tiled {

// kR * kP = 16 pairs per step (the transposing butterfly leaves pair q = r * kP + pp in lanes 2q, 2q+1).  Two anchors
// x eight partners instead of 4 x 4 doubles the number of warps a batch provides (B / kR: 2048 at B = 4096, where
// the sweeps are bound by how many warps an SM has in flight), and 16-row tiles let four CTAs share an SM.
#ifndef VK_TC_R
#define VK_TC_R 4
#define VK_TC_PSH 2
#define VK_TC_TR 32
#define VK_TC_TW 4
#define VK_TC_MINB 2
#endif
#ifndef VK_TC_MINB_COLS
#define VK_TC_MINB_COLS 2      // (3 CTAs per SM fit the packed column sweep at 168 registers; measured r02z: 2.81 ms against 2.58)
#endif
constexpr int kR = VK_TC_R;    // anchors per warp
constexpr int kPSh = VK_TC_PSH;   // log2 kP
constexpr int kP = 1 << kPSh;  // partners per step
constexpr int kTW = VK_TC_TW;  // warps per CTA
constexpr int kTR = VK_TC_TR;  // partner rows staged per tile
static_assert(kR * kP == 16 && (1 << kPSh) == kP && kTR % kP == 0, "pair layout");
constexpr int kThreads = 32 * kTW;
constexpr float kLn2 = 0.6931471805599453f;

// hs = 0.5 exp(-lv) log2e ; cs = (-0.5 (log 2pi + lv) + lw_j) log2e, lw_j = the generic-row log importance weight of column j
__global__ void prep_kernel(const float* __restrict__ lv, TcParams p, float* __restrict__ hs, float* __restrict__ cs) {
  const int64_t n = (int64_t)p.B * p.D;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = lv[i];
    const int j = (int)(i / p.D);
    const float lw = j == 0 ? p.lw_ds : (j == 1 ? p.lw_strat : p.lw_bm1);
    hs[i] = 0.5f * expf(-v) * kLog2e;
    cs[i] = (-0.5f * (kLog2Pi + v) + lw) * kLog2e;
  }
}

// Cb_j = sum_d cs_jd in double: the part of  sum_d (ld_ijd + lw_j) log2e  that does not depend on the anchor.  Keeping it
// out of the fp32 pair sums (which then hold only the positive quadratic part, a few tens) and combining per pair in
// double keeps the softmax weights p_ij = exp(A_ij - log q(z_i)) at ~1e-6 instead of ulp(250).
__global__ void cbar_kernel(const float* __restrict__ cs, TcParams p, double* __restrict__ cb) {
  const int j = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (j >= p.B) return;
  double t = 0.0;
  for (int d = lane; d < p.D; d += 32) t += (double)cs[(size_t)j * p.D + d];
  t = warp_sum(t);
  if (lane == 0) cb[j] = t;
}

// 16 per-pair partial sums (index q = r * kP + pp) in every lane -> the complete sum of pair q = lane >> 1
__device__ __forceinline__ float pair_reduce(const float (&a)[kR][kP], int lane) {
  float v8[8], v4[4], v2[2];
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    const float lo = a[t >> kPSh][t & (kP - 1)], hi = a[(t + 8) >> kPSh][(t + 8) & (kP - 1)];
    v8[t] = (b4 ? hi : lo) + __shfl_xor_sync(0xFFFFFFFFu, b4 ? lo : hi, 16);
  }
#pragma unroll
  for (int t = 0; t < 4; ++t) v4[t] = (b3 ? v8[t + 4] : v8[t]) + __shfl_xor_sync(0xFFFFFFFFu, b3 ? v8[t] : v8[t + 4], 8);
#pragma unroll
  for (int t = 0; t < 2; ++t) v2[t] = (b2 ? v4[t + 2] : v4[t]) + __shfl_xor_sync(0xFFFFFFFFu, b2 ? v4[t] : v4[t + 2], 4);
  float v = (b1 ? v2[1] : v2[0]) + __shfl_xor_sync(0xFFFFFFFFu, b1 ? v2[0] : v2[1], 2);
  v += __shfl_xor_sync(0xFFFFFFFFu, v, 1);
  return v;
}

// Complete pair sums when NS = 2 warps split the latent dims of one anchor group: each warp reduces its half, the halves
// meet in shared memory (slot alternates with the step parity so one 64-thread named barrier per step is enough).
template <int NS>
__device__ __forceinline__ float pair_sum(const float (&a)[kR][kP], int lane, int grp, int slice, float* xch, int step_no) {
  float v = pair_reduce(a, lane);
  if (NS == 2) {
    float* slot = xch + (((step_no & 1) * (kTW / 2) + grp) * 2) * 16;
    if ((lane & 1) == 0) slot[slice * 16 + (lane >> 1)] = v;
    asm volatile("bar.sync %0, 64;" ::"r"(1 + grp) : "memory");
    v += slot[(slice ^ 1) * 16 + (lane >> 1)];
  }
  return v;
}

// stage kTR rows of a row-major [B, D] array into a [kTR][DP] tile with cp.async (the copy of tile t+1 overlaps the
// sweep of tile t): rows >= B get `dead`, padded dims of live rows 0
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async4(float* dst, const float* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8(double* dst, const double* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void cp_async16(float* dst, const float* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}

template <int DP>
__device__ __forceinline__ void stage(float* t, const float* __restrict__ src, int B, int D, int j0, float dead) {
  if (D == DP && j0 + kTR <= B && ((uintptr_t)src & 15) == 0) {
    // full tile of unpadded rows: one contiguous block, 16-byte copies, no index arithmetic
    const float* s0 = src + (size_t)j0 * D;
    for (int e = threadIdx.x * 4; e < kTR * DP; e += kThreads * 4) cp_async16(t + e, s0 + e);
    return;
  }
  for (int e = threadIdx.x; e < kTR * DP; e += kThreads) {
    const int r = e / DP, c = e % DP;
    const int j = j0 + r;
    if (j < B && c < D) cp_async4(t + e, src + (size_t)j * D + c);
    else t[e] = j < B ? 0.f : dead;
  }
}

// ---- forward -----------------------------------------------------------------------------------------------------
template <int KD, int NS>
__global__ void __launch_bounds__(kThreads, VK_TC_MINB)
fwd_kernel(const float* __restrict__ Z, const float* __restrict__ MU, const float* __restrict__ LV,
           const float* __restrict__ Hs, const float* __restrict__ Cs, const double* __restrict__ Cb, TcParams p,
           float* __restrict__ lqz_out, float* __restrict__ lse_out, double* __restrict__ partials, unsigned int* ticket,
           float* __restrict__ out, float2* __restrict__ part_ms, double2* __restrict__ part_r) {
  // gridDim.y > 1 (row block of a multi-GPU shard): partner tiles split over blockIdx.y; every split leaves its online
  // logsumexp states (max, sum) in part_ms[y][row - r0][d] / part_r[y][row - r0] and fwd_merge_kernel finishes.
  constexpr int DP = 32 * KD * NS;                   // staged tile width: NS warps share an anchor group, KD dims per lane each
  extern __shared__ __align__(16) float sm[];
  __shared__ float xch_s[2 * (kTW / 2) * 2 * 16];
  float* xch = xch_s;
  int step_no = 0;
  constexpr int kBuf = 2 * kTR + 3 * kTR * DP;       // floats per staging buffer: Cb (double) | mu | hs | cs
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int grp = w / NS, dbase = (w % NS) * 32 * KD;
  const int i0 = p.r0 + (blockIdx.x * (kTW / NS) + grp) * kR;   // anchors [p.r0, p.r1): a rank's row block
  float z[kR][KD], m[kR][KD], s[kR][KD];
#pragma unroll
  for (int r = 0; r < kR; ++r)
#pragma unroll
    for (int k = 0; k < KD; ++k) {
      const int d = dbase + lane + 32 * k;
      z[r][k] = (i0 + r < p.r1 && d < p.D) ? Z[(size_t)(i0 + r) * p.D + d] : 0.f;
      m[r][k] = -INFINITY;
      s[r][k] = 0.f;
    }
  double rm = -INFINITY;                             // online row logsumexp of the pair slot q = lane >> 1
  float rs = 0.f;
  const float dcorr = (p.lw_strat - p.lw_ds) * kLog2e;  // pair (B-2, 0) uses lw_strat where column 0 folds lw_ds
  const int r_special = p.B - 2 - i0;                // anchor slot of row B-2 in this warp (if in [0, kR))
  const int pq = (lane >> 1) & (kP - 1);                    // partner slot of this lane's pair
  auto issue = [&](int t, int buf) {
    float* base = sm + buf * kBuf;
    const int j0 = t * kTR;
    stage<DP>(base + 2 * kTR, MU, p.B, p.D, j0, 0.f);
    stage<DP>(base + 2 * kTR + kTR * DP, Hs, p.B, p.D, j0, 0.f);
    stage<DP>(base + 2 * kTR + 2 * kTR * DP, Cs, p.B, p.D, j0, -INFINITY);   // dead partners: every density -inf
    if (threadIdx.x < kTR) {
      double* tb = reinterpret_cast<double*>(base) + threadIdx.x;
      if (j0 + threadIdx.x < p.B) cp_async8(tb, Cb + j0 + threadIdx.x); else *tb = -(double)INFINITY;
    }
    cp_async_commit();
  };
  const int ntiles_all = (p.B + kTR - 1) / kTR;
  const int tile0 = (int)((long long)ntiles_all * blockIdx.y / gridDim.y), ntiles = (int)((long long)ntiles_all * (blockIdx.y + 1) / gridDim.y);
  if (tile0 < ntiles) issue(tile0, 0);
  for (int tile = tile0; tile < ntiles; ++tile) {
    const int j0 = tile * kTR;
    if (tile + 1 < ntiles) { issue(tile + 1, (tile + 1 - tile0) & 1); cp_async_wait<1>(); } else cp_async_wait<0>();
    __syncthreads();
    const float* base = sm + ((tile - tile0) & 1) * kBuf;
    const double* Tb = reinterpret_cast<const double*>(base);   // [kTR] Cb of the staged partners (-inf for dead rows)
    const float* Tm = base + 2 * kTR;
    const float* Th = Tm + kTR * DP;
    const float* Tc = Th + kTR * DP;
    for (int jb = 0; jb < kTR && j0 + jb < p.B; jb += kP) {
      auto step = [&](auto special_c) {
        constexpr bool SPECIAL = decltype(special_c)::value;
        float a[kR][kP];
#pragma unroll
        for (int r = 0; r < kR; ++r)
#pragma unroll
          for (int pp = 0; pp < kP; ++pp) a[r][pp] = 0.f;
#pragma unroll
        for (int k = 0; k < KD; ++k) {
          float pm[kP], ph[kP], pc[kP];
#pragma unroll
          for (int pp = 0; pp < kP; ++pp) {
            const int o = (jb + pp) * DP + dbase + lane + 32 * k;
            pm[pp] = Tm[o]; ph[pp] = Th[o]; pc[pp] = Tc[o];
          }
#pragma unroll
          for (int r = 0; r < kR; ++r) {
            float e[kP];
#pragma unroll
            for (int pp = 0; pp < kP; ++pp) {
              const float df = z[r][k] - pm[pp];
              const float d2 = df * df;
              e[pp] = fmaf(-d2, ph[pp], pc[pp]);                // (log N(z_id; mu_jd, var_jd) + lw_j) log2e
              if (SPECIAL && pp == 0 && r == r_special && dbase + lane + 32 * k < p.D) e[pp] += dcorr;
              a[r][pp] = fmaf(d2, ph[pp], a[r][pp]);            // quadratic part of the pair sum
            }
            float bm[kP / 2];
#pragma unroll
            for (int pp = 0; pp < kP / 2; ++pp) bm[pp] = fmaxf(e[2 * pp], e[2 * pp + 1]);
#pragma unroll
            for (int h = kP / 4; h >= 1; h >>= 1)
#pragma unroll
              for (int pp = 0; pp < h; ++pp) bm[pp] = fmaxf(bm[pp], bm[pp + h]);
            const float mn = fmaxf(fmaxf(m[r][k], bm[0]), -1e30f);
            float sm2[kP / 2];
#pragma unroll
            for (int pp = 0; pp < kP / 2; ++pp) sm2[pp] = ex2f(e[2 * pp] - mn) + ex2f(e[2 * pp + 1] - mn);
#pragma unroll
            for (int h = kP / 4; h >= 1; h >>= 1)
#pragma unroll
              for (int pp = 0; pp < h; ++pp) sm2[pp] += sm2[pp + h];
            const float sum = sm2[0];
            s[r][k] = fmaf(s[r][k], ex2f(m[r][k] - mn), sum);
            m[r][k] = mn;
          }
        }
        // A_ij + D lw_ij complete for the 16 pairs of this step: online logsumexp over j (betatc_vae.py:185-187)
        double val = Tb[jb + pq] - (double)pair_sum<NS>(a, lane, grp, w % NS, xch, step_no);
        if (SPECIAL && pq == 0 && (lane >> (kPSh + 1)) == r_special) val += (double)p.D * (double)dcorr;
        const double mn = fmax(fmax(rm, val), -1e300);
        rs = fmaf(rs, ex2f((float)(rm - mn)), ex2f((float)(val - mn)));
        rm = mn;
      };
      if (j0 + jb == 0 && r_special >= 0 && r_special < kR) step(std::true_type{}); else step(std::false_type{});
      ++step_no;
    }
    __syncthreads();                                  // everyone is done with this buffer before it is refilled
  }
  // merge the kP partner slots of every anchor (lanes 2 kP r .. 2 kP (r+1) - 1; the xor-1 neighbour is a duplicate)
#pragma unroll
  for (int o = 2; o <= kP; o <<= 1) {
    const double om = __shfl_xor_sync(0xFFFFFFFFu, rm, o);
    const float os = __shfl_xor_sync(0xFFFFFFFFu, rs, o);
    const double mn = fmax(fmax(rm, om), -1e300);
    rs = rs * ex2f((float)(rm - mn)) + os * ex2f((float)(om - mn));
    rm = mn;
  }
  if (gridDim.y > 1) {
    const size_t rows = (size_t)(p.r1 - p.r0);
#pragma unroll
    for (int r = 0; r < kR; ++r) {
      const int i = i0 + r;
      const double rmr = __shfl_sync(0xFFFFFFFFu, rm, 2 * kP * r);
      const float rsr = __shfl_sync(0xFFFFFFFFu, rs, 2 * kP * r);
      if (i >= p.r1) continue;
      const size_t ro = (size_t)blockIdx.y * rows + (size_t)(i - p.r0);
      if (dbase == 0 && lane == 0) part_r[ro] = make_double2(rmr, (double)rsr);
#pragma unroll
      for (int k = 0; k < KD; ++k) {
        const int d = dbase + lane + 32 * k;
        if (d < p.D) part_ms[ro * p.D + d] = make_float2(m[r][k], s[r][k]);
      }
    }
    return;
  }
  // log q(z_i), i = i0 + (lane >> (kPSh + 1)), as an fp32 value plus its fp32 residual (the backward sweeps use both)
  const double lqz_d = (rm + (double)log2f(rs)) * 0.6931471805599453;
  const float lqz_mine = (float)lqz_d, lqz_res = (float)(lqz_d - (double)lqz_mine);
  double t_qzx = 0.0, t_qz = 0.0, t_pq = 0.0, t_pz = 0.0;
#pragma unroll
  for (int r = 0; r < kR; ++r) {
    const int i = i0 + r;
    const float lqz = __shfl_sync(0xFFFFFFFFu, lqz_mine, 2 * kP * r);
    const float lqr = __shfl_sync(0xFFFFFFFFu, lqz_res, 2 * kP * r);
    float lpq = 0.f, lqzx = 0.f, lpz = 0.f;
#pragma unroll
    for (int k = 0; k < KD; ++k) {
      const int d = dbase + lane + 32 * k;
      if (i < p.r1 && d < p.D) {
        const size_t o = (size_t)i * p.D + d;
        const float l = (m[r][k] + log2f(s[r][k])) * kLn2;
        lse_out[o] = l;
        lpq += l;
        const float v = LV[o], df = z[r][k] - MU[o];
        lqzx += fmaf(-df * df, 0.5f * expf(-v), -0.5f * (kLog2Pi + v));
        lpz += -0.5f * kLog2Pi - 0.5f * z[r][k] * z[r][k];
      }
    }
    lpq = warp_sum(lpq); lqzx = warp_sum(lqzx); lpz = warp_sum(lpz);
    if (i < p.r1) {
      if (dbase == 0) {                              // one of the NS warps of the group owns the row value
        if (lane == 0) { lqz_out[i] = lqz; lqz_out[p.B + i] = lqr; }
        t_qz += lqz;
      }
      t_qzx += lqzx; t_pq += lpq; t_pz += lpz;
    }
  }
  __shared__ double red[kTW * 4];
  if (lane == 0) { red[w * 4 + 0] = t_qzx; red[w * 4 + 1] = t_qz; red[w * 4 + 2] = t_pq; red[w * 4 + 3] = t_pz; }
  __syncthreads();
  if (threadIdx.x < 4) {
    double t = 0.0;
    for (int k = 0; k < kTW; ++k) t += red[k * 4 + threadIdx.x];
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

// ---- backward, row sweep: d/dz_i -------------------------------------------------------------------------------------
// L = (1/B) sum_i [ alpha lqzx_i + (beta - alpha) lqz_i + (ga - beta) lpq_i - ga lpz_i ],  ga = anneal * gamma
//   d/d ld_ijd = [ (beta - alpha) p_ij + (ga - beta) r_ijd ] / B,   p = softmax_j(A), r = softmax_j(ld + lw) per d
//   d ld_ijd / d z_id = -2 h_jd (z_id - mu_jd)
template <int KD, int NS>
__global__ void __launch_bounds__(kThreads, VK_TC_MINB)
bwd_rows_kernel(const float* __restrict__ Z, const float* __restrict__ MU, const float* __restrict__ LV,
                const float* __restrict__ Hs, const float* __restrict__ Cs, const double* __restrict__ Cb,
                const float* __restrict__ lqz, const float* __restrict__ lse, TcParams p, const float* __restrict__ g_loss,
                float* __restrict__ g_z, float* __restrict__ part) {
  // gridDim.y > 1: partner tiles are split over blockIdx.y (a row block of a multi-GPU shard has too few anchor groups
  // to fill the GPU); each split writes its raw sweep sums to part[y][row - r0][d] and bwd_merge_kernel finishes.
  constexpr int DP = 32 * KD * NS;                   // staged tile width: NS warps share an anchor group, KD dims per lane each
  extern __shared__ __align__(16) float sm[];
  __shared__ float xch_s[2 * (kTW / 2) * 2 * 16];
  float* xch = xch_s;
  int step_no = 0;
  constexpr int kBuf = 2 * kTR + 3 * kTR * DP;
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int grp = w / NS, dbase = (w % NS) * 32 * KD;
  const int i0 = p.r0 + (blockIdx.x * (kTW / NS) + grp) * kR;   // anchors [p.r0, p.r1): a rank's row block
  const float g = g_loss ? *g_loss : 1.f;
  const float invB = 1.f / (float)p.B;
  const float cq = (p.beta - p.alpha) * invB * g, cp = (p.gamma_anneal - p.beta) * invB * g;
  float z[kR][KD], l2[kR][KD], gz[kR][KD];
#pragma unroll
  for (int r = 0; r < kR; ++r)
#pragma unroll
    for (int k = 0; k < KD; ++k) {
      const int d = dbase + lane + 32 * k;
      const bool ok = i0 + r < p.r1 && d < p.D;
      z[r][k] = ok ? Z[(size_t)(i0 + r) * p.D + d] : 0.f;
      l2[r][k] = ok ? lse[(size_t)(i0 + r) * p.D + d] * kLog2e : 0.f;
      gz[r][k] = 0.f;
    }
  // log q(z_i) (log2 units) of the anchor of this lane's pair slot q = lane >> 1
  const int iq = i0 + (lane >> (kPSh + 1));
  const int pq = (lane >> 1) & (kP - 1);
  const double lqz2 = iq < p.r1 ? ((double)lqz[iq] + (double)lqz[p.B + iq]) * 1.4426950408889634 : (double)INFINITY;
  const float dcorr = (p.lw_strat - p.lw_ds) * kLog2e;
  const int r_special = p.B - 2 - i0;
  auto issue = [&](int t, int buf) {
    float* base = sm + buf * kBuf;
    const int j0 = t * kTR;
    stage<DP>(base + 2 * kTR, MU, p.B, p.D, j0, 0.f);
    stage<DP>(base + 2 * kTR + kTR * DP, Hs, p.B, p.D, j0, 0.f);
    stage<DP>(base + 2 * kTR + 2 * kTR * DP, Cs, p.B, p.D, j0, -INFINITY);
    if (threadIdx.x < kTR) {
      double* tb = reinterpret_cast<double*>(base) + threadIdx.x;
      if (j0 + threadIdx.x < p.B) cp_async8(tb, Cb + j0 + threadIdx.x); else *tb = -(double)INFINITY;
    }
    cp_async_commit();
  };
  const int ntiles_all = (p.B + kTR - 1) / kTR;
  const int tile0 = (int)((long long)ntiles_all * blockIdx.y / gridDim.y), ntiles = (int)((long long)ntiles_all * (blockIdx.y + 1) / gridDim.y);
  if (tile0 < ntiles) issue(tile0, 0);
  for (int tile = tile0; tile < ntiles; ++tile) {
    const int j0 = tile * kTR;
    if (tile + 1 < ntiles) { issue(tile + 1, (tile + 1 - tile0) & 1); cp_async_wait<1>(); } else cp_async_wait<0>();
    __syncthreads();
    const float* base = sm + ((tile - tile0) & 1) * kBuf;
    const double* Tb = reinterpret_cast<const double*>(base);
    const float* Tm = base + 2 * kTR;
    const float* Th = Tm + kTR * DP;
    const float* Tc = Th + kTR * DP;
    for (int jb = 0; jb < kTR && j0 + jb < p.B; jb += kP) {
      auto step = [&](auto special_c) {
        constexpr bool SPECIAL = decltype(special_c)::value;
        // the staged partner values are read from shared memory in both phases rather than held across the pair
        // reduction: 3 KD kP registers would cost a resident CTA per SM
        float a[kR][kP];
#pragma unroll
        for (int r = 0; r < kR; ++r)
#pragma unroll
          for (int pp = 0; pp < kP; ++pp) a[r][pp] = 0.f;
#pragma unroll
        for (int k = 0; k < KD; ++k) {
#pragma unroll
          for (int pp = 0; pp < kP; ++pp) {
            const int o = (jb + pp) * DP + dbase + lane + 32 * k;
            const float pm = Tm[o], ph = Th[o];
#pragma unroll
            for (int r = 0; r < kR; ++r) {
              const float df = z[r][k] - pm;
              a[r][pp] = fmaf(df * df, ph, a[r][pp]);                // quadratic part of the pair sum
            }
          }
        }
        // softmax weight of (i, j) in log q(z_i), times its coefficient; dead partners / anchors give exactly 0
        double x = (Tb[jb + pq] - (double)pair_sum<NS>(a, lane, grp, w % NS, xch, step_no)) - lqz2;
        if (SPECIAL && pq == 0 && (lane >> (kPSh + 1)) == r_special) x += (double)p.D * (double)dcorr;
        const float wq = cq * ex2f((float)x);
        float cqp[kR][kP];
#pragma unroll
        for (int r = 0; r < kR; ++r)
#pragma unroll
          for (int pp = 0; pp < kP; ++pp) cqp[r][pp] = __shfl_sync(0xFFFFFFFFu, wq, 2 * (r * kP + pp));
#pragma unroll
        for (int k = 0; k < KD; ++k)
#pragma unroll
          for (int pp = 0; pp < kP; ++pp) {
            const int o = (jb + pp) * DP + dbase + lane + 32 * k;
            const float pm = Tm[o], ph = Th[o], pc = Tc[o];
#pragma unroll
            for (int r = 0; r < kR; ++r) {
              const float df = z[r][k] - pm;
              const float t = ph * df;                               // h (z - mu) log2e
              float e = fmaf(-df, t, pc);
              if (SPECIAL && pp == 0 && r == r_special && dbase + lane + 32 * k < p.D) e += dcorr;
              const float rr = ex2f(e - l2[r][k]);                   // softmax weight inside LSE_j[i, d]
              const float cf = fmaf(cp, rr, cqp[r][pp]);             // dL / d ld_ijd
              gz[r][k] = fmaf(-cf, t, gz[r][k]);
            }
          }
      };
      if (j0 + jb == 0 && r_special >= 0 && r_special < kR) step(std::true_type{}); else step(std::false_type{});
      ++step_no;
    }
    __syncthreads();                                  // everyone is done with this buffer before it is refilled
  }
  const float ca = p.alpha * invB * g, cz = p.gamma_anneal * invB * g;
#pragma unroll
  for (int r = 0; r < kR; ++r)
#pragma unroll
    for (int k = 0; k < KD; ++k) {
      const int i = i0 + r, d = dbase + lane + 32 * k;
      if (gridDim.y > 1) {
        if (i < p.r1 && d < p.D) part[((size_t)blockIdx.y * (p.r1 - p.r0) + (i - p.r0)) * p.D + d] = gz[r][k];
        continue;
      }
      if (i < p.r1 && d < p.D) {
        const size_t o = (size_t)i * p.D + d;
        const float df = z[r][k] - MU[o];
        const float t = expf(-LV[o]) * df;                           // 2 h (z - mu)
        // off-diagonal sweep (t carried the factor log2e / 2) + diagonal terms: alpha/B d log q(z_i|x_i)/dz and
        // -ga/B d log p(z_i)/dz = +ga/B z
        g_z[o] = 2.f * kLn2 * gz[r][k] - ca * t + cz * z[r][k];
      }
    }
}

// ---- backward, column sweep: d/dmu_j, d/dlog_var_j -------------------------------------------------------------------
//   d ld_ijd / d mu_jd = 2 h_jd (z_id - mu_jd) ;  d ld_ijd / d lv_jd = h_jd (z_id - mu_jd)^2 - 1/2
template <int KD, int NS>
__global__ void __launch_bounds__(kThreads, VK_TC_MINB_COLS)
bwd_cols_kernel(const float* __restrict__ Z, const float* __restrict__ MU, const float* __restrict__ LV,
                const float* __restrict__ Hs, const float* __restrict__ Cs, const double* __restrict__ Cb,
                const float* __restrict__ lqz, const float* __restrict__ lse, TcParams p, const float* __restrict__ g_loss,
                float* __restrict__ g_mu, float* __restrict__ g_lv, float* __restrict__ part_mu, float* __restrict__ part_lv) {
  constexpr int DP = 32 * KD * NS;                   // staged tile width: NS warps share an anchor group, KD dims per lane each
  extern __shared__ __align__(16) float sm[];
  __shared__ float xch_s[2 * (kTW / 2) * 2 * 16];
  float* xch = xch_s;
  int step_no = 0;
  constexpr int kBuf = 2 * kTR + 3 * kTR * DP;       // (same size as the other sweeps; the third tile is unused here)
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int grp = w / NS, dbase = (w % NS) * 32 * KD;
  const int c0 = p.r0 + (blockIdx.x * (kTW / NS) + grp) * kR;   // first anchor column of this warp, in [p.r0, p.r1)
  const float g = g_loss ? *g_loss : 1.f;
  const float invB = 1.f / (float)p.B;
  const float cq = (p.beta - p.alpha) * invB * g, cp = (p.gamma_anneal - p.beta) * invB * g;
  // Anchor columns are carried in PAIRS (r = 2 rp, 2 rp + 1) as packed fp32 (FADD2 / FMUL2 / FFMA2: one issue slot for the
  // two anchors that share a staged partner value): 6 packed + 1 MUFU instead of 12 + 1 instructions per density.  Signs
  // are folded into the operands (-mu, -h; -g0 accumulated) so every element sees exactly the scalar version's operations.
  static_assert(kR % 2 == 0, "anchor pairs");
  constexpr int RP = kR / 2;
  uint64_t nmu2[RP][KD], nhs2[RP][KD], cs2[RP][KD], g0n2[RP][KD], g12[RP][KD];
#pragma unroll
  for (int rp = 0; rp < RP; ++rp)
#pragma unroll
    for (int k = 0; k < KD; ++k) {
      const int d = dbase + lane + 32 * k;
      float m_[2], h_[2], c_[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int r = 2 * rp + e;
        const bool ok = c0 + r < p.r1 && d < p.D;
        const size_t o = (size_t)(c0 + r) * p.D + d;
        m_[e] = ok ? MU[o] : 0.f; h_[e] = ok ? Hs[o] : 0.f; c_[e] = ok ? Cs[o] : 0.f;
      }
      nmu2[rp][k] = pk2(-m_[0], -m_[1]); nhs2[rp][k] = pk2(-h_[0], -h_[1]); cs2[rp][k] = pk2(c_[0], c_[1]);
      g0n2[rp][k] = pk2(0.f, 0.f); g12[rp][k] = pk2(0.f, 0.f);
    }
  const uint64_t cp2 = pk2(cp, cp), log2e2 = pk2(kLog2e, kLog2e), nln2_2 = pk2(-kLn2, -kLn2), nhalf2 = pk2(-0.5f, -0.5f);
  const float dcorr = (p.lw_strat - p.lw_ds) * kLog2e;
  const int pq = (lane >> 1) & (kP - 1);                    // partner slot of this lane's pair
  const bool anchor_dead = c0 + (lane >> (kPSh + 1)) >= p.r1;
  const double cbq = anchor_dead ? 0.0 : Cb[c0 + (lane >> (kPSh + 1))];
  auto issue = [&](int t, int buf) {
    float* base = sm + buf * kBuf;
    const int i0 = t * kTR;
    stage<DP>(base + 2 * kTR, Z, p.B, p.D, i0, 0.f);
    stage<DP>(base + 2 * kTR + kTR * DP, lse, p.B, p.D, i0, INFINITY);          // dead partners: softmax weight exactly 0
    if (threadIdx.x < kTR) {                                                     // log q(z_i): value | residual
      const int i = i0 + threadIdx.x;
      if (i < p.B) { cp_async4(base + threadIdx.x, lqz + i); cp_async4(base + kTR + threadIdx.x, lqz + p.B + i); }
      else { base[threadIdx.x] = INFINITY; base[kTR + threadIdx.x] = 0.f; }
    }
    cp_async_commit();
  };
  const int ntiles_all = (p.B + kTR - 1) / kTR;
  const int tile0 = (int)((long long)ntiles_all * blockIdx.y / gridDim.y), ntiles = (int)((long long)ntiles_all * (blockIdx.y + 1) / gridDim.y);
  if (tile0 < ntiles) issue(tile0, 0);
  for (int tile = tile0; tile < ntiles; ++tile) {
    const int i0 = tile * kTR;
    if (tile + 1 < ntiles) { issue(tile + 1, (tile + 1 - tile0) & 1); cp_async_wait<1>(); } else cp_async_wait<0>();
    __syncthreads();
    const float* base = sm + ((tile - tile0) & 1) * kBuf;
    const float* Tq = base;                          // partner rows i: log q(z_i) value [kTR] | residual [kTR]
    const float* Tz = base + 2 * kTR;                // z
    const float* Tl = Tz + kTR * DP;                 // LSE_j[i, d] (natural log units)
    for (int ib = 0; ib < kTR && i0 + ib < p.B; ib += kP) {
      const int pp_special = p.B - 2 - (i0 + ib);     // partner slot of row B-2 in this step (if in [0, kP))
      auto step = [&](auto special_c) {
        constexpr bool SPECIAL = decltype(special_c)::value;
        uint64_t an2[RP][kP];                                        // MINUS the quadratic part of the pair sums
#pragma unroll
        for (int rp = 0; rp < RP; ++rp)
#pragma unroll
          for (int pp = 0; pp < kP; ++pp) an2[rp][pp] = pk2(0.f, 0.f);
#pragma unroll
        for (int k = 0; k < KD; ++k) {
#pragma unroll
          for (int pp = 0; pp < kP; ++pp) {
            const float pz = Tz[(ib + pp) * DP + dbase + lane + 32 * k];
            const uint64_t pz2 = pk2(pz, pz);
#pragma unroll
            for (int rp = 0; rp < RP; ++rp) {
              const uint64_t df2 = add2(pz2, nmu2[rp][k]);
              an2[rp][pp] = fma2(mul2(df2, df2), nhs2[rp][k], an2[rp][pp]);
            }
          }
        }
        float a[kR][kP];
#pragma unroll
        for (int rp = 0; rp < RP; ++rp)
#pragma unroll
          for (int pp = 0; pp < kP; ++pp) upk2(an2[rp][pp], a[2 * rp][pp], a[2 * rp + 1][pp]);
        // (the sum of the negated values is the negated sum, bit for bit)
        double x = (cbq + (double)pair_sum<NS>(a, lane, grp, w % NS, xch, step_no)) -
                   ((double)Tq[ib + pq] + (double)Tq[kTR + ib + pq]) * 1.4426950408889634;
        if (SPECIAL && (lane >> (kPSh + 1)) == 0 && pq == pp_special) x += (double)p.D * (double)dcorr;
        const float wq = anchor_dead ? 0.f : cq * ex2f((float)x);
        uint64_t cqp2[RP][kP];
#pragma unroll
        for (int rp = 0; rp < RP; ++rp)
#pragma unroll
          for (int pp = 0; pp < kP; ++pp)
            cqp2[rp][pp] = pk2(__shfl_sync(0xFFFFFFFFu, wq, 2 * (2 * rp * kP + pp)), __shfl_sync(0xFFFFFFFFu, wq, 2 * ((2 * rp + 1) * kP + pp)));
#pragma unroll
        for (int k = 0; k < KD; ++k)
#pragma unroll
          for (int pp = 0; pp < kP; ++pp) {
            const int o = (ib + pp) * DP + dbase + lane + 32 * k;
            const float pz = Tz[o], pl = Tl[o];
            const uint64_t pz2 = pk2(pz, pz), npl2 = pk2(-pl, -pl);
#pragma unroll
            for (int rp = 0; rp < RP; ++rp) {
              const uint64_t df2 = add2(pz2, nmu2[rp][k]);
              const uint64_t tn2 = mul2(nhs2[rp][k], df2);           // -h (z - mu) log2e
              const uint64_t w2 = mul2(df2, tn2);                    // -h (z - mu)^2 log2e
              uint64_t e2 = add2(cs2[rp][k], w2);
              if (SPECIAL && rp == 0 && pp == pp_special && dbase + lane + 32 * k < p.D) e2 = add2(e2, pk2(dcorr, 0.f));
              float a0, a1;
              upk2(fma2(npl2, log2e2, e2), a0, a1);
              const uint64_t rr2 = pk2(ex2f(a0), ex2f(a1));          // softmax weights inside LSE_j[i, d]
              const uint64_t cf2 = fma2(cp2, rr2, cqp2[rp][pp]);     // dL / d ld_ijd
              g0n2[rp][k] = fma2(cf2, tn2, g0n2[rp][k]);
              g12[rp][k] = fma2(cf2, fma2(w2, nln2_2, nhalf2), g12[rp][k]);   // h (z - mu)^2 - 1/2
            }
          }
      };
      if (c0 == 0 && pp_special >= 0 && pp_special < kP) step(std::true_type{}); else step(std::false_type{});
      ++step_no;
    }
    __syncthreads();
  }
  const float ca = p.alpha * invB * g;
#pragma unroll
  for (int r = 0; r < kR; ++r)
#pragma unroll
    for (int k = 0; k < KD; ++k) {
      float mu[kR][KD], g0[kR][KD], g1[kR][KD];                    // (this element only; the compiler keeps the pairs)
      {
        float q0, q1;
        upk2(nmu2[r / 2][k], q0, q1); mu[r][k] = -((r & 1) ? q1 : q0);
        upk2(g0n2[r / 2][k], q0, q1); g0[r][k] = -((r & 1) ? q1 : q0);
        upk2(g12[r / 2][k], q0, q1); g1[r][k] = (r & 1) ? q1 : q0;
      }
      const int j = c0 + r, d = dbase + lane + 32 * k;
      if (gridDim.y > 1) {
        if (j < p.r1 && d < p.D) {
          const size_t po = ((size_t)blockIdx.y * (p.r1 - p.r0) + (j - p.r0)) * p.D + d;
          part_mu[po] = g0[r][k];
          part_lv[po] = g1[r][k];
        }
        continue;
      }
      if (j < p.r1 && d < p.D) {
        const size_t o = (size_t)j * p.D + d;
        const float df = Z[o] - mu[r][k];
        const float h2 = expf(-LV[o]);                               // 2 h
        g_mu[o] = 2.f * kLn2 * g0[r][k] + ca * (h2 * df);
        g_lv[o] = g1[r][k] + ca * (0.5f * h2 * df * df - 0.5f);
      }
    }
}

// Finishes a partner-split forward: merges the splits' online-logsumexp states in a fixed order, then does what the
// epilogue of fwd_kernel does (per-dimension LSE rows, log q(z_i) with its residual, the four batch sums).  One warp per
// anchor row, lanes over latent dims.
__global__ void __launch_bounds__(256)
fwd_merge_kernel(const float* __restrict__ Z, const float* __restrict__ MU, const float* __restrict__ LV, TcParams p,
                 int nsplit, const float2* __restrict__ part_ms, const double2* __restrict__ part_r,
                 float* __restrict__ lqz_out, float* __restrict__ lse_out, double* __restrict__ partials,
                 unsigned int* ticket, float* __restrict__ out) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const size_t rows = (size_t)(p.r1 - p.r0);
  double t_qzx = 0.0, t_qz = 0.0, t_pq = 0.0, t_pz = 0.0;
  for (size_t row = (size_t)blockIdx.x * 8 + w; row < rows; row += (size_t)gridDim.x * 8) {
    const size_t i = (size_t)p.r0 + row;
    float lpq = 0.f, lqzx = 0.f, lpz = 0.f;
    for (int d = lane; d < p.D; d += 32) {
      float mn = -1e30f;
      for (int y = 0; y < nsplit; ++y) mn = fmaxf(mn, part_ms[((size_t)y * rows + row) * p.D + d].x);
      float sum = 0.f;
      for (int y = 0; y < nsplit; ++y) {
        const float2 ms = part_ms[((size_t)y * rows + row) * p.D + d];
        sum = fmaf(ms.y, ex2f(ms.x - mn), sum);
      }
      const size_t o = i * p.D + d;
      const float l = (mn + log2f(sum)) * kLn2;
      lse_out[o] = l;
      lpq += l;
      const float v = LV[o], zz = Z[o], df = zz - MU[o];
      lqzx += fmaf(-df * df, 0.5f * expf(-v), -0.5f * (kLog2Pi + v));
      lpz += -0.5f * kLog2Pi - 0.5f * zz * zz;
    }
    lpq = warp_sum(lpq); lqzx = warp_sum(lqzx); lpz = warp_sum(lpz);
    double rm = -1e300;
    for (int y = 0; y < nsplit; ++y) rm = fmax(rm, part_r[(size_t)y * rows + row].x);
    float rs = 0.f;
    for (int y = 0; y < nsplit; ++y) {
      const double2 r2 = part_r[(size_t)y * rows + row];
      rs = fmaf((float)r2.y, ex2f((float)(r2.x - rm)), rs);
    }
    const double lqz_d = (rm + (double)log2f(rs)) * 0.6931471805599453;
    const float lqz = (float)lqz_d, lqr = (float)(lqz_d - (double)lqz);
    if (lane == 0) { lqz_out[i] = lqz; lqz_out[p.B + i] = lqr; }
    t_qz += lqz; t_qzx += lqzx; t_pq += lpq; t_pz += lpz;
  }
  __shared__ double red[8 * 4];
  if (lane == 0) { red[w * 4 + 0] = t_qzx; red[w * 4 + 1] = t_qz; red[w * 4 + 2] = t_pq; red[w * 4 + 3] = t_pz; }
  __syncthreads();
  if (threadIdx.x < 4) {
    double t = 0.0;
    for (int k = 0; k < 8; ++k) t += red[k * 4 + threadIdx.x];
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
      out[0] = p.alpha * mi + (p.beta * tc + p.gamma_anneal * kld);
      out[1] = kld;
      out[2] = tc;
      out[3] = mi;
    }
  }
}

// Finishes a partner-split backward: sums the splits' raw sweep sums in a fixed order and adds the diagonal terms (the
// epilogues of bwd_rows_kernel / bwd_cols_kernel).
__global__ void __launch_bounds__(256)
bwd_merge_kernel(const float* __restrict__ Z, const float* __restrict__ MU, const float* __restrict__ LV, TcParams p,
                 const float* __restrict__ g_loss, int nsplit, const float* __restrict__ part_z,
                 const float* __restrict__ part_mu, const float* __restrict__ part_lv, float* __restrict__ g_z,
                 float* __restrict__ g_mu, float* __restrict__ g_lv) {
  const float g = g_loss ? *g_loss : 1.f;
  const float invB = 1.f / (float)p.B;
  const float ca = p.alpha * invB * g, cz = p.gamma_anneal * invB * g;
  const size_t rows = (size_t)(p.r1 - p.r0), n = rows * p.D;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
    float sz = 0.f, s0 = 0.f, s1 = 0.f;
    for (int y = 0; y < nsplit; ++y) { sz += part_z[y * n + e]; s0 += part_mu[y * n + e]; s1 += part_lv[y * n + e]; }
    const size_t o = (size_t)p.r0 * p.D + e;
    const float zv = Z[o], df = zv - MU[o], h2 = expf(-LV[o]);     // 2 h
    g_z[o] = 2.f * kLn2 * sz - ca * (h2 * df) + cz * zv;
    g_mu[o] = 2.f * kLn2 * s0 + ca * (h2 * df);
    g_lv[o] = s1 + ca * (0.5f * h2 * df * df - 0.5f);
  }
}

inline size_t smem_bytes(int DP) { return (size_t)2 * (3 * kTR * DP + 2 * kTR) * sizeof(float); }   // two staging buffers of width DP

}  // namespace tiled
