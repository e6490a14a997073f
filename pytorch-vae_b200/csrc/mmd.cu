// K4: WAE-MMD  compute_kernel / compute_rbf / compute_inv_mult_quad / compute_mmd
// (reference: models/wae_mmd.py:125-204), forward value and d/dz in ONE sweep over the three Gram matrices.
//
// The reference materialises x1.expand(N,N,D) - x2.expand(N,N,D) (2.2 TB per temporary at N = 65536).  Here the
// pairwise squared distance is  |x|^2 + |y|^2 - 2 x.y  with x.y on tcgen05 tensor cores:
//
//   stats kernels   column means of z and prior -> the common centre m (distances are translation invariant, so the
//                   whole sweep runs on x - m: smaller norms, less cancellation in |x|^2 + |y|^2 - 2 x.y), column sums of
//                   the centred data, and the coefficient offsets c0 (the coefficient at the mean squared distance)
//   pack kernels    z - m, prior - m (fp32 [N,D]) -> bf16 hi/lo operand images, 128-row blocks in the UMMA no-swizzle
//                   canonical layout, + fp32 row norms.  The same image is a K-major operand (K = latent dim)
//                   for the distance GEMM and an MN-major operand (K = sample index) for the gradient GEMM.
//   main kernel     persistent, warp-specialised, one CTA per SM:
//     TMA warp      1-D bulk copies (cp.async.bulk -> mbarrier) of 128-column B stages
//     MMA warp      GEMM1  S[128,128] = X_i . Y_j^T         (G1P passes, fp32 accumulate in TMEM)
//                   GEMM2  O[128,D]  += C'_ij . Y_j         (G2P passes; centred coefficient tile written by the epilogue)
//     epilogue 16w  tcgen05.ld S -> d^2 -> k (ex2 / rcp) -> running sum of k, c' = dk/dd^2 - c0 (signed) ->
//                   row sums of c', bf16 (hi / lo) of c' back into the S columns just read = the A operand of GEMM2
//   finalize        dL/dz_i (unscaled) = z_i * sum_j c'_ij - O'_i + [c0 part from the column sums, float64];
//                   per-CTA partial sums -> three kernel sums
//   sample check    (auto mode, reduced passes) float64 SIMT gradient of 32 sample rows against all columns; if the
//                   measured error of the tensor sweep exceeds the budget a device flag makes the parity sweep run
//
// Gradient identity:  sum_j c_ij (z_i - y_j) = z_i * (sum_j c_ij) - (C Y)_i,  with c = c' + c0:
//   = z_i * sum_j c'_ij - (C' Y)_i + c0 * (n z_i - sum_j y_j).
// The tcgen05 fp32 accumulator truncates; accumulating C' (values ~ the SPREAD of the kernel over the batch) instead of
// C (values ~ the kernel itself) shrinks the accumulated magnitude -- and with it the truncation bias and the bf16
// rounding of the coefficients -- by 1 / (relative spread): 2000x for the RBF at the benchmark size.
// The prior-prior Gram needs no gradient: it is evaluated on a circulant half (each unordered tile pair once,
// weight 2), which is also balanced when row blocks are sharded across GPUs.
// Reference quirks kept: RBF = exp(-|x-y|^2 / (2 D^2 z_var)) averaged over the full N x N including the diagonal
// (wae_mmd.py:164-167); IMQ = C / (1e-7 + C + |x-y|^2), C = 2 D z_var, summed with the diagonal removed from
// all three terms, including k(prior_i, z_i) (wae_mmd.py:185-191).
// Latent dims > 128 do not fit the tile kernel's tensor-memory map: they take the SIMT kernel at the end of this file
// (the reference's literal (x - y)^2 form, float64 accumulation) -- slower, same results, no exception.
#include <stdlib.h>
#include <type_traits>
#include "common.cuh"
#include "tc.cuh"

namespace vk {
using namespace tc;

namespace {

constexpr int kRows = 128;       // row block (MMA M)
constexpr int kCols = 128;       // column block (GEMM1 N, GEMM2 K): one B stage, one S buffer
constexpr int kMaxBStages = 6;   // B ring: 6 hi-only stages, or 3 hi + lo stages when a pass reads the lo image
constexpr int kNormRing = 8;     // >= B stages + 2 (prior-prior items free a stage before the epilogue has read its norms)
// mbarriers
constexpr int kBarAFull = 0, kBarAEmpty = 1, kBarBFull = 2, kBarBEmpty = kBarBFull + kMaxBStages,
              kBarSFull = kBarBEmpty + kMaxBStages, kBarKFull = kBarSFull + 2, kBarOFull = kBarKFull + 2,
              kBarOEmpty = kBarOFull + 1, kNumBars = kBarOEmpty + 1;
// tensor memory map (512 columns): S / coefficient buffers 2 x 128 | O <= 128 | row-block operand hi, lo (Dp/2 columns
// each, two bf16 per column).  The coefficient block (bf16 hi [+ lo]: 16 [+ 16] columns per 32-column slice) overwrites
// the S columns its warp has just read, so it needs no columns -- and no barriers -- of its own.
constexpr uint32_t kTmS = 0, kTmO = 256, kTmAhi = 384, kTmAlo = 448;
constexpr int kEpiWarps = 16, kColGroups = kEpiWarps / 4, kTmaWarp = 16, kMmaWarp = 17;
constexpr int kEpiThreads = 32 * kEpiWarps;
constexpr int kMmdThreads = 32 * 18;  // warps 0-15 epilogue, warp 16 TMA, warp 17 MMA
// Column chunks per row block.  Two rules pick Q: enough gradient items for the static schedule to balance (up to
// kBalanceQ), and -- when a chain limit is set -- a bound on how many 128-column blocks are accumulated into one gradient
// tile in tensor memory: the tcgen05 fp32 accumulator truncates, a bias that grows with the length of the chain AND with
// the magnitude of the accumulated values.  Round 1 accumulated the raw coefficients (|O| ~ N k |y|: 2.3e-4 of the
// gradient's maximum at N = 65 536 with one 683-tile chain, 1.8e-5 with 32-tile chains at 17 ms per step).  Round 2
// accumulates CENTRED coefficients (see the head of the file), so the RBF needs no chain limit at all; the partial tiles
// of a limited chain are summed in fp32 round-to-nearest by the finalize kernel.
constexpr int kMaxQ = 64, kBalanceQ = 16;
constexpr int64_t kMaxGradItems = 65536;
// rows of the verification sample (auto mode).  16 rows x 128 dims = 2048 gradient entries against all N columns: the
// rounding error of a reduced-precision sweep is statistical and spread evenly over the rows (measured r02b/r02g: the
// sample's error tracks the full gradient's within 2x on every data set tried).  mmd_sample_kernel takes them eight at a
// time in registers (r02v: 2 x 248 us for the one-row-per-CTA version it replaced).
constexpr int kSampleRows = 16;
constexpr float kVerifyBudget = 4e-6f;                  // accepted sample error relative to max |grad| (north star: 1e-5)
constexpr int64_t kAutoMinN = 8192;                     // below this the parity sweep costs < 0.3 ms: nothing to gain

// Device-side per-call state at the start of the workspace (after the ticket header).
struct MmdStats {
  float m[128];             // centre of the data: mean of the z and prior column means
  double Sz[128], Sp[128];  // column sums of the centred z / prior over all N rows
  float c0[4];              // coefficient offsets of the zz and pz tiles ([2], [3] unused)
  unsigned int gmax_bits;   // max |grad_unscaled| written by the last finalize (float bit pattern)
  int flag;                 // 1: the sampled check of the reduced-precision sweep failed -> the parity sweep runs
  float verify_rel;         // measured sample error / max |grad|
  int mode_used;            // 10 * g1p + g2p of the sweep whose result stands
  double gref[kSampleRows * 128];   // float64 SIMT gradient sums of the sample rows (before dk_scale)
};

struct MmdShape {
  int64_t N;
  int D, Dp, nkc;       // latent dim, padded to 16, number of 8-element K chunks
  int T;                // 128-row tiles of the full matrix
  int rb0, rb1;         // row-block range owned by this call
  int Q;                // column chunks per row block (gradient items)
  int P;                // pieces of the prior-prior item of a row block
  int qmajor;           // item enumeration: 1 q-major, 0 row-block-major
  int kernel_type;      // 0 rbf, 1 imq
  int g1p, g2p;         // tensor passes of the distance GEMM (3: both operands hi/lo, 2: row operand hi/lo, 1) / gradient GEMM (3, 1)
  float a_exp;          // rbf: -log2(e) / (2 D^2 z_var)
  float dk_scale;       // dk/dd^2 = dk_scale * coefficient: rbf -1 / (2 D^2 z_var) (coef k) ; imq -C (coef (k/C)^2)
  float C, Ceps;        // imq: C, fl32(1e-7 + C)
};

__host__ __device__ inline int rup(int x, int m) { return (x + m - 1) / m * m; }
__host__ __device__ inline uint32_t tile_img_bytes(int Dp) { return 2u * kRows * Dp * 2u; }   // hi + lo of 128 rows

struct MmdWs {
  size_t stats, colpart, cta_sums, imgZb, imgPb, normZ, normP, Opart, rspart, total;
};
constexpr int kStatBlocks = 296;
inline MmdWs mmd_ws(int64_t N, int D, int rows_local_blocks, int Q) {
  const int Dp = rup(D, 16);
  const int T = (int)((N + kRows - 1) / kRows);
  MmdWs w;
  size_t o = kWsHeader;
  w.stats = o; o += sizeof(MmdStats);
  o = (o + 255) & ~(size_t)255;
  w.colpart = o; o += (size_t)kStatBlocks * 2 * 130 * sizeof(double);
  w.cta_sums = o; o += (size_t)256 * 3 * sizeof(double);
  o = (o + 255) & ~(size_t)255;
  w.imgZb = o; o += (size_t)T * tile_img_bytes(Dp);      // bf16 hi | lo images of 128-row blocks (column stages: one bulk copy each)
  w.imgPb = o; o += (size_t)T * tile_img_bytes(Dp);
  w.normZ = o; o += (size_t)T * kRows * 4;
  w.normP = o; o += (size_t)T * kRows * 4;
  w.Opart = o; o += (size_t)Q * rows_local_blocks * kRows * Dp * 4;
  w.rspart = o; o += (size_t)kColGroups * Q * rows_local_blocks * kRows * 4;   // one partial per column group and chunk
  w.total = o;
  return w;
}
inline int q_cap(int row_blocks) {
  int64_t c = kMaxGradItems / (row_blocks > 0 ? row_blocks : 1);
  if (c > kMaxQ) c = kMaxQ;
  if (c < kBalanceQ) c = kBalanceQ;
  return (int)c;
}
// pieces of a row block's prior-prior item: the nearest number of gradient-item lengths
inline int pick_P(int T, int Q) {
  const int len = (2 * T + Q - 1) / Q, noff = (T % 2 == 0) ? T / 2 + 1 : (T + 1) / 2;
  int P = (noff + len / 2) / len;
  if (P < 1) P = 1;
  if (P > noff) P = noff;
  return P;
}
// Makespan (in 128-column blocks) of the static schedule: items enumerated q-major, dealt round-robin to the grid.  An
// item boundary costs about as much as kItemCost blocks (row operand refill, O read-out, pipeline drain and restart:
// r02x sweep, ~18k cycles), a prior-prior block ~3/4 of a gradient block (no gradient GEMM).
constexpr int kItemCost = 10;
inline int64_t schedule_makespan(int T, int row_blocks, int Q) {
  const int P = pick_P(T, Q), n = 2 * T, len = (n + Q - 1) / Q;
  const int noff = (T % 2 == 0) ? T / 2 + 1 : (T + 1) / 2, plen = (noff + P - 1) / P;
  const int64_t items = (int64_t)row_blocks * (Q + P);
  const int grid = (int)(items < sm_count() ? items : sm_count());
  int64_t worst = 0;
  for (int b = 0; b < grid; ++b) {
    int64_t load = 0;
    for (int64_t idx = b; idx < items; idx += grid) {
      const int q = (int)(idx / row_blocks);
      int blocks;
      if (q < Q) { const int u0 = q * len; blocks = (u0 + len < n ? u0 + len : n) - u0; if (blocks < 0) blocks = 0; load += 4 * (blocks + kItemCost); }
      else { const int u0 = (q - Q) * plen; blocks = (u0 + plen < noff ? u0 + plen : noff) - u0; if (blocks < 0) blocks = 0; load += 3 * blocks + 4 * kItemCost; }
    }
    if (load > worst) worst = load;
  }
  return worst;
}
// chain_tiles: longest accumulation chain in 64-column units (0 = no limit); the sequence of a gradient item is made of
// 2 T blocks of 128 columns ([z blocks | prior blocks]).  Q = the chunk count in [chain bound, kBalanceQ] whose static
// schedule has the shortest makespan (one GPU, N = 65 536: every Q from 3 to 12 measures the same within noise; a shard
// of 64 row blocks on 148 SMs is where the quantisation bites: 640 items are 5 for some CTAs and 4 for others).
inline int pick_Q(int T, int row_blocks, int chain_tiles) {
  static const int q_env = [] { const char* e = getenv("VK_MMD_Q"); return e ? atoi(e) : 0; }();
  int qmin = 1, qmax = kBalanceQ;
  if (chain_tiles > 0) {
    qmin = (4 * T + chain_tiles - 1) / chain_tiles;
    if (qmin > q_cap(row_blocks)) qmin = q_cap(row_blocks);
    if (qmax < qmin) qmax = qmin;
  }
  if (qmax > 2 * T) qmax = 2 * T;
  if (qmin > qmax) qmin = qmax;
  if (qmin < 1) qmin = 1;
  // one-entry memo: a training loop calls with the same shape every step
  static thread_local int mT = -1, mR = -1, mC = -1, mQ = 1;
  int q;
  if (q_env > 0) q = q_env < qmin ? qmin : q_env;
  else if (mT == T && mR == row_blocks && mC == chain_tiles) q = mQ;
  else {
    q = qmin;
    int64_t best = -1;
    for (int c = qmin; c <= qmax; ++c) {
      const int len = (2 * T + c - 1) / c;
      if ((2 * T + len - 1) / len != c) continue;           // would leave an empty chunk
      const int64_t m = schedule_makespan(T, row_blocks, c);
      if (best < 0 || m < best) { best = m; q = c; }
    }
    mT = T; mR = row_blocks; mC = chain_tiles; mQ = q;
  }
  const int n = 2 * T, len = (n + q - 1) / q;
  return (n + len - 1) / len;                 // no empty chunk
}
// upper bound of Q * row_blocks over every row range of an N-row problem (what the workspace query must cover)
inline int64_t max_grad_items(int T) {
  int64_t a = (int64_t)kMaxQ * T;
  if (a > kMaxGradItems) a = kMaxGradItems;
  const int64_t b = (int64_t)kBalanceQ * T;
  return a > b ? a : b;
}

struct MmdSmem {
  uint32_t B, norms, bars, tmem, red, total;
};
__host__ __device__ inline MmdSmem mmd_smem(int Dp) {
  MmdSmem m;
  uint32_t o = 0;
  m.B = o; o += 3u * 2u * kCols * Dp * 2u;                   // column stages: 3 x (hi + lo image of 128 rows) or 6 x hi
  m.norms = o; o += kNormRing * kCols * 4u;   // deeper than the B ring: read by the epilogue after b_empty may have fired
  m.bars = o; o += kNumBars * 8;
  m.tmem = o; o += 16;
  m.red = o; o += kEpiWarps * 3 * 8;
  m.total = o;
  return m;
}

// ---- stats: column sums and squared norms of z and prior (double), then centre, centred sums, c0 ---------------
// part[block][mat][0..127] column sums, [128] sum of squared row norms; one warp per row, lanes over latent dims.
__global__ void __launch_bounds__(256)
mmd_colsum_kernel(const float* __restrict__ Z, const float* __restrict__ P, int64_t N, int D, double* __restrict__ part) {
  __shared__ double sh[8][2][130];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double az[4] = {0, 0, 0, 0}, ap[4] = {0, 0, 0, 0}, nz = 0.0, np = 0.0;
  // four rows in flight per warp (independent loads), fp32 -> double accumulation per element
  const int64_t stride = (int64_t)gridDim.x * 8;
  for (int64_t r0 = (int64_t)blockIdx.x * 8 + warp; r0 < N; r0 += 4 * stride) {
    float a[4][4], b[4][4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t r = r0 + u * stride;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int d = lane + 32 * k;
        const bool ok = r < N && d < D;
        a[u][k] = ok ? __ldg(Z + r * D + d) : 0.f;
        b[u][k] = ok ? __ldg(P + r * D + d) : 0.f;
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const double x = (double)a[u][k], y = (double)b[u][k];
        az[k] += x; ap[k] += y; nz += x * x; np += y * y;
      }
    }
  }
  nz = warp_sum(nz); np = warp_sum(np);
#pragma unroll
  for (int k = 0; k < 4; ++k) { sh[warp][0][lane + 32 * k] = az[k]; sh[warp][1][lane + 32 * k] = ap[k]; }
  if (lane == 0) { sh[warp][0][128] = nz; sh[warp][1][128] = np; }
  __syncthreads();
  for (int t = threadIdx.x; t < 2 * 130; t += blockDim.x) {
    const int mat = t / 130, c = t % 130;
    double v = 0.0;
    if (c <= 128) for (int w = 0; w < 8; ++w) v += sh[w][mat][c];
    part[((size_t)blockIdx.x * 2 + mat) * 130 + c] = v;
  }
}

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(1024)
mmd_stats_finalize_kernel(const double* __restrict__ part, int blocks, int64_t N, int D, int kernel_type, float a_exp,
                          float Ceps, MmdStats* __restrict__ st) {
  __shared__ double red[4][128];
  __shared__ double part_s[8][2][130];
  // 8 slices of the block partials per column (fixed order inside a slice and across slices: deterministic)
  {
    const int c = threadIdx.x & 127, sl = threadIdx.x >> 7;
    const int per = (blocks + 7) / 8, b0 = sl * per, b1 = b0 + per < blocks ? b0 + per : blocks;
    double a0 = 0.0, a1 = 0.0, n0 = 0.0, n1 = 0.0;
    // eight block partials in flight per thread (a dependent add per L2 load made this single-CTA kernel 24 us); the
    // summation order is unchanged
    for (int b = b0; b < b1; b += 8) {
      double t0[8], t1[8], u0[8], u1[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const bool ok = b + e < b1;
        t0[e] = ok ? part[((size_t)(b + e) * 2 + 0) * 130 + c] : 0.0;
        t1[e] = ok ? part[((size_t)(b + e) * 2 + 1) * 130 + c] : 0.0;
        u0[e] = (ok && c == 0) ? part[((size_t)(b + e) * 2 + 0) * 130 + 128] : 0.0;
        u1[e] = (ok && c == 0) ? part[((size_t)(b + e) * 2 + 1) * 130 + 128] : 0.0;
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) { a0 += t0[e]; a1 += t1[e]; n0 += u0[e]; n1 += u1[e]; }
    }
    part_s[sl][0][c] = a0; part_s[sl][1][c] = a1;
    if (c == 0) { part_s[sl][0][128] = n0; part_s[sl][1][128] = n1; }
  }
  __syncthreads();
  if (threadIdx.x >= 128) return;
  const int d = threadIdx.x;
  double sz = 0.0, sp = 0.0;
  for (int sl = 0; sl < 8; ++sl) { sz += part_s[sl][0][d]; sp += part_s[sl][1][d]; }
  if (d >= D) { sz = 0.0; sp = 0.0; }
  const double n = (double)N;
  const float m = d < D ? (float)(0.5 * (sz / n + sp / n)) : 0.f;
  st->m[d] = m;
  const double szc = sz - n * (double)m, spc = sp - n * (double)m;     // column sums of the centred data
  st->Sz[d] = d < D ? szc : 0.0;
  st->Sp[d] = d < D ? spc : 0.0;
  // per-dim pieces of  sum |x - m|^2 = sum |x|^2 - 2 m . sum x + n |m|^2  and of  mean(z - m) . mean(p - m)
  red[0][d] = d < D ? -2.0 * (double)m * sz + n * (double)m * (double)m : 0.0;
  red[1][d] = d < D ? -2.0 * (double)m * sp + n * (double)m * (double)m : 0.0;
  red[2][d] = d < D ? (szc / n) * (szc / n) : 0.0;
  red[3][d] = d < D ? (szc / n) * (spc / n) : 0.0;
  for (int i = d; i < kSampleRows * 128; i += 128) st->gref[i] = 0.0;
  asm volatile("bar.sync 1, 128;" ::: "memory");           // the first 128 threads only (the others have returned)
  if (d == 0) {
    double nz = 0.0, np = 0.0, a = 0.0, b = 0.0, mz2 = 0.0, mzp = 0.0;
    for (int sl = 0; sl < 8; ++sl) { nz += part_s[sl][0][128]; np += part_s[sl][1][128]; }
    for (int i = 0; i < 128; ++i) { a += red[0][i]; b += red[1][i]; mz2 += red[2][i]; mzp += red[3][i]; }
    const double ez = (nz + a) / n, ep = (np + b) / n;            // mean squared norms of the centred rows
    double dzz = 2.0 * (ez - mz2), dpz = ez + ep - 2.0 * mzp;    // mean squared distances z-z, z-prior
    if (dzz < 0.0) dzz = 0.0;
    if (dpz < 0.0) dpz = 0.0;
    float c0zz, c0pz;
    if (kernel_type == 0) {
      // -log2(e) / (2 D^2 z_var) = a_exp: exp(-d^2 / (2 D^2 z_var)) = 2^(a_exp d^2)
      c0zz = exp2f(a_exp * (float)dzz); c0pz = exp2f(a_exp * (float)dpz);
    } else {
      const float kz = 1.0f / (Ceps + (float)dzz), kp = 1.0f / (Ceps + (float)dpz);
      c0zz = kz * kz; c0pz = kp * kp;
    }
    st->c0[0] = c0zz; st->c0[1] = c0pz; st->c0[2] = 0.f; st->c0[3] = 0.f;
    st->gmax_bits = 0u; st->flag = 0; st->verify_rel = 0.f; st->mode_used = 0;
  }
}

// ---- pack: fp32 [N,D] - m -> image tiles [T][hi/lo][Dp/8][128][8] bf16 ; norms [T*128] fp32 ---------------------
// pack + norms of BOTH matrices in one launch (gridDim.y: 0 z, 1 prior): half a warp per row, lanes over the 8-element K
// chunks.  Each lane centres its 8 values, adds its share of the row norm, writes the hi / lo image chunks; a 16-lane
// shuffle reduction (fixed order) finishes the norm.  Replaces 2 pack + 2 norm launches: the fixed cost matters once the
// sweep is sharded over 8 GPUs.
__global__ void __launch_bounds__(256)
mmd_pack_norm_kernel(const float* __restrict__ Z, const float* __restrict__ P, const MmdStats* __restrict__ st, int64_t N,
                     int D, int Dp, float scale, uint8_t* __restrict__ imgZ, uint8_t* __restrict__ imgP,
                     float* __restrict__ normZ, float* __restrict__ normP, int T) {
  const float* X = blockIdx.y ? P : Z;
  uint8_t* imgB = blockIdx.y ? imgP : imgZ;
  float* norms = blockIdx.y ? normP : normZ;
  const int nkc = Dp / 8;
  const int kc = threadIdx.x & 15;
  const int64_t rows = (int64_t)T * kRows;
  const bool vec = (D & 3) == 0 && (reinterpret_cast<uintptr_t>(X) & 15) == 0;
  for (int64_t row = (int64_t)blockIdx.x * 16 + (threadIdx.x >> 4); row < rows; row += (int64_t)gridDim.x * 16) {
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = 0.f;
    if (row < N && kc < nkc) {
      const int d0 = kc * 8;
      if (vec && d0 + 8 <= D) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(X + row * D + d0));
        const float4 b = __ldg(reinterpret_cast<const float4*>(X + row * D + d0 + 4));
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] -= st->m[d0 + e];
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) if (d0 + e < D) v[e] = X[row * D + d0 + e] - st->m[d0 + e];
      }
    }
    float sq = 0.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) sq = fmaf(v[e], v[e], sq);
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);      // stays inside the 16-lane half
    if (kc == 0) norms[row] = sq * scale;              // RBF: |x|^2 * a_exp (the epilogue's ex2 argument), IMQ: |x|^2
    if (kc < nkc) {
      uint32_t ph[4], pl[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) split_pair(v[2 * e], v[2 * e + 1], ph[e], pl[e]);
      // 128-row blocks: [block][hi/lo][Dp/8][128][8]
      uint8_t* gb = imgB + (row / kCols) * tile_img_bytes(Dp);
      const int rb = (int)(row % kCols);
      *reinterpret_cast<uint4*>(gb + ((size_t)kc * kCols + rb) * 16) = make_uint4(ph[0], ph[1], ph[2], ph[3]);
      *reinterpret_cast<uint4*>(gb + ((size_t)(nkc + kc) * kCols + rb) * 16) = make_uint4(pl[0], pl[1], pl[2], pl[3]);
    }
  }
}

// ---- work items ---------------------------------------------------------------------------------------------
// Items: Q gradient chunks and P prior-prior pieces per owned row block, enumerated q-major (see get_item).
// The column sequence of an item is made of 128-column blocks; a block that reaches beyond N is not skipped but takes
// the masked epilogue (its operand rows and norms are zero-filled by the pack kernel).
struct Item {
  int i;          // row block (global index)
  int q;          // item index within the row block: [0, Q) gradient chunks, [Q, Q + P) prior-prior pieces
  int grad;       // 1: rows = z, columns = [z blocks | prior blocks] chunk q ; 0: rows = prior, circulant half of prior
  int u0, u1;     // block sequence range
};
struct TileD {
  int bm;         // column matrix: 0 z, 1 prior
  int jb;         // 128-column block index
  int kind;       // 0 zz, 1 zp, 2 pp
  float weight;
  int diag;       // block on the main diagonal
};
__device__ __forceinline__ Item get_item(const MmdShape& s, int idx) {
  Item it;
  // q-major: consecutive items are the same column chunk of consecutive row blocks, so (a) a CTA's items, a grid stride
  // apart, run through every chunk kind in proportion -- with the row-block-major order and a grid that is a multiple
  // of Q + P every CTA drew one kind only (r02w: 8 % imbalance at one GPU, 16 % at eight) -- and (b) the CTAs of a
  // wave read the same column stages (L2); the cheaper prior-prior pieces come last
  const int nrb = s.rb1 - s.rb0;
  const int q = s.qmajor ? idx / nrb : idx % (s.Q + s.P);
  it.i = s.rb0 + (s.qmajor ? idx % nrb : idx / (s.Q + s.P));
  it.q = q;
  if (q < s.Q) {
    it.grad = 1;
    const int n = 2 * s.T, len = (n + s.Q - 1) / s.Q;
    it.u0 = q * len;
    it.u1 = it.u0 + len < n ? it.u0 + len : n;
    if (it.u0 > n) it.u0 = n;
  } else {
    // prior-prior: the circulant half of this row block, cut into P pieces about as long as a gradient item so that the
    // static round-robin schedule deals equal work (one piece of T/2 + 1 blocks was twice a gradient item at Q = 8)
    it.grad = 0;
    const int noff = (s.T % 2 == 0) ? s.T / 2 + 1 : (s.T + 1) / 2, len = (noff + s.P - 1) / s.P;
    it.u0 = (q - s.Q) * len;
    it.u1 = it.u0 + len < noff ? it.u0 + len : noff;
    if (it.u0 > noff) it.u0 = noff;
  }
  return it;
}
__device__ __forceinline__ TileD get_tile(const MmdShape& s, const Item& it, int u) {
  TileD t;
  if (it.grad) {
    const bool zside = u < s.T;
    t.bm = zside ? 0 : 1;
    t.jb = zside ? u : u - s.T;
    t.kind = zside ? 0 : 1;
    t.weight = 1.f;
    t.diag = t.jb == it.i;
  } else {
    int jb = it.i + u;                                    // u <= T/2 + 1: one conditional wrap instead of a modulo
    if (jb >= s.T) jb -= s.T;
    t.bm = 1;
    t.jb = jb;
    t.kind = 2;
    t.weight = (u == 0 || (s.T % 2 == 0 && u == s.T / 2)) ? 1.f : 2.f;
    t.diag = u == 0;
  }
  return t;
}
// two fp32 -> packed bf16 pair (round to nearest), no lo part
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}

// ---- main kernel ----------------------------------------------------------------------------------------------
// G1P / G2P: tensor passes of the distance / gradient GEMM.  run_if >= 0: the kernel only runs when st->flag == run_if
// (the parity sweep behind a failed sample check).
//
// Pipeline per 128-column block c (buffer sb = c & 1 of tensor memory):
//   TMA   B stage c (hi [+ lo] image of the 128 column rows, their norms)            -> B_full[slot]
//   MMA   G1(c): S[sb] = X_i . Y_c^T            (after B_full)                        -> S_full[sb]
//   epi   S[sb] -> registers -> k, c' ; bf16 c' back INTO the S columns just read     -> K_full[sb]   (16 warp arrivals)
//   MMA   G2(c): O += C'[sb] . Y_c              (after K_full)                        -> B_empty[slot]
// The MMA warp issues G1(c+1) before it waits for K_full(c): the tensor pipe works on the next block's distances
// while the epilogue is busy, and because tcgen05.mma executes in issue order, G1(c+2) -- issued after G2(c) -- may
// overwrite buffer sb without another barrier (K_full(c) already says every warp has S(c) in registers).
// Round 2 measurements behind this shape (profiles/role_prof_mmd_r02r.log, trace_mmd_r02r.log): with 64-column tiles
// the MMA warp's own instruction stream (~250 uniform-datapath instructions per tile at ~5 cycles) took 1260 cycles per
// tile against 768 cycles of tensor work, and every epilogue warp walked the loop header of both tile groups.
template <int G1P, int G2P, int KT>
// (18 warps = 5 on two of the four SM sub-partitions: 16 384 / 5 / 32 = 102 -> 96 registers per thread is the most a
// 576-thread CTA can have; __maxnreg__(112) compiles with fewer spills but cannot launch)
__global__ void __launch_bounds__(kMmdThreads, 1)
mmd_main_kernel(const float* __restrict__ Zf, const float* __restrict__ Pf, const uint8_t* __restrict__ imgZb,
                const uint8_t* __restrict__ imgPb, const float* __restrict__ normZ,
                const float* __restrict__ normP, MmdShape s, int n_items, float* __restrict__ Opart,
                float* __restrict__ rspart, double* __restrict__ cta_sums, const MmdStats* __restrict__ st, int run_if) {
  if (run_if >= 0 && st->flag != run_if) return;          // uniform over the grid: decided before the kernel started
  extern __shared__ __align__(128) uint8_t smem[];
  const MmdSmem L = mmd_smem(s.Dp);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(smem);
  auto bar = [&](int i) { return sbase + L.bars + 8u * i; };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L.tmem);
  constexpr bool kNeedLo = G1P == 3 || G2P == 3;             // the lo image of a column stage is read by some pass
  constexpr uint32_t NST = kNeedLo ? 3u : 6u;                // B ring slots (same bytes either way)
  if (threadIdx.x == 0) {
    mbar_init(bar(kBarAFull), kEpiWarps); mbar_init(bar(kBarAEmpty), 1);
    for (int i = 0; i < kMaxBStages; ++i) { mbar_init(bar(kBarBFull + i), 1); mbar_init(bar(kBarBEmpty + i), 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar(kBarSFull + i), 1); mbar_init(bar(kBarKFull + i), kEpiWarps); }
    mbar_init(bar(kBarOFull), 1); mbar_init(bar(kBarOEmpty), kEpiWarps);
    mbar_fence_init();
  }
  if (warp == kMmaWarp) tmem_alloc(smem_u32(tmem_slot), 512);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tmem_S = tmem + kTmS, tmem_O = tmem + kTmO;
  const uint32_t b_half = (uint32_t)kCols * s.Dp * 2u;      // the hi (or lo) image of a 128-row stage
  const uint32_t b_bytes = kNeedLo ? 2u * b_half : b_half;   // bytes of one ring slot = what one bulk copy brings

  if (warp == kTmaWarp) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t nb = 0;
      for (int idx = blockIdx.x; idx < n_items; idx += gridDim.x) {
        const Item it = get_item(s, idx);
        for (int u = it.u0; u < it.u1; ++u, ++nb) {
          const TileD t = get_tile(s, it, u);
          const uint32_t slot = nb % NST;
          if (nb >= NST) mbar_wait(bar(kBarBEmpty + slot), ((nb / NST) - 1) & 1);
          const uint8_t* bimg = (t.bm ? imgPb : imgZb) + (size_t)t.jb * (2 * b_half);
          const float* nrm = (t.bm ? normP : normZ) + (size_t)t.jb * kCols;
          mbar_expect_tx(bar(kBarBFull + slot), b_bytes + kCols * 4);
          bulk_g2s(sbase + L.B + slot * b_bytes, bimg, b_bytes, bar(kBarBFull + slot));      // hi (+ lo) image, one copy
          bulk_g2s(sbase + L.norms + (nb % kNormRing) * kCols * 4, nrm, kCols * 4, bar(kBarBFull + slot));
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // ===== MMA issuer: the whole warp runs the loop converged (operands stay in uniform registers), one elected
    // lane issues =====
    {
      const uint32_t leader = elect_one();
      const uint32_t idesc1 = idesc_bf16(kRows, kCols, 0, 0);       // S = A(K-major) . B(K-major)^T
      const uint32_t idesc2 = idesc_bf16(kRows, s.Dp, 0, 1);        // O += Cmat(K-major) . B(MN-major)
      const uint32_t b_lbo = kCols * 16;
      const uint64_t b_step = (2u * b_lbo) >> 4, b2_step = 256u >> 4;
      const uint32_t tmem_Ahi = tmem + kTmAhi, tmem_Alo = tmem + kTmAlo;
      const int ksteps1 = s.Dp / 16;
      uint32_t cnt = 0, n_grad_items = 0;                      // cnt: blocks whose G1 has been issued
      int iter = 0;
      PROF_DECL
      TRACE_DECL(4)
      auto gemm1 = [&](bool release_b) {
        const uint32_t slot = cnt % NST, sb = cnt & 1u;
        TRACE(0, cnt)
        PROF_LAP(3)
        mbar_wait(bar(kBarBFull + slot), (cnt / NST) & 1);
        PROF_LAP(0)
        fence_after_sync();
        const uint32_t b_hi = sbase + L.B + slot * b_bytes;
        const uint64_t db_hi = smem_desc(b_hi, b_lbo, 128), db_lo = smem_desc(b_hi + b_half, b_lbo, 128);   // lo: kNeedLo modes only
        const uint32_t s_dst = tmem_S + sb * (uint32_t)kCols;
        uint32_t acc = 0;
#pragma unroll
        for (int pass = 0; pass < G1P; ++pass) {
          // passes: lo*hi, hi*lo, then hi*hi LAST -- the tensor core's fp32 accumulation truncates (~2^-23 |acc| per
          // MMA), so the correction terms are summed while the accumulator is still small (gemm3x.cuh, measured 8x)
          // G1P == 2: lo*hi + hi*hi -- the ROW operand keeps its 16 bits, the column operand is plain bf16 (see the policy)
          uint32_t ta = (G1P >= 2 && pass == 0) ? tmem_Alo : tmem_Ahi;   // operand from tensor memory: 8 columns per K=16
          uint64_t db = (G1P == 3 && pass == 1) ? db_lo : db_hi;
#pragma unroll 4
          for (int k = 0; k < ksteps1; ++k) {
            mma_bf16_ts_if(leader, s_dst, ta, db, idesc1, acc);
            acc = 1;
            ta += 8;
            db += b_step;
          }
        }
        mma_commit_if(leader, bar(kBarSFull + sb));
        if (release_b) mma_commit_if(leader, bar(kBarBEmpty + slot));   // no gradient GEMM will read this stage
        TRACE(1, cnt)
        ++cnt;
      };
      for (int idx = blockIdx.x; idx < n_items; idx += gridDim.x, ++iter) {
        const Item it = get_item(s, idx);
        mbar_wait(bar(kBarAFull), iter & 1);
        const int n = it.u1 - it.u0;
        const uint32_t c0 = cnt;
        // z-side blocks of this chunk come first (block sequence: [z blocks | prior blocks])
        int nz = 0;
        if (it.grad) {
          const int zend = it.u1 < s.T ? it.u1 : s.T;
          nz = zend > it.u0 ? zend - it.u0 : 0;
        }
        if (n > 0) gemm1(!it.grad);
        uint32_t oacc = 0;
        for (int j = 0; j < n; ++j) {
          // G1(j+1) goes first: the distance GEMM of the next block runs while the epilogue works on block j
          if (j + 1 < n) gemm1(!it.grad);
          const uint32_t c = c0 + (uint32_t)j, sb = c & 1u;
          PROF_LAP(3)
          mbar_wait(bar(kBarKFull + sb), (c >> 1) & 1);       // block j is in the epilogue's registers (+ coefficients written)
          if (it.grad && j == 0 && n_grad_items > 0) mbar_wait(bar(kBarOEmpty), (n_grad_items - 1) & 1);
          PROF_LAP(2)
          if (it.grad) {
            TRACE(2, c)
            fence_after_sync();
            const uint32_t slot = c % NST;
            const uint32_t neg_a = j >= nz ? (1u << 13) : 0u;             // k(z, prior) enters with a minus sign
            // A: coefficient block in tensor memory, per 32-column slice 16 columns of bf16 pairs hi, then 16 lo (G2P == 3)
            const uint32_t tk = tmem_S + sb * (uint32_t)kCols;
            const uint32_t b_hi = sbase + L.B + slot * b_bytes;
            // B: MN-major (N = latent dims, K = rows of the stage)
            const uint64_t db_hi = smem_desc(b_hi, 128, b_lbo), db_lo = smem_desc(b_hi + b_half, 128, b_lbo);
#pragma unroll
            for (int pass = 0; pass < G2P; ++pass) {
              // G2P == 3: lo*hi, hi*lo, hi*hi ; G2P == 2: lo*hi, hi*hi (coefficient split, column operand plain bf16)
              const uint32_t lo_off = (G2P >= 2 && pass == 0) ? 16u : 0u;
              uint64_t db = (G2P == 3 && pass == 1) ? db_lo : db_hi;
#pragma unroll
              for (int k = 0; k < kCols / 16; ++k) {
                mma_bf16_ts_if(leader, tmem_O, tk + (uint32_t)((k >> 1) * 32 + (k & 1) * 8) + lo_off, db, idesc2 | neg_a, oacc);
                oacc = 1;
                db += b2_step;
              }
            }
            mma_commit_if(leader, bar(kBarBEmpty + slot));
            TRACE(3, c)
          }
        }
        if (it.grad && n > 0) { mma_commit_if(leader, bar(kBarOFull)); ++n_grad_items; }
        mma_commit_if(leader, bar(kBarAEmpty));
      }
      PROF_LAP(3)
      PROF_OUT("mma wait_b/-/wait_k_full/issue", leader)
      TRACE_OUT("mma g1_start/g1_issued/g2_kfull/g2_issued", 4, leader)
    }
  } else {
    // ===== epilogue: 16 warps = 4 TMEM lane quadrants x 4 slices of 32 columns; every warp takes its slice of every
    // block (four warps per scheduler hide the TMEM / MUFU / shared-memory latencies) =====
    const int ew = warp & 3;                                  // TMEM lane quadrant (must equal warp id % 4)
    const int cg = warp >> 2;                                 // 32-column slice of the block; slot of this warp's partial row sums / O chunk
    const int t = ew * 32 + lane;                             // row within the block
    double acc_zz = 0.0, acc_pz = 0.0, acc_pp = 0.0;
    // sum of the current term (run_kind) as a compensated fp32 pair: exact to ~2^-45 of the run, no conversion and no
    // DADD per block (the DADD alone held 5 % of the r02g stall samples)
    float run_hi = 0.f, run_lo = 0.f;
    int run_kind = -1;
    auto flush_run = [&]() {
      const double r = (double)run_hi + (double)run_lo;
      if (run_kind == 0) acc_zz += r; else if (run_kind == 1) acc_pz += r; else if (run_kind == 2) acc_pp += r;
      run_hi = 0.f; run_lo = 0.f;
    };
    uint32_t cnt = 0, n_grad_items = 0, item_iter = 0;
    // RBF: norms arrive pre-scaled by a_exp, so ex2's argument is one FADD + one FFMA; IMQ: t = C + eps + d^2
    const float s_coef = KT == 0 ? -2.f * s.a_exp : -2.f;
    const float c0zz = st->c0[0], c0pz = st->c0[1];          // coefficient offsets (written by the stats kernel)
    const int n32 = (int)s.N;                                  // N <= 2^24: 32-bit index math in the block loop
    PROF_DECL
    TRACE_DECL(3)
    for (int idx = blockIdx.x; idx < n_items; idx += gridDim.x) {
      const Item it = get_item(s, idx);
      const int grow = it.i * kRows + t;
      const bool rvalid = grow < n32;
      const bool row_tail = (it.i + 1) * kRows > n32;     // this row block has rows beyond N
      const float nx = (it.grad ? normZ : normP)[grow] + (KT == 1 ? s.Ceps : 0.f);
      float rs = 0.f;
      {
        // this row block's operand -> tensor memory: dims [cg*32, cg*32+32) of row t as bf16 hi / lo pairs, two per
        // column.  The previous item's MMAs must be finished with the old block first.
        if (item_iter > 0) mbar_wait(bar(kBarAEmpty), (item_iter - 1) & 1);
        fence_after_sync();
        // The pack kernel has already split x - m into the bf16 hi / lo images ([block][hi/lo][Dp/8][128 rows][8]): chunk
        // kc of this row is 16 bytes, and the 32 lanes of a warp (32 consecutive rows) read 512 contiguous bytes -- four
        // coalesced 16-byte loads per image instead of the fp32 row (lanes on 32 different rows: ~16k LSU cycles per
        // item with scalar loads, r02y / r02z role counters).  Rows >= N and padded dims are zero in the image.
        const uint8_t* img = (it.grad ? imgZb : imgPb) + (size_t)it.i * (2 * b_half);
        const uint32_t ta = tmem + ((uint32_t)(ew * 32) << 16) + cg * 16;
        uint32_t ah[16], al[16];
#pragma unroll
        for (int m = 0; m < 4; ++m) {
          const int kc = cg * 4 + m;
          uint4 h = make_uint4(0u, 0u, 0u, 0u), l = h;
          if (kc < s.nkc) {
            h = __ldg(reinterpret_cast<const uint4*>(img + ((size_t)kc * kRows + t) * 16));
            if (G1P >= 2) l = __ldg(reinterpret_cast<const uint4*>(img + b_half + ((size_t)kc * kRows + t) * 16));
          }
          ah[4 * m] = h.x; ah[4 * m + 1] = h.y; ah[4 * m + 2] = h.z; ah[4 * m + 3] = h.w;
          al[4 * m] = l.x; al[4 * m + 1] = l.y; al[4 * m + 2] = l.z; al[4 * m + 3] = l.w;
        }
#pragma unroll
        for (int j8 = 0; j8 < 16; j8 += 8) {
          uint32_t h8[8], l8[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) { h8[j] = ah[j8 + j]; l8[j] = al[j8 + j]; }
          tmem_st8(ta + kTmAhi + j8, h8);
          if (G1P >= 2) tmem_st8(ta + kTmAlo + j8, l8);
        }
        tmem_st_wait();
        fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(kBarAFull));
        ++item_iter;
      }
      for (int u = it.u0; u < it.u1; ++u, ++cnt) {
        const TileD td = get_tile(s, it, u);
        const uint32_t sb = cnt & 1u;
        PROF_LAP(1)
        mbar_wait(bar(kBarSFull + sb), (cnt >> 1) & 1);
        PROF_LAP(0)
        TRACE(0, cnt)
        fence_after_sync();
        uint32_t rr[2][16];
        // this warp's 32 columns of S; its coefficients go back into the same columns (hi: +0..15, lo: +16..31)
        const uint32_t taddr = tmem_S + ((uint32_t)(ew * 32) << 16) + sb * (uint32_t)kCols + cg * 32;
        tmem_ld16(taddr, rr[0]);
        tmem_ld16(taddr + 16, rr[1]);
        const float* ny = reinterpret_cast<const float*>(smem + L.norms + (cnt % kNormRing) * kCols * 4) + cg * 32;
        const int gcol0 = td.jb * kCols + cg * 32;
        // IMQ leaves the diagonal out: only the slice that meets this warp's rows has diagonal elements
        const bool diag_excl = KT == 1 && td.diag && cg == ew;
        // uniform across the warp (the two instances below contain .sync.aligned tcgen05 instructions)
        const bool slow = diag_excl || gcol0 + 32 > n32 || row_tail;
        float ksum = 0.f, csum = 0.f;                       // sum of k (value) and of the centred coefficient (gradient)
        const float c0k = td.kind == 0 ? c0zz : c0pz;       // (unused by the prior-prior blocks)
        // interior slices take the branch-free instance; diagonal / tail slices the masked one
        auto tile_math = [&](auto masked_c, auto grad_c) {
          constexpr bool MASKED = decltype(masked_c)::value, GRAD = decltype(grad_c)::value;
          // interior: two elements per instruction (FADD2 / FFMA2), one running sum where one is enough
          const uint64_t nx2 = pk2(nx, nx), sc2 = pk2(s_coef, s_coef), nc0 = pk2(-c0k, -c0k);
          uint64_t ksum2 = pk2(0.f, 0.f), csum2 = pk2(0.f, 0.f);
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            tmem_ld_wait16(rr[hf]);                           // (tcgen05.wait::ld covers both loads: S is in registers)
            if (!GRAD && hf == 1) {                           // no coefficients to write: the buffer is free already
              fence_before_sync();
              __syncwarp();
              if (lane == 0) mbar_arrive(bar(kBarKFull + sb));
            }
            uint32_t kh[8], kl[8];                            // 16 coefficients: 8 hi + 8 lo columns of bf16 pairs
#pragma unroll
            for (int c8 = 0; c8 < 16; c8 += 8) {
              float cv[8];
              const float4 n0 = *reinterpret_cast<const float4*>(ny + hf * 16 + c8);
              const float4 n1 = *reinterpret_cast<const float4*>(ny + hf * 16 + c8 + 4);
              if constexpr (!MASKED) {
                const uint64_t nyp[4] = {pk2(n0.x, n0.y), pk2(n0.z, n0.w), pk2(n1.x, n1.y), pk2(n1.z, n1.w)};
#pragma unroll
                for (int e = 0; e < 8; e += 2) {
                  // no clamp of d^2 at 0: a rounding-level negative distance moves k by < 1e-9 relative
                  const uint64_t a2 = fma2(pk2u(rr[hf][c8 + e], rr[hf][c8 + e + 1]), sc2, add2(nx2, nyp[e >> 1]));
                  float a0, a1, k0, k1;
                  upk2(a2, a0, a1);
                  if (KT == 0) { k0 = ex2_approx(a0); k1 = ex2_approx(a1); }      // exp(-d^2 / (2 D^2 z_var))
                  else { k0 = rcp_approx(a0); k1 = rcp_approx(a1); }              // k / C = 1 / (C + eps + d^2)
                  const uint64_t k2 = pk2(k0, k1);
                  if (GRAD) {
                    // centred coefficient: rbf dk/dd^2 = k * dk_scale, imq -C (k/C)^2 (the constants are applied in finalize)
                    const uint64_t c2 = KT == 0 ? add2(k2, nc0) : fma2(k2, k2, nc0);
                    csum2 = add2(csum2, c2);
                    upk2(c2, cv[e], cv[e + 1]);
                  }
                  if (!GRAD || KT == 1) ksum2 = add2(ksum2, k2);                  // rbf gradient blocks: sum k = sum c' + 32 c0
                }
              } else {
                const float nyv[8] = {n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, n1.z, n1.w};
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                  const int c = hf * 16 + c8 + e;
                  const float arg = fmaf(__uint_as_float(rr[hf][c8 + e]), s_coef, nx + nyv[e]);
                  float k;
                  if (KT == 0) { k = ex2_approx(arg); cv[e] = k - c0k; }
                  else { k = rcp_approx(arg); cv[e] = fmaf(k, k, -c0k); }
                  // diagonal exclusion, tail rows / columns
                  const bool ok = rvalid && (gcol0 + c < n32) && !(diag_excl && gcol0 + c == grow);
                  k = ok ? k : 0.f;
                  cv[e] = ok ? cv[e] : 0.f;
                  ksum += k;
                  if (GRAD) csum += cv[e];
                }
              }
              if (GRAD) {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  if (G2P >= 2) split_pair(cv[2 * e], cv[2 * e + 1], kh[(c8 >> 1) + e], kl[(c8 >> 1) + e]);
                  else kh[(c8 >> 1) + e] = pack_bf16x2(cv[2 * e], cv[2 * e + 1]);
                }
              }
            }
            if (GRAD) {
              tmem_st8(taddr + hf * 8, kh);
              if (G2P >= 2) tmem_st8(taddr + 16 + hf * 8, kl);
            }
          }
          if constexpr (!MASKED) {
            float a, b;
            if (GRAD) { upk2(csum2, a, b); csum = a + b; }
            if (GRAD && KT == 0) ksum = fmaf(32.f, c0k, csum);
            else { upk2(ksum2, a, b); ksum = a + b; }
          }
        };
        if (it.grad) {
          if (slow) tile_math(std::true_type{}, std::true_type{}); else tile_math(std::false_type{}, std::true_type{});
          tmem_st_wait();
          fence_before_sync();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar(kBarKFull + sb));
        } else {
          if (slow) tile_math(std::true_type{}, std::false_type{}); else tile_math(std::false_type{}, std::false_type{});
        }
        rs += td.kind == 1 ? -csum : csum;                  // GEMM2 applies the same sign with the negate-A bit
        if (td.kind != run_kind) {                          // (warp-uniform, once or twice per item)
          flush_run();
          run_kind = td.kind;
        }
        {
          const float x = ksum * td.weight, sum = run_hi + x, bp = sum - run_hi;      // TwoSum
          run_lo += (run_hi - (sum - bp)) + (x - bp);
          run_hi = sum;
        }
        TRACE(2, cnt)
      }
      PROF_LAP(1)
      if (it.grad) {
        const int q = it.q;
        const size_t lrow = (size_t)(it.i - s.rb0) * kRows + t;
        const size_t nloc = (size_t)(s.rb1 - s.rb0) * kRows;
        float* op = Opart + ((size_t)q * nloc + lrow) * s.Dp;
        rspart[((size_t)q * kColGroups + cg) * nloc + lrow] = rs;
        const int c0 = cg * 32;                               // this warp's 32-column chunk of O (Dp <= 128)
        if (it.u1 <= it.u0) {                                 // empty chunk (cannot happen with pick_Q's Q; kept safe)
          for (int e = c0; e < c0 + 32 && e < s.Dp; ++e) op[e] = 0.f;
          continue;
        }
        // O complete -> partial gradient rows to the workspace
        mbar_wait(bar(kBarOFull), n_grad_items & 1);
        fence_after_sync();
        if (c0 < s.Dp) {
          uint32_t o[32];
          tmem_ld32(tmem_O + ((uint32_t)(ew * 32) << 16) + c0, o);
          tmem_ld_wait32(o);
          if (c0 + 32 <= s.Dp) {
#pragma unroll
            for (int e = 0; e < 32; e += 4)
              *reinterpret_cast<float4*>(op + c0 + e) = make_float4(__uint_as_float(o[e]), __uint_as_float(o[e + 1]),
                                                                     __uint_as_float(o[e + 2]), __uint_as_float(o[e + 3]));
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) if (c0 + e < s.Dp) op[c0 + e] = __uint_as_float(o[e]);
          }
        }
        fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(kBarOEmpty));
        ++n_grad_items;
      }
      PROF_LAP(3)
    }
    PROF_OUT("epilogue wait_s_full/work/-/item_end", threadIdx.x == 0)
    TRACE_OUT("epi s_full/-/done", 3, threadIdx.x == 0)
    flush_run();
    if (KT == 1) { acc_zz *= (double)s.C; acc_pz *= (double)s.C; acc_pp *= (double)s.C; }   // IMQ sums were of k / C
    // per-CTA sums (fixed order) -> workspace
    double* red = reinterpret_cast<double*>(smem + L.red);
    acc_zz = warp_sum(acc_zz); acc_pz = warp_sum(acc_pz); acc_pp = warp_sum(acc_pp);
    if (lane == 0) { red[warp * 3 + 0] = acc_zz; red[warp * 3 + 1] = acc_pz; red[warp * 3 + 2] = acc_pp; }
    asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
    if (threadIdx.x < 3) {
      double tot = 0.0;
      for (int w = 0; w < kEpiWarps; ++w) tot += red[w * 3 + threadIdx.x];
      cta_sums[blockIdx.x * 3 + threadIdx.x] = tot;
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem, 512);
}

// ---- finalize: G_i = z_i * sum_q rs' - sum_q O' + c0 part ; the three kernel sums -----------------------------
// The c0 part of the gradient, from the column sums (float64):  rbf (every column carries a coefficient)
//   (c0zz - c0pz) N z_i - c0zz Sz + c0pz Sp ;  imq (column i carries none):
//   (c0zz - c0pz) (N - 1) z_i - c0zz (Sz - z_i) + c0pz (Sp - p_i)         (z, p, Sz, Sp of the centred data)
__global__ void mmd_finalize_kernel(const float* __restrict__ Z, const float* __restrict__ P, MmdShape s,
                                    const float* __restrict__ Opart, const float* __restrict__ rspart,
                                    const double* __restrict__ cta_sums, int grid_main, float* __restrict__ grad,
                                    double* __restrict__ sums, MmdStats* __restrict__ st, int run_if) {
  if (run_if >= 0 && st->flag != run_if) return;
  const int64_t row0 = (int64_t)s.rb0 * kRows;
  const int64_t nloc_pad = (int64_t)(s.rb1 - s.rb0) * kRows;
  int64_t nloc = s.N - row0;
  if (nloc > nloc_pad) nloc = nloc_pad;
  const double c0zz = (double)st->c0[0], c0pz = (double)st->c0[1];
  const double n = (double)s.N;
  const int lane = threadIdx.x & 31;
  float gmax = 0.f;
  // one warp per row: the lanes first sum the row's Q x 4 partial row sums (fixed order), then take latent dims
  for (int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < nloc; r += ((int64_t)gridDim.x * blockDim.x) >> 5) {
    float rs = 0.f;
    for (int e = lane; e < s.Q * kColGroups; e += 32) rs += rspart[(size_t)e * nloc_pad + r];
    rs = warp_sum(rs);
    for (int d = lane; d < s.D; d += 32) {
      float o = 0.f;
      for (int q = 0; q < s.Q; ++q) o += Opart[((size_t)q * nloc_pad + r) * s.Dp + d];
      const float zc = Z[(row0 + r) * s.D + d] - st->m[d];
      double ana;
      if (s.kernel_type == 0) {
        ana = (c0zz - c0pz) * n * (double)zc - c0zz * st->Sz[d] + c0pz * st->Sp[d];
      } else {
        const float pc = P[(row0 + r) * s.D + d] - st->m[d];
        ana = (c0zz - c0pz) * (n - 1.0) * (double)zc - c0zz * (st->Sz[d] - (double)zc) + c0pz * (st->Sp[d] - (double)pc);
      }
      const float gv = s.dk_scale * (float)((double)fmaf(zc, rs, -o) + ana);   // dk/dd^2 constant: -1/(2 D^2 z_var) | -C
      grad[r * s.D + d] = gv;
      gmax = fmaxf(gmax, fabsf(gv));
    }
  }
  gmax = warp_max(gmax);
  if ((threadIdx.x & 31) == 0 && gmax > 0.f) atomicMax(&st->gmax_bits, __float_as_uint(gmax));
  if (blockIdx.x == 0 && threadIdx.x < 3) {
    double t = 0.0;
    for (int c = 0; c < grid_main; ++c) t += cta_sums[c * 3 + threadIdx.x];
    sums[threadIdx.x] = t;                       // order: zz, pz, pp
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) st->mode_used = 10 * s.g1p + s.g2p;
}

// ---- SIMT rows: the reference's literal (x - y)^2 form, float64 accumulation --------------------------------------
// One CTA (8 warps) per row; a warp takes one column at a time, lanes over latent dims.  The whole path for latent dims
// > 128 (every row, gridDim.y == 1, deterministic); the sampled check of the tensor sweeps is mmd_sample_kernel below.
//   X rows, Y columns [N, D]; rows i = row0 + blockIdx.x * row_step; sign +1 / -1 of the gradient contribution;
//   gacc [rows, D] float64 accumulators (NULL: no gradient), krow [rows]: each row's kernel sum (NULL: none; needs
//   gridDim.y == 1).
template <int KT>
__global__ void __launch_bounds__(256)
mmd_simt_rows_kernel(const float* __restrict__ X, const float* __restrict__ Y, int64_t N, int D, int64_t row0,
                     int64_t row_step, int64_t n_rows, float a_exp, float Ceps, float C, double sign,
                     double* __restrict__ gacc, bool g_atomic, double* __restrict__ krow) {
  extern __shared__ double sh[];                 // [8][D] gradient partials | [8] k partials
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if ((int64_t)blockIdx.x >= n_rows) return;
  const int64_t i = row0 + (int64_t)blockIdx.x * row_step;
  const float* xi = X + i * D;
  double* acc = sh + (size_t)warp * D;
  double ks = 0.0;
  const int64_t span = (N + gridDim.y - 1) / gridDim.y;
  const int64_t j0 = (int64_t)blockIdx.y * span, j1 = j0 + span < N ? j0 + span : N;
  {
    for (int d = lane; d < D; d += 32) acc[d] = 0.0;
    for (int64_t j = j0 + warp; j < j1; j += 8) {
      const float* yj = Y + j * D;
      float s2 = 0.f;
      for (int d = lane; d < D; d += 32) { const float t = xi[d] - yj[d]; s2 = fmaf(t, t, s2); }
      s2 = warp_sum(s2);
      float k, coef;
      if (KT == 0) { k = exp2f(a_exp * s2); coef = k; }
      else { const float r = 1.0f / (Ceps + s2); k = C * r; coef = r * r; }
      if (KT == 1 && j == i) continue;                           // imq: the diagonal is removed from every term
      ks += (double)k;
      if (gacc) for (int d = lane; d < D; d += 32) acc[d] += (double)(coef * (xi[d] - yj[d]));
    }
  }
  __syncthreads();
  double* kred = sh + (size_t)8 * D;
  if (lane == 0) kred[warp] = ks;
  __syncthreads();
  if (gacc) {
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
      double t = 0.0;
      for (int w = 0; w < 8; ++w) t += sh[(size_t)w * D + d];
      double* dst = gacc + (size_t)blockIdx.x * D + d;
      if (g_atomic) atomicAdd(dst, sign * t); else *dst += sign * t;
    }
  }
  if (krow && threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += kred[w];
    krow[blockIdx.x] = t;
  }
}

// ---- sampled check of a reduced-precision sweep: float64 gradient sums of a few rows against ALL columns -----------
// A CTA keeps kSampleGroup sample rows in registers and walks a column range, so every column is read once per row group
// (the first version gave a CTA one row: 16 x 33.5 MB of L2 reads per launch and 32 lanes computing the same exponential).
// A warp takes four columns at a time, lanes over latent dims: the 8 x 4 squared distances of a step are summed with one
// transposing butterfly (31 shuffles; lane q = 4 r + c ends with pair (row r, column c)), that lane alone evaluates the
// kernel, and the coefficients go back to every lane by shuffle for the float64 accumulation of coef * (x - y).
// Terms in fp32 exactly as the SIMT path computes them (literal (x - y)^2 form), sums in float64, atomics into gref.
constexpr int kSampleGroup = 8;
template <int KT>
__global__ void __launch_bounds__(256)
mmd_sample_kernel(const float* __restrict__ X, const float* __restrict__ Y, int64_t N, int D, int64_t row0,
                  int64_t row_step, int n_rows, float a_exp, float Ceps, double sign, double* __restrict__ gref) {
  constexpr int NR = kSampleGroup;
  __shared__ double sh[NR][128];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rbase = blockIdx.y * NR;
  float xr[NR][4];
  double ar[NR][4];
#pragma unroll
  for (int r = 0; r < NR; ++r) {
    const bool rok = rbase + r < n_rows;
    const float* xi = X + (row0 + (int64_t)(rbase + r) * row_step) * D;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      xr[r][k] = (rok && lane + 32 * k < D) ? xi[lane + 32 * k] : 0.f;
      ar[r][k] = 0.0;
    }
  }
  const int my_r = lane >> 2, my_c = lane & 3;                 // the pair this lane finishes
  const bool my_rok = rbase + my_r < n_rows;
  const int64_t my_i = row0 + (int64_t)(rbase + my_r) * row_step;
  const int64_t span = (N + gridDim.x - 1) / gridDim.x;
  const int64_t j0 = (int64_t)blockIdx.x * span, j1 = j0 + span < N ? j0 + span : N;
  for (int64_t j = j0 + warp * 4; j < j1; j += 32) {
    float y[4][4];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
      for (int k = 0; k < 4; ++k)
        y[c][k] = (j + c < j1 && lane + 32 * k < D) ? __ldg(Y + (j + c) * D + lane + 32 * k) : 0.f;
    float v[32];
#pragma unroll
    for (int r = 0; r < NR; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float s2 = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) { const float df = xr[r][k] - y[c][k]; s2 = fmaf(df, df, s2); }
        v[r * 4 + c] = s2;
      }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
      const bool up = lane & o;
#pragma unroll
      for (int t = 0; t < o; ++t) v[t] = (up ? v[t + o] : v[t]) + __shfl_xor_sync(0xFFFFFFFFu, up ? v[t] : v[t + o], o);
    }
    float coef;
    if (KT == 0) coef = exp2f(a_exp * v[0]);
    else { const float rr = 1.0f / (Ceps + v[0]); coef = rr * rr; }
    if (!my_rok || j + my_c >= j1 || (KT == 1 && j + my_c == my_i)) coef = 0.f;      // imq: no diagonal in any term
#pragma unroll
    for (int r = 0; r < NR; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float cf = __shfl_sync(0xFFFFFFFFu, coef, r * 4 + c);
#pragma unroll
        for (int k = 0; k < 4; ++k) ar[r][k] += (double)(cf * (xr[r][k] - y[c][k]));
      }
  }
  // the eight warps add their sums in turn (fixed order), then one atomic per (row, dim) and CTA
  for (int w = 0; w < 8; ++w) {
    if (warp == w) {
#pragma unroll
      for (int r = 0; r < NR; ++r)
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          double* p = &sh[r][lane + 32 * k];
          *p = (w == 0 ? 0.0 : *p) + ar[r][k];
        }
    }
    __syncthreads();
  }
  for (int t = threadIdx.x; t < NR * 128; t += blockDim.x) {
    const int r = t >> 7, d = t & 127;
    if (rbase + r < n_rows && d < D) atomicAdd(gref + (size_t)(rbase + r) * D + d, sign * sh[r][d]);
  }
}

// sample check: max |grad - dk_scale * gref| over the sample rows against the budget -> st->flag
__global__ void __launch_bounds__(256)
mmd_verify_kernel(const float* __restrict__ grad, int D, int64_t row_step, int n_rows, float dk_scale, float budget,
                  MmdStats* __restrict__ st) {
  __shared__ float red[8];
  float e = 0.f;
  for (int t = threadIdx.x; t < n_rows * D; t += blockDim.x) {
    const int r = t / D, d = t % D;
    const float ref = dk_scale * (float)st->gref[(size_t)r * D + d];
    e = fmaxf(e, fabsf(grad[(size_t)r * row_step * D + d] - ref));
  }
  e = warp_max(e);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = e;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) e = fmaxf(e, red[w]);
    const float gmax = __uint_as_float(st->gmax_bits);
    const float rel = gmax > 0.f ? e / gmax : 0.f;
    st->verify_rel = rel;
    st->flag = (rel <= budget) ? 0 : 1;                       // NaN compares false: the parity sweep runs
  }
}

// general-D path: float64 accumulators -> grad_unscaled; per-row kernel sums -> the three sums in a fixed order
__global__ void __launch_bounds__(256)
mmd_simt_finish_kernel(const double* __restrict__ gacc, int64_t count, float dk_scale, float* __restrict__ grad,
                       const double* __restrict__ krow, int64_t rows, double* __restrict__ sums) {
  __shared__ double red[3 * 8];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x)
    grad[i] = dk_scale * (float)gacc[i];
  if (blockIdx.x == 0) {
    double v[3] = {0.0, 0.0, 0.0};
    for (int64_t r = threadIdx.x; r < rows; r += blockDim.x) { v[0] += krow[r]; v[1] += krow[rows + r]; v[2] += krow[2 * rows + r]; }
    block_sum<3>(v, red);
    if (threadIdx.x == 0) { sums[0] = v[0]; sums[1] = v[1]; sums[2] = v[2]; }
  }
}

// out[4] = { mmd, term_pp, term_zz, term_pz } from the (all-reduced) sums {zz, pz, pp}; wae_mmd.py:197-203
__global__ void mmd_combine_kernel(const double* __restrict__ sums, double n, int kernel_type, float reg, float* out) {
  const double m = kernel_type == 0 ? 1.0 / (n * n) : 1.0;
  const float zz = (float)(sums[0] * m), pz = (float)(sums[1] * m), pp = (float)(sums[2] * m);
  out[0] = reg * pp + reg * zz - 2 * reg * pz;
  out[1] = pp;
  out[2] = zz;
  out[3] = pz;
}

__global__ void scale_by_scalar_kernel(const float* __restrict__ in, const float* __restrict__ g, float factor, int64_t n,
                                       float* __restrict__ out) {
  const float f = (g ? *g : 1.f) * factor;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = in[i] * f;
}

__global__ void mmd_diag_kernel(const MmdStats* __restrict__ st, float* __restrict__ diag) {
  diag[0] = (float)st->mode_used;
  diag[1] = st->verify_rel;
  diag[2] = (float)st->flag;
  diag[3] = __uint_as_float(st->gmax_bits);
}

int chain_limit(int kernel_type, int64_t D) {
  // longest accumulation chain in 64-column units (two per block), 0 = none.  The tcgen05 accumulator truncates, a bias
  // ~ chain length x magnitude of the accumulated CENTRED coefficients, i.e. x the relative spread of the kernel over the
  // batch: IMQ ~8 % at the benchmark size -> 128; RBF ~ sqrt(2) / (D^1.5 z_var): 1/2000 at D = 128 (no limit needed),
  // but 1 % at D = 16 and 25 % at D = 2 (N = 9000, D = 2 in the parity sweep: 6.4e-6 with 36-unit chains, 1.5e-5 with
  // 142) -> narrow latent dims get short chains.  VK_MMD_CHAIN overrides.
  static const int env = [] { const char* e = getenv("VK_MMD_CHAIN"); return e ? atoi(e) : -1; }();
  if (env >= 0) return env;
  // (IMQ: the spread of d^2 relative to C + d^2 goes like sqrt(2 / D): N = 9000, D = 2 measures 6.9e-6 with 36-unit
  // chains and 1.55e-5 with 95)
  if (kernel_type != 0) return D <= 4 ? 16 : D <= 16 ? 32 : D <= 64 ? 64 : 128;
  return D <= 4 ? 16 : D <= 16 ? 64 : D <= 32 ? 128 : 0;
}
constexpr int kFallbackChain = 32;      // the parity sweep behind a failed sample check: short chains, it is the slow path anyway

void make_shape(int64_t N, int64_t D, int kernel_type, float z_var, int64_t row_begin, int64_t row_end, MmdShape& s) {
  s.N = N; s.D = (int)D; s.Dp = rup((int)D, 16); s.nkc = s.Dp / 8;
  s.T = (int)((N + kRows - 1) / kRows);
  s.rb0 = (int)(row_begin / kRows);
  s.rb1 = (int)((row_end + kRows - 1) / kRows);
  s.Q = pick_Q(s.T, s.rb1 - s.rb0, chain_limit(kernel_type, D));
  { static const int qm = [] { const char* e = getenv("VK_MMD_QMAJOR"); return e ? atoi(e) : 1; }(); s.qmajor = qm; }
  s.P = pick_P(s.T, s.Q);
  s.kernel_type = kernel_type; s.g1p = 3; s.g2p = 3;
  const double sp = 2.0 * (double)D * (double)D * (double)z_var;
  s.a_exp = (float)(-1.4426950408889634 / sp);
  s.C = (float)(2.0 * (double)D * (double)z_var);
  s.Ceps = (float)(1e-7 + (double)s.C);
  s.dk_scale = kernel_type == 0 ? (float)(-1.0 / sp) : -s.C;      // applied once, in finalize
}

template <int G1P, int G2P>
int launch_sweep(const float* z, const float* prior, const MmdShape& s, const MmdWs& w, uint8_t* ws, double* sums_out,
                 float* grad_unscaled, int run_if, cudaStream_t st) {
  const MmdSmem L = mmd_smem(s.Dp);
  const int n_items = (s.rb1 - s.rb0) * (s.Q + s.P);
  const int grid = n_items < sm_count() ? n_items : sm_count();
  auto kern = s.kernel_type == 0 ? mmd_main_kernel<G1P, G2P, 0> : mmd_main_kernel<G1P, G2P, 1>;
  static bool attr_done[2] = {false, false};                  // once per instantiation and kernel type
  if (!attr_done[s.kernel_type]) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mmd_smem(128).total) != cudaSuccess)
      return check_launch("mmd_fwd (shared memory attribute)");
    attr_done[s.kernel_type] = true;
  }
  MmdStats* stats = (MmdStats*)(ws + w.stats);
  kern<<<grid, kMmdThreads, L.total, st>>>(z, prior, ws + w.imgZb, ws + w.imgPb, (const float*)(ws + w.normZ),
                                           (const float*)(ws + w.normP), s, n_items, (float*)(ws + w.Opart),
                                           (float*)(ws + w.rspart), (double*)(ws + w.cta_sums), stats, run_if);
  int rc = check_launch("mmd_main");
  if (rc != VK_OK) return rc;
  mmd_finalize_kernel<<<sm_count() * 8, 256, 0, st>>>(z, prior, s, (const float*)(ws + w.Opart), (const float*)(ws + w.rspart),
                                                       (const double*)(ws + w.cta_sums), grid, grad_unscaled, sums_out, stats,
                                                       run_if);
  return check_launch("mmd_finalize");
}

int dispatch_sweep(int g1p, int g2p, const float* z, const float* prior, MmdShape s, const MmdWs& w, uint8_t* ws,
                   double* sums_out, float* grad_unscaled, int run_if, cudaStream_t st) {
  s.g1p = g1p; s.g2p = g2p;
  if (g1p == 3 && g2p == 3) return launch_sweep<3, 3>(z, prior, s, w, ws, sums_out, grad_unscaled, run_if, st);
  if (g1p == 2 && g2p == 1) return launch_sweep<2, 1>(z, prior, s, w, ws, sums_out, grad_unscaled, run_if, st);
  if (g1p == 2 && g2p == 2) return launch_sweep<2, 2>(z, prior, s, w, ws, sums_out, grad_unscaled, run_if, st);
  if (g1p == 3 && g2p == 1) return launch_sweep<3, 1>(z, prior, s, w, ws, sums_out, grad_unscaled, run_if, st);
  if (g1p == 3 && g2p == 2) return launch_sweep<3, 2>(z, prior, s, w, ws, sums_out, grad_unscaled, run_if, st);
  return launch_sweep<1, 1>(z, prior, s, w, ws, sums_out, grad_unscaled, run_if, st);
}

struct SimtWs { size_t gacc, krow, total; };
inline SimtWs simt_ws(int64_t rows, int64_t D) {
  SimtWs w;
  size_t o = kWsHeader;
  w.krow = o; o += (size_t)3 * rows * sizeof(double);
  o = (o + 255) & ~(size_t)255;
  w.gacc = o; o += (size_t)rows * D * sizeof(double);
  w.total = o;
  return w;
}

}  // namespace
}  // namespace vk

using namespace vk;

extern "C" int64_t vk_mmd_workspace_bytes(int64_t N, int64_t D) {
  if (N < 1 || D < 1) return -1;
  if (D > 128) return (int64_t)simt_ws(N, D).total;           // SIMT path: float64 gradient accumulators
  const int T = (int)((N + kRows - 1) / kRows);
  const int64_t items = max_grad_items(T);
  VK_REQUIRE(items < (int64_t)1 << 31, -1, "mmd_workspace_bytes: N too large");
  return (int64_t)mmd_ws(N, (int)D, (int)items, 1).total;
}

extern "C" int vk_mmd_fwd(const float* z, const float* prior, int64_t N, int64_t D, int kernel_type, float z_var,
                          int passes, int64_t row_begin, int64_t row_end, double* sums_out, float* grad_unscaled,
                          float* diag, void* workspace, int64_t workspace_bytes, void* stream) {
  VK_REQUIRE(z && prior && sums_out && grad_unscaled && workspace, VK_ERR_BAD_ARG, "mmd_fwd: null pointer");
  VK_REQUIRE(N >= 1 && D >= 1, VK_ERR_BAD_ARG, "mmd: empty z [%lld,%lld]", (long long)N, (long long)D);
  VK_REQUIRE(N <= (int64_t)1 << 24, VK_ERR_UNSUPPORTED, "mmd: N=%lld too large", (long long)N);
  VK_REQUIRE(kernel_type == 0 || kernel_type == 1, VK_ERR_BAD_ARG, "mmd: unknown kernel type %d", kernel_type);
  VK_REQUIRE((passes >= 0 && passes <= 3) || passes == 22 || passes == 31 || passes == 32, VK_ERR_BAD_ARG,
             "mmd: passes must be 0 (auto), 1, 2, 3 or one of the explicit pairs 22, 31, 32");
  VK_REQUIRE(row_begin >= 0 && row_begin < row_end && row_end <= N && row_begin % kRows == 0 &&
             (row_end % kRows == 0 || row_end == N), VK_ERR_BAD_ARG,
             "mmd: row range [%lld,%lld) must be 128-aligned inside [0,%lld)", (long long)row_begin, (long long)row_end, (long long)N);
  cudaStream_t st = (cudaStream_t)stream;
  uint8_t* ws = (uint8_t*)workspace;
  MmdShape s;
  make_shape(N, D, kernel_type, z_var, row_begin, row_end, s);
  if (D > 128) {
    // ---- general latent dim: SIMT rows, every term in the reference's (x - y)^2 form ----
    const int64_t rows = row_end - row_begin;
    const SimtWs w = simt_ws(rows, D);
    VK_REQUIRE(workspace_bytes >= (int64_t)w.total, VK_ERR_WORKSPACE, "mmd_fwd: workspace %lld < %lld",
               (long long)workspace_bytes, (long long)w.total);
    VK_REQUIRE(D <= 2048, VK_ERR_UNSUPPORTED, "mmd: latent dim %lld > 2048", (long long)D);
    double* ks = (double*)(ws + w.krow);
    double* gacc = (double*)(ws + w.gacc);
    cudaMemsetAsync(gacc, 0, (size_t)rows * D * sizeof(double), st);
    const size_t shb = ((size_t)8 * D + 8) * sizeof(double);
    auto kern = kernel_type == 0 ? mmd_simt_rows_kernel<0> : mmd_simt_rows_kernel<1>;
    if (shb > 48 * 1024 && cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shb) != cudaSuccess)
      return check_launch("mmd_fwd (SIMT shared memory attribute)");
    const dim3 grid((unsigned)rows, 1);
    kern<<<grid, 256, shb, st>>>(z, z, N, (int)D, row_begin, 1, rows, s.a_exp, s.Ceps, s.C, 1.0, gacc, false, ks);
    kern<<<grid, 256, shb, st>>>(z, prior, N, (int)D, row_begin, 1, rows, s.a_exp, s.Ceps, s.C, -1.0, gacc, false, ks + rows);
    kern<<<grid, 256, shb, st>>>(prior, prior, N, (int)D, row_begin, 1, rows, s.a_exp, s.Ceps, s.C, 0.0, nullptr, false, ks + 2 * rows);
    int rc = check_launch("mmd_simt_rows");
    if (rc != VK_OK) return rc;
    mmd_simt_finish_kernel<<<sm_count() * 2, 256, 0, st>>>(gacc, rows * D, s.dk_scale, grad_unscaled, ks, rows, sums_out);
    if (diag) cudaMemsetAsync(diag, 0, 4 * sizeof(float), st);
    return check_launch("mmd_simt_finish");
  }
  const MmdWs w = mmd_ws(N, (int)D, s.rb1 - s.rb0, s.Q);
  VK_REQUIRE(workspace_bytes >= (int64_t)w.total, VK_ERR_WORKSPACE, "mmd_fwd: workspace %lld < %lld",
             (long long)workspace_bytes, (long long)w.total);
  MmdStats* stats = (MmdStats*)(ws + w.stats);
  double* colpart = (double*)(ws + w.colpart);
  // ---- statistics, centre, operand images ----
  const int sblocks = (int)((N + 7) / 8 < kStatBlocks ? (N + 7) / 8 : kStatBlocks);
  mmd_colsum_kernel<<<sblocks, 256, 0, st>>>(z, prior, N, (int)D, colpart);
  mmd_stats_finalize_kernel<<<1, 1024, 0, st>>>(colpart, sblocks, N, (int)D, kernel_type, s.a_exp, s.Ceps, stats);
  const int64_t pack_rows = (int64_t)s.T * kRows;
  const int64_t pack_cap = (int64_t)sm_count() * 8;
  const int pack_blocks = (int)((pack_rows + 15) / 16 < pack_cap ? (pack_rows + 15) / 16 : pack_cap);
  const float nscale = kernel_type == 0 ? s.a_exp : 1.f;
  mmd_pack_norm_kernel<<<dim3(pack_blocks, 2), 256, 0, st>>>(z, prior, stats, N, (int)D, s.Dp, nscale, ws + w.imgZb, ws + w.imgPb,
                                                             (float*)(ws + w.normZ), (float*)(ws + w.normP), s.T);
  int rc = check_launch("mmd_pack");
  if (rc != VK_OK) return rc;
  // ---- precision policy ----
  int g1p = 3, g2p = 3;
  bool verify = false;
  if (passes == 1) { g1p = 1; g2p = 1; }
  else if (passes == 2) { g1p = 2; g2p = 1; }
  else if (passes > 3) { g1p = passes / 10; g2p = passes % 10; }
  else if (passes == 0 && N >= kAutoMinN && kernel_type == 0) {
    // RBF at large N: |dk/dd^2| = k / (2 D^2 z_var) makes k insensitive to the bf16 rounding of the COLUMN operand (an
    // error that is random over the columns and averages out of a row's gradient) and the centred coefficients are tiny,
    // so the gradient GEMM runs in one pass and the distance GEMM in two (row operand hi + lo, column operand hi).  The
    // rounding of the ROW operand does not average -- it moves the whole row's kernel values coherently (measured at
    // N = 65 536: 8.7e-6 with one pass, 2.8e-7 with the row operand split) -- so that half of the split stays.
    // Whether the data agree is checked below, not assumed.
    g1p = 2; g2p = 1; verify = true;
  }
  rc = dispatch_sweep(g1p, g2p, z, prior, s, w, ws, sums_out, grad_unscaled, -1, st);
  if (rc != VK_OK) return rc;
  if (verify) {
    const int64_t nloc = (row_end - row_begin);
    // a rank that sweeps 1/G of the rows checks 1/G of the sample (at least 2 rows): the SIMT passes cost rows x N and
    // would otherwise be a fixed cost that does not shrink with the shard
    int64_t want = (kSampleRows * nloc + N - 1) / N;
    if (want < 2) want = 2;
    if (want > kSampleRows) want = kSampleRows;
    const int n_rows = (int)(nloc < want ? nloc : want);
    const int64_t row_step = nloc / n_rows;
    auto kern = kernel_type == 0 ? mmd_sample_kernel<0> : mmd_sample_kernel<1>;
    const dim3 grid((unsigned)(sm_count() * 2), (unsigned)((n_rows + kSampleGroup - 1) / kSampleGroup));
    kern<<<grid, 256, 0, st>>>(z, z, N, (int)D, row_begin, row_step, n_rows, s.a_exp, s.Ceps, 1.0, stats->gref);
    kern<<<grid, 256, 0, st>>>(z, prior, N, (int)D, row_begin, row_step, n_rows, s.a_exp, s.Ceps, -1.0, stats->gref);
    mmd_verify_kernel<<<1, 256, 0, st>>>(grad_unscaled, (int)D, row_step, n_rows, s.dk_scale, kVerifyBudget, stats);
    rc = check_launch("mmd_verify");
    if (rc != VK_OK) return rc;
    MmdShape sf = s;                                      // the parity sweep that runs only if the check failed
    const int ch = chain_limit(kernel_type, D);
    sf.Q = pick_Q(s.T, s.rb1 - s.rb0, ch > 0 && ch < kFallbackChain ? ch : kFallbackChain);
    sf.P = pick_P(s.T, sf.Q);
    const MmdWs wf = mmd_ws(N, (int)D, s.rb1 - s.rb0, sf.Q);
    VK_REQUIRE(workspace_bytes >= (int64_t)wf.total, VK_ERR_WORKSPACE, "mmd_fwd: workspace %lld < %lld (parity sweep)",
               (long long)workspace_bytes, (long long)wf.total);
    rc = dispatch_sweep(3, 3, z, prior, sf, wf, ws, sums_out, grad_unscaled, 1, st);
    if (rc != VK_OK) return rc;
  }
  if (diag) {
    mmd_diag_kernel<<<1, 1, 0, st>>>(stats, diag);
    return check_launch("mmd_diag");
  }
  return VK_OK;
}

extern "C" int vk_mmd_combine(const double* sums, int64_t N, int kernel_type, float reg_weight, float* out, void* stream) {
  VK_REQUIRE(sums && out && N >= 1, VK_ERR_BAD_ARG, "mmd_combine: bad argument");
  mmd_combine_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(sums, (double)N, kernel_type, reg_weight, out);
  return check_launch("mmd_combine");
}

extern "C" int vk_scale_by_scalar(const float* in, const float* g, float factor, int64_t n, float* out, void* stream) {
  VK_REQUIRE(in && out && n >= 0, VK_ERR_BAD_ARG, "scale_by_scalar: bad argument");
  if (n == 0) return VK_OK;
  const int64_t cap = (int64_t)sm_count() * 8;
  const int blocks = (int)((n + 1023) / 1024 < cap ? (n + 1023) / 1024 : cap);
  scale_by_scalar_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(in, g, factor, n, out);
  return check_launch("scale_by_scalar");
}
