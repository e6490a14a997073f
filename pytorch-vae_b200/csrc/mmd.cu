// K4: WAE-MMD  compute_kernel / compute_rbf / compute_inv_mult_quad / compute_mmd
// (reference: models/wae_mmd.py:125-204), forward value and d/dz in ONE sweep over the three Gram matrices.
//
// The reference materialises x1.expand(N,N,D) - x2.expand(N,N,D) (2.2 TB per temporary at N = 65536).  Here the
// pairwise squared distance is  |x|^2 + |y|^2 - 2 x.y  with x.y on tcgen05 tensor cores:
//
//   pack kernels    z, prior_z (fp32 [N,D]) -> bf16 hi/lo operand images, 128-row tiles in the UMMA no-swizzle
//                   canonical layout, + fp32 row norms.  The same image is a K-major operand (K = latent dim)
//                   for the distance GEMM and an MN-major operand (K = sample index) for the gradient GEMM.
//   main kernel     persistent, warp-specialised, one CTA per SM:
//     TMA warp      1-D bulk copies (cp.async.bulk -> mbarrier) of the A row block and of 64-column B stages
//     MMA warp      GEMM1  S[128,64]  = X_i . Y_j^T         (hi*hi + lo*hi + hi*lo, fp32 accumulate in TMEM)
//                   GEMM2  O[128,D]  += C_ij . Y_j          (coefficient tile C written by the epilogue)
//     epilogue (4w) tcgen05.ld S -> d^2 -> k (ex2 / rcp) -> running sum of k, c = dk/dd^2 (signed) ->
//                   row sums of c, bf16 hi/lo of c into shared memory as the A operand of GEMM2
//   finalize        dL/dz_i (unscaled) = z_i * sum_j c_ij - O_i ; per-CTA partial sums -> three kernel sums
//
// Gradient identity:  sum_j c_ij (z_i - y_j) = z_i * (sum_j c_ij) - (C Y)_i.
// The prior-prior Gram needs no gradient: it is evaluated on a circulant half (each unordered tile pair once,
// weight 2), which is also balanced when row blocks are sharded across GPUs.
// Reference quirks kept: RBF = exp(-|x-y|^2 / (2 D^2 z_var)) averaged over the full N x N including the diagonal
// (wae_mmd.py:164-167); IMQ = C / (1e-7 + C + |x-y|^2), C = 2 D z_var, summed with the diagonal removed from
// all three terms, including k(prior_i, z_i) (wae_mmd.py:185-191).
#include <type_traits>
#include "common.cuh"
#include "tc.cuh"

namespace vk {
using namespace tc;

namespace {

constexpr int kRows = 128;       // row block (MMA M)
constexpr int kCols = 64;        // column tile (GEMM1 N, GEMM2 K)
constexpr int kBStages = 4;
constexpr int kNormRing = 8;
// mbarriers
constexpr int kBarAFull = 0, kBarAEmpty = 1, kBarBFull = 2, kBarBEmpty = kBarBFull + kBStages, kBarSFull = kBarBEmpty + kBStages,
              kBarSEmpty = kBarSFull + 2, kBarKFull = kBarSEmpty + 2, kBarKEmpty = kBarKFull + 2, kBarOFull = kBarKEmpty + 2,
              kBarOEmpty = kBarOFull + 1, kNumBars = kBarOEmpty + 1;
// tensor memory map (512 columns): S 2 x 64 | O <= 128 | row-block operand hi, lo (Dp/2 columns each, two bf16 per
// column) | coefficient tile hi, lo (32 columns each), double-buffered
constexpr uint32_t kTmS = 0, kTmO = 128, kTmAhi = 256, kTmAlo = 320, kTmK = 384;
constexpr int kEpiWarps = 16, kColGroups = kEpiWarps / 4, kTmaWarp = 16, kMmaWarp = 17;
constexpr int kEpiThreads = 32 * kEpiWarps;
constexpr int kMmdThreads = 32 * 18;  // warps 0-15 epilogue, warp 16 TMA, warp 17 MMA
// Column chunks per row block.  Two rules pick Q: enough gradient items for the static schedule to balance (up to
// kBalanceQ), and a bound on how many 64-column tiles are accumulated into one gradient tile in tensor memory: the
// tcgen05 fp32 accumulator truncates, a bias that grows with the length of the chain (measured at N = 65 536 with 683
// tiles per chain: 2.3e-4 of the gradient's maximum against float64; 16-tile chains at N = 4096 stay below 2e-5).
// Chains are therefore cut at kChainTiles and the partial tiles are summed in fp32 round-to-nearest by the finalize
// kernel; kMaxGradItems bounds the workspace that costs (4 GiB of partial tiles at D = 128).
constexpr int kMaxQ = 64, kBalanceQ = 8, kChainTiles = 32;
constexpr int64_t kMaxGradItems = 65536;
// -DVK_MMD_INNER_DRAIN=1 (NOT validated on a GPU yet): cut the chains INSIDE a scheduler item instead --
// the MMA warp commits o_full every s.chain live tiles, the epilogue drains O into sub-slot (q * subs + sub) while the
// distance GEMMs of the next tiles run, and the row block's A operand stays in tensor memory.  Q follows the balance
// rule only.  With the switch off the preprocessed source is the validated r01z kernel, token for token.
#ifndef VK_MMD_INNER_DRAIN
#define VK_MMD_INNER_DRAIN 0
#endif

struct MmdShape {
  int64_t N;
  int D, Dp, nkc;       // latent dim, padded to 16, number of 8-element K chunks
  int T;                // 128-row tiles of the full matrix
  int lz;               // live 64-column tiles per side: ceil(N / 64)
  int rb0, rb1;         // row-block range owned by this call
  int Q;                // column chunks per row block (gradient items)
#if VK_MMD_INNER_DRAIN
  int subs, chain;      // partial-gradient sub-slots per item, live tiles per accumulation chain
#endif
  int kernel_type;      // 0 rbf, 1 imq
  int passes;           // 3 (hi/lo split, fp32-grade) or 1 (plain bf16)
  float a_exp;          // rbf: -log2(e) / (2 D^2 z_var)
  float dk_scale;       // dk/dd^2 = dk_scale * coefficient: rbf -1 / (2 D^2 z_var) (coef k) ; imq -C (coef (k/C)^2)
  float C, Ceps;        // imq: C, fl32(1e-7 + C)
};

__host__ __device__ inline int rup(int x, int m) { return (x + m - 1) / m * m; }
__host__ __device__ inline uint32_t tile_img_bytes(int Dp) { return 2u * kRows * Dp * 2u; }   // hi + lo of 128 rows

struct MmdWs {
  size_t cta_sums, imgZb, imgPb, normZ, normP, Opart, rspart, total;
};
inline MmdWs mmd_ws(int64_t N, int D, int rows_local_blocks, int Q) {
  const int Dp = rup(D, 16);
  const int T = (int)((N + kRows - 1) / kRows);
  MmdWs w;
  size_t o = kWsHeader;
  w.cta_sums = o; o += (size_t)256 * 3 * sizeof(double);
  o = (o + 255) & ~(size_t)255;
  w.imgZb = o; o += (size_t)T * tile_img_bytes(Dp);      // bf16 hi/lo images in 64-row granules (column stages: one bulk copy each)
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
inline int pick_Q(int T, int row_blocks) {
  // aim for >= 8 gradient items per SM so the static round-robin schedule balances
  int q = (148 * 8 + row_blocks - 1) / row_blocks;
  if (q > kBalanceQ) q = kBalanceQ;
  const int chain = (4 * T + kChainTiles - 1) / kChainTiles;      // accumulation chains of at most kChainTiles tiles
  if (chain > q) q = chain;
  if (q > q_cap(row_blocks)) q = q_cap(row_blocks);
  if (q > 4 * T) q = 4 * T;
  if (q < 1) q = 1;
  const int n = 4 * T, len = (n + q - 1) / q;
  return (n + len - 1) / len;                 // no empty chunk
}
#if VK_MMD_INNER_DRAIN
inline int pick_Q_balance(int T, int row_blocks) {
  int q = (148 * 8 + row_blocks - 1) / row_blocks;
  if (q > kBalanceQ) q = kBalanceQ;
  if (q > 4 * T) q = 4 * T;
  if (q < 1) q = 1;
  const int n = 4 * T, len = (n + q - 1) / q;
  return (n + len - 1) / len;
}
// sub-slots per item and live tiles per chain: kChainTiles unless that needs more than kMaxGradItems partial tiles
inline void pick_chain(int T, int row_blocks, int Q, int& subs, int& chain) {
  const int n = 4 * T, len = (n + Q - 1) / Q;
  int64_t per_rb = kMaxGradItems / (row_blocks > 0 ? row_blocks : 1);
  if (per_rb < Q) per_rb = Q;
  const int subs_cap = (int)(per_rb / Q);                        // >= 1
  chain = kChainTiles;
  subs = (len + chain - 1) / chain;
  if (subs > subs_cap) {
    chain = (len + subs_cap - 1) / subs_cap;
    subs = (len + chain - 1) / chain;
  }
  if (subs < 1) subs = 1;
}
// bound of Q * subs * row_blocks over every row range: Q * subs <= T / 8 + 9, and never more than the cap allows
inline int64_t max_grad_slots(int T) {
  const int64_t a = (int64_t)T * (T / 8 + 9);
  int64_t b = kMaxGradItems;
  if (b < (int64_t)kBalanceQ * T) b = (int64_t)kBalanceQ * T;
  return a < b ? a : b;
}
#endif
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
  m.B = o; o += kBStages * 2u * kCols * Dp * 2u;             // column stages: hi + lo image of 64 rows each
  m.norms = o; o += kNormRing * kCols * 4u;   // deeper than the B ring: read by the epilogue after b_empty may have fired
  m.bars = o; o += kNumBars * 8;
  m.tmem = o; o += 16;
  m.red = o; o += kEpiWarps * 3 * 8;
  m.total = o;
  return m;
}

// ---- pack: fp32 [N,D] -> image tiles [T][hi/lo][Dp/8][128][8] bf16 ; norms [T*128] fp32 ----------------------
__global__ void mmd_pack_kernel(const float* __restrict__ X, int64_t N, int D, int Dp, uint8_t* __restrict__ imgB, int T) {
  const int nkc = Dp / 8;
  const int64_t total = (int64_t)T * kRows * nkc;
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = g / nkc;
    const int kc = (int)(g % nkc);
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int d = kc * 8 + e;
      v[e] = (row < N && d < D) ? X[row * D + d] : 0.f;
    }
    uint32_t ph[4], pl[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      __nv_bfloat16 h0, l0, h1, l1;
      split_bf16(v[2 * e], h0, l0);
      split_bf16(v[2 * e + 1], h1, l1);
      ph[e] = pack2(h0, h1);
      pl[e] = pack2(l0, l1);
    }
    // 64-row granules: [granule][hi/lo][Dp/8][64][8]
    uint8_t* gb = imgB + (row / kCols) * (tile_img_bytes(Dp) / 2);
    const int rb = (int)(row % kCols);
    *reinterpret_cast<uint4*>(gb + ((size_t)kc * kCols + rb) * 16) = make_uint4(ph[0], ph[1], ph[2], ph[3]);
    *reinterpret_cast<uint4*>(gb + ((size_t)(nkc + kc) * kCols + rb) * 16) = make_uint4(pl[0], pl[1], pl[2], pl[3]);
  }
}
// one warp per row, fixed summation order
__global__ void mmd_norm_kernel(const float* __restrict__ X, int64_t N, int D, float scale, float* __restrict__ norms, int T) {
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= (int64_t)T * kRows) return;
  float s = 0.f;
  if (row < N) for (int d = lane; d < D; d += 32) { const float v = X[row * D + d]; s = fmaf(v, v, s); }
  s = warp_sum(s);
  if (lane == 0) norms[row] = s * scale;                 // RBF: |x|^2 * a_exp (the epilogue's ex2 argument), IMQ: |x|^2
}

// ---- work items ---------------------------------------------------------------------------------------------
// Item order: for each owned row block i: Q gradient items (i, q), then the prior-prior item of row block i.
struct Item {
  int i;          // row block (global index)
  int grad;       // 1: rows = z, columns = [z tiles | prior tiles] chunk q ; 0: rows = prior, circulant half of prior
  int u0, u1;     // tile sequence range
};
struct TileD {
  int bm;         // column matrix: 0 z, 1 prior
  int jt;         // 64-column tile index
  int kind;       // 0 zz, 1 zp, 2 pp
  float weight;
  int diag;       // tile touches the main diagonal
};
__device__ __forceinline__ Item get_item(const MmdShape& s, int idx) {
  Item it;
  const int per = s.Q + 1;
  it.i = s.rb0 + idx / per;
  const int q = idx % per;
  if (q < s.Q) {
    it.grad = 1;
    const int n = 4 * s.T, len = (n + s.Q - 1) / s.Q;
    it.u0 = q * len;
    it.u1 = it.u0 + len < n ? it.u0 + len : n;
    if (it.u0 > n) it.u0 = n;
  } else {
    it.grad = 0;
    const int noff = (s.T % 2 == 0) ? s.T / 2 + 1 : (s.T + 1) / 2;
    it.u0 = 0;
    it.u1 = 2 * noff;
  }
  return it;
}
__device__ __forceinline__ TileD get_tile(const MmdShape& s, const Item& it, int u) {
  TileD t;
  if (it.grad) {
    const bool zside = u < 2 * s.T;
    t.bm = zside ? 0 : 1;
    t.jt = zside ? u : u - 2 * s.T;
    t.kind = zside ? 0 : 1;
    t.weight = 1.f;
    t.diag = (t.jt >> 1) == it.i;
  } else {
    const int o = u >> 1;
    int j128 = it.i + o;                                  // o <= T/2 + 1: one conditional wrap instead of a modulo
    if (j128 >= s.T) j128 -= s.T;
    t.bm = 1;
    t.jt = 2 * j128 + (u & 1);
    t.kind = 2;
    t.weight = (o == 0 || (s.T % 2 == 0 && o == s.T / 2)) ? 1.f : 2.f;
    t.diag = o == 0;
  }
  return t;
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
__device__ __forceinline__ bool tile_live(const MmdShape& s, const TileD& t) { return t.jt < s.lz; }

// ---- main kernel ----------------------------------------------------------------------------------------------
template <int PASSES, int KT>
__global__ void __launch_bounds__(kMmdThreads, 1)
mmd_main_kernel(const float* __restrict__ Zf, const float* __restrict__ Pf, const uint8_t* __restrict__ imgZb,
                const uint8_t* __restrict__ imgPb, const float* __restrict__ normZ,
                const float* __restrict__ normP, MmdShape s, int n_items, float* __restrict__ Opart,
                float* __restrict__ rspart, double* __restrict__ cta_sums) {
  extern __shared__ __align__(128) uint8_t smem[];
  const MmdSmem L = mmd_smem(s.Dp);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(smem);
  auto bar = [&](int i) { return sbase + L.bars + 8u * i; };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L.tmem);
  if (threadIdx.x == 0) {
    mbar_init(bar(kBarAFull), kEpiThreads); mbar_init(bar(kBarAEmpty), 1);
    for (int i = 0; i < kBStages; ++i) { mbar_init(bar(kBarBFull + i), 1); mbar_init(bar(kBarBEmpty + i), 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(kBarSFull + i), 1); mbar_init(bar(kBarSEmpty + i), kEpiThreads / 2);      // one tile group per S / Kmat buffer use
      mbar_init(bar(kBarKFull + i), kEpiThreads / 2); mbar_init(bar(kBarKEmpty + i), 1);
    }
    mbar_init(bar(kBarOFull), 1); mbar_init(bar(kBarOEmpty), kEpiThreads);
    mbar_fence_init();
  }
  if (warp == kMmaWarp) tmem_alloc(smem_u32(tmem_slot), 512);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tmem_S = tmem + kTmS, tmem_O = tmem + kTmO;
  const uint32_t b_half = (uint32_t)kCols * s.Dp * 2u;      // one hi (or lo) image of a 64-row stage

  if (warp == kTmaWarp) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t nb = 0;
      int iter = 0;
      for (int idx = blockIdx.x; idx < n_items; idx += gridDim.x, ++iter) {
        const Item it = get_item(s, idx);
        for (int u = it.u0; u < it.u1; ++u) {
          const TileD t = get_tile(s, it, u);
          if (!tile_live(s, t)) continue;
          const uint32_t slot = nb % kBStages;
          if (nb >= kBStages) mbar_wait(bar(kBarBEmpty + slot), ((nb / kBStages) - 1) & 1);
          const uint8_t* bimg = (t.bm ? imgPb : imgZb) + (size_t)t.jt * (2 * b_half);
          const float* nrm = (t.bm ? normP : normZ) + (size_t)t.jt * kCols;
          mbar_expect_tx(bar(kBarBFull + slot), 2 * b_half + kCols * 4);
          bulk_g2s(sbase + L.B + slot * (2u * b_half), bimg, 2 * b_half, bar(kBarBFull + slot));      // hi + lo images of 64 rows, one copy
          bulk_g2s(sbase + L.norms + (nb % kNormRing) * kCols * 4, nrm, kCols * 4, bar(kBarBFull + slot));
          ++nb;
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // ===== MMA issuer: the whole warp runs the loop converged (operands stay in uniform registers), one elected
    // lane issues =====
    {
      const uint32_t leader = elect_one();
      const uint32_t idesc1 = idesc_bf16(kRows, kCols, 0, 0);       // S = A(K-major) . B(K-major)^T
      const uint32_t idesc2 = idesc_bf16(kRows, s.Dp, 0, 1);        // O += Kmat(K-major) . B(MN-major)
      const uint32_t b_lbo = kCols * 16;
      const uint64_t b_step = (2u * b_lbo) >> 4, b2_step = 256u >> 4;
      const uint32_t tmem_Ahi = tmem + kTmAhi, tmem_Alo = tmem + kTmAlo;
      const int ksteps1 = s.Dp / 16;
      uint32_t nb = 0, ns = 0, nk = 0, n_grad_items = 0;
      int iter = 0;
      PROF_DECL
      TRACE_DECL(4)
      auto gemm1 = [&](uint32_t slot) {
        TRACE(0, ns)
        PROF_LAP(3)
        mbar_wait(bar(kBarBFull + slot), (nb / kBStages) & 1);
        PROF_LAP(0)
        const uint32_t sb = ns & 1;
        if (ns >= 2) mbar_wait(bar(kBarSEmpty + sb), ((ns >> 1) - 1) & 1);
        PROF_LAP(1)
        fence_after_sync();
        const uint32_t b_hi = sbase + L.B + slot * (2u * b_half);
        const uint64_t db_hi = smem_desc(b_hi, b_lbo, 128), db_lo = smem_desc(b_hi + b_half, b_lbo, 128);
        uint32_t acc = 0;
#pragma unroll
        for (int pass = 0; pass < PASSES; ++pass) {
          // passes: lo*hi, hi*lo, then hi*hi LAST -- the tensor core's fp32 accumulation truncates (~2^-23 |acc| per
          // MMA), so the correction terms are summed while the accumulator is still small (gemm3x.cuh, measured 8x)
          uint32_t ta = (PASSES == 3 && pass == 0) ? tmem_Alo : tmem_Ahi;   // operand from tensor memory: 8 columns per K=16
          uint64_t db = (PASSES == 3 && pass == 1) ? db_lo : db_hi;
#pragma unroll 4
          for (int k = 0; k < ksteps1; ++k) {
            mma_bf16_ts_if(leader, tmem_S + sb * kCols, ta, db, idesc1, acc);
            acc = 1;
            ta += 8;
            db += b_step;
          }
        }
        mma_commit_if(leader, bar(kBarSFull + sb));
        TRACE(1, ns)
        ++ns;
      };
      for (int idx = blockIdx.x; idx < n_items; idx += gridDim.x, ++iter) {
        const Item it = get_item(s, idx);
        mbar_wait(bar(kBarAFull), iter & 1);
        fence_after_sync();
        // live tiles of this item, in order
        int live = 0;
        for (int u = it.u0; u < it.u1; ++u) live += tile_live(s, get_tile(s, it, u)) ? 1 : 0;
        const uint32_t nb0 = nb;
        // software pipeline: G1(0), G1(1), then for each j: G1(j+2), G2(j)
        int issued1 = 0;
        for (; issued1 < 2 && issued1 < live; ++issued1) {
          gemm1((nb0 + issued1) % kBStages);
          if (!it.grad) mma_commit_if(leader, bar(kBarBEmpty + (nb0 + issued1) % kBStages));
          ++nb;
        }
        uint32_t oacc = 0;
#if VK_MMD_INNER_DRAIN
        bool o_wait = true;                                  // the next G2 is the first one into a fresh accumulator
#endif
        // live z-side tiles of this chunk come first (tile sequence: [z tiles | prior tiles], dead tiles = columns >= N)
        int nz_live = 0;
        if (it.grad) {
          const int zend = it.u1 < 2 * s.T ? it.u1 : 2 * s.T;
          const int hi = zend < s.lz ? zend : s.lz;
          nz_live = hi > it.u0 ? hi - it.u0 : 0;
        }
        for (int j = 0; j < live; ++j) {
          // G1(j+2) goes first: its S buffer was freed as soon as the epilogue had tile j in registers, so the distance
          // GEMM of tile j+2 runs while the epilogue is still computing the coefficients G2(j) waits for
          if (issued1 < live) {
            gemm1((nb0 + issued1) % kBStages);
            if (!it.grad) mma_commit_if(leader, bar(kBarBEmpty + (nb0 + issued1) % kBStages));
            ++nb;
            ++issued1;
          }
          if (it.grad) {
            const uint32_t slot = (nb0 + j) % kBStages, kb = nk & 1;
            const uint32_t neg_a = j >= nz_live ? (1u << 13) : 0u;           // k(z, prior) enters with a minus sign
            PROF_LAP(3)
            mbar_wait(bar(kBarKFull + kb), (nk >> 1) & 1);
#if VK_MMD_INNER_DRAIN
            if (o_wait && n_grad_items > 0) mbar_wait(bar(kBarOEmpty), (n_grad_items - 1) & 1);
            o_wait = false;
#else
            if (j == 0 && n_grad_items > 0) mbar_wait(bar(kBarOEmpty), (n_grad_items - 1) & 1);
#endif
            PROF_LAP(2)
            TRACE(2, nk)
            fence_after_sync();
            const uint32_t tk_hi = tmem + kTmK + kb * 64, tk_lo = tk_hi + 32;
            const uint32_t b_hi = sbase + L.B + slot * (2u * b_half);
            // A: coefficient tile in tensor memory (K = column index) ; B: MN-major (N = latent dims, K = rows of the stage)
            const uint64_t db_hi = smem_desc(b_hi, 128, b_lbo), db_lo = smem_desc(b_hi + b_half, 128, b_lbo);
#pragma unroll
            for (int pass = 0; pass < PASSES; ++pass) {
              uint32_t ta = (PASSES == 3 && pass == 0) ? tk_lo : tk_hi;
              uint64_t db = (PASSES == 3 && pass == 1) ? db_lo : db_hi;
#pragma unroll
              for (int k = 0; k < kCols / 16; ++k) {
                mma_bf16_ts_if(leader, tmem_O, ta, db, idesc2 | neg_a, oacc);
                oacc = 1;
                ta += 8;
                db += b2_step;
              }
            }
            mma_commit_if(leader, bar(kBarBEmpty + slot));
            mma_commit_if(leader, bar(kBarKEmpty + kb));
            TRACE(3, nk)
            ++nk;
#if VK_MMD_INNER_DRAIN
            if ((j + 1) % s.chain == 0 && j + 1 < live) {      // chain complete: hand O to the epilogue, start a new one
              mma_commit_if(leader, bar(kBarOFull));
              ++n_grad_items;
              oacc = 0;
              o_wait = true;
            }
#endif
          }
        }
        if (it.grad && live > 0) { mma_commit_if(leader, bar(kBarOFull)); ++n_grad_items; }
        mma_commit_if(leader, bar(kBarAEmpty));
      }
      PROF_LAP(3)
      PROF_OUT("mma wait_b/wait_s_empty/wait_k_full/issue", leader)
      TRACE_OUT("mma g1_start/g1_issued/g2_kfull/g2_issued", 4, leader)
    }
  } else {
    // ===== epilogue: 16 warps = 2 tile groups x 4 TMEM lane quadrants x 2 column halves.  Group g takes the tiles
    // that land in S buffer g, so two tiles are in the epilogue at once (four warps per scheduler to hide the TMEM /
    // MUFU / shared-memory latencies) while every warp still amortises its per-tile bookkeeping over 32 columns =====
    const int ew = warp & 3;                                  // TMEM lane quadrant (must equal warp id % 4)
    const int h = (warp >> 2) & 1;                            // column half of the tile
    const uint32_t grp = warp >> 3;                           // tile group == S buffer
    const int cg = warp >> 2;                                 // slot of this warp's partial row sums / O chunk
    const int t = ew * 32 + lane;                             // row within the block
    double acc_zz = 0.0, acc_pz = 0.0, acc_pp = 0.0, acc_run = 0.0;   // acc_run: sums of the term run_kind, not yet flushed
    int run_kind = -1;
    uint32_t ns = 0, nk = 0, nb = 0, n_grad_items = 0, item_iter = 0;
    // RBF: norms arrive pre-scaled by a_exp, so ex2's argument is one FADD + one FFMA; IMQ: t = C + eps + d^2
    const float s_coef = KT == 0 ? -2.f * s.a_exp : -2.f;
    PROF_DECL
    TRACE_DECL(3)
    for (int idx = blockIdx.x; idx < n_items; idx += gridDim.x) {
      const Item it = get_item(s, idx);
      const int n32 = (int)s.N;                                // N <= 2^24: 32-bit index math in the tile loop
      const int grow = it.i * kRows + t;
      const bool rvalid = grow < n32;
      const bool row_tail = (it.i + 1) * kRows > n32;     // this row block has rows beyond N
      const float nx = (it.grad ? normZ : normP)[grow] + (KT == 1 ? s.Ceps : 0.f);
      float rs = 0.f;
      int live = 0;
#if VK_MMD_INNER_DRAIN
      const int q_item = idx % (s.Q + 1);
      const size_t lrow_i = (size_t)(it.i - s.rb0) * kRows + t;
      const size_t nloc_i = (size_t)(s.rb1 - s.rb0) * kRows;
      int nlive = 0;
      if (it.grad)
        for (int u = it.u0; u < it.u1; ++u) nlive += tile_live(s, get_tile(s, it, u)) ? 1 : 0;
      // partial slot (q * subs + sub) <- this warp's row sums of the chain and its 32 x 32 chunk of O (zeros when there
      // is no accumulator to read); every epilogue warp calls it at the same live-tile count (.sync.aligned inside)
      auto drain_o = [&](int sub, bool have_o) {
        const size_t slot = (size_t)q_item * s.subs + sub;
        float* op = Opart + (slot * nloc_i + lrow_i) * s.Dp;
        rspart[(slot * kColGroups + cg) * nloc_i + lrow_i] = rs;
        rs = 0.f;
        const int c0 = cg * 32;
        if (!have_o) {
          for (int e = c0; e < c0 + 32 && e < s.Dp; ++e) op[e] = 0.f;
          return;
        }
        mbar_wait(bar(kBarOFull), n_grad_items & 1);
        fence_after_sync();
        if (c0 < s.Dp) {
          uint32_t o[32];
          tmem_ld32(tmem_O + ((uint32_t)(ew * 32) << 16) + c0, o);
          tmem_ld_wait();
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
        mbar_arrive(bar(kBarOEmpty));
        ++n_grad_items;
      };
#endif
      {
        // this row block's operand -> tensor memory: fp32 row slice [cg*32, cg*32+32) split into bf16 hi / lo, two per
        // column.  The previous item's MMAs must be finished with the old block first.
        if (item_iter > 0) mbar_wait(bar(kBarAEmpty), (item_iter - 1) & 1);
        fence_after_sync();
        const float* xr = (it.grad ? Zf : Pf) + (int64_t)grow * s.D + cg * 32;
        const uint32_t ta = tmem + ((uint32_t)(ew * 32) << 16) + cg * 16;
#pragma unroll
        for (int j8 = 0; j8 < 16; j8 += 8) {
          uint32_t ah[8], al[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int d = cg * 32 + 2 * (j8 + j);
            const float a = (rvalid && d < s.D) ? __ldg(xr + 2 * (j8 + j)) : 0.f;
            const float b = (rvalid && d + 1 < s.D) ? __ldg(xr + 2 * (j8 + j) + 1) : 0.f;
            split_pair(a, b, ah[j], al[j]);
          }
          tmem_st8(ta + kTmAhi + j8, ah);
          tmem_st8(ta + kTmAlo + j8, al);
        }
        tmem_st_wait();
        fence_before_sync();
        mbar_arrive(bar(kBarAFull));
        ++item_iter;
      }
      for (int u = it.u0; u < it.u1; ++u) {
        const TileD td = get_tile(s, it, u);
        if (!tile_live(s, td)) continue;
        if ((ns & 1) == grp) {
          PROF_LAP(1)
          mbar_wait(bar(kBarSFull + grp), (ns >> 1) & 1);
          PROF_LAP(0)
          TRACE(0, ns)
          fence_after_sync();
          uint32_t rr[2][16];
          const uint32_t taddr = tmem_S + ((uint32_t)(ew * 32) << 16) + grp * kCols + h * 32;
          tmem_ld16(taddr, rr[0]);
          tmem_ld16(taddr + 16, rr[1]);
          const float* ny = reinterpret_cast<const float*>(smem + L.norms + (nb % kNormRing) * kCols * 4) + h * 32;
          const int gcol0 = td.jt * kCols + h * 32;
          // uniform across the warp (the two instances below contain .sync.aligned tcgen05 instructions)
          const bool slow = td.diag || gcol0 + 32 > n32 || row_tail;
          float ksum = 0.f, csum = 0.f;                       // sum of k (value) and of the coefficient (gradient)
          const uint32_t tk = tmem + ((uint32_t)(ew * 32) << 16) + kTmK + (nk & 1) * 64 + h * 16;
          const bool diag_excl = KT == 1 && td.diag;
          // interior tiles take the branch-free instance; diagonal / tail tiles the masked one
          auto tile_math = [&](auto masked_c, auto grad_c) {
            constexpr bool MASKED = decltype(masked_c)::value, GRAD = decltype(grad_c)::value;
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            tmem_ld_wait16(rr[hf]);
            if (hf == 1) {                                    // both halves are in registers: the next GEMM1 into this
              fence_before_sync();                            // S buffer may start
              mbar_arrive(bar(kBarSEmpty + grp));
            }
            uint32_t kh[8], kl[8];                            // 16 coefficients: 8 hi + 8 lo columns of the tile in TMEM
#pragma unroll
            for (int c8 = 0; c8 < 16; c8 += 8) {
              float cv[8];
              const float4 n0 = *reinterpret_cast<const float4*>(ny + hf * 16 + c8);
              const float4 n1 = *reinterpret_cast<const float4*>(ny + hf * 16 + c8 + 4);
              const float nyv[8] = {n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, n1.z, n1.w};
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                const int c = hf * 16 + c8 + e;
                // no clamp of d^2 at 0: a rounding-level negative distance moves k by < 1e-9 relative
                const float arg = fmaf(__uint_as_float(rr[hf][c8 + e]), s_coef, nx + nyv[e]);
                float k;
                if (KT == 0) {
                  k = ex2_approx(arg);                        // exp(-d^2 / (2 D^2 z_var))
                  cv[e] = k;                                  // dk/dd^2 = k * dk_scale: the constant is applied in finalize
                } else {
                  k = rcp_approx(arg);                        // k / C = 1 / (C + eps + d^2)
                  cv[e] = k * k;                              // dk/dd^2 = -C * (k/C)^2
                }
                if (MASKED) {                                 // diagonal exclusion, tail rows / columns
                  const bool ok = rvalid && (gcol0 + c < n32) && !(diag_excl && gcol0 + c == grow);
                  k = ok ? k : 0.f;
                  cv[e] = ok ? cv[e] : 0.f;
                }
                ksum += k;
                if (KT == 1) csum += cv[e];
              }
              if (GRAD) {
#pragma unroll
                for (int e = 0; e < 4; ++e) split_pair(cv[2 * e], cv[2 * e + 1], kh[(c8 >> 1) + e], kl[(c8 >> 1) + e]);
              }
            }
            if (GRAD) {
              if (hf == 0) {                                  // the coefficient buffer is needed only now: G2 of the tile
                PROF_LAP(1)                                   // two back has had the first half of this tile's math to finish
                if (nk >= 2) mbar_wait(bar(kBarKEmpty + (nk & 1)), ((nk >> 1) - 1) & 1);
                PROF_LAP(2)
                TRACE(1, ns)
              }
              tmem_st8(tk + hf * 8, kh);
              tmem_st8(tk + 32 + hf * 8, kl);
            }
          }
          };
          if (it.grad) {
            if (slow) tile_math(std::true_type{}, std::true_type{}); else tile_math(std::false_type{}, std::true_type{});
          } else {
            if (slow) tile_math(std::true_type{}, std::false_type{}); else tile_math(std::false_type{}, std::false_type{});
          }
          if (KT == 0) csum = ksum;
          rs += td.kind == 1 ? -csum : csum;                  // GEMM2 applies the same sign with the negate-A bit
          if (td.kind != run_kind) {                          // (warp-uniform, once or twice per item)
            if (run_kind == 0) acc_zz += acc_run; else if (run_kind == 1) acc_pz += acc_run; else if (run_kind == 2) acc_pp += acc_run;
            acc_run = 0.0;
            run_kind = td.kind;
          }
          acc_run += (double)(ksum * td.weight);
          if (it.grad) {
            tmem_st_wait();
            fence_before_sync();
            mbar_arrive(bar(kBarKFull + (nk & 1)));
          }
          TRACE(2, ns)
        }
        ++ns;
        ++live;
        if (it.grad) ++nk;
        ++nb;
#if VK_MMD_INNER_DRAIN
        if (it.grad && live % s.chain == 0 && live < nlive) drain_o(live / s.chain - 1, true);
#endif
      }
      PROF_LAP(1)
#if VK_MMD_INNER_DRAIN
      if (it.grad) {
        const int last = live == 0 ? 0 : (live - 1) / s.chain;
        drain_o(last, live > 0);
        for (int sub = last + 1; sub < s.subs; ++sub) drain_o(sub, false);     // unused sub-slots of a short item
      }
#else
      if (it.grad) {
        const int q = (idx % (s.Q + 1));
        const size_t lrow = (size_t)(it.i - s.rb0) * kRows + t;
        const size_t nloc = (size_t)(s.rb1 - s.rb0) * kRows;
        float* op = Opart + ((size_t)q * nloc + lrow) * s.Dp;
        rspart[((size_t)q * kColGroups + cg) * nloc + lrow] = rs;
        const int c0 = cg * 32;                               // this warp's 32-column chunk of O (Dp <= 128)
        if (live == 0) {                                      // chunk made only of column tiles beyond N
          for (int e = c0; e < c0 + 32 && e < s.Dp; ++e) op[e] = 0.f;
          continue;
        }
        // O complete -> partial gradient rows to the workspace
        mbar_wait(bar(kBarOFull), n_grad_items & 1);
        fence_after_sync();
        if (c0 < s.Dp) {
          uint32_t o[32];
          tmem_ld32(tmem_O + ((uint32_t)(ew * 32) << 16) + c0, o);
          tmem_ld_wait();
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
        mbar_arrive(bar(kBarOEmpty));
        ++n_grad_items;
      }
#endif
      PROF_LAP(3)
    }
    PROF_OUT("epilogue wait_s_full/work/wait_k_empty/item_end", threadIdx.x == 0)
    TRACE_OUT("epi s_full/k_empty/done", 3, threadIdx.x == 0 || threadIdx.x == 256)
    if (run_kind == 0) acc_zz += acc_run; else if (run_kind == 1) acc_pz += acc_run; else if (run_kind == 2) acc_pp += acc_run;
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

// ---- finalize: G_i = z_i * sum_q rs - sum_q O ; the three kernel sums ----------------------------------------
__global__ void mmd_finalize_kernel(const float* __restrict__ Z, MmdShape s, const float* __restrict__ Opart,
                                    const float* __restrict__ rspart, const double* __restrict__ cta_sums, int grid_main,
                                    float* __restrict__ grad, double* __restrict__ sums) {
  const int64_t row0 = (int64_t)s.rb0 * kRows;
  const int64_t nloc_pad = (int64_t)(s.rb1 - s.rb0) * kRows;
  int64_t nloc = s.N - row0;
  if (nloc > nloc_pad) nloc = nloc_pad;
  const int64_t total = nloc * s.D;
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = g / s.D;
    const int d = (int)(g % s.D);
    float rs = 0.f, o = 0.f;
#if VK_MMD_INNER_DRAIN
    const int n_slots = s.Q * s.subs;
#else
    const int n_slots = s.Q;
#endif
    for (int q = 0; q < n_slots; ++q) {
      for (int c = 0; c < kColGroups; ++c) rs += rspart[(size_t)(q * kColGroups + c) * nloc_pad + r];
      o += Opart[((size_t)q * nloc_pad + r) * s.Dp + d];
    }
    grad[g] = s.dk_scale * fmaf(Z[(row0 + r) * s.D + d], rs, -o);     // dk/dd^2 constant: -1/(2 D^2 z_var) | -C
  }
  if (blockIdx.x == 0 && threadIdx.x < 3) {
    double t = 0.0;
    for (int c = 0; c < grid_main; ++c) t += cta_sums[c * 3 + threadIdx.x];
    sums[threadIdx.x] = t;                       // order: zz, pz, pp
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

int make_shape(int64_t N, int64_t D, int kernel_type, float z_var, int passes, int64_t row_begin, int64_t row_end,
               MmdShape& s) {
  VK_REQUIRE(N >= 1 && D >= 1, VK_ERR_BAD_ARG, "mmd: empty z [%lld,%lld]", (long long)N, (long long)D);
  VK_REQUIRE(D <= 128, VK_ERR_UNSUPPORTED, "mmd: latent dim %lld > 128 is not supported by the Gram tile kernel", (long long)D);
  VK_REQUIRE(N <= (int64_t)1 << 24, VK_ERR_UNSUPPORTED, "mmd: N=%lld too large", (long long)N);
  VK_REQUIRE(kernel_type == 0 || kernel_type == 1, VK_ERR_BAD_ARG, "mmd: unknown kernel type %d", kernel_type);
  VK_REQUIRE(passes == 1 || passes == 3, VK_ERR_BAD_ARG, "mmd: passes must be 1 or 3");
  VK_REQUIRE(row_begin >= 0 && row_begin < row_end && row_end <= N && row_begin % kRows == 0 &&
             (row_end % kRows == 0 || row_end == N), VK_ERR_BAD_ARG,
             "mmd: row range [%lld,%lld) must be 128-aligned inside [0,%lld)", (long long)row_begin, (long long)row_end, (long long)N);
  s.N = N; s.D = (int)D; s.Dp = rup((int)D, 16); s.nkc = s.Dp / 8;
  s.T = (int)((N + kRows - 1) / kRows);
  s.lz = (int)((N + kCols - 1) / kCols);
  s.rb0 = (int)(row_begin / kRows);
  s.rb1 = (int)((row_end + kRows - 1) / kRows);
#if VK_MMD_INNER_DRAIN
  s.Q = pick_Q_balance(s.T, s.rb1 - s.rb0);
  pick_chain(s.T, s.rb1 - s.rb0, s.Q, s.subs, s.chain);
#else
  s.Q = pick_Q(s.T, s.rb1 - s.rb0);
#endif
  s.kernel_type = kernel_type; s.passes = passes;
  const double sp = 2.0 * (double)D * (double)D * (double)z_var;
  s.a_exp = (float)(-1.4426950408889634 / sp);
  s.C = (float)(2.0 * (double)D * (double)z_var);
  s.Ceps = (float)(1e-7 + (double)s.C);
  s.dk_scale = kernel_type == 0 ? (float)(-1.0 / sp) : -s.C;      // applied once, in finalize
  return VK_OK;
}

}  // namespace
}  // namespace vk

using namespace vk;

extern "C" int64_t vk_mmd_workspace_bytes(int64_t N, int64_t D) {
  if (N < 1 || D < 1 || D > 128) return -1;
  const int T = (int)((N + kRows - 1) / kRows);
#if VK_MMD_INNER_DRAIN
  const int64_t items = max_grad_slots(T);
#else
  const int64_t items = max_grad_items(T);
#endif
  VK_REQUIRE(items < (int64_t)1 << 31, -1, "mmd_workspace_bytes: N too large");
  return (int64_t)mmd_ws(N, (int)D, (int)items, 1).total;
}

extern "C" int vk_mmd_fwd(const float* z, const float* prior, int64_t N, int64_t D, int kernel_type, float z_var,
                          int passes, int64_t row_begin, int64_t row_end, double* sums_out, float* grad_unscaled,
                          void* workspace, int64_t workspace_bytes, void* stream) {
  VK_REQUIRE(z && prior && sums_out && grad_unscaled && workspace, VK_ERR_BAD_ARG, "mmd_fwd: null pointer");
  MmdShape s;
  int rc = make_shape(N, D, kernel_type, z_var, passes, row_begin, row_end, s);
  if (rc != VK_OK) return rc;
#if VK_MMD_INNER_DRAIN
  const MmdWs w = mmd_ws(N, (int)D, s.rb1 - s.rb0, s.Q * s.subs);
#else
  const MmdWs w = mmd_ws(N, (int)D, s.rb1 - s.rb0, s.Q);
#endif
  VK_REQUIRE(workspace_bytes >= (int64_t)w.total, VK_ERR_WORKSPACE, "mmd_fwd: workspace %lld < %lld",
             (long long)workspace_bytes, (long long)w.total);
  cudaStream_t st = (cudaStream_t)stream;
  uint8_t* ws = (uint8_t*)workspace;
  const int64_t pack_work = (int64_t)s.T * kRows * s.nkc;
  const int pack_blocks = (int)((pack_work + 255) / 256 < 148 * 16 ? (pack_work + 255) / 256 : 148 * 16);
  const int norm_blocks = (int)(((int64_t)s.T * kRows * 32 + 255) / 256);
  mmd_pack_kernel<<<pack_blocks, 256, 0, st>>>(z, N, (int)D, s.Dp, ws + w.imgZb, s.T);
  mmd_pack_kernel<<<pack_blocks, 256, 0, st>>>(prior, N, (int)D, s.Dp, ws + w.imgPb, s.T);
  const float nscale = kernel_type == 0 ? s.a_exp : 1.f;
  mmd_norm_kernel<<<norm_blocks, 256, 0, st>>>(z, N, (int)D, nscale, (float*)(ws + w.normZ), s.T);
  mmd_norm_kernel<<<norm_blocks, 256, 0, st>>>(prior, N, (int)D, nscale, (float*)(ws + w.normP), s.T);
  const MmdSmem L = mmd_smem(s.Dp);
  const int n_items = (s.rb1 - s.rb0) * (s.Q + 1);
  const int grid = n_items < sm_count() ? n_items : sm_count();
  auto kern = passes == 3 ? (kernel_type == 0 ? mmd_main_kernel<3, 0> : mmd_main_kernel<3, 1>)
                          : (kernel_type == 0 ? mmd_main_kernel<1, 0> : mmd_main_kernel<1, 1>);
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total) != cudaSuccess)
    return check_launch("mmd_fwd (shared memory attribute)");
  kern<<<grid, kMmdThreads, L.total, st>>>(z, prior, ws + w.imgZb, ws + w.imgPb, (const float*)(ws + w.normZ),
                                           (const float*)(ws + w.normP), s, n_items, (float*)(ws + w.Opart),
                                           (float*)(ws + w.rspart), (double*)(ws + w.cta_sums));
  rc = check_launch("mmd_main");
  if (rc != VK_OK) return rc;
  mmd_finalize_kernel<<<148 * 4, 256, 0, st>>>(z, s, (const float*)(ws + w.Opart), (const float*)(ws + w.rspart),
                                               (const double*)(ws + w.cta_sums), grid, grad_unscaled, sums_out);
  return check_launch("mmd_finalize");
}

extern "C" int vk_mmd_combine(const double* sums, int64_t N, int kernel_type, float reg_weight, float* out, void* stream) {
  VK_REQUIRE(sums && out && N >= 1, VK_ERR_BAD_ARG, "mmd_combine: bad argument");
  mmd_combine_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(sums, (double)N, kernel_type, reg_weight, out);
  return check_launch("mmd_combine");
}

extern "C" int vk_scale_by_scalar(const float* in, const float* g, float factor, int64_t n, float* out, void* stream) {
  VK_REQUIRE(in && out && n >= 0, VK_ERR_BAD_ARG, "scale_by_scalar: bad argument");
  if (n == 0) return VK_OK;
  const int blocks = (int)((n + 1023) / 1024 < 148 * 8 ? (n + 1023) / 1024 : 148 * 8);
  scale_by_scalar_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(in, g, factor, n, out);
  return check_launch("scale_by_scalar");
}
