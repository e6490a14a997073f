// fp32-grade GEMM on tcgen05 for the two SWAE matrix products (swae.py:164-166 projection, and its backward):
//     C(n, j) = sum_k A(n, k) * B(j, k)        n: up to millions of rows, j and k: a few hundred
//
// Every fp32 operand is split three ways into bf16 (x = h + m + l, 24 significant bits) and the product is accumulated
// in fp32 in tensor memory from six kind::f16 passes  hh' + hm' + mh' + mm' + hl' + lh'  (the dropped terms are below
// 2^-24 |x||y|).  A 3-pass / 16-bit split is NOT enough here: the projections are sort keys, and 2^-17 errors reorder
// neighbours of a 10^6-key segment whose gaps are ~1e-6.
//
//   B (the small matrix: theta) is packed once per call into three UMMA K-major images [Kp/8][Jp][8] that stay resident
//   in shared memory for the whole persistent CTA.
//   A (the tall matrix) never touches shared memory: four producer warps read 128-row tiles with their natural strides
//   (software-pipelined one 64-k group ahead), split in registers and tcgen05.st the three images straight into tensor memory (lane = row, 32-bit column = two
//   consecutive k), from where the MMAs take their A operand.
//   The accumulator is two column halves in tensor memory, so the epilogue of one half (tcgen05.ld -> strided global
//   stores, coalesced along whichever of n / j is contiguous in C) overlaps the MMAs of the other.
#pragma once
#include "common.cuh"
#include "tc.cuh"

namespace vk {
namespace g3 {
using namespace tc;

constexpr int kRows = 128;
constexpr int kProducerWarps = 4, kEpilogueWarps = 4, kMmaWarp = 8;
constexpr int kThreads = 32 * (kMmaWarp + 1);

struct Args {
  const float* A; float* C;
  int64_t N;                 // rows
  int J, K, Jp, Kp;          // output columns / contraction length, padded to 16
  int64_t a_sn, a_sk;        // element strides of A(n, k)
  int64_t c_sn, c_sj;        // element strides of C(n, j)
  int JH0;                   // columns of the first accumulator half (multiple of 16); the second has Jp - JH0
  int64_t tiles;
  // Optional column tables (device arrays): a_tab[k] = address of A(0, k), c_tab[j] = address of C(0, j).  A column may
  // live in ANOTHER GPU's memory (peer-mapped): the multi-GPU SWD stores each projection straight into the segment buffer
  // of the rank that sorts it and reads d/dlat columns straight from their owners -- the exchange is fused into the GEMM.
  const float* const* a_tab;
  float* const* c_tab;
};

__host__ __device__ inline int rup16(int x) { return (x + 15) / 16 * 16; }
inline size_t b_image_bytes(int Jp, int Kp) { return (size_t)Jp * Kp * 2; }
inline size_t smem_bytes(int Jp, int Kp) { return 3 * b_image_bytes(Jp, Kp) + 256; }
// tensor memory: three A images of Kp/2 columns, then the accumulator (Jp columns)
inline bool fits(int J, int K) {
  const int Jp = rup16(J), Kp = rup16(K);
  return Jp <= 256 && Kp <= 256 && 3 * (Kp / 2) + Jp <= 512 && smem_bytes(Jp, Kp) <= 200 * 1024;
}

// three-way bf16 split of a pair: packed {hi: b, lo: a} for each of the three terms
__device__ __forceinline__ void split3_pair(float a, float b, uint32_t& h2, uint32_t& m2, uint32_t& l2) {
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h2) : "f"(b), "f"(a));
  const float ra = a - __uint_as_float(h2 << 16), rb = b - __uint_as_float(h2 & 0xFFFF0000u);
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(m2) : "f"(rb), "f"(ra));
  const float sa = ra - __uint_as_float(m2 << 16), sb = rb - __uint_as_float(m2 & 0xFFFF0000u);
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(l2) : "f"(sb), "f"(sa));
}

// B(j, k) = src[j * b_sj + k * b_sk]  ->  images h | m | l, each [Kp/8][Jp][8] bf16 (zero padded)
__global__ void pack_b_kernel(const float* __restrict__ src, int J, int K, int Jp, int Kp, int64_t b_sj, int64_t b_sk,
                              uint8_t* __restrict__ img) {
  const int total = Jp * (Kp / 8);
  const size_t ib = (size_t)Jp * Kp * 2;
  for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < total; g += gridDim.x * blockDim.x) {
    const int j = g % Jp, kc = g / Jp;
    uint32_t h[4], m[4], l[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int k0 = kc * 8 + 2 * e;
      const float a = (j < J && k0 < K) ? src[j * b_sj + k0 * b_sk] : 0.f;
      const float b = (j < J && k0 + 1 < K) ? src[j * b_sj + (k0 + 1) * b_sk] : 0.f;
      split3_pair(a, b, h[e], m[e], l[e]);
    }
    const size_t o = ((size_t)kc * Jp + j) * 16;
    *reinterpret_cast<uint4*>(img + o) = make_uint4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<uint4*>(img + ib + o) = make_uint4(m[0], m[1], m[2], m[3]);
    *reinterpret_cast<uint4*>(img + 2 * ib + o) = make_uint4(l[0], l[1], l[2], l[3]);
  }
}

// barriers: 0 b_full | 1 a_full | 2 a_empty | 3,4 acc_full | 5,6 acc_empty
__global__ void __launch_bounds__(kThreads, 1)
gemm3x_kernel(Args g, const uint8_t* __restrict__ bimg) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[8];
  __shared__ uint32_t tmem_slot;
  __shared__ const float* a_tab_s[256];                // column tables (K, J <= 256), staged once per CTA
  __shared__ float* c_tab_s[256];
  if (g.a_tab) for (int i = threadIdx.x; i < g.K; i += kThreads) a_tab_s[i] = g.a_tab[i];
  if (g.c_tab) for (int i = threadIdx.x; i < g.J; i += kThreads) c_tab_s[i] = g.c_tab[i];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(smem);
  auto bar = [&](int i) { return smem_u32(&bars[i]); };
  const uint32_t ib = (uint32_t)g.Jp * g.Kp * 2u;      // bytes of one B image
  const int kch = g.Kp / 16;                           // K steps of 16
  const uint32_t acol = (uint32_t)g.Kp / 2;            // tensor-memory columns of one A image
  if (threadIdx.x == 0) {
    mbar_init(bar(0), 1);
    mbar_init(bar(1), 32 * kProducerWarps); mbar_init(bar(2), 1);
    mbar_init(bar(3), 1); mbar_init(bar(4), 1);
    mbar_init(bar(5), 32 * kEpilogueWarps); mbar_init(bar(6), 32 * kEpilogueWarps);
    mbar_fence_init();
  }
  if (warp == kMmaWarp) tmem_alloc(smem_u32(&tmem_slot), 512);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const uint32_t tA = tmem, tAcc = tmem + 3 * acol;
  const int JH[2] = {g.JH0, g.Jp - g.JH0};

  if (warp < kProducerWarps) {
    // ===== producers: A rows -> three bf16 images in tensor memory =====
    // Work unit = one group of kGroup K-steps (64 k) of one tile.  The loads of unit u+1 (which may belong to the
    // next tile) are issued before unit u is split and stored, so every lane keeps 2 x 256 B of A in flight --
    // across the wait for the MMAs of the previous tile too; without this the loop exposed one DRAM round trip per
    // 64 bytes and the whole GEMM ran at 1.2 TB/s.
    constexpr int kGroup = 4;
    const bool vec = !g.a_tab && g.a_sk == 1 && (g.K & 3) == 0 && (g.a_sn & 3) == 0 && ((uintptr_t)g.A & 15) == 0;
    const int groups = (kch + kGroup - 1) / kGroup;
    auto load_unit = [&](float (&v)[16 * kGroup], int64_t tile, int grp) {
      const int64_t n = tile * kRows + warp * 32 + lane;
      const bool live = tile < g.tiles && n < g.N;
      const float* ap = g.A + (live ? n : 0) * g.a_sn;
      const int k0 = grp * (16 * kGroup);
      if (g.a_tab) {                                  // per-column bases (possibly peer memory): plain loads
        const int64_t no = (live ? n : 0) * g.a_sn;
#pragma unroll
        for (int e = 0; e < 16 * kGroup; ++e) v[e] = (live && k0 + e < g.K) ? a_tab_s[k0 + e][no] : 0.f;
        return;
      }
      if (live && k0 + 16 * kGroup <= g.K) {          // whole group in range: no predicates, pointer bumps only
        if (vec) {
          const float4* q = reinterpret_cast<const float4*>(ap + k0);
#pragma unroll
          for (int e = 0; e < 4 * kGroup; ++e) {
            const float4 f = __ldg(q + e);
            v[4 * e] = f.x; v[4 * e + 1] = f.y; v[4 * e + 2] = f.z; v[4 * e + 3] = f.w;
          }
        } else {
          const float* q = ap + (int64_t)k0 * g.a_sk;
#pragma unroll
          for (int e = 0; e < 16 * kGroup; ++e) { v[e] = __ldg(q); q += g.a_sk; }
        }
        return;
      }
      // ragged group (tail of K, or rows past N): per 16-k step, skipping steps that lie wholly in the padding
#pragma unroll
      for (int kk = 0; kk < kGroup; ++kk) {
        const int k1 = k0 + 16 * kk;
        if (!live || k1 >= g.K) {
#pragma unroll
          for (int e = 0; e < 16; ++e) v[16 * kk + e] = 0.f;
        } else if (vec) {
#pragma unroll
          for (int e = 0; e < 16; e += 4) {
            const float4 f = k1 + e < g.K ? __ldg(reinterpret_cast<const float4*>(ap + k1 + e)) : make_float4(0.f, 0.f, 0.f, 0.f);
            v[16 * kk + e] = f.x; v[16 * kk + e + 1] = f.y; v[16 * kk + e + 2] = f.z; v[16 * kk + e + 3] = f.w;
          }
        } else {
          const float* q = ap + (int64_t)k1 * g.a_sk;
#pragma unroll
          for (int e = 0; e < 16; ++e) { v[16 * kk + e] = k1 + e < g.K ? __ldg(q) : 0.f; q += g.a_sk; }
        }
      }
    };
    // The tile two rounds ahead is pulled into L2 by bulk prefetches (one row or one k-slice per thread), so the
    // register prefetch one group ahead only has to cover an L2 hit, not a DRAM round trip.
    const bool pf_rows = vec && !g.a_tab;                                                       // row-major A: K * 4 bytes per row
    const bool pf_cols = !vec && !g.a_tab && g.a_sn == 1 && (g.a_sk & 3) == 0 && ((uintptr_t)g.A & 15) == 0;   // k-major A
    auto prefetch_tile = [&](int64_t tile) {
      if (tile >= g.tiles) return;
      const int64_t n0 = tile * kRows;
      if (pf_rows) {
        const int64_t n = n0 + warp * 32 + lane;
        if (n < g.N) l2_prefetch_bulk(g.A + n * g.a_sn, (uint32_t)g.K * 4u);
      } else if (pf_cols) {
        const int64_t rows = g.N - n0 < kRows ? g.N - n0 : kRows;
        const uint32_t bytes = (uint32_t)(rows * 4) & ~15u;
        if (bytes)
          for (int k = warp * 32 + lane; k < g.K; k += 32 * kProducerWarps) l2_prefetch_bulk(g.A + k * g.a_sk + n0, bytes);
      }
    };
    float cur[16 * kGroup], nxt[16 * kGroup];
    int it = 0;
    if ((int64_t)blockIdx.x < g.tiles) load_unit(cur, blockIdx.x, 0);
    prefetch_tile((int64_t)blockIdx.x + gridDim.x);
    for (int64_t tile = blockIdx.x; tile < g.tiles; tile += gridDim.x, ++it) {
      const uint32_t ta = tA + ((uint32_t)(warp * 32) << 16);
      prefetch_tile(tile + 2 * (int64_t)gridDim.x);
      for (int grp = 0; grp < groups; ++grp) {
        const bool last = grp + 1 == groups;
        load_unit(nxt, last ? tile + gridDim.x : tile, last ? 0 : grp + 1);
        if (grp == 0) {
          if (it > 0) mbar_wait(bar(2), (it - 1) & 1);         // the MMAs of the previous tile have consumed A
          fence_after_sync();
        }
#pragma unroll
        for (int kk = 0; kk < kGroup; ++kk) {
          const int kc = grp * kGroup + kk;
          if (kc < kch) {
            uint32_t h[8], m[8], l[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) split3_pair(cur[16 * kk + 2 * e], cur[16 * kk + 2 * e + 1], h[e], m[e], l[e]);
            tmem_st8(ta + kc * 8, h);
            tmem_st8(ta + acol + kc * 8, m);
            tmem_st8(ta + 2 * acol + kc * 8, l);
          }
        }
#pragma unroll
        for (int e = 0; e < 16 * kGroup; ++e) cur[e] = nxt[e];
      }
      tmem_st_wait();
      fence_before_sync();
      mbar_arrive(bar(1));
    }
  } else if (warp == kMmaWarp) {
    // ===== MMA issuer (converged warp, elected lane) =====
    const uint32_t leader = elect_one();
    if (leader) {
      mbar_expect_tx(bar(0), 3 * ib);
      bulk_g2s(sbase, bimg, 3 * ib, bar(0));
    }
    __syncwarp();
    mbar_wait(bar(0), 0);
    const uint32_t b_lbo = (uint32_t)g.Jp * 16;
    const uint64_t b_step = (2u * b_lbo) >> 4;
    int it = 0;
    for (int64_t tile = blockIdx.x; tile < g.tiles; tile += gridDim.x, ++it) {
      mbar_wait(bar(1), it & 1);
      fence_after_sync();
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int nh = JH[half];
        if (nh == 0) continue;
        if (it > 0) mbar_wait(bar(5 + half), (it - 1) & 1);
        fence_after_sync();
        const uint32_t idesc = idesc_bf16(kRows, nh, 0, 0);
        const uint32_t acc = tAcc + (half ? g.JH0 : 0);
        const uint32_t rowoff = (half ? (uint32_t)g.JH0 : 0u) * 16u;
        uint32_t accum = 0;
        // (A image, B image):  hl' + lh' + mm' + hm' + mh' first, hh' LAST.  The tensor core's fp32 accumulation
        // truncates (measured: ~2^-23 |acc| lost per MMA), so the 5/6 of the MMAs that carry the small correction
        // terms run while the accumulator is still ~2^-8 of its final size; only the K/16 MMAs of hh' see it large.
#pragma unroll
        for (int pass = 0; pass < 6; ++pass) {
          const int pa = pass == 1 ? 2 : (pass == 2 || pass == 4 ? 1 : 0);
          const int pb = pass == 0 ? 2 : (pass == 2 || pass == 3 ? 1 : 0);
          uint32_t ta = tA + pa * acol;
          uint64_t db = smem_desc(sbase + pb * ib + rowoff, b_lbo, 128);
          for (int k = 0; k < kch; ++k) {
            mma_bf16_ts_if(leader, acc, ta, db, idesc, accum);
            accum = 1;
            ta += 8;
            db += b_step;
          }
        }
        mma_commit_if(leader, bar(3 + half));
      }
      mma_commit_if(leader, bar(2));
    }
  } else {
    // ===== epilogue: accumulator half -> C with its strides =====
    const int q = warp & 3;
    const bool vec = !g.c_tab && g.c_sj == 1 && (g.J & 3) == 0 && (g.c_sn & 3) == 0 && ((uintptr_t)g.C & 15) == 0;
    int it = 0;
    for (int64_t tile = blockIdx.x; tile < g.tiles; tile += gridDim.x, ++it) {
      const int64_t n = tile * kRows + q * 32 + lane;
      float* cp = g.C + n * g.c_sn;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int nh = JH[half];
        if (nh == 0) continue;
        const int hoff = half ? g.JH0 : 0;
        mbar_wait(bar(3 + half), it & 1);
        fence_after_sync();
        for (int c0 = 0; c0 < nh; c0 += 16) {
          uint32_t r[16];
          tmem_ld16(tAcc + ((uint32_t)(q * 32) << 16) + hoff + c0, r);
          tmem_ld_wait16(r);
          const int j0 = hoff + c0;
          if (g.c_tab) {                              // per-column bases (possibly peer memory)
            if (n < g.N) {
              const int64_t no = n * g.c_sn;
#pragma unroll
              for (int e = 0; e < 16; ++e)
                if (j0 + e < g.J) c_tab_s[j0 + e][no] = __uint_as_float(r[e]);
            }
          } else if (n < g.N && j0 + 16 <= g.J) {            // whole block in range: no predicates, pointer bumps only
            if (vec) {
              float4* q = reinterpret_cast<float4*>(cp + j0);
#pragma unroll
              for (int e = 0; e < 4; ++e)
                q[e] = make_float4(__uint_as_float(r[4 * e]), __uint_as_float(r[4 * e + 1]), __uint_as_float(r[4 * e + 2]),
                                   __uint_as_float(r[4 * e + 3]));
            } else {
              float* q = cp + (int64_t)j0 * g.c_sj;
#pragma unroll
              for (int e = 0; e < 16; ++e) { *q = __uint_as_float(r[e]); q += g.c_sj; }
            }
          } else if (n < g.N) {
            if (vec) {
#pragma unroll
              for (int e = 0; e < 16; e += 4)
                if (j0 + e < g.J)
                  *reinterpret_cast<float4*>(cp + j0 + e) = make_float4(__uint_as_float(r[e]), __uint_as_float(r[e + 1]),
                                                                        __uint_as_float(r[e + 2]), __uint_as_float(r[e + 3]));
            } else {
#pragma unroll
              for (int e = 0; e < 16; ++e)
                if (j0 + e < g.J) cp[(int64_t)(j0 + e) * g.c_sj] = __uint_as_float(r[e]);
            }
          }
        }
        fence_before_sync();
        mbar_arrive(bar(5 + half));
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem, 512);
}

// C(n, j) = sum_k A(n, k) B(j, k); `bimg` is scratch for the packed B images (3 * Jp * Kp * 2 bytes)
inline int launch(const float* A, int64_t a_sn, int64_t a_sk, const float* B, int64_t b_sj, int64_t b_sk, float* C,
                  int64_t c_sn, int64_t c_sj, int64_t N, int J, int K, uint8_t* bimg, cudaStream_t st, const char* what,
                  const float* const* a_tab = nullptr, float* const* c_tab = nullptr) {
  Args g;
  g.a_tab = a_tab; g.c_tab = c_tab;
  g.A = A; g.C = C; g.N = N; g.J = J; g.K = K; g.Jp = rup16(J); g.Kp = rup16(K);
  g.a_sn = a_sn; g.a_sk = a_sk; g.c_sn = c_sn; g.c_sj = c_sj;
  g.JH0 = rup16(g.Jp / 2);
  g.tiles = (N + kRows - 1) / kRows;
  const int pk = g.Jp * (g.Kp / 8);
  pack_b_kernel<<<(pk + 255) / 256, 256, 0, st>>>(B, J, K, g.Jp, g.Kp, b_sj, b_sk, bimg);
  const size_t sm = smem_bytes(g.Jp, g.Kp);
  if (cudaFuncSetAttribute(gemm3x_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm) != cudaSuccess)
    return check_launch(what);
  const int64_t grid = g.tiles < sm_count() ? g.tiles : sm_count();
  gemm3x_kernel<<<(unsigned)grid, kThreads, sm, st>>>(g, bimg);
  return check_launch(what);
}

}  // namespace g3
}  // namespace vk
