// K7: VQ-VAE VectorQuantizer forward / backward (reference: models/vq_vae.py:24-55).
//
// The reference permutes NCHW -> NHWC, builds the [R,K] distance matrix with two SGEMM-class calls,
// argmin, a [R,K] one-hot, one-hot @ codebook, two mse_loss, the straight-through add and permutes back
// (about 1 GB of temporaries at R = 262144, K = 512).  Here one persistent warp-specialised kernel does all
// of it in a single pass over the latents:
//
//   producer warps (4)  read a 128-row tile of x straight from NCHW (coalesced along h*w), compute |x|^2,
//                       split every value into bf16 hi + lo and write the UMMA operand image to shared memory
//   MMA warp (1 thread) tcgen05.mma  dot[128, 256-code chunk] = x . e^T  as hi*hi + lo*hi + hi*lo
//                       (3 bf16 passes, fp32 accumulate in TMEM: ~2^-17 relative, enough for the argmin to
//                       agree with the fp32 reference wherever its best-vs-second gap exceeds 1e-5)
//   epilogue warps (4)  tcgen05.ld the accumulators, dist = (|x|^2 + |e|^2) - 2 dot  in the reference's
//                       association (vq_vae.py:30-32), running first-index argmin, then gather the fp32 code,
//                       q_out = x + (q - x) (vq_vae.py:51), indices, and the squared-error partial sum
//
// The bf16 hi/lo image of the codebook (built once per call by a tiny pack kernel) is brought into shared
// memory with one 1-D bulk TMA copy and stays resident; TMEM holds two 256-column accumulator buffers so the
// MMA of the next chunk overlaps the argmin of the current one.
#include "common.cuh"
#include "tc.cuh"

namespace vk {
using namespace tc;

namespace {

constexpr int kTileRows = 128;
constexpr int kProducerWarps = 8, kEpilogueWarps = 8, kOutputWarps = 8, kMmaWarp = 24, kParts = kEpilogueWarps / 4;
constexpr int kVqThreads = 32 * (kMmaWarp + 1);   // warps 0-7 producers, 8-15 output, 16-23 argmin, 24 MMA

struct VqShape {
  int64_t B, D, HW, K, R;
  int Dp;       // D padded to a multiple of 16 (MMA K step)
  int NC;       // codes per MMA chunk (<= 256, multiple of 32)
  int Kp;       // K padded to a multiple of NC
  int nch;      // chunks
  int64_t tiles;
};

__host__ __device__ inline int round_up(int x, int m) { return (x + m - 1) / m * m; }

// workspace layout (after the 256-byte ticket header)
struct VqWs {
  size_t partials, image, ee, total;
};
inline VqWs vq_ws(const VqShape& s) {
  VqWs w;
  w.partials = kWsHeader;
  w.image = w.partials + (size_t)kMaxReduceBlocks * sizeof(double);
  w.ee = w.image + (size_t)2 * s.Kp * s.Dp * 2;
  w.total = w.ee + (size_t)s.Kp * 4;
  return w;
}

// smem layout of the forward kernel
struct VqSmem {
  uint32_t cb, x[2], ee, cand, bars, tmem, total;
};
__host__ __device__ inline VqSmem vq_smem(int Dp, int Kp) {
  VqSmem m;
  uint32_t o = 0;
  m.cb = o;            o += 2u * Kp * Dp * 2;            // hi image then lo image
  m.x[0] = o;          o += 2u * kTileRows * Dp * 2;
  m.x[1] = o;          o += 2u * kTileRows * Dp * 2;
  m.ee = o;            o += Kp * 4u;
  m.cand = o;          o += 2u * 4u * kTileRows * 8 + 2u * kTileRows * 4;   // per-part argmin candidates (dist, index), double-buffered; idx hand-over
  m.bars = o;          o += 16 * 8;
  m.tmem = o;          o += 16;
  m.total = o;
  return m;
}

// ---- codebook pack: fp32 [K,D] -> bf16 hi/lo operand images [Dp/8][Kp][8] + |e|^2 ---------------------------
__global__ void vq_pack_codebook_kernel(const float* __restrict__ E, int K, int D, int Kp, int Dp,
                                        __nv_bfloat16* __restrict__ img, float* __restrict__ ee) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= Kp) return;
  __nv_bfloat16* hi = img;
  __nv_bfloat16* lo = img + (size_t)Kp * Dp;
  float s = 0.f;
  for (int d = 0; d < Dp; ++d) {
    const float v = (k < K && d < D) ? E[(size_t)k * D + d] : 0.f;
    s += v * v;
    __nv_bfloat16 h, l;
    split_bf16(v, h, l);
    const size_t a = ((size_t)(d >> 3) * Kp + k) * 8 + (d & 7);
    hi[a] = h;
    lo[a] = l;
  }
  ee[k] = (k < K) ? s : INFINITY;          // padded codes can never win the argmin
}

__device__ __forceinline__ float min3f(float a, float b, float c) {
  float d;
  asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

// ---- forward -------------------------------------------------------------------------------------------------
// warps 0-7 producers (4 row quadrants x 2 halves of D), warps 16-23 argmin (4 TMEM lane quadrants x 2 column parts),
// warps 8-15 output (gather / straight-through store / loss, decoupled so its global-memory latency is off the
// argmin -> MMA critical path), warp 24 MMA issuer.
__global__ void __launch_bounds__(kVqThreads, 1)
vq_fwd_kernel(const float* __restrict__ X, const float* __restrict__ E, const uint8_t* __restrict__ cb_image,
              const float* __restrict__ ee_g, VqShape s, float beta, float* __restrict__ q_out,
              int32_t* __restrict__ idx_out, float* __restrict__ out, unsigned int* ticket, double* partials) {
  extern __shared__ __align__(128) uint8_t smem[];
  const VqSmem L = vq_smem(s.Dp, s.Kp);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sbase = smem_u32(smem);
  // barriers: 0 cb_full | 1,2 x_full | 3,4 x_empty | 5,6 tmem_full | 7,8 tmem_empty | 9,10 idx_full | 11,12 idx_empty
  auto bar = [&](int i) { return sbase + L.bars + 8u * i; };
  float* ee_s = reinterpret_cast<float*>(smem + L.ee);
  float* cand_d = reinterpret_cast<float*>(smem + L.cand);      // [2][4 parts][128]
  int* cand_i = reinterpret_cast<int*>(smem + L.cand) + 2 * 4 * kTileRows;
  int* idx_s = cand_i + 2 * 4 * kTileRows;                      // [2][128] winning code per row
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L.tmem);

  if (threadIdx.x == 0) {
    mbar_init(bar(0), 1);
    mbar_init(bar(1), 32 * kProducerWarps); mbar_init(bar(2), 32 * kProducerWarps);
    mbar_init(bar(3), 1); mbar_init(bar(4), 1);
    mbar_init(bar(5), 1); mbar_init(bar(6), 1);
    mbar_init(bar(7), 32 * kEpilogueWarps); mbar_init(bar(8), 32 * kEpilogueWarps);
    mbar_init(bar(9), kTileRows); mbar_init(bar(10), kTileRows);
    mbar_init(bar(11), 32 * kOutputWarps); mbar_init(bar(12), 32 * kOutputWarps);
    mbar_fence_init();
  }
  if (warp == kMmaWarp) tmem_alloc(smem_u32(tmem_slot), 512);
  for (int k = threadIdx.x; k < s.Kp; k += blockDim.x) ee_s[k] = ee_g[k];
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = *tmem_slot;
  const uint32_t img_bytes = 2u * kTileRows * s.Dp;        // one x image (hi or lo)
  const uint32_t cb_img_bytes = 2u * s.Kp * s.Dp;          // one codebook image (hi or lo)

  if (warp < kProducerWarps) {
    // ===== producers: NCHW fp32 -> bf16 hi/lo operand image, |x|^2 (each warp: 32 rows x half of the latent dims) =====
    const int t = (warp & 3) * 32 + lane;                    // row within the tile
    const int half = warp >> 2;
    const int dbeg = half * (s.Dp / 2), dend = dbeg + s.Dp / 2;
    const uint32_t hw32 = (uint32_t)s.HW;
    const bool fast = (s.D & 63) == 0;                       // each half of the dims is whole 32-dim groups
    int it = 0;
    PROF_DECL
    for (int64_t tile = blockIdx.x; tile < s.tiles; tile += gridDim.x, ++it) {
      const int st = it & 1;
      PROF_LAP(1)
      if (it >= 2) mbar_wait_relaxed(bar(3 + st), ((it >> 1) - 1) & 1);
      PROF_LAP(0)
      const uint32_t r = (uint32_t)tile * kTileRows + t;     // host guarantees R, D*HW < 2^31: 32-bit index math
      const bool live = r < (uint32_t)s.R;
      const uint32_t b = live ? r / hw32 : 0u, hw = live ? r - b * hw32 : 0u;
      const float* xp = X + ((size_t)b * s.D + dbeg) * hw32 + hw;
      uint8_t* hi = smem + L.x[0] + (uint32_t)st * 2u * img_bytes;
      uint8_t* lo = hi + img_bytes;
      if (fast) {
        // whole 32-dim groups, live or not: 32 strided loads in flight, one pointer bump each
        for (int d0 = dbeg; d0 < dend; d0 += 32) {
          float v[32];
#pragma unroll
          for (int e = 0; e < 32; ++e) { v[e] = live ? __ldg(xp) : 0.f; xp += hw32; }
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t ph[4], pl[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) split_pair(v[8 * c + 2 * e], v[8 * c + 2 * e + 1], ph[e], pl[e]);
            const uint32_t off = ((uint32_t)((d0 >> 3) + c) * kTileRows + t) * 16;
            *reinterpret_cast<uint4*>(hi + off) = make_uint4(ph[0], ph[1], ph[2], ph[3]);
            *reinterpret_cast<uint4*>(lo + off) = make_uint4(pl[0], pl[1], pl[2], pl[3]);
          }
        }
      } else {
        for (int d0 = dbeg; d0 < dend; d0 += 8) {
          float v[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) { v[e] = (live && d0 + e < s.D) ? __ldg(xp) : 0.f; xp += hw32; }
          uint32_t ph[4], pl[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) split_pair(v[2 * e], v[2 * e + 1], ph[e], pl[e]);
          const uint32_t off = ((uint32_t)(d0 >> 3) * kTileRows + t) * 16;
          *reinterpret_cast<uint4*>(hi + off) = make_uint4(ph[0], ph[1], ph[2], ph[3]);
          *reinterpret_cast<uint4*>(lo + off) = make_uint4(pl[0], pl[1], pl[2], pl[3]);
        }
      }
      fence_proxy_async();
      mbar_arrive(bar(1 + st));
    }
    PROF_LAP(1)
    PROF_OUT("producer wait_empty/work", threadIdx.x == 0)
  } else if (warp == kMmaWarp) {
    // ===== MMA issuer: the warp stays converged (uniform-register operands), one elected lane issues =====
    {
      const uint32_t leader = elect_one();
      if (leader) {
        mbar_expect_tx(bar(0), 2 * cb_img_bytes);
        bulk_g2s(sbase + L.cb, cb_image, 2 * cb_img_bytes, bar(0));
      }
      __syncwarp();
      mbar_wait(bar(0), 0);
      const uint32_t idesc = idesc_bf16(128, s.NC, 0, 0);
      const uint32_t a_lbo = kTileRows * 16, b_lbo = (uint32_t)s.Kp * 16;
      const int ksteps = s.Dp / 16;
      int it = 0;
      uint32_t c = 0;                                       // global chunk counter -> TMEM buffer c & 1
      PROF_DECL
      for (int64_t tile = blockIdx.x; tile < s.tiles; tile += gridDim.x, ++it) {
        const int st = it & 1;
        PROF_LAP(2)
        mbar_wait(bar(1 + st), (it >> 1) & 1);
        PROF_LAP(0)
        fence_after_sync();
        const uint32_t a_hi = sbase + L.x[0] + (uint32_t)st * 2u * img_bytes, a_lo = a_hi + img_bytes;
        for (int ch = 0; ch < s.nch; ++ch, ++c) {
          const uint32_t buf = c & 1;
          PROF_LAP(2)
          if (c >= 2) mbar_wait(bar(7 + buf), ((c >> 1) - 1) & 1);
          PROF_LAP(1)
          fence_after_sync();
          const uint32_t b_hi = sbase + L.cb + (uint32_t)ch * s.NC * 16, b_lo = b_hi + cb_img_bytes;
          const uint32_t d = tmem + buf * 256;
          // descriptors advance by a constant per K step (start-address field, 16-byte units): no per-MMA rebuild
          const uint64_t da_hi = smem_desc(a_hi, a_lbo, 128), da_lo = smem_desc(a_lo, a_lbo, 128);
          const uint64_t db_hi = smem_desc(b_hi, b_lbo, 128), db_lo = smem_desc(b_lo, b_lbo, 128);
          const uint64_t a_step = (2u * a_lbo) >> 4, b_step = (2u * b_lbo) >> 4;
          uint32_t acc = 0;
#pragma unroll
          for (int pass = 0; pass < 3; ++pass) {
            // lo*hi, hi*lo, then hi*hi last: the truncating fp32 accumulation sees the small terms first
            uint64_t da = pass == 0 ? da_lo : da_hi, db = pass == 1 ? db_lo : db_hi;
#pragma unroll 4
            for (int k = 0; k < ksteps; ++k) {
              mma_bf16_if(leader, d, da, db, idesc, acc);
              acc = 1;
              da += a_step;
              db += b_step;
            }
          }
          mma_commit_if(leader, bar(5 + buf));
        }
        mma_commit_if(leader, bar(3 + st));                            // x stage free once every MMA of this tile has run
      }
      PROF_LAP(2)
      PROF_OUT("mma wait_x/wait_tmem/issue", leader)
    }
  } else if (warp >= kProducerWarps + kOutputWarps) {
    // ===== argmin warps (highest warp ids: the scheduler favours them): first-index argmin over half of the codes, merge =====
    const int ew = warp & 3;                                 // TMEM lanes 32*ew .. 32*ew+31 (hardware: warp id % 4)
    const int part = (warp - kProducerWarps - kOutputWarps) >> 2;   // 0..1: which 32-column blocks of a chunk
    const int t = ew * 32 + lane;
    const int nblk = s.NC / 32;
    int it = 0;
    uint32_t c = 0;
    PROF_DECL
    for (int64_t tile = blockIdx.x; tile < s.tiles; tile += gridDim.x, ++it) {
      // Running minimum at the granularity of 8-code groups: a three-input-min tree finds a group's minimum in 4
      // instructions, one compare / select updates (best, group base, the group's 8 distances), and the exact first
      // index inside the winning group is resolved once per row at the end.  ~2.5 instructions per distance instead of
      // 5.2 for the per-distance (compare, select index, min) chains this replaces (r02k role profile: the argmin warps
      // were the critical role, 100k of 112k cycles).
      float best = INFINITY, sv[8];
      int bk = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) sv[i] = INFINITY;
      for (int ch = 0; ch < s.nch; ++ch, ++c) {
        const uint32_t buf = c & 1;
        PROF_LAP(1)
        mbar_wait(bar(5 + buf), (c >> 1) & 1);
        PROF_LAP(0)
        fence_after_sync();
        // this part's 32-column blocks, one tcgen05.ld.x32 each: tcgen05.wait::ld waits for EVERY outstanding load, so two
        // ping-ponged x16 loads never were two loads in flight -- one x32 per wait halves the exposed waits
        const uint32_t tbase = tmem + ((uint32_t)(ew * 32) << 16) + buf * 256;
        const int nb32 = (nblk - part + kParts - 1) / kParts;
        auto consume = [&](const uint32_t* r, int kb) {       // 16 distances
          float ee[16];
#pragma unroll
          for (int i = 0; i < 16; i += 4)
            *reinterpret_cast<float4*>(ee + i) = *reinterpret_cast<const float4*>(ee_s + kb + i);
#pragma unroll
          for (int g = 0; g < 16; g += 8) {
            // |e|^2 - 2 x.e: the row-constant |x|^2 of vq_vae.py:35-37 does not move the argmin (and its
            // fp32 rounding is far below the split-bf16 error of the dot product), so it is left out
            float d[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) d[i] = fmaf(-2.f, __uint_as_float(r[g + i]), ee[g + i]);
            float m = min3f(d[0], d[1], d[2]);
            m = min3f(m, d[3], d[4]);
            m = min3f(m, d[5], d[6]);
            m = fminf(m, d[7]);
            if (m < best) {                                   // strict: the earlier group keeps a tie (first index wins)
              best = m;
              bk = kb + g;
#pragma unroll
              for (int i = 0; i < 8; ++i) sv[i] = d[i];
            }
          }
        };
        uint32_t ra[32];
        for (int b = 0; b < nb32; ++b) {
          const int col = (part + b * kParts) << 5;
          tmem_ld32(tbase + col, ra);
          tmem_ld_wait32(ra);
          consume(ra, ch * s.NC + col);
          consume(ra + 16, ch * s.NC + col + 16);
        }
        fence_before_sync();
        mbar_arrive(bar(7 + buf));
      }
      int bi = bk + 7;
#pragma unroll
      for (int i = 6; i >= 0; --i) if (sv[i] == best) bi = bk + i;
      // merge the column parts of this row (first index wins ties, like torch.argmin) and hand the index over
      const int cbuf = (it & 1) * kParts * kTileRows;
      cand_d[cbuf + part * kTileRows + t] = best;
      cand_i[cbuf + part * kTileRows + t] = bi;
      PROF_LAP(1)
      asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpilogueWarps) : "memory");
      PROF_LAP(2)
      if (part == 0) {
#pragma unroll
        for (int q = 1; q < kParts; ++q) {
          const float dq = cand_d[cbuf + q * kTileRows + t];
          const int iq = cand_i[cbuf + q * kTileRows + t];
          if (dq < best || (dq == best && iq < bi)) { best = dq; bi = iq; }
        }
        const int sl = it & 1;
        PROF_LAP(1)
        if (it >= 2) mbar_wait(bar(11 + sl), ((it >> 1) - 1) & 1);
        PROF_LAP(3)
        idx_s[sl * kTileRows + t] = bi;
        mbar_arrive(bar(9 + sl));
      }
    }
    PROF_LAP(1)
    PROF_OUT("argmin wait_tmem/work/merge_bar/wait_idx_empty", threadIdx.x == 32 * (kProducerWarps + kOutputWarps))
  } else {
    // ===== output warps: gather the fp32 code, straight-through value back to NCHW, indices, squared error =====
    const int ot = threadIdx.x - 32 * kProducerWarps;        // 0..255
    const int t = ot & 127, dh = ot >> 7;
    double sq = 0.0;
    int it = 0;
    PROF_DECL
    for (int64_t tile = blockIdx.x; tile < s.tiles; tile += gridDim.x, ++it) {
      const int sl = it & 1;
      PROF_LAP(1)
      mbar_wait_relaxed(bar(9 + sl), (it >> 1) & 1);
      PROF_LAP(0)
      const int bi = idx_s[sl * kTileRows + t];
      mbar_arrive(bar(11 + sl));
      const uint32_t r = (uint32_t)tile * kTileRows + t;
      if (r < (uint32_t)s.R) {
        if (dh == 0) idx_out[r] = bi;
        const uint32_t hw32 = (uint32_t)s.HW;
        const uint32_t b = r / hw32, hw = r - b * hw32;
        float acc = 0.f;
        if ((s.D & 31) == 0) {
          const int dq = (int)s.D / 2;
          const size_t o = ((size_t)b * s.D + (size_t)dh * dq) * hw32 + hw;
          const float* xp = X + o;
          float* qp = q_out + o;
          const float4* ep = reinterpret_cast<const float4*>(E + (size_t)bi * s.D + dh * dq);
          for (int d = 0; d < dq; d += 16) {
            const float4 e0 = __ldg(ep), e1 = __ldg(ep + 1), e2 = __ldg(ep + 2), e3 = __ldg(ep + 3);
            ep += 4;
            const float ev[16] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w,
                                  e2.x, e2.y, e2.z, e2.w, e3.x, e3.y, e3.z, e3.w};
            float xv[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) { xv[j] = __ldg(xp); xp += hw32; }
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float df = ev[j] - xv[j];
              *qp = xv[j] + df;                              // latents + (quantized - latents).detach()
              qp += hw32;
              acc = fmaf(df, df, acc);
            }
          }
        } else {
          const size_t o = ((size_t)b * s.D + dh) * hw32 + hw;
          const float* xp = X + o;
          float* qp = q_out + o;
          const float* ep = E + (size_t)bi * s.D;
          for (int d = dh; d < s.D; d += 2) {
            const float x = __ldg(xp);
            const float df = __ldg(ep + d) - x;
            *qp = x + df;
            xp += 2 * hw32;
            qp += 2 * hw32;
            acc = fmaf(df, df, acc);
          }
        }
        sq += (double)acc;
      }
    }
    PROF_LAP(1)
    PROF_OUT("output wait_idx/work", ot == 0)
    // block partial of sum (q-x)^2 over the output warps, then the last-block combine
    __shared__ double red[kOutputWarps];
    sq = warp_sum(sq);
    if (lane == 0) red[ot >> 5] = sq;
    asm volatile("bar.sync 2, %0;" ::"n"(32 * kOutputWarps) : "memory");
    if (ot == 0) {
      double tot = 0;
      for (int i = 0; i < kOutputWarps; ++i) tot += red[i];
      partials[blockIdx.x] = tot;
      __threadfence();
      const unsigned int tk = atomicAdd(ticket, 1u);
      if (tk == gridDim.x - 1) {
        __threadfence();
        double all = 0;
        for (unsigned int i = 0; i < gridDim.x; ++i) all += ((volatile double*)partials)[i];
        const float m = (float)(all / ((double)s.R * (double)s.D));   // mse_loss(q, x): commitment == embedding in value
        out[0] = m * beta + m;                                         // vq_vae.py:50
        out[1] = m;
        *ticket = 0u;
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem, 512);
}

// ---- backward ------------------------------------------------------------------------------------------------
// g_x = g_q + g_loss * (2 beta / numel) (x - q);   g_E[idx] += g_loss * (2 / numel) (q - x)   (vq_vae.py:47-51)
//
// One CTA (1024 threads) per SM walks 128-row tiles.  Loads / stores run with lanes over rows (coalesced along h*w in
// NCHW; thread = row x slice of the latent dims) and the NEXT tile's x, g_q and indices are fetched into registers while
// the current tile's codebook-gradient phase runs.  The codebook gradient is accumulated in a per-CTA [K][D] shared
// tile WITHOUT atomics (a shared fp32 atomicAdd is an ATOMS.CAST.SPIN loop at >= 2 cycles per lane, which bounded the
// previous version at > 100 us): the tile's x is transposed through shared memory, rows with equal codes are grouped
// with match.any inside each 32-row group and a per-code table of the four groups' member masks; the first row of the
// tile with a given code (its warp has lanes over latent dims, bank-conflict free) adds its members' (e - x) and does
// one plain read-modify-write per (code, dim), so a code has exactly one writer per tile and the phase needs no global
// loads and a single barrier.  The accumulator is flushed with 128-bit global reductions.
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

constexpr int kBwdThreads = 1024;
struct VqBwdSmem { uint32_t acc, xs, idx, tab, total; int ldx; };
inline VqBwdSmem vq_bwd_smem(int64_t D, int64_t K, bool want_e) {
  VqBwdSmem m;
  m.ldx = (int)D | 1;
  uint32_t o = 0;
  m.acc = o;  o += want_e ? (uint32_t)(K * D * 4) : 0u;
  m.xs = o;   o += want_e ? (uint32_t)(kTileRows * m.ldx * 4) : 0u;
  m.idx = o;  o += kTileRows * 4;
  m.tab = o;  o += want_e ? (uint32_t)(K * 16) : 0u;          // per code: member masks of the tile's four 32-row groups
  m.total = o;
  return m;
}

// DPER > 0: D == 8 * DPER, thread (row, slice) owns DPER consecutive dims, register prefetch of the next tile.
// DPER == 0: any D, dims strided by 8, no prefetch.
template <int DPER>
__global__ void __launch_bounds__(kBwdThreads, 1)
vq_bwd_kernel(const float* __restrict__ g_q, const float* __restrict__ g_loss, const float* __restrict__ X,
              const float* __restrict__ E, const int32_t* __restrict__ idx, VqShape s, float beta, VqBwdSmem L,
              float* __restrict__ g_x, float* __restrict__ g_E) {
  extern __shared__ __align__(16) uint8_t bsm[];
  float* acc_s = reinterpret_cast<float*>(bsm + L.acc);      // [K][D] sum over this CTA's rows of (e - x)
  float* x_s = reinterpret_cast<float*>(bsm + L.xs);         // [128][ldx] the tile's (e - x), row-major
  int* idx_s = reinterpret_cast<int*>(bsm + L.idx);
  uint32_t* tab_s = reinterpret_cast<uint32_t*>(bsm + L.tab);
  const bool want_e = g_E != nullptr;
  const int D = (int)s.D, ldx = L.ldx;
  if (want_e) {
    for (int i = threadIdx.x; i < s.K * D; i += kBwdThreads) acc_s[i] = 0.f;
    for (int i = threadIdx.x; i < s.K * 4; i += kBwdThreads) tab_s[i] = 0u;
  }
  const float gl = g_loss ? *g_loss : 0.f;
  const float cx = gl * 2.f * beta / (float)((double)s.R * (double)s.D);
  const float ce = gl * 2.f / (float)((double)s.R * (double)s.D);
  const int rl = threadIdx.x & 127, dc = threadIdx.x >> 7;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t hw32 = (uint32_t)s.HW;
  constexpr int NP = DPER > 0 ? DPER : 1;
  float xr[NP], gr[NP];
  int kr = -1;
  size_t base = 0;                                            // element offset of (row, first dim of the slice)

  auto fetch = [&](int64_t tile) {
    const uint32_t r = (uint32_t)tile * kTileRows + rl;
    kr = -1;
    if (tile < s.tiles && r < (uint32_t)s.R) {
      const uint32_t b = r / hw32, hw = r - b * hw32;
      kr = __ldg(idx + r);
      if (DPER > 0) {
        base = ((size_t)b * D + (size_t)(dc * DPER)) * hw32 + hw;
#pragma unroll
        for (int j = 0; j < NP; ++j) xr[j] = __ldg(X + base + (uint32_t)j * hw32);
        if (g_q) {
#pragma unroll
          for (int j = 0; j < NP; ++j) gr[j] = __ldg(g_q + base + (uint32_t)j * hw32);
        } else {
#pragma unroll
          for (int j = 0; j < NP; ++j) gr[j] = 0.f;
        }
      } else {
        base = (size_t)b * D * hw32 + hw;
      }
    }
  };

  fetch(blockIdx.x);
  __syncthreads();
  for (int64_t tile = blockIdx.x; tile < s.tiles; tile += gridDim.x) {
    // ---- consume the registers: straight-through + commitment gradient, x into the transposed tile
    const int k = kr;
    if (dc == 0) idx_s[rl] = k;
    if (k >= 0) {
      const float* ep = E + (size_t)k * D;
      if (DPER > 0) {
        float ev[NP];
        if (DPER % 4 == 0) {
#pragma unroll
          for (int j = 0; j < NP; j += 4) {
            const float4 e4 = __ldg(reinterpret_cast<const float4*>(ep + dc * DPER + j));
            ev[j] = e4.x; ev[j + 1 < NP ? j + 1 : 0] = e4.y; ev[j + 2 < NP ? j + 2 : 0] = e4.z; ev[j + 3 < NP ? j + 3 : 0] = e4.w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < NP; ++j) ev[j] = __ldg(ep + dc * DPER + j);
        }
#pragma unroll
        for (int j = 0; j < NP; ++j) {
          const float df = ev[j] - xr[j];                     // q - x
          if (g_x) g_x[base + (uint32_t)j * hw32] = gr[j] - cx * df;
          if (want_e) x_s[rl * ldx + dc * DPER + j] = df;
        }
      } else {
        for (int d = dc; d < D; d += 8) {
          const float df = __ldg(ep + d) - __ldg(X + base + (uint32_t)d * hw32);
          if (g_x) g_x[base + (uint32_t)d * hw32] = (g_q ? __ldg(g_q + base + (uint32_t)d * hw32) : 0.f) - cx * df;
          if (want_e) x_s[rl * ldx + d] = df;
        }
      }
    }
    fetch(tile + gridDim.x);                                  // in flight during the codebook-gradient phase
    if (!want_e) continue;
    __syncthreads();
    // rows with equal codes: match.any inside each 32-row group, the group's first such row publishes the member mask
    uint32_t my_mask = 0u;
    int my_code = -1;
    if (warp < kTileRows / 32) {
      const int kk = idx_s[warp * 32 + lane];
      const uint32_t m = __match_any_sync(0xffffffffu, kk);
      if (kk >= 0 && (__ffs(m) - 1) == lane) { my_mask = m; my_code = kk; tab_s[kk * 4 + warp] = m; }
    }
    __syncthreads();
    // warp w looks at rows w, w + 32, w + 64, w + 96; the FIRST row of the tile with a given code sums all its members
    // (lanes over latent dims) and is the only writer of that code's accumulator row in this tile
#pragma unroll 1
    for (int g = 0; g < kTileRows / 32; ++g) {
      const int kk = idx_s[g * 32 + warp];
      if (kk < 0) continue;
      const uint4 t4 = *reinterpret_cast<const uint4*>(tab_s + kk * 4);
      const uint32_t m[4] = {t4.x, t4.y, t4.z, t4.w};
      bool lead = (__ffs(m[g]) - 1) == warp;
      for (int h = 0; h < g; ++h) lead = lead && m[h] == 0u;
      if (!lead) continue;
      for (int d = lane; d < D; d += 32) {
        float sum = 0.f;
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          uint32_t mm = m[h];
          while (mm) {
            const int j = __ffs(mm) - 1;
            mm &= mm - 1;
            sum += x_s[(h * 32 + j) * ldx + d];
          }
        }
        acc_s[kk * D + d] += sum;
      }
    }
    __syncthreads();
    if (my_code >= 0) tab_s[my_code * 4 + warp] = 0u;         // (next written after the next tile's first barrier)
    (void)my_mask;
  }
  if (want_e) {
    __syncthreads();
    if ((D & 3) == 0 && ((reinterpret_cast<uintptr_t>(g_E) & 15) == 0)) {
      for (int i = threadIdx.x; i < s.K * D / 4; i += kBwdThreads) {
        const float4 v = *reinterpret_cast<const float4*>(acc_s + 4 * i);
        if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f) red_add_v4(g_E + 4 * i, ce * v.x, ce * v.y, ce * v.z, ce * v.w);
      }
    } else {
      for (int i = threadIdx.x; i < s.K * D; i += kBwdThreads) {
        const float v = acc_s[i];
        if (v != 0.f) atomicAdd(&g_E[i], ce * v);
      }
    }
  }
}

__global__ void zero_kernel(float* p, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = 0.f;
}


// ---- general shapes: SIMT kernels (any D, any K) -------------------------------------------------------------------
// The tensor-core forward keeps the whole bf16 hi/lo image of the codebook resident in shared memory and the backward a
// [K][D] fp32 accumulator tile; codebooks that do not fit (e.g. K = 1024, D = 128: 512 KB) take these kernels instead:
// same results (fp32 dot products in the reference's association (|x|^2 + |e|^2) - 2 x.e, vq_vae.py:35-37; first index
// wins ties), several times slower, no exception.
constexpr int kSimtRows = 128, kSimtCodes = 32, kSimtDims = 128;   // thread = row; 32 codes x <= 128 dims staged per step

__global__ void vq_code_norms_kernel(const float* __restrict__ E, int64_t K, int64_t D, float* __restrict__ ee) {
  const int64_t k = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (k >= K) return;
  float s = 0.f;
  for (int64_t d = lane; d < D; d += 32) { const float v = E[k * D + d]; s = fmaf(v, v, s); }
  s = warp_sum(s);
  if (lane == 0) ee[k] = s;
}

__global__ void __launch_bounds__(kSimtRows)
vq_fwd_simt_kernel(const float* __restrict__ X, const float* __restrict__ E, const float* __restrict__ ee, VqShape s,
                   float beta, float* __restrict__ q_out, int32_t* __restrict__ idx_out, float* __restrict__ out,
                   unsigned int* ticket, double* partials) {
  __shared__ __align__(16) float es[kSimtCodes][kSimtDims];
  __shared__ double red[kSimtRows / 32];
  const int t = threadIdx.x;
  const uint32_t hw32 = (uint32_t)s.HW;
  const int D = (int)s.D;
  double sq = 0.0;
  for (int64_t tile = blockIdx.x; tile < s.tiles; tile += gridDim.x) {
    const uint32_t r = (uint32_t)tile * kSimtRows + t;
    const bool live = r < (uint32_t)s.R;
    const uint32_t b = live ? r / hw32 : 0u, hw = live ? r - b * hw32 : 0u;
    const float* xp = X + (size_t)b * D * hw32 + hw;
    float xx = 0.f;
    if (live) for (int d = 0; d < D; ++d) { const float v = __ldg(xp + (size_t)d * hw32); xx = fmaf(v, v, xx); }
    float best = INFINITY;
    int bi = 0;
    for (int64_t k0 = 0; k0 < s.K; k0 += kSimtCodes) {
      float dot[kSimtCodes];
#pragma unroll
      for (int c = 0; c < kSimtCodes; ++c) dot[c] = 0.f;
      for (int d0 = 0; d0 < D; d0 += kSimtDims) {
        const int dn = D - d0 < kSimtDims ? D - d0 : kSimtDims;
        __syncthreads();
        for (int i = t; i < kSimtCodes * kSimtDims; i += kSimtRows) {
          const int c = i / kSimtDims, d = i % kSimtDims;
          es[c][d] = (k0 + c < s.K && d < dn) ? __ldg(E + (size_t)(k0 + c) * D + d0 + d) : 0.f;
        }
        __syncthreads();
        for (int dd = 0; dd < dn; dd += 4) {
          float xv[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) xv[j] = (live && dd + j < dn) ? __ldg(xp + (size_t)(d0 + dd + j) * hw32) : 0.f;
#pragma unroll
          for (int c = 0; c < kSimtCodes; ++c) {
            const float4 e4 = *reinterpret_cast<const float4*>(&es[c][dd]);
            dot[c] = fmaf(xv[0], e4.x, dot[c]); dot[c] = fmaf(xv[1], e4.y, dot[c]);
            dot[c] = fmaf(xv[2], e4.z, dot[c]); dot[c] = fmaf(xv[3], e4.w, dot[c]);
          }
        }
      }
#pragma unroll
      for (int c = 0; c < kSimtCodes; ++c) {
        if (k0 + c < s.K) {
          const float dist = (xx + __ldg(ee + k0 + c)) - 2.f * dot[c];      // vq_vae.py:35-37
          if (dist < best) { best = dist; bi = (int)(k0 + c); }             // strict: the first minimum wins
        }
      }
    }
    if (live) {
      idx_out[r] = bi;
      float* qp = q_out + (size_t)b * D * hw32 + hw;
      const float* ep = E + (size_t)bi * D;
      float acc = 0.f;
      for (int d = 0; d < D; ++d) {
        const float x = __ldg(xp + (size_t)d * hw32);
        const float df = __ldg(ep + d) - x;
        qp[(size_t)d * hw32] = x + df;                       // latents + (quantized - latents).detach()
        acc = fmaf(df, df, acc);
      }
      sq += (double)acc;
    }
  }
  sq = warp_sum(sq);
  if ((t & 31) == 0) red[t >> 5] = sq;
  __syncthreads();
  if (t == 0) {
    double tot = 0;
    for (int i = 0; i < kSimtRows / 32; ++i) tot += red[i];
    partials[blockIdx.x] = tot;
    __threadfence();
    const unsigned int tk = atomicAdd(ticket, 1u);
    if (tk == gridDim.x - 1) {
      __threadfence();
      double all = 0;
      for (unsigned int i = 0; i < gridDim.x; ++i) all += ((volatile double*)partials)[i];
      const float m = (float)(all / ((double)s.R * (double)s.D));
      out[0] = m * beta + m;
      out[1] = m;
      *ticket = 0u;
    }
  }
}

// g_x = g_q - cx (q - x);  g_E[idx] += ce (q - x) with global atomics (summation order varies run to run at rounding level)
__global__ void __launch_bounds__(kSimtRows)
vq_bwd_simt_kernel(const float* __restrict__ g_q, const float* __restrict__ g_loss, const float* __restrict__ X,
                   const float* __restrict__ E, const int32_t* __restrict__ idx, VqShape s, float beta,
                   float* __restrict__ g_x, float* __restrict__ g_E) {
  const float gl = g_loss ? *g_loss : 0.f;
  const float cx = gl * 2.f * beta / (float)((double)s.R * (double)s.D);
  const float ce = gl * 2.f / (float)((double)s.R * (double)s.D);
  const uint32_t hw32 = (uint32_t)s.HW;
  const int D = (int)s.D;
  for (int64_t tile = blockIdx.x; tile < s.tiles; tile += gridDim.x) {
    const uint32_t r = (uint32_t)tile * kSimtRows + threadIdx.x;
    if (r >= (uint32_t)s.R) continue;
    const uint32_t b = r / hw32, hw = r - b * hw32;
    const size_t base = (size_t)b * D * hw32 + hw;
    const int k = __ldg(idx + r);
    const float* ep = E + (size_t)k * D;
    for (int d = 0; d < D; ++d) {
      const size_t o = base + (size_t)d * hw32;
      const float df = __ldg(ep + d) - __ldg(X + o);
      if (g_x) g_x[o] = (g_q ? __ldg(g_q + o) : 0.f) - cx * df;
      if (g_E && gl != 0.f) atomicAdd(g_E + (size_t)k * D + d, ce * df);
    }
  }
}

int make_shape(int64_t B, int64_t D, int64_t HW, int64_t K, VqShape& s) {
  VK_REQUIRE(B > 0 && D > 0 && HW > 0 && K > 0, VK_ERR_BAD_ARG, "vq: empty shape B=%lld D=%lld HW=%lld K=%lld",
             (long long)B, (long long)D, (long long)HW, (long long)K);
  VK_REQUIRE(D <= (1 << 20) && K <= (1 << 24), VK_ERR_UNSUPPORTED, "vq: D=%lld or K=%lld beyond the index math of the kernels",
             (long long)D, (long long)K);
  VK_REQUIRE(B * HW < (1ll << 31) - 128 && HW * D < (1ll << 31), VK_ERR_UNSUPPORTED,
             "vq: B*HW=%lld rows or D*HW=%lld beyond the 32-bit index math of the kernels", (long long)(B * HW),
             (long long)(D * HW));
  s.B = B; s.D = D; s.HW = HW; s.K = K; s.R = B * HW;
  s.Dp = round_up((int)D, 16);
  s.NC = K >= 256 ? 256 : round_up((int)K, 32);
  s.Kp = round_up((int)K, s.NC);
  s.nch = s.Kp / s.NC;
  s.tiles = (s.R + kTileRows - 1) / kTileRows;
  return VK_OK;
}

}  // namespace
}  // namespace vk

using namespace vk;

// the tensor-core forward fits: codebook image + two x stages in shared memory, MMA N <= 256 per chunk
static bool vq_tile_fwd_fits(const VqShape& s) {
  return s.D <= 256 && s.K <= 4096 && vq_smem(s.Dp, s.Kp).total + 256 <= 227 * 1024;
}

extern "C" int64_t vk_vq_workspace_bytes(int64_t D, int64_t K) {
  VqShape s;
  if (make_shape(1, D, 1, K, s) != VK_OK) return -1;
  return (int64_t)vq_ws(s).total;
}

extern "C" int vk_vq_fwd(const float* latents, const float* codebook, int64_t B, int64_t D, int64_t HW, int64_t K,
                         float beta, float* q_out, int32_t* idx_out, float* out, void* workspace,
                         int64_t workspace_bytes, void* stream) {
  VK_REQUIRE(latents && codebook && q_out && idx_out && out && workspace, VK_ERR_BAD_ARG, "vq_fwd: null pointer");
  VqShape s;
  int rc = make_shape(B, D, HW, K, s);
  if (rc != VK_OK) return rc;
  const VqWs w = vq_ws(s);
  VK_REQUIRE(workspace_bytes >= (int64_t)w.total, VK_ERR_WORKSPACE, "vq_fwd: workspace %lld < %lld",
             (long long)workspace_bytes, (long long)w.total);
  cudaStream_t st = (cudaStream_t)stream;
  uint8_t* ws = (uint8_t*)workspace;
  if (!vq_tile_fwd_fits(s)) {
    // codebook image beyond shared memory (or D / K beyond the tile kernel's map): SIMT kernel, same results
    float* ee = (float*)(ws + w.ee);
    vq_code_norms_kernel<<<(unsigned)((K * 32 + 255) / 256), 256, 0, st>>>(codebook, K, D, ee);
    const int64_t cap = (int64_t)sm_count() * 8;
    const int grid = (int)(s.tiles < cap ? s.tiles : cap);
    vq_fwd_simt_kernel<<<grid, kSimtRows, 0, st>>>(latents, codebook, ee, s, beta, q_out, idx_out, out, (unsigned int*)ws,
                                                    (double*)(ws + w.partials));
    return check_launch("vq_fwd (SIMT)");
  }
  const VqSmem L = vq_smem(s.Dp, s.Kp);
  vq_pack_codebook_kernel<<<(s.Kp + 127) / 128, 128, 0, st>>>(codebook, (int)K, (int)D, s.Kp, s.Dp,
                                                              (__nv_bfloat16*)(ws + w.image), (float*)(ws + w.ee));
  if (cudaFuncSetAttribute(vq_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total) != cudaSuccess)
    return check_launch("vq_fwd (shared memory attribute)");
  const int grid = (int)(s.tiles < sm_count() ? s.tiles : sm_count());
  vq_fwd_kernel<<<grid, kVqThreads, L.total, st>>>(latents, codebook, ws + w.image, (const float*)(ws + w.ee), s,
                                                          beta, q_out, idx_out, out, (unsigned int*)ws,
                                                          (double*)(ws + w.partials));
  return check_launch("vq_fwd");
}

extern "C" int vk_vq_bwd(const float* g_q, const float* g_loss, const float* latents, const float* codebook,
                         const int32_t* idx, int64_t B, int64_t D, int64_t HW, int64_t K, float beta,
                         float* g_latents, float* g_codebook, void* stream) {
  VK_REQUIRE(latents && codebook && idx, VK_ERR_BAD_ARG, "vq_bwd: null pointer");
  VqShape s;
  int rc = make_shape(B, D, HW, K, s);
  if (rc != VK_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const VqBwdSmem L = vq_bwd_smem(D, K, g_codebook != nullptr);
  if ((uint64_t)K * (uint64_t)D * 4 + (uint64_t)kTileRows * ((uint64_t)D + 1) * 4 + (uint64_t)K * 16 + 1024 > 226 * 1024) {
    // the [K][D] accumulator tile does not fit shared memory: SIMT kernel with global atomics
    if (g_codebook) zero_kernel<<<64, 256, 0, st>>>(g_codebook, K * D);
    const int64_t cap = (int64_t)sm_count() * 8;
    const int64_t blocks = s.tiles < cap ? s.tiles : cap;
    vq_bwd_simt_kernel<<<(int)blocks, kSimtRows, 0, st>>>(g_q, g_loss, latents, codebook, idx, s, beta, g_latents, g_codebook);
    return check_launch("vq_bwd (SIMT)");
  }
  auto kern = D == 32 ? vq_bwd_kernel<4> : D == 64 ? vq_bwd_kernel<8> : D == 128 ? vq_bwd_kernel<16> : vq_bwd_kernel<0>;
  if (L.total > 48 * 1024 && cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total) != cudaSuccess)
    return check_launch("vq_bwd (shared memory attribute)");
  if (g_codebook) zero_kernel<<<64, 256, 0, st>>>(g_codebook, K * D);
  int64_t blocks = s.tiles;
  if (blocks > sm_count()) blocks = sm_count();
  kern<<<(int)blocks, kBwdThreads, L.total, st>>>(g_q, g_loss, latents, codebook, idx, s, beta, L, g_latents, g_codebook);
  return check_launch("vq_bwd");
}
