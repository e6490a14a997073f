// Segmented LSD radix sort of fp32 keys (ascending), key-only or key + uint32 payload, written from scratch for
// the SWAE sliced-Wasserstein path (reference: torch.sort(dim=1) at models/swae.py:176-177).
//
// Segments are equal-length rows of a [segments, n] matrix.  8-bit digits, 4 passes; every pass is
//   hist    : each CTA counts the digits of its chunk of one segment            (read keys)
//   scan    : one CTA per segment turns [chunk][digit] counts into global offsets (digit-major, chunk-minor)
//   scatter : each CTA re-reads its chunk, ranks it stably (ballot peer masks + per-warp digit counters), sorts it by
//             digit inside shared memory and writes digit runs to their global place (coalesced runs)
// The float -> ordered-uint transform is folded into the first pass and undone in the last one.
#pragma once
#include "common.cuh"

namespace vk {
namespace segsort {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kItems = 16;                       // keys per thread
constexpr int kChunk = kThreads * kItems;        // keys per CTA: 4096
constexpr int kBins = 256;

__device__ __forceinline__ uint32_t f2ord(uint32_t u) { return u ^ ((u >> 31) ? 0xFFFFFFFFu : 0x80000000u); }
__device__ __forceinline__ uint32_t ord2f(uint32_t u) { return u ^ ((u >> 31) ? 0x80000000u : 0xFFFFFFFFu); }

// Lanes of the warp that hold the same 8-bit digit (among lanes with ok == true): eight ballots, fixed cost.
// (match.any costs one pass per distinct value in the warp -- ~30 for uniform digits -- and made ranking the
// bottleneck of the whole sort.)
__device__ __forceinline__ uint32_t digit_peers(uint32_t d, bool ok) {
  uint32_t peers = __ballot_sync(0xFFFFFFFFu, ok);
#pragma unroll
  for (int b = 0; b < 8; ++b) {
    const bool bit = (d >> b) & 1u;
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, bit);
    peers &= bit ? m : ~m;
  }
  return peers;
}

// peers & (lanes whose key has the same value of the masked bit as mine), in four instructions: bit test straight
// into a predicate, ballot, all-ones/zero select, one LOP3 computing peers & (ballot ^ select).  (Written in C this
// compiles to six.)
__device__ __forceinline__ uint32_t same_bit_peers(uint32_t peers, uint32_t key, uint32_t mask) {
  asm volatile(
      "{\n .reg .pred p, q;\n .reg .b32 t, nb;\n setp.ne.u32 q, 0, 0;\n"
      " lop3.or.b32 t|p, %1, %2, 0, 0xC0, q;\n"
      " vote.sync.ballot.b32 t, p, 0xffffffff;\n"
      " selp.b32 nb, 0, -1, p;\n"
      " lop3.b32 %0, %0, t, nb, 0x60;\n}"
      : "+r"(peers) : "r"(key), "r"(mask));
  return peers;
}

inline int chunks_for(int64_t n) { return (int)((n + kChunk - 1) / kChunk); }
// bytes of scratch for the per-chunk digit counts of `segments` segments of length n
inline size_t counts_bytes(int64_t segments, int64_t n) { return (size_t)segments * chunks_for(n) * kBins * sizeof(uint32_t); }

// ---- hist: counts[seg][chunk][digit] ---------------------------------------------------------------------------
template <bool FIRST>
__global__ void __launch_bounds__(kThreads)
hist_kernel(const uint32_t* __restrict__ keys, int64_t n, int shift, uint32_t* __restrict__ counts) {
  __shared__ uint32_t h[kWarps][kBins];          // warp-private counters
  const int seg = blockIdx.y, chunk = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < kWarps * kBins; i += kThreads) (&h[0][0])[i] = 0;
  __syncthreads();
  const uint32_t* k = keys + (size_t)seg * n;
  const int64_t base = (int64_t)chunk * kChunk + (int64_t)warp * (32 * kItems);
  uint32_t key[kItems];
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const int64_t e = base + i * 32 + lane;
    uint32_t u = e < n ? k[e] : 0xFFFFFFFFu;
    if (FIRST && e < n) u = f2ord(u);
    key[i] = u;
  }
  // counting needs no order: plain shared-memory atomics on the warp-private counters (a few bank-conflict replays
  // per warp instruction) are ~4x cheaper than building peer masks with nine ballots per key
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    if (base + i * 32 + lane < n) atomicAdd(&h[warp][(key[i] >> shift) & 0xFFu], 1u);
  }
  __syncthreads();
  uint32_t t = 0;
#pragma unroll
  for (int w = 0; w < kWarps; ++w) t += h[w][threadIdx.x];
  counts[((size_t)seg * gridDim.x + chunk) * kBins + threadIdx.x] = t;
}

// ---- scan: counts -> exclusive global offsets, in place ----------------------------------------------------------
// One CTA per segment, 4 chunk ranges x 256 digits: every thread totals its range (batches of independent loads),
// the digit totals are scanned across digits, then the range is rewritten with its running prefix.
constexpr int kScanQ = 4;
constexpr int kScanBatch = 8;
__global__ void __launch_bounds__(kScanQ * kBins)
scan_kernel(uint32_t* __restrict__ counts, int chunks) {
  __shared__ uint32_t qtot[kScanQ][kBins];
  __shared__ uint32_t part[kScanQ][8];
  const int seg = blockIdx.x, d = threadIdx.x & (kBins - 1), q = threadIdx.x >> 8;
  uint32_t* c = counts + (size_t)seg * chunks * kBins + d;
  const int per = (chunks + kScanQ - 1) / kScanQ;
  const int b0 = q * per, b1 = b0 + per < chunks ? b0 + per : chunks;
  uint32_t sum = 0;
  for (int b = b0; b < b1; b += kScanBatch) {          // coalesced across digits
    uint32_t v[kScanBatch];
#pragma unroll
    for (int u = 0; u < kScanBatch; ++u) v[u] = b + u < b1 ? c[(size_t)(b + u) * kBins] : 0u;
#pragma unroll
    for (int u = 0; u < kScanBatch; ++u) sum += v[u];
  }
  qtot[q][d] = sum;
  __syncthreads();
  uint32_t total = 0, below = 0;
#pragma unroll
  for (int i = 0; i < kScanQ; ++i) { const uint32_t t = qtot[i][d]; total += t; if (i < q) below += t; }
  // exclusive scan of the 256 digit totals (one warp-shuffle scan per 32, then 8 partials), once per chunk range
  const int lane = d & 31, w = d >> 5;
  uint32_t incl = total;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) part[q][w] = incl;
  __syncthreads();
  uint32_t run = incl - total + below;
  for (int i = 0; i < w; ++i) run += part[q][i];
  for (int b = b0; b < b1; b += kScanBatch) {
    uint32_t v[kScanBatch];
#pragma unroll
    for (int u = 0; u < kScanBatch; ++u) v[u] = b + u < b1 ? c[(size_t)(b + u) * kBins] : 0u;
#pragma unroll
    for (int u = 0; u < kScanBatch; ++u) {
      if (b + u < b1) c[(size_t)(b + u) * kBins] = run;
      run += v[u];
    }
  }
}

// ---- scatter -----------------------------------------------------------------------------------------------------
// PASS 0: input keys are raw floats (transform on load) and the payload is the element index; PASS 3: keys are
// transformed back to floats on store.  FULL: the chunk has kChunk live keys (no bounds predicates at all).
// The kernel is bound by ALU-pipe instructions per key (ncu r01j: 131 instructions per key, ALU pipe 69 % busy,
// DRAM 15 %), so everything around the nine ballots is kept minimal: immediate digit masks (PASS is a template
// parameter), 32-bit chunk-local indices, the peer masks of all 16 keys built before the serial counter chain
// (every peer reads the warp's counter, the highest peer lane bumps it: no leader election, no shuffle), and
// precombined tables for the two passes over the chunk (counter := position of the warp's run, gdelta := global
// run start minus chunk-sorted run start).
//
// FUSE (last pass of a key + index sort only): instead of storing the sorted keys and the permutation, the pass
// consumes them on the spot for the sliced-Wasserstein loss (swae.py:176-179): with `other` = the already sorted
// keys of the second operand, element r of the sorted order contributes |key_r - other_r|^p to the CTA's partial
// sum and its derivative goes straight to the source position of the key (g_out[perm_r]).  That removes one write
// and one read of the sorted keys and of the permutation (3.4 GB at config 5) and a kernel.
struct LossFuse {
  const float* other;      // [segments, n] sorted ascending
  float* g_out;            // [segments, n] in source order, may be NULL
  double* partials;        // [segments * chunks]
  float p;
};

// kScatterMinBlocks CTAs per SM: ncu (profiles/ncu_swae_r01v_summary.txt) shows the pass latency-bound (issue slots 57 %
// busy, long-scoreboard stalls) at 3 resident CTAs -- the 70 registers per thread were the occupancy limit.
#ifndef VK_SCATTER_MINB
#define VK_SCATTER_MINB 4
#endif
#ifndef VK_SCATTER_MINB_KV
#define VK_SCATTER_MINB_KV 4
#endif
template <int PASS, bool PAYLOAD, bool FULL, bool FUSE>
__global__ void __launch_bounds__(kThreads, (PAYLOAD || FUSE) ? VK_SCATTER_MINB_KV : VK_SCATTER_MINB)
scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, int64_t n,
               const uint32_t* __restrict__ offsets, uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
               int chunk0, int chunks, LossFuse fuse) {
  static_assert(!FUSE || (PASS == 3 && PAYLOAD), "the fused loss replaces the stores of the last key+index pass");
  constexpr int shift = 8 * PASS;
  constexpr bool FIRST = PASS == 0, LAST = PASS == 3;
  __shared__ uint32_t wcount[kWarps][kBins];     // per-warp digit counters -> position of the warp's run of the digit
  __shared__ uint32_t gdelta[kBins];             // global offset of the chunk's digit run - its start inside the chunk
  __shared__ uint32_t part[8];
  __shared__ uint32_t skey[kChunk];
  __shared__ uint32_t sval[PAYLOAD ? kChunk : 1];
  const int seg = blockIdx.y, chunk = chunk0 + blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < kWarps * kBins; i += kThreads) (&wcount[0][0])[i] = 0;
  const uint32_t my_gofs = offsets[((size_t)seg * chunks + chunk) * kBins + threadIdx.x];
  __syncthreads();
  const size_t sbase = (size_t)seg * n;
  const uint32_t cbase = (uint32_t)chunk * kChunk;                      // n < 2^32 (checked by the host)
  const uint32_t* kin = keys_in + sbase + cbase;
  const int live = FULL ? kChunk : (int)((n - cbase) < kChunk ? (n - cbase) : kChunk);
  const int wbase = warp * (32 * kItems) + lane;                       // warp-blocked, lane-striped: stable order
  uint32_t key[kItems], rank[kItems];
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const int idx = wbase + i * 32;
    uint32_t u = (FULL || idx < live) ? kin[idx] : 0xFFFFFFFFu;
    if (FIRST && (FULL || idx < live)) u = f2ord(u);
    key[i] = u;
  }
#pragma unroll
  for (int i = 0; i < kItems; ++i) {                                   // peer masks: 16 independent ballot chains
    uint32_t peers = FULL ? 0xFFFFFFFFu : __ballot_sync(0xFFFFFFFFu, wbase + i * 32 < live);
#pragma unroll
    for (int b = 0; b < 8; ++b) peers = same_bit_peers(peers, key[i], 1u << (shift + b));
    rank[i] = peers;
  }
#pragma unroll
  for (int i = 0; i < kItems; ++i) {                                   // serial chain through the warp's counters
    const bool ok = FULL || wbase + i * 32 < live;
    const uint32_t peers = rank[i];
    uint32_t* c = &wcount[warp][(key[i] >> shift) & 0xFFu];
    const uint32_t r = *c + __popc(peers & lt);
    __syncwarp();
    if (ok && (peers >> lane) == 1u) *c = r + 1u;                       // highest peer lane: count so far + all peers
    __syncwarp();
    rank[i] = r;
  }
  __syncthreads();
  {   // thread d: digit totals, exclusive scan over digits, then counter[w][d] := chunk position of warp w's run
    const int d = threadIdx.x;
    uint32_t cw[kWarps], run = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) { cw[w] = wcount[w][d]; run += cw[w]; }
    const int l = d & 31, wq = d >> 5;
    uint32_t incl = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (l >= o) incl += t;
    }
    if (l == 31) part[wq] = incl;
    __syncthreads();
    uint32_t pos = incl - run;
    for (int i = 0; i < wq; ++i) pos += part[i];
    gdelta[d] = my_gofs - pos;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) { wcount[w][d] = pos; pos += cw[w]; }
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const int idx = wbase + i * 32;
    if (FULL || idx < live) {
      const uint32_t pos = wcount[warp][(key[i] >> shift) & 0xFFu] + rank[i];
      skey[pos] = key[i];
      if (PAYLOAD) sval[pos] = FIRST ? cbase + (uint32_t)idx : vals_in[sbase + cbase + idx];
    }
  }
  __syncthreads();
  uint32_t* kout = keys_out + sbase;
  uint32_t* vout = PAYLOAD ? vals_out + sbase : nullptr;
  float acc = 0.f;
  const bool p2 = fuse.p == 2.0f;
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const int pos = i * kThreads + threadIdx.x;
    if (FULL || pos < live) {
      const uint32_t u = skey[pos];
      const uint32_t g = gdelta[(u >> shift) & 0xFFu] + (uint32_t)pos;
      if (FUSE) {
        const float w = __uint_as_float(ord2f(u)) - fuse.other[sbase + g];
        float v, dv;
        if (p2) { v = w * w; dv = 2.f * w; }
        else { v = powf(w, fuse.p); dv = fuse.p * powf(w, fuse.p - 1.f); }
        acc += v;
        if (fuse.g_out) fuse.g_out[sbase + sval[pos]] = dv;
      } else {
        kout[g] = LAST ? ord2f(u) : u;
        if (PAYLOAD) vout[g] = sval[pos];
      }
    }
  }
  if (FUSE) {
    __syncthreads();
    double* red = reinterpret_cast<double*>(skey);      // the sorted chunk is consumed
    double v[1] = {(double)acc};
    block_sum<1>(v, red);
    if (threadIdx.x == 0) fuse.partials[(size_t)seg * chunks + chunk] = v[0];
  }
}

// deterministic sum of the per-chunk partials of the fused loss
__global__ void __launch_bounds__(1024)
sum_partials_kernel(const double* __restrict__ partials, int64_t count, double* __restrict__ out) {
  __shared__ double red[32];
  double t = 0.0;
  for (int64_t i = threadIdx.x; i < count; i += blockDim.x) t += partials[i];
  double v[1] = {t};
  block_sum<1>(v, red);
  if (threadIdx.x == 0) *out = v[0];
}

template <int PASS, bool PAYLOAD, bool FUSE = false>
inline void scatter_pass(dim3 grid, int64_t n, const uint32_t* ki, const uint32_t* vi, const uint32_t* counts,
                         uint32_t* ko, uint32_t* vo, cudaStream_t st, LossFuse fuse = LossFuse{nullptr, nullptr, nullptr, 0.f}) {
  // all chunks but the last are full; when n is a multiple of kChunk the whole grid takes the predicate-free path
  const unsigned full = (unsigned)(n / kChunk);
  if (full > 0)
    scatter_kernel<PASS, PAYLOAD, true, FUSE><<<dim3(full, grid.y), kThreads, 0, st>>>(ki, vi, n, counts, ko, vo, 0,
                                                                                      (int)grid.x, fuse);
  if (full < grid.x)      // the ragged last chunk of every segment: the bounds-checked instance
    scatter_kernel<PASS, PAYLOAD, false, FUSE><<<dim3(1, grid.y), kThreads, 0, st>>>(ki, vi, n, counts, ko, vo,
                                                                                     (int)full, (int)grid.x, fuse);
}

// Sorts `segments` rows of length n.  keys: input floats in buf0; result (floats) lands in buf0 again after 4
// passes (buf0 -> buf1 -> buf0 -> buf1 -> buf0).  vals (PAYLOAD): element indices, result in val0.
// With `fuse` (PAYLOAD only) the last pass stores nothing: buf0 / val0 are left untouched by it and the loss partials
// / gradient are produced instead (fuse->g_out may alias buf0 / val0, not buf1 / val1 which the pass reads).
template <bool PAYLOAD>
inline int sort(uint32_t* key0, uint32_t* key1, uint32_t* val0, uint32_t* val1, int64_t segments, int64_t n,
                uint32_t* counts, cudaStream_t st, const LossFuse* fuse = nullptr) {
  const int chunks = chunks_for(n);
  const dim3 grid(chunks, (unsigned)segments);
  uint32_t *ki = key0, *ko = key1, *vi = val0, *vo = val1;
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 8 * pass;
    if (pass == 0) hist_kernel<true><<<grid, kThreads, 0, st>>>(ki, n, shift, counts);
    else hist_kernel<false><<<grid, kThreads, 0, st>>>(ki, n, shift, counts);
    scan_kernel<<<(unsigned)segments, kScanQ * kBins, 0, st>>>(counts, chunks);
    if (pass == 0) scatter_pass<0, PAYLOAD>(grid, n, ki, vi, counts, ko, vo, st);
    else if (pass == 1) scatter_pass<1, PAYLOAD>(grid, n, ki, vi, counts, ko, vo, st);
    else if (pass == 2) scatter_pass<2, PAYLOAD>(grid, n, ki, vi, counts, ko, vo, st);
    else if (PAYLOAD && fuse) scatter_pass<3, true, true>(grid, n, ki, vi, counts, ko, vo, st, *fuse);
    else scatter_pass<3, PAYLOAD>(grid, n, ki, vi, counts, ko, vo, st);
    uint32_t* t = ki; ki = ko; ko = t;
    t = vi; vi = vo; vo = t;
  }
  return check_launch("seg_radix_sort");
}

}  // namespace segsort
}  // namespace vk
