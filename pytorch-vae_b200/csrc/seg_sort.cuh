// Segmented LSD radix sort of fp32 keys (ascending), key-only or key + uint32 payload, written from scratch for
// the SWAE sliced-Wasserstein path (reference: torch.sort(dim=1) at models/swae.py:176-177).
//
// Segments are equal-length rows of a [segments, n] matrix.  8-bit digits, 4 passes; every pass is
//   hist    : each CTA counts the digits of its chunk of one segment            (read keys)
//   scan    : one CTA per segment turns [chunk][digit] counts into global offsets (digit-major, chunk-minor)
//   scatter : each CTA re-reads its chunk, ranks it stably (warp match + per-warp digit counters), sorts it by
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
__global__ void __launch_bounds__(kBins)
scan_kernel(uint32_t* __restrict__ counts, int chunks) {
  __shared__ uint32_t tot[kBins];
  const int seg = blockIdx.x, d = threadIdx.x;
  uint32_t* c = counts + (size_t)seg * chunks * kBins;
  uint32_t run = 0;
  for (int b = 0; b < chunks; ++b) {          // coalesced across digits
    const uint32_t v = c[(size_t)b * kBins + d];
    c[(size_t)b * kBins + d] = run;
    run += v;
  }
  tot[d] = run;
  __syncthreads();
  // exclusive scan of the 256 digit totals (one warp-shuffle scan per 32, then 8 partials)
  __shared__ uint32_t part[8];
  const int lane = d & 31, w = d >> 5;
  uint32_t v = tot[d], incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) part[w] = incl;
  __syncthreads();
  uint32_t pre = 0;
  for (int i = 0; i < w; ++i) pre += part[i];
  const uint32_t start = pre + incl - v;
  for (int b = 0; b < chunks; ++b) c[(size_t)b * kBins + d] += start;
}

// ---- scatter -----------------------------------------------------------------------------------------------------
// FIRST: input keys are raw floats (transform on load) and the payload is the element index; LAST: keys are
// transformed back to floats on store.
template <bool FIRST, bool LAST, bool PAYLOAD>
__global__ void __launch_bounds__(kThreads)
scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, int64_t n, int shift,
               const uint32_t* __restrict__ offsets, uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out) {
  __shared__ uint32_t wcount[kWarps][kBins];     // per-warp digit counters -> exclusive prefix across warps
  __shared__ uint32_t dstart[kBins];             // start of each digit inside the CTA-sorted chunk
  __shared__ uint32_t gofs[kBins];               // global offset of this chunk's run of each digit
  __shared__ uint32_t skey[kChunk];
  __shared__ uint32_t sval[PAYLOAD ? kChunk : 1];
  const int seg = blockIdx.y, chunk = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < kWarps * kBins; i += kThreads) (&wcount[0][0])[i] = 0;
  gofs[threadIdx.x] = offsets[((size_t)seg * gridDim.x + chunk) * kBins + threadIdx.x];
  __syncthreads();
  const size_t sbase = (size_t)seg * n;
  const int64_t base = (int64_t)chunk * kChunk + (int64_t)warp * (32 * kItems);   // warp-blocked, lane-striped: stable order
  uint32_t key[kItems], rank[kItems];
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const int64_t e = base + i * 32 + lane;
    uint32_t u = e < n ? keys_in[sbase + e] : 0xFFFFFFFFu;
    if (FIRST && e < n) u = f2ord(u);
    key[i] = u;
  }
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const int64_t e = base + i * 32 + lane;
    const bool ok = e < n;
    const uint32_t d = (key[i] >> shift) & 0xFFu;
    const uint32_t peers = digit_peers(d, ok);
    const int leader = ok ? __ffs(peers) - 1 : lane;
    uint32_t prev = 0;
    if (ok && lane == leader) { prev = wcount[warp][d]; wcount[warp][d] = prev + __popc(peers); }
    prev = __shfl_sync(0xFFFFFFFFu, prev, leader);
    rank[i] = prev + __popc(peers & lt);
    __syncwarp();
  }
  __syncthreads();
  {   // thread d: exclusive prefix of digit d over warps, digit totals
    const int d = threadIdx.x;
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) { const uint32_t v = wcount[w][d]; wcount[w][d] = run; run += v; }
    // exclusive scan of totals over digits -> dstart
    const int l = d & 31, wq = d >> 5;
    uint32_t incl = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (l >= o) incl += t;
    }
    __shared__ uint32_t part[8];
    if (l == 31) part[wq] = incl;
    __syncthreads();
    uint32_t pre = 0;
    for (int i = 0; i < wq; ++i) pre += part[i];
    dstart[d] = pre + incl - run;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const int64_t e = base + i * 32 + lane;
    if (e < n) {
      const uint32_t d = (key[i] >> shift) & 0xFFu;
      const uint32_t pos = dstart[d] + wcount[warp][d] + rank[i];
      skey[pos] = key[i];
      if (PAYLOAD) sval[pos] = FIRST ? (uint32_t)e : vals_in[sbase + e];
    }
  }
  __syncthreads();
  const int64_t cbase = (int64_t)chunk * kChunk;
  const int live = (int)((n - cbase) < kChunk ? (n - cbase) : kChunk);
#pragma unroll 4
  for (int i = 0; i < kItems; ++i) {
    const int pos = i * kThreads + threadIdx.x;
    if (pos < live) {
      const uint32_t u = skey[pos];
      const uint32_t d = (u >> shift) & 0xFFu;
      const size_t g = sbase + gofs[d] + (pos - dstart[d]);
      keys_out[g] = LAST ? ord2f(u) : u;
      if (PAYLOAD) vals_out[g] = sval[pos];
    }
  }
}

// Sorts `segments` rows of length n.  keys: input floats in buf0; result (floats) lands in buf0 again after 4
// passes (buf0 -> buf1 -> buf0 -> buf1 -> buf0).  vals (PAYLOAD): element indices, result in val0.
template <bool PAYLOAD>
inline int sort(uint32_t* key0, uint32_t* key1, uint32_t* val0, uint32_t* val1, int64_t segments, int64_t n,
                uint32_t* counts, cudaStream_t st) {
  const int chunks = chunks_for(n);
  const dim3 grid(chunks, (unsigned)segments);
  uint32_t *ki = key0, *ko = key1, *vi = val0, *vo = val1;
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 8 * pass;
    if (pass == 0) hist_kernel<true><<<grid, kThreads, 0, st>>>(ki, n, shift, counts);
    else hist_kernel<false><<<grid, kThreads, 0, st>>>(ki, n, shift, counts);
    scan_kernel<<<(unsigned)segments, kBins, 0, st>>>(counts, chunks);
    if (pass == 0) scatter_kernel<true, false, PAYLOAD><<<grid, kThreads, 0, st>>>(ki, vi, n, shift, counts, ko, vo);
    else if (pass == 3) scatter_kernel<false, true, PAYLOAD><<<grid, kThreads, 0, st>>>(ki, vi, n, shift, counts, ko, vo);
    else scatter_kernel<false, false, PAYLOAD><<<grid, kThreads, 0, st>>>(ki, vi, n, shift, counts, ko, vo);
    uint32_t* t = ki; ki = ko; ko = t;
    t = vi; vi = vo; vo = t;
  }
  return check_launch("seg_radix_sort");
}

}  // namespace segsort
}  // namespace vk
