// Shared host/device helpers for libvaekit (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/vaekit.h"

namespace vk {

// ---- error plumbing (thread-local message, C ABI never throws) -------------
char* err_buf();
int fail(int code, const char* fmt, ...);
int check_launch(const char* what);

#define VK_REQUIRE(cond, code, ...) \
  do { if (!(cond)) return ::vk::fail((code), __VA_ARGS__); } while (0)

int sm_count();

// ---- workspace layout for the two-stage deterministic reductions ------------
// [0,256)   header: uint32 ticket counters (zero on entry, zero on exit)
// [256, ..) partial sums
constexpr int64_t kWsHeader = VK_WS_HEADER_BYTES;
constexpr int kMaxReduceBlocks = 148 * 16;
constexpr int64_t kWsBytes = kWsHeader + (int64_t)kMaxReduceBlocks * 8 * sizeof(double) + (1 << 20);

// ---- device helpers ---------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Streaming 128-bit load that does not pollute L1 (data is touched once).
__device__ __forceinline__ float4 ldg_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void stg_stream(float4* p, const float4& v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// Block-wide sum of NV doubles per thread -> valid in thread 0.  `red` must hold
// NV * (blockDim.x/32) doubles of shared memory.
template <int NV>
__device__ __forceinline__ void block_sum(double (&v)[NV], double* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
  for (int i = 0; i < NV; ++i) v[i] = warp_sum(v[i]);
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < NV; ++i) red[i * nw + wid] = v[i];
  }
  __syncthreads();
  if (wid == 0) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      double x = (lane < nw) ? red[i * nw + lane] : 0.0;
      v[i] = warp_sum(x);
    }
  }
}

// "Last block done" ticket: returns true in exactly one block (all its threads),
// after every other block's global writes are visible.  Resets the counter.
__device__ __forceinline__ bool last_block_ticket(unsigned int* counter, unsigned int nblocks) {
  __shared__ bool is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int t = atomicAdd(counter, 1u);
    is_last = (t == nblocks - 1);
    if (is_last) *counter = 0u;  // self-reset for the next call on this stream
  }
  __syncthreads();
  if (is_last) __threadfence();
  return is_last;
}

// packed fp32 pairs (sm_100: FADD2 / FMUL2 / FFMA2 process two values per issue slot; every element sees the same IEEE
// operation as the scalar instruction)
__device__ __forceinline__ uint64_t pk2(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ uint64_t pk2u(uint32_t a, uint32_t b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ void upk2(uint64_t v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}

__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}

}  // namespace vk

// Per-role cycle counters (debug builds only: VK_NVCC_EXTRA=-DVK_ROLE_PROF): CTA 0 prints, for the first thread of
// each role, the cycles spent in each lap class (waits vs work).
#ifdef VK_ROLE_PROF
#define PROF_DECL long long p_[4] = {0, 0, 0, 0}; long long t_ = clock64();
#define PROF_LAP(i) { const long long n_ = clock64(); p_[i] += n_ - t_; t_ = n_; }
#define PROF_OUT(name, cond) if (blockIdx.x == 0 && (cond)) printf("%s: %lld %lld %lld %lld\n", name, p_[0], p_[1], p_[2], p_[3]);
// event trace: clock of event `ev` of pipeline step `j` for steps [kTraceLo, kTraceLo + 8), printed by CTA 0
#define TRACE_DECL(nev) long long tr_[nev][8]; for (int a_ = 0; a_ < nev; ++a_) for (int b_ = 0; b_ < 8; ++b_) tr_[a_][b_] = 0;
#define TRACE(ev, j) { const int j_ = (int)(j) - 40; if (j_ >= 0 && j_ < 8) tr_[ev][j_] = clock64(); }
#define TRACE_OUT(name, nev, cond) if (blockIdx.x == 0 && (cond)) for (int a_ = 0; a_ < nev; ++a_) \
  printf("trace %s ev%d: %lld %lld %lld %lld %lld %lld %lld %lld\n", name, a_, tr_[a_][0], tr_[a_][1], tr_[a_][2], tr_[a_][3], \
         tr_[a_][4], tr_[a_][5], tr_[a_][6], tr_[a_][7]);
#else
#define PROF_DECL
#define PROF_LAP(i)
#define PROF_OUT(name, cond)
#define TRACE_DECL(nev)
#define TRACE(ev, j)
#define TRACE_OUT(name, nev, cond)
#endif
