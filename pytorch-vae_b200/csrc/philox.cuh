// Philox4x32-10 counter RNG + Box-Muller (the RNG contract of include/vaekit.h).
// CPU statement of the same stream: oracle/vae_oracle.py philox_uint32 / philox_normal.
#pragma once
#include <stdint.h>

namespace vk {

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
    const uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += W0;
    k.y += W1;
  }
  return c;
}

// Words of quad q (elements 4q .. 4q+3) of stream (seed, offset).
__device__ __forceinline__ uint4 philox_quad(uint64_t seed, uint64_t offset, uint64_t q) {
  const uint64_t ctr = q + offset;
  return philox4x32_10(make_uint4((uint32_t)ctr, (uint32_t)(ctr >> 32), 0u, 0u),
                       make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
}

// 23-bit uniform in (0,1), exactly representable in fp32.
__device__ __forceinline__ float u23(uint32_t w) {
  return ((float)(w >> 9) + 0.5f) * 1.1920928955078125e-07f;  // 2^-23
}

__device__ __forceinline__ void box_muller(uint32_t wa, uint32_t wb, float& n0, float& n1) {
  const float r = sqrtf(-2.0f * logf(u23(wa)));
  float s, c;
  sincospif(2.0f * u23(wb), &s, &c);
  n0 = r * c;
  n1 = r * s;
}

__device__ __forceinline__ float4 philox_normal4(uint64_t seed, uint64_t offset, uint64_t q) {
  const uint4 w = philox_quad(seed, offset, q);
  float4 n;
  box_muller(w.x, w.y, n.x, n.y);
  box_muller(w.z, w.w, n.z, n.w);
  return n;
}

// Single element e of the stream (slow path for rows whose length is not a multiple of 4).
__device__ __forceinline__ float philox_normal1(uint64_t seed, uint64_t offset, uint64_t e) {
  const float4 n = philox_normal4(seed, offset, e >> 2);
  const int c = (int)(e & 3);
  return c == 0 ? n.x : c == 1 ? n.y : c == 2 ? n.z : n.w;
}

}  // namespace vk
