// K1: fused reparameterisation + Gaussian-KL rows, forward and backward.
// Replaces the 5+4 ATen launches of models/vanilla_vae.py:115-117 (and the
// identical bodies in beta_vae / logcosh_vae / iwae / miwae / betatc_vae) plus
// the KL row sums of vanilla_vae.py:143 / iwae.py:150.  HBM-streaming kernel:
// one 128-bit load of mu and log_var per lane, one 128-bit store of z; noise
// comes from in-register Philox (no eps tensor in HBM, regenerated in backward).
#include "common.cuh"
#include "philox.cuh"

namespace vk {

struct RowMap {
  int64_t n0, n1, n2;        // logical outer grid
  int64_t ms0, ms1, ms2;     // mu row strides (elements)
  int64_t ls0, ls1, ls2;     // log_var row strides (elements)
  int64_t rows, D;
};

__device__ __forceinline__ void row_offsets(const RowMap& m, int64_t r, int64_t& mo, int64_t& lo) {
  const int64_t i2 = r % m.n2;
  const int64_t t = r / m.n2;
  const int64_t i1 = t % m.n1;
  const int64_t i0 = t / m.n1;
  mo = i0 * m.ms0 + i1 * m.ms1 + i2 * m.ms2;
  lo = i0 * m.ls0 + i1 * m.ls1 + i2 * m.ls2;
}

__device__ __forceinline__ float group_sum(float v, int gs) {
  for (int o = gs >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float kl_term(float mu, float lv, float elv) {
  // 1 + lv - mu^2 - e^lv, same association as the reference expression
  return ((1.0f + lv) - mu * mu) - elv;
}

// VEC: D % 4 == 0, strides % 4 == 0, 16-byte aligned bases -> float4 path, one Philox quad per lane step.
template <bool VEC>
__global__ void __launch_bounds__(256)
reparam_kl_fwd_kernel(const float* __restrict__ mu, const float* __restrict__ lv, RowMap m, int gs,
                      uint64_t seed, uint64_t offset, const int64_t* __restrict__ key, float* __restrict__ z,
                      float* __restrict__ kl_rows, float* __restrict__ eps_out) {
  if (key) { seed = (uint64_t)key[0]; offset = (uint64_t)key[1]; }      // (seed, offset) held on the device: graph-replayable
  const int lane = threadIdx.x & 31;
  const int rpw = 32 / gs;                      // rows per warp
  const int sub = lane % gs;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t r = warp * rpw + lane / gs;
  const bool active = r < m.rows;
  const bool need_noise = (z != nullptr) || (eps_out != nullptr);
  float acc = 0.0f;
  if (active) {
    int64_t mo, lo;
    row_offsets(m, r, mo, lo);
    if (VEC) {
      const int64_t nq = m.D >> 2;
      for (int64_t j = sub; j < nq; j += gs) {
        const float4 a = *reinterpret_cast<const float4*>(mu + mo + 4 * j);
        const float4 b = *reinterpret_cast<const float4*>(lv + lo + 4 * j);
        if (kl_rows) {
          acc += kl_term(a.x, b.x, expf(b.x)) + kl_term(a.y, b.y, expf(b.y)) +
                 kl_term(a.z, b.z, expf(b.z)) + kl_term(a.w, b.w, expf(b.w));
        }
        if (need_noise) {
          const int64_t q = r * nq + j;
          const float4 e = philox_normal4(seed, offset, (uint64_t)q);
          if (z) {
            float4 o;
            o.x = e.x * expf(0.5f * b.x) + a.x;
            o.y = e.y * expf(0.5f * b.y) + a.y;
            o.z = e.z * expf(0.5f * b.z) + a.z;
            o.w = e.w * expf(0.5f * b.w) + a.w;
            *reinterpret_cast<float4*>(z + 4 * q) = o;
          }
          if (eps_out) *reinterpret_cast<float4*>(eps_out + 4 * q) = e;
        }
      }
    } else {
      for (int64_t d = sub; d < m.D; d += gs) {
        const float a = mu[mo + d], b = lv[lo + d];
        if (kl_rows) acc += kl_term(a, b, expf(b));
        if (need_noise) {
          const int64_t e_idx = r * m.D + d;
          const float e = philox_normal1(seed, offset, (uint64_t)e_idx);
          if (z) z[e_idx] = e * expf(0.5f * b) + a;
          if (eps_out) eps_out[e_idx] = e;
        }
      }
    }
  }
  if (kl_rows) {
    acc = group_sum(acc, gs);
    if (active && sub == 0) kl_rows[r] = -0.5f * acc;
  }
}

template <bool VEC>
__global__ void __launch_bounds__(256)
reparam_kl_bwd_kernel(const float* __restrict__ g_z, const float* __restrict__ g_kl,
                      const float* __restrict__ mu, const float* __restrict__ lv, RowMap m, int gs,
                      uint64_t seed, uint64_t offset, const int64_t* __restrict__ key, float* __restrict__ g_mu,
                      float* __restrict__ g_lv) {
  if (key) { seed = (uint64_t)key[0]; offset = (uint64_t)key[1]; }
  const int lane = threadIdx.x & 31;
  const int rpw = 32 / gs;
  const int sub = lane % gs;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t r = warp * rpw + lane / gs;
  if (r >= m.rows) return;
  int64_t mo, lo;
  row_offsets(m, r, mo, lo);
  const float gk = g_kl ? g_kl[r] : 0.0f;
  if (VEC) {
    const int64_t nq = m.D >> 2;
    for (int64_t j = sub; j < nq; j += gs) {
      const float4 a = *reinterpret_cast<const float4*>(mu + mo + 4 * j);
      const float4 b = *reinterpret_cast<const float4*>(lv + lo + 4 * j);
      const int64_t q = r * nq + j;
      float4 gm = make_float4(gk * a.x, gk * a.y, gk * a.z, gk * a.w);
      float4 gl = make_float4(-0.5f * gk * (1.0f - expf(b.x)), -0.5f * gk * (1.0f - expf(b.y)),
                              -0.5f * gk * (1.0f - expf(b.z)), -0.5f * gk * (1.0f - expf(b.w)));
      if (g_z) {
        const float4 g = *reinterpret_cast<const float4*>(g_z + 4 * q);
        const float4 e = philox_normal4(seed, offset, (uint64_t)q);
        gm.x += g.x; gm.y += g.y; gm.z += g.z; gm.w += g.w;
        gl.x += g.x * e.x * 0.5f * expf(0.5f * b.x);
        gl.y += g.y * e.y * 0.5f * expf(0.5f * b.y);
        gl.z += g.z * e.z * 0.5f * expf(0.5f * b.z);
        gl.w += g.w * e.w * 0.5f * expf(0.5f * b.w);
      }
      *reinterpret_cast<float4*>(g_mu + 4 * q) = gm;
      *reinterpret_cast<float4*>(g_lv + 4 * q) = gl;
    }
  } else {
    for (int64_t d = sub; d < m.D; d += gs) {
      const float a = mu[mo + d], b = lv[lo + d];
      const int64_t e_idx = r * m.D + d;
      float gm = gk * a, gl = -0.5f * gk * (1.0f - expf(b));
      if (g_z) {
        const float g = g_z[e_idx];
        const float e = philox_normal1(seed, offset, (uint64_t)e_idx);
        gm += g;
        gl += g * e * 0.5f * expf(0.5f * b);
      }
      g_mu[e_idx] = gm;
      g_lv[e_idx] = gl;
    }
  }
}

__global__ void __launch_bounds__(256)
philox_normal_kernel(float* __restrict__ out, int64_t numel, uint64_t seed, uint64_t offset) {
  const int64_t nq = (numel + 3) >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < nq; q += stride) {
    const float4 n = philox_normal4(seed, offset, (uint64_t)q);
    if (4 * q + 3 < numel && ((reinterpret_cast<uintptr_t>(out) & 15) == 0)) {
      *reinterpret_cast<float4*>(out + 4 * q) = n;
    } else {
      const float v[4] = {n.x, n.y, n.z, n.w};
      for (int c = 0; c < 4; ++c)
        if (4 * q + c < numel) out[4 * q + c] = v[c];
    }
  }
}

// Same stream, with (seed, offset) read from device memory: lets a process group share one rank's generator state with a
// broadcast and no host synchronisation (vaekit/dist.py).  key[0] = seed, key[1] = offset (both as int64 bit patterns).
__global__ void __launch_bounds__(256)
philox_normal_dev_kernel(float* __restrict__ out, int64_t numel, const int64_t* __restrict__ key, int64_t quad_shift) {
  const uint64_t seed = (uint64_t)key[0], offset = (uint64_t)key[1] + (uint64_t)quad_shift;
  const int64_t nq = (numel + 3) >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < nq; q += stride) {
    const float4 n = philox_normal4(seed, offset, (uint64_t)q);
    if (4 * q + 3 < numel && ((reinterpret_cast<uintptr_t>(out) & 15) == 0)) {
      *reinterpret_cast<float4*>(out + 4 * q) = n;
    } else {
      const float v[4] = {n.x, n.y, n.z, n.w};
      for (int c = 0; c < 4; ++c)
        if (4 * q + c < numel) out[4 * q + c] = v[c];
    }
  }
}

__global__ void philox_raw_kernel(uint32_t* __restrict__ out, int64_t nq, uint64_t seed, uint64_t offset) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q < nq) {
    const uint4 w = philox_quad(seed, offset, (uint64_t)q);
    out[4 * q + 0] = w.x; out[4 * q + 1] = w.y; out[4 * q + 2] = w.z; out[4 * q + 3] = w.w;
  }
}

static int make_rowmap(const int64_t* outer, const int64_t* ms, const int64_t* ls, int64_t D, RowMap& m) {
  VK_REQUIRE(outer && ms && ls, VK_ERR_BAD_ARG, "reparam: null shape/stride array");
  VK_REQUIRE(D > 0 && outer[0] >= 0 && outer[1] >= 0 && outer[2] >= 0, VK_ERR_BAD_ARG, "reparam: bad sizes");
  m.n0 = outer[0]; m.n1 = outer[1]; m.n2 = outer[2];
  m.ms0 = ms[0]; m.ms1 = ms[1]; m.ms2 = ms[2];
  m.ls0 = ls[0]; m.ls1 = ls[1]; m.ls2 = ls[2];
  m.rows = m.n0 * m.n1 * m.n2;
  m.D = D;
  return VK_OK;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

static bool vec_ok(const RowMap& m, const void* a, const void* b, const void* c, const void* d, const void* e) {
  return (m.D % 4 == 0) && (m.ms0 % 4 == 0) && (m.ms1 % 4 == 0) && (m.ms2 % 4 == 0) && (m.ls0 % 4 == 0) &&
         (m.ls1 % 4 == 0) && (m.ls2 % 4 == 0) && aligned16(a) && aligned16(b) && aligned16(c) && aligned16(d) &&
         aligned16(e);
}

static int group_size(int64_t work_items) {  // smallest power of two >= work_items, capped at 32
  int gs = 1;
  while (gs < 32 && gs < work_items) gs <<= 1;
  return gs;
}

}  // namespace vk

using namespace vk;

static int reparam_fwd(const float* mu, const float* log_var, const int64_t* outer_host,
                       const int64_t* mu_strides_host, const int64_t* lv_strides_host, int64_t D,
                       uint64_t seed, uint64_t offset, const int64_t* key, float* z, float* kl_rows, float* eps_out,
                       void* stream) {
  VK_REQUIRE(mu && log_var, VK_ERR_BAD_ARG, "reparam_kl_fwd: null mu/log_var");
  VK_REQUIRE(z || kl_rows || eps_out, VK_ERR_BAD_ARG, "reparam_kl_fwd: no output requested");
  RowMap m;
  int rc = make_rowmap(outer_host, mu_strides_host, lv_strides_host, D, m);
  if (rc) return rc;
  if (m.rows == 0) return VK_OK;
  const bool vec = vec_ok(m, mu, log_var, z, eps_out, nullptr);
  const int gs = group_size(vec ? D / 4 : D);
  const int64_t warps = (m.rows + (32 / gs) - 1) / (32 / gs);
  const int64_t blocks = (warps + 7) / 8;
  VK_REQUIRE(blocks < (1ll << 31), VK_ERR_UNSUPPORTED, "reparam_kl_fwd: too many rows");
  cudaStream_t s = (cudaStream_t)stream;
  if (vec)
    reparam_kl_fwd_kernel<true><<<(unsigned)blocks, 256, 0, s>>>(mu, log_var, m, gs, seed, offset, key, z, kl_rows, eps_out);
  else
    reparam_kl_fwd_kernel<false><<<(unsigned)blocks, 256, 0, s>>>(mu, log_var, m, gs, seed, offset, key, z, kl_rows, eps_out);
  return check_launch("reparam_kl_fwd");
}

extern "C" int vk_reparam_kl_fwd(const float* mu, const float* log_var, const int64_t* outer_host,
                                 const int64_t* mu_strides_host, const int64_t* lv_strides_host, int64_t D,
                                 uint64_t seed, uint64_t offset, float* z, float* kl_rows, float* eps_out,
                                 void* stream) {
  return reparam_fwd(mu, log_var, outer_host, mu_strides_host, lv_strides_host, D, seed, offset, nullptr, z, kl_rows,
                     eps_out, stream);
}

extern "C" int vk_reparam_kl_fwd_dev(const float* mu, const float* log_var, const int64_t* outer_host,
                                     const int64_t* mu_strides_host, const int64_t* lv_strides_host, int64_t D,
                                     const int64_t* seed_offset, float* z, float* kl_rows, float* eps_out,
                                     void* stream) {
  VK_REQUIRE(seed_offset, VK_ERR_BAD_ARG, "reparam_kl_fwd_dev: null (seed, offset) pointer");
  return reparam_fwd(mu, log_var, outer_host, mu_strides_host, lv_strides_host, D, 0, 0, seed_offset, z, kl_rows,
                     eps_out, stream);
}

static int reparam_bwd(const float* g_z, const float* g_kl, const float* mu, const float* log_var,
                       const int64_t* outer_host, const int64_t* mu_strides_host,
                       const int64_t* lv_strides_host, int64_t D, uint64_t seed, uint64_t offset, const int64_t* key,
                       float* g_mu, float* g_lv, void* stream) {
  VK_REQUIRE(mu && log_var && g_mu && g_lv, VK_ERR_BAD_ARG, "reparam_kl_bwd: null pointer");
  RowMap m;
  int rc = make_rowmap(outer_host, mu_strides_host, lv_strides_host, D, m);
  if (rc) return rc;
  if (m.rows == 0) return VK_OK;
  const bool vec = vec_ok(m, mu, log_var, g_z, g_mu, g_lv);
  const int gs = group_size(vec ? D / 4 : D);
  const int64_t warps = (m.rows + (32 / gs) - 1) / (32 / gs);
  const int64_t blocks = (warps + 7) / 8;
  VK_REQUIRE(blocks < (1ll << 31), VK_ERR_UNSUPPORTED, "reparam_kl_bwd: too many rows");
  cudaStream_t s = (cudaStream_t)stream;
  if (vec)
    reparam_kl_bwd_kernel<true><<<(unsigned)blocks, 256, 0, s>>>(g_z, g_kl, mu, log_var, m, gs, seed, offset, key, g_mu, g_lv);
  else
    reparam_kl_bwd_kernel<false><<<(unsigned)blocks, 256, 0, s>>>(g_z, g_kl, mu, log_var, m, gs, seed, offset, key, g_mu, g_lv);
  return check_launch("reparam_kl_bwd");
}

extern "C" int vk_reparam_kl_bwd(const float* g_z, const float* g_kl, const float* mu, const float* log_var,
                                 const int64_t* outer_host, const int64_t* mu_strides_host,
                                 const int64_t* lv_strides_host, int64_t D, uint64_t seed, uint64_t offset,
                                 float* g_mu, float* g_lv, void* stream) {
  return reparam_bwd(g_z, g_kl, mu, log_var, outer_host, mu_strides_host, lv_strides_host, D, seed, offset, nullptr, g_mu,
                     g_lv, stream);
}

extern "C" int vk_reparam_kl_bwd_dev(const float* g_z, const float* g_kl, const float* mu, const float* log_var,
                                     const int64_t* outer_host, const int64_t* mu_strides_host,
                                     const int64_t* lv_strides_host, int64_t D, const int64_t* seed_offset,
                                     float* g_mu, float* g_lv, void* stream) {
  VK_REQUIRE(seed_offset, VK_ERR_BAD_ARG, "reparam_kl_bwd_dev: null (seed, offset) pointer");
  return reparam_bwd(g_z, g_kl, mu, log_var, outer_host, mu_strides_host, lv_strides_host, D, 0, 0, seed_offset, g_mu, g_lv,
                     stream);
}

extern "C" int vk_philox_normal(float* out, int64_t numel, uint64_t seed, uint64_t offset, void* stream) {
  VK_REQUIRE(out && numel >= 0, VK_ERR_BAD_ARG, "philox_normal: bad argument");
  if (numel == 0) return VK_OK;
  const int64_t nq = (numel + 3) / 4;
  int64_t blocks = (nq + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  philox_normal_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(out, numel, seed, offset);
  return check_launch("philox_normal");
}

extern "C" int vk_philox_normal_dev(float* out, int64_t numel, const int64_t* seed_offset, int64_t quad_shift,
                                    void* stream) {
  VK_REQUIRE(out && seed_offset && numel >= 0 && quad_shift >= 0, VK_ERR_BAD_ARG, "philox_normal_dev: bad argument");
  if (numel == 0) return VK_OK;
  const int64_t nq = (numel + 3) / 4;
  int64_t blocks = (nq + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  philox_normal_dev_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(out, numel, seed_offset, quad_shift);
  return check_launch("philox_normal_dev");
}

extern "C" int vk_philox_raw(uint32_t* out, int64_t n_quads, uint64_t seed, uint64_t offset, void* stream) {
  VK_REQUIRE(out && n_quads >= 0, VK_ERR_BAD_ARG, "philox_raw: bad argument");
  if (n_quads == 0) return VK_OK;
  philox_raw_kernel<<<(unsigned)((n_quads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(out, n_quads, seed, offset);
  return check_launch("philox_raw");
}
