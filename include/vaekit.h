/*
 * vaekit.h -- C ABI of libvaekit.so: the B200 (sm_100a) kernels behind the
 * stochastic-latent + loss hot path of bplaut/PyTorch-VAE.
 *
 * The reference has no FFI of its own (it is pure Python over torch); the
 * boundary it offers is the method surface of its BaseVAE subclasses
 * (models/base.py:5-28).  Each entry point below replaces the ATen call
 * sequence of the cited reference lines; the Python host layer
 * (pytorch-vae_b200/vaekit) binds them with ctypes and exposes the reference's
 * method names on top.
 *
 * Conventions (all entry points):
 *   - plain C types only; every pointer is a DEVICE pointer to fp32 data
 *     unless the parameter name ends in _host;
 *   - returns 0 on success, a negative VK_ERR_* otherwise; never throws,
 *     never allocates, never synchronises the stream;
 *   - `stream` is a cudaStream_t passed as void*; work is enqueued on it;
 *   - workspaces are caller-provided; vk_workspace_bytes() gives the size.  The
 *     first VK_WS_HEADER_BYTES of a workspace must be ZERO on first use and are
 *     left zero by every call (self-resetting tickets), so one zero-initialised
 *     buffer per stream can be reused for all calls;
 *   - tensors are dense row-major unless strides are passed explicitly.
 *
 * RNG contract: noise is a pure function of (seed, offset, element index):
 *   quad q = element/4 draws Philox4x32-10 with counter (lo32(q+offset),
 *   hi32(q+offset), 0, 0) and key (lo32(seed), hi32(seed)); the four words give
 *   four normals by Box-Muller on 23-bit uniforms (oracle/vae_oracle.py
 *   philox_normal is the CPU statement of the same stream).  The result does
 *   not depend on launch geometry or on how a batch is sharded over GPUs.
 */
#ifndef VAEKIT_H_
#define VAEKIT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VK_OK                 0
#define VK_ERR_BAD_ARG       -1   /* null pointer, negative size, misaligned, unknown mode */
#define VK_ERR_UNSUPPORTED   -2   /* shape outside what the kernel was built for */
#define VK_ERR_WORKSPACE     -3   /* workspace too small */
#define VK_ERR_CUDA          -4   /* launch / runtime error, see vk_last_error_string() */

#define VK_WS_HEADER_BYTES   256

int         vk_version(void);             /* major*10000 + minor*100 + patch */
const char* vk_last_error_string(void);   /* thread-local, valid until the next failing call */
/* Generic upper bound for the reduction workspaces of the streaming ops. */
int64_t     vk_workspace_bytes(void);

/* ---------------------------------------------------------------------------
 * K1  reparameterize + Gaussian KL          (reference: models/vanilla_vae.py:107-117,
 *     beta_vae.py:112-122, logcosh_vae.py:108-118, iwae.py:111-119, miwae.py:114-122,
 *     betatc_vae.py:115-125; KL rows: vanilla_vae.py:143 inner sum, iwae.py:150)
 *
 * mu / log_var are addressed as rows of D contiguous floats; row r of the
 * logical [n0,n1,n2] outer grid (r = (i0*n1 + i1)*n2 + i2) starts at element
 * i0*s[0] + i1*s[1] + i2*s[2]  (this covers the IWAE/MIWAE repeat().permute()
 * views, iwae.py:123-124, without materialising them).
 *   z        [rows, D] dense, or NULL      z = eps * exp(0.5*log_var) + mu
 *   kl_rows  [rows]    or NULL             -0.5 * sum_d(1 + lv - mu^2 - exp(lv))
 *   eps_out  [rows, D] dense, or NULL      the noise used (for oracle injection)
 * If z and eps_out are both NULL no noise is generated (KL-only mode).
 * -------------------------------------------------------------------------*/
int vk_reparam_kl_fwd(const float* mu, const float* log_var,
                      const int64_t* outer_host /*[3]*/, const int64_t* mu_strides_host /*[3]*/,
                      const int64_t* lv_strides_host /*[3]*/, int64_t D,
                      uint64_t seed, uint64_t offset,
                      float* z, float* kl_rows, float* eps_out, void* stream);

/* Backward of the above.  g_z [rows,D] dense or NULL; g_kl [rows] or NULL.
 *   g_mu = g_z + g_kl * mu ;  g_lv = g_z * eps * 0.5 * exp(0.5 lv) - 0.5 * g_kl * (1 - exp(lv))
 * eps is regenerated from (seed, offset).  g_mu / g_lv are dense [rows, D]. */
int vk_reparam_kl_bwd(const float* g_z, const float* g_kl, const float* mu, const float* log_var,
                      const int64_t* outer_host, const int64_t* mu_strides_host,
                      const int64_t* lv_strides_host, int64_t D,
                      uint64_t seed, uint64_t offset,
                      float* g_mu, float* g_lv, void* stream);

/* The same two entry points with (seed, offset) read from DEVICE memory (seed_offset[0], [1] as int64 bit patterns,
 * like vk_philox_normal_dev): the pair is no launch argument, so a CUDA graph that captured these launches draws fresh
 * noise on every replay once the caller advances seed_offset[1] on the device between replays (config 1 at
 * batch 64 is launch-latency-bound and wants the whole step in one graph). */
int vk_reparam_kl_fwd_dev(const float* mu, const float* log_var,
                          const int64_t* outer_host, const int64_t* mu_strides_host,
                          const int64_t* lv_strides_host, int64_t D, const int64_t* seed_offset,
                          float* z, float* kl_rows, float* eps_out, void* stream);
int vk_reparam_kl_bwd_dev(const float* g_z, const float* g_kl, const float* mu, const float* log_var,
                          const int64_t* outer_host, const int64_t* mu_strides_host,
                          const int64_t* lv_strides_host, int64_t D, const int64_t* seed_offset,
                          float* g_mu, float* g_lv, void* stream);

/* Standard-normal fill with the contract stream (prior_z of wae_mmd.py:195 / swae.py:165). */
int vk_philox_normal(float* out, int64_t numel, uint64_t seed, uint64_t offset, void* stream);
/* The same stream with (seed, offset) read from DEVICE memory (seed_offset[0], [1] as int64 bit patterns) and started
 * quad_shift counters later: a process group shares one rank's generator state by broadcast without a host sync, and a
 * rank generates only its row block of a global [N, D] sample (quad_shift = first_row * D / 4). */
int vk_philox_normal_dev(float* out, int64_t numel, const int64_t* seed_offset, int64_t quad_shift, void* stream);
/* Raw Philox words, [n_quads,4] uint32 -- for the bit-exact stream test. */
int vk_philox_raw(uint32_t* out, int64_t n_quads, uint64_t seed, uint64_t offset, void* stream);

/* ---------------------------------------------------------------------------
 * K2  reconstruction term + KL combine       (reference: vanilla_vae.py:134-146,
 *     beta_vae.py:129-152, logcosh_vae.py:125-155, betatc_vae.py:163,
 *     swae.py:121-122, wae_mmd.py:118, vq_vae.py:205, info_vae.py:127-133)
 *
 * recon_mode: what out[1] is
 *   VK_RECON_MSE_MEAN     mean((r-x)^2)
 *   VK_RECON_MSE_SUM      sum((r-x)^2)
 *   VK_RECON_LOGCOSH      (1/alpha) * mean(alpha*t + log(1+exp(-2 alpha t)) - log 2), t = r-x
 *   VK_RECON_MSE_L1_MEAN  mean((r-x)^2) + mean(|r-x|)
 *   VK_RECON_BCE_MEAN / VK_RECON_BCE_SUM   the "MSE/BCE" of the north star: the reference itself has no BCE term
 *                         (SURVEY 0.5), so the contract is torch's F.binary_cross_entropy(recons, input, reduction=
 *                         'mean' | 'sum'): -(x * max(log r, -100) + (1 - x) * max(log1p(-r), -100)); backward
 *                         (r - x) / max((1 - r) r, 1e-12).  recons outside [0, 1] give NaN (torch raises).
 * combine_mode: what out[0] is (kld = mean(kl_rows); ignored when kl_rows is NULL)
 *   VK_COMBINE_NONE       out[0] = out[1]
 *   VK_COMBINE_LINEAR     out[0] = out[1] + w * kld
 *   VK_COMBINE_ABS        out[0] = out[1] + w * |kld - C|
 *   VK_COMBINE_SCALED     out[0] = C * out[1] + w * kld        (info_vae.py:131-132; applies without kl_rows too)
 * out[4] = { loss, recon_loss, kld, -kld }.
 * -------------------------------------------------------------------------*/
enum { VK_RECON_MSE_MEAN = 0, VK_RECON_MSE_SUM = 1, VK_RECON_LOGCOSH = 2, VK_RECON_MSE_L1_MEAN = 3, VK_RECON_BCE_MEAN = 4,
       VK_RECON_BCE_SUM = 5 };
enum { VK_COMBINE_NONE = 0, VK_COMBINE_LINEAR = 1, VK_COMBINE_ABS = 2, VK_COMBINE_SCALED = 3 };

int vk_elbo_fwd(const float* recons, const float* input, int64_t numel,
                const float* kl_rows, int64_t rows,
                int recon_mode, float alpha, int combine_mode, float w, float C,
                float* out /*[4]*/, void* workspace, int64_t workspace_bytes, void* stream);

/* Backward.  g_loss/g_recon/g_kld/g_negkld: device pointers to the upstream
 * scalar gradients of out[0..3], each may be NULL (= 0).  `out` is the forward
 * result (sign of kld - C).  g_recons [numel] and g_kl_rows [rows] may be NULL. */
int vk_elbo_bwd(const float* recons, const float* input, int64_t numel, int64_t rows,
                int recon_mode, float alpha, int combine_mode, float w, float C,
                const float* out, const float* g_loss, const float* g_recon,
                const float* g_kld, const float* g_negkld,
                float* g_recons, float* g_kl_rows, void* stream);

/* Per-image reconstruction error for the trainer (reference: experiment.py:58-62 recomputes
 * F.mse_loss(recons, input, reduction='none').mean(dim=[1,2,3]) and reads it back image by image):
 * out[b] = mean over the P = C*H*W elements of image b of (recons - input)^2, one streaming pass. */
int vk_per_image_mse(const float* recons, const float* input, int64_t B, int64_t P, float* out, void* stream);

/* ---------------------------------------------------------------------------
 * K3  IWAE / MIWAE importance-weighted bound   (reference: iwae.py:129-160,
 *     miwae.py:132-164; closed-form backward: SURVEY.md A.2)
 *
 * recons [B, G, K, P] dense (G = num_estimates, 1 for IWAE), input [B, P] is
 * broadcast over (G, K) in-kernel (the reference materialises the repeat,
 * iwae.py:145).  kl [B*G*K] are the KL rows of K1.
 *   lp[b,g,k] = mean_P (r - x)^2 ;  lw = lp + M_N*kl ;  w = softmax_k(lw)
 *   out[4] = { mean_{b,g} sum_k w*lw,  mean(lp),  mean(kl),  -mean(kl) }
 * lp_out [B*G*K] is saved for the backward.
 * -------------------------------------------------------------------------*/
int vk_iw_bound_fwd(const float* recons, const float* input, const float* kl,
                    int64_t B, int64_t G, int64_t K, int64_t P, float M_N,
                    float* lp_out, float* out /*[4]*/,
                    void* workspace, int64_t workspace_bytes, void* stream);

int vk_iw_bound_bwd(const float* recons, const float* input, const float* lp, const float* kl,
                    int64_t B, int64_t G, int64_t K, int64_t P, float M_N,
                    const float* g_loss, const float* g_recon, const float* g_kld, const float* g_negkld,
                    float* g_recons /*[B,G,K,P] or NULL*/, float* g_kl /*[B*G*K] or NULL*/, void* stream);

/* Allocation-free variant of vk_iw_bound_fwd.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments and results as vk_iw_bound_fwd without the workspace: the scratch (a
 * 256-byte ticket header + 3 doubles per batch item) lives in a file-scope `static __device__` array of the library,
 * 256 + 24 * VK_IW_STATIC_MAX_BATCH = 393 472 bytes, which every GPU that loads the library holds until the process
 * exits.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine); it enqueues one
 * 256-byte cudaMemsetAsync (the header must start at zero; the array keeps the previous call's contents) and the kernel.
 *   B > VK_IW_STATIC_MAX_BATCH -> VK_ERR_UNSUPPORTED with a message naming the limit (use vk_iw_bound_fwd instead).
 *   The array's device address is resolved with cudaGetSymbolAddress on the first call per device (that device current)
 *   and cached: make that first call outside a stream capture.
 *   Every call on a device shares the one array: calls that OVERLAP on the GPU (different streams or host threads)
 *   race on it and give wrong results; calls on one stream are ordered and safe. */
#define VK_IW_STATIC_MAX_BATCH 16384
int vk_iw_bound_fwd_static(const float* recons, const float* input, const float* kl,
                           int64_t B, int64_t G, int64_t K, int64_t P, float M_N,
                           float* lp_out, float* out /*[4]*/, void* stream);

/* ---------------------------------------------------------------------------
 * K7  VQ-VAE VectorQuantizer               (reference: models/vq_vae.py:24-55)
 *
 * latents [B, D, H*W] (NCHW, read in place -- no NHWC copy), codebook [K, D].
 *   idx[r]  = argmin_k (|x_r|^2 + |e_k|^2) - 2 x_r.e_k   (first index on ties), r = b*HW + hw
 *   q_out   = x + (e_idx - x)                            (straight-through value, vq_vae.py:51), NCHW
 *   out[2]  = { vq_loss = beta*mse + mse,  mse = mean (e_idx - x)^2 }
 * The distance GEMM runs on tcgen05 tensor cores as three bf16 passes (hi*hi + lo*hi + hi*lo, fp32
 * accumulation in TMEM): |error| on x.e is ~2^-17 |x||e| sqrt(D).
 * Backward (g_q upstream of q_out or NULL, g_loss upstream of vq_loss or NULL):
 *   g_latents  = g_q + g_loss * (2 beta / numel) (x - q)
 *   g_codebook[k] = sum_{r: idx_r = k} g_loss * (2 / numel) (q_r - x_r)      (overwritten, not accumulated)
 * -------------------------------------------------------------------------*/
int64_t vk_vq_workspace_bytes(int64_t D, int64_t K);
int vk_vq_fwd(const float* latents, const float* codebook, int64_t B, int64_t D, int64_t HW, int64_t K,
              float beta, float* q_out, int32_t* idx_out, float* out /*[2]*/,
              void* workspace, int64_t workspace_bytes, void* stream);
int vk_vq_bwd(const float* g_q, const float* g_loss, const float* latents, const float* codebook,
              const int32_t* idx, int64_t B, int64_t D, int64_t HW, int64_t K, float beta,
              float* g_latents /*or NULL*/, float* g_codebook /*or NULL*/, void* stream);

/* ---------------------------------------------------------------------------
 * K4  WAE-MMD                              (reference: models/wae_mmd.py:125-204)
 *
 * z, prior [N, D] fp32 (D <= 128).  kernel_type: VK_KERNEL_RBF  exp(-|x-y|^2 / (2 D^2 z_var)), full N x N sums
 * (wae_mmd.py:164-167); VK_KERNEL_IMQ  C/(1e-7 + C + |x-y|^2), C = 2 D z_var, diagonal removed from all three
 * sums (wae_mmd.py:185-191).  One sweep produces, for the row range [row_begin, row_end) (128-aligned; the
 * whole matrix on one GPU, a row block per rank when sharded after an all-gather of z):
 *   sums_out[3] (double, device) = { sum k(z_i,z_j), sum k(z_i,prior_j), sum k(prior_i,prior_j) } over the rows
 *   grad_unscaled [(row_end-row_begin), D]:  G_i = sum_j dk/dd^2(z_i,z_j) (z_i-z_j) - sum_j dk/dd^2(z_i,p_j) (z_i-p_j)
 *                 so that  d mmd / d z_i = 4 * reg * m * G_i,  m = 1/N^2 (rbf) or 1 (imq)
 * x.y runs on tcgen05 tensor cores in bf16 with fp32 accumulation; `passes` selects the precision of the two GEMMs
 * (distance GEMM / gradient GEMM):
 *   3  both in three passes of a bf16 hi/lo split (hi*hi + lo*hi + hi*lo, ~2^-17 relative): the parity sweep
 *   2  distance GEMM two passes (row operand hi/lo, column operand plain bf16), gradient GEMM one pass
 *   1  both in one plain bf16 pass
 *   22, 31, 32  an explicit (distance GEMM, gradient GEMM) pair; a two-pass gradient GEMM splits the coefficients only
 *   0  auto: the library picks (rbf at N >= 8192: mode 2; otherwise the parity sweep); a reduced-precision sweep
 *      is CHECKED against a float64 SIMT evaluation of 16 sample rows over all columns, and when its error exceeds
 *      4e-6 of max |grad| a device flag makes the parity sweep run behind it -- no host synchronisation either way.
 * The data are centred on the device (distances are translation invariant) and the gradient GEMM accumulates CENTRED
 * coefficients c - c0 (c0 = the coefficient at the mean squared distance; its share of the gradient is exact float64
 * arithmetic on column sums), which is what keeps the tcgen05 accumulator's truncation out of the result.
 * Latent dims > 128 take a SIMT kernel in the reference's literal (x - y)^2 form (slower, any D <= 2048).
 * diag (device, 4 floats, may be NULL): { 10 * g1 + g2 passes of the sweep whose result stands, measured sample error /
 * max |grad| (0 when nothing was checked), 1 if the parity sweep had to run, max |grad_unscaled| }.
 * vk_mmd_combine: out[4] = { reg*pp + reg*zz - 2*reg*pz, pp, zz, pz } with pp = mean (rbf) / sum (imq) of the
 * (all-reduced) sums, in the reference's fp32 association (wae_mmd.py:201-203).
 * -------------------------------------------------------------------------*/
enum { VK_KERNEL_RBF = 0, VK_KERNEL_IMQ = 1 };
int64_t vk_mmd_workspace_bytes(int64_t N, int64_t D);
int vk_mmd_fwd(const float* z, const float* prior, int64_t N, int64_t D, int kernel_type, float z_var, int passes,
               int64_t row_begin, int64_t row_end, double* sums_out /*[3]*/, float* grad_unscaled,
               float* diag /*[4] or NULL*/, void* workspace, int64_t workspace_bytes, void* stream);
int vk_mmd_combine(const double* sums /*[3]*/, int64_t N, int kernel_type, float reg_weight, float* out /*[4]*/, void* stream);
/* out[i] = in[i] * (*g) * factor  (g may be NULL = 1): backward of an op whose gradient was produced in the forward sweep */
int vk_scale_by_scalar(const float* in, const float* g, float factor, int64_t n, float* out, void* stream);

/* ---------------------------------------------------------------------------
 * K4b  the kernel matrix on its own   (reference: models/wae_mmd.py:125-191, models/info_vae.py:150-201 --
 *      compute_kernel / compute_rbf / compute_inv_mult_quad called directly; compute_mmd uses vk_mmd_fwd instead)
 *
 * x1, x2: the reference's [N1, N2, D] operands as base pointer + element strides s[3] = {i, j, d}; the expand() views
 * of wae_mmd.py:140-141 have stride 0 on the broadcast axis and are read in place.
 *   rbf: kmat[i, j] = exp(-(mean_d (x1 - x2)^2) / (2 D z_var))                                (wae_mmd.py:164-167)
 *   imq: kmat[i, j] = C / (eps + C + sum_d (x1 - x2)^2), C = 2 D z_var, and
 *        imq_out[0] = sum kmat - sum diag kmat   (the scalar the reference returns, wae_mmd.py:185-189);
 *        imq_out must be NULL for rbf and non-NULL for imq.
 * Backward, g = upstream gradient read as g[i * g_sr + j * g_sc] (rbf: the [N1, N2] gradient of kmat; imq: the scalar
 * gradient with both strides 0, the diagonal is excluded):
 *   vk_pair_kernel_bwd_rows : x1 [N1, D], x2 [N2, D] (the compute_kernel case) -> g_x1 [N1, D], g_x2 [N2, D] (either
 *                             may be NULL), i.e. the expanded gradients already summed over the broadcast axis;
 *   vk_pair_kernel_bwd_dense: general [N1, N2, D] operands -> g_x1, g_x2 [N1, N2, D] contiguous (either may be NULL).
 * workspace: vk_workspace_bytes().
 * -------------------------------------------------------------------------*/
int vk_pair_kernel_fwd(const float* x1, const int64_t* s1, const float* x2, const int64_t* s2, int64_t N1, int64_t N2,
                       int64_t D, int kernel_type, float z_var, float eps, float* kmat, float* imq_out, void* workspace,
                       int64_t workspace_bytes, void* stream);
int vk_pair_kernel_bwd_rows(const float* x1, int64_t x1_sr, int64_t x1_sd, const float* x2, int64_t x2_sr, int64_t x2_sd,
                            int64_t N1, int64_t N2, int64_t D, int kernel_type, float z_var, float eps, const float* kmat,
                            const float* g, int64_t g_sr, int64_t g_sc, float* g_x1, float* g_x2, void* stream);
int vk_pair_kernel_bwd_dense(const float* x1, const int64_t* s1, const float* x2, const int64_t* s2, int64_t N1, int64_t N2,
                             int64_t D, int kernel_type, float z_var, float eps, const float* kmat, const float* g,
                             int64_t g_sr, int64_t g_sc, float* g_x1, float* g_x2, void* stream);

/* ---------------------------------------------------------------------------
 * K6  SWAE sliced-Wasserstein distance       (reference: models/swae.py:151-179)
 *
 * z, prior [N, D]; proj [S, D] unit-norm projection directions (swae.py:129-148 builds them on the CPU generator;
 * S may be one rank's shard of the model's projections).  One call produces
 *   sum_out (double, device)      = sum_{s,i} (sort_i(z.theta_s) - sort_i(prior.theta_s))^p
 *   grad_unscaled [N, D] or NULL  = d sum_out / d z          (through the sort permutation)
 * so that  swd = reg * sum / (S_total N)  and  d swd / dz = reg / (S_total N) * grad_unscaled.
 * Steps: fp32 projection GEMM stored projection-major, segmented LSD radix sort (8-bit digits; key + sample index
 * for z, key-only for the prior), loss + scatter of d/dlat, fp32 GEMM back to [N, D].
 * vk_mean_scale: out[0] = scale * (float)(sum / count)     (reg_weight * w_dist.mean(), swae.py:179)
 * vk_swd_segments_fwd: the middle of vk_swd_fwd on its own, for the multi-GPU decomposition in which ranks project
 * their own rows and exchange projected values (all-to-all): lat, pri [S, N] projection-major segments (swae.py:176-177
 * sorts exactly these rows).  On return *sum_out = sum (sort(lat) - sort(pri))^p, `lat` holds d sum / d lat in SAMPLE
 * order when want_grad (else garbage) and `pri` holds the sorted prior projections.
 * vk_seg_sort: the sort on its own -- keys [segments, n] fp32 ascending in place, perm_out (uint32 source index
 * of every sorted element) or NULL.
 * -------------------------------------------------------------------------*/
int64_t vk_swd_workspace_bytes(int64_t N, int64_t S);
int vk_swd_fwd(const float* z, const float* prior, const float* proj, int64_t N, int64_t D, int64_t S, float p,
               double* sum_out, float* grad_unscaled, void* workspace, int64_t workspace_bytes, void* stream);
int64_t vk_swd_segments_workspace_bytes(int64_t N, int64_t S);
int vk_swd_segments_fwd(float* lat, float* pri, int64_t N, int64_t S, float p, double* sum_out, int want_grad,
                        void* workspace, int64_t workspace_bytes, void* stream);
int vk_mean_scale(const double* sum, double count, float scale, float* out, void* stream);
int64_t vk_seg_sort_workspace_bytes(int64_t segments, int64_t n);
int vk_seg_sort(float* keys, uint32_t* perm_out, int64_t segments, int64_t n, void* workspace,
                int64_t workspace_bytes, void* stream);

/* vk_gemm_nt: the fp32-grade tensor-core product used by the two SWAE GEMMs, on its own:
 *   C(n, j) = sum_k A(n, k) * B(j, k)     A(n, k) = A[n*a_sn + k*a_sk], B(j, k) = B[j*b_sj + k*b_sk], C likewise
 * J and K small (J, K <= 256 after padding to 16, 3*K/2 + J <= 512 tensor-memory columns), N arbitrary.  Operands are
 * split three ways into bf16 and accumulated from six tcgen05 passes in fp32 (error ~ 2^-23 |A||B|).
 * Returns VK_ERR_UNSUPPORTED when the shape does not fit (vk_swd_fwd then uses its fp32 FFMA GEMM).
 * workspace: vk_gemm_nt_workspace_bytes() bytes. */
int64_t vk_gemm_nt_workspace_bytes(void);
int vk_gemm_nt(const float* A, int64_t a_sn, int64_t a_sk, const float* B, int64_t b_sj, int64_t b_sk, float* C,
               int64_t c_sn, int64_t c_sj, int64_t N, int64_t J, int64_t K, void* workspace, int64_t workspace_bytes,
               void* stream);

/* vk_gemm_nt_tab: the same GEMM with optional per-column base addresses (device arrays of device pointers):
 *   a_cols[k] = address of A(0, k) (then A(n, k) = a_cols[k][n * a_sn]; A and a_sk are ignored),
 *   c_cols[j] = address of C(0, j) (then C(n, j) = c_cols[j][n * c_sn]; C and c_sj are ignored).
 * A column may live in another GPU's peer-mapped memory: the multi-GPU SWD (vaekit/dist.py, exchange="peer") stores
 * every projection straight into the segment buffer of the rank that sorts it and reads d/dlat columns straight from
 * their owners over NVLink, so the exchange that follows / precedes the GEMM is fused into it. */
int vk_gemm_nt_tab(const float* const* a_cols, const float* A, int64_t a_sn, int64_t a_sk, const float* B, int64_t b_sj,
                   int64_t b_sk, float* const* c_cols, float* C, int64_t c_sn, int64_t c_sj, int64_t N, int64_t J,
                   int64_t K, void* workspace, int64_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------
 * K5  Beta-TC-VAE latent terms             (reference: models/betatc_vae.py:132-142, 165-211)
 *
 * z, mu, log_var [B, D] (B >= 2).  ld_ijd = log N(z_id; mu_jd, exp(lv_jd)); the [B,B] log importance weights
 * (columns 0 and 1 and entry [B-2,0] special, betatc_vae.py:179-182) are added to every latent dim before both
 * logsumexps (betatc_vae.py:185).  gamma_anneal = anneal_rate * gamma (host state, betatc_vae.py:196-200).
 *   out[4]   = { alpha*mi + (beta*tc + gamma_anneal*kld),  kld,  tc,  mi }       (means over the batch)
 *   lqz_out [2B], lse_out [B, D]: log_q_z[i] (first B floats; the next B hold its fp32 rounding residual, so the
 *   backward can rebuild the softmax weights exp(A_ij - log_q_z[i]) without ulp(|log_q_z|) noise) and
 *   LSE_j(ld_ijd + lw_ij), saved for the backward
 * Backward: g_loss upstream of out[0] (NULL = 1); writes g_z, g_mu, g_lv [B, D] (overwritten).
 * The B^2 D densities each need their own exp (per-dimension LSE), so the kernels are MUFU/FP32 bound.
 * -------------------------------------------------------------------------*/
int64_t vk_tc_workspace_bytes(int64_t B, int64_t D);
int vk_tc_fwd(const float* z, const float* mu, const float* log_var, int64_t B, int64_t D, float M_N,
              float alpha, float beta, float gamma_anneal, float* lqz_out, float* lse_out, float* out /*[4]*/,
              void* workspace, int64_t workspace_bytes, void* stream);
int vk_tc_bwd(const float* z, const float* mu, const float* log_var, const float* lqz, const float* lse,
              int64_t B, int64_t D, float M_N, float alpha, float beta, float gamma_anneal, const float* g_loss,
              float* g_z, float* g_mu, float* g_lv, void* workspace, int64_t workspace_bytes, void* stream);
/* Row-block forms for the multi-GPU decomposition (all-gather z, mu, log_var; rank r owns anchor rows
 * [row_begin, row_end) against all B partner columns).  All pointers address the FULL [B, ...] arrays; only the rows of
 * the range are written (lqz_out, lse_out, g_z, g_mu, g_lv), and out[4] holds the range's share of the batch means, so
 * the shares of disjoint ranges add up to vk_tc_fwd's out.  The backward needs lqz / lse of ALL rows (all-gathered):
 * g_mu / g_lv of a partner column sum over every anchor row.  Latent dim <= 128. */
int vk_tc_fwd_rows(const float* z, const float* mu, const float* log_var, int64_t B, int64_t D, int64_t row_begin,
                   int64_t row_end, float M_N, float alpha, float beta, float gamma_anneal, float* lqz_out, float* lse_out,
                   float* out /*[4]*/, void* workspace, int64_t workspace_bytes, void* stream);
int vk_tc_bwd_rows(const float* z, const float* mu, const float* log_var, const float* lqz, const float* lse, int64_t B,
                   int64_t D, int64_t row_begin, int64_t row_end, float M_N, float alpha, float beta, float gamma_anneal,
                   const float* g_loss, float* g_z, float* g_mu, float* g_lv, void* workspace, int64_t workspace_bytes,
                   void* stream);

/* ---------------------------------------------------------------------------
 * K8  Gumbel-softmax reparameterisation and categorical KL   (reference: models/cat_vae.py:118-131, 133-173; the same
 *     trick in joint_vae.py:142-167) -- SURVEY 8(f) rank 4
 *
 * q [R, C] logits, one categorical row per (sample, latent dim).
 *   vk_gumbel_softmax_fwd: u ~ U(0,1) from the contract Philox stream (one 23-bit uniform per element, keyed by the
 *     element index) or read from u_in when non-NULL; g = -log(-log(u + eps) + eps); s_out = softmax((q + g) / temp, -1).
 *     u_out (optional) receives the uniforms so the oracle can be evaluated on the same noise.
 *   vk_gumbel_softmax_bwd: g_q = s (g_s - sum_c g_s s) / temp.
 *   vk_cat_kl_rows_fwd: rows_out[b] = sum_{l, c} p (log(p + eps) - log(1/C + eps)), p = softmax(q[b, l, :]); the batch
 *     mean and the loss combine happen in vk_elbo_fwd (rows_out goes in as its kl_rows).
 *   vk_cat_kl_rows_bwd: g_q from the per-sample upstream gradients g_rows [B].
 * -------------------------------------------------------------------------*/
int vk_gumbel_softmax_fwd(const float* q, int64_t R, int64_t C, float temp, float eps, uint64_t seed, uint64_t offset,
                          const float* u_in, float* s_out, float* u_out, void* stream);
int vk_gumbel_softmax_bwd(const float* s, const float* g_s, int64_t R, int64_t C, float temp, float* g_q, void* stream);
int vk_cat_kl_rows_fwd(const float* q, int64_t B, int64_t L, int64_t C, float eps, float* rows_out, void* stream);
int vk_cat_kl_rows_bwd(const float* q, const float* g_rows, int64_t B, int64_t L, int64_t C, float eps, float* g_q,
                       void* stream);

/* ---------------------------------------------------------------------------
 * K9  VampPrior term   (reference: models/vampvae.py:124-160) -- SURVEY 8(f) rank 4
 *
 * z [B, D] samples; (pmu, plv) [K, D] the encodings of the K pseudo-inputs; (mu, log_var) [B, D] the posterior.
 *   vk_gauss_logq_rows_fwd: rows_out[i] = sum_d -0.5 (lv_id + (z_id - mu_id)^2) / exp(lv_id)   (vampvae.py:136-138; the
 *     batch mean is E_log_q_z).  _bwd: (g_z, g_mu, g_lv) from the per-row upstream gradients g_rows [B].
 *   vk_vamp_prior_fwd: t_out[i, k] = sum_d -0.5 (plv_kd + (z_id - pmu_kd)^2) / exp(plv_kd) - log K, lse_out[i] =
 *     logsumexp_k t[i, k]   (vampvae.py:153-158; the batch mean is E_log_p_z).  The [B, K, D] broadcast of the reference
 *     is never materialised.
 *   vk_vamp_prior_bwd: g_z [B, D], g_pmu [K, D], g_plv [K, D] from g_rows [B] (gradient w.r.t. lse_out), with the
 *     mixture weights recomputed from (t, lse).  Fixed summation order, no atomics.
 * K, D <= 8192.
 * -------------------------------------------------------------------------*/
int vk_gauss_logq_rows_fwd(const float* z, const float* mu, const float* log_var, int64_t B, int64_t D, float* rows_out,
                           void* stream);
int vk_gauss_logq_rows_bwd(const float* z, const float* mu, const float* log_var, const float* g_rows, int64_t B, int64_t D,
                           float* g_z, float* g_mu, float* g_lv, void* stream);
int vk_vamp_prior_fwd(const float* z, const float* pmu, const float* plv, int64_t B, int64_t K, int64_t D, float* t_out,
                      float* lse_out, void* stream);
int vk_vamp_prior_bwd(const float* z, const float* pmu, const float* plv, const float* t, const float* lse,
                      const float* g_rows, int64_t B, int64_t K, int64_t D, float* g_z, float* g_pmu, float* g_plv,
                      void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VAEKIT_H_ */
