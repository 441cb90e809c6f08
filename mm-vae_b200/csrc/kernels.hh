// Internal launcher interface between the C-ABI layer (model.cu) and the
// sm_100a kernels.  Plain structs of device pointers; every launcher returns 0
// or nonzero (message via mmv::get_error()).
#pragma once

#include <cuda_runtime.h>
#include <cstddef>
#include <cstdint>

namespace mmv {

// ---- K2 fp32 GEMM (gemm_f32.cu)
struct gemm_args_t {
    const float *A = nullptr, *B = nullptr;
    float *C = nullptr;
    const float *bias = nullptr;   // [N], added after alpha*AB + beta*C
    int M = 0, N = 0, K = 0;
    int lda = 0, ldb = 0, ldc = 0;
    int transA = 0, transB = 0;
    float alpha = 1.f, beta = 0.f;
    int relu = 0;
};
int launch_gemm_f32(const gemm_args_t &g, cudaStream_t stream);

// ---- K2 bf16 tcgen05 GEMM (gemm_tc.cu).  A, B are bf16 (device), each K-major or MN-major:
//   a_mn = 0: A stored [M x K] (lda) ; a_mn = 1: A stored [K x M] (lda)
//   b_mn = 0: B stored [N x K] (ldb) ; b_mn = 1: B stored [K x N] (ldb)
struct gemm_tc_args_t {
    const void *A = nullptr, *B = nullptr;
    // fp32-tolerance mode: residual copies (same layout) of both operands.  The kernel then walks K three times into one
    // accumulator — A B_lo, A_lo B, A B — i.e. A and B stand for hi + lo pairs (both null: the plain bf16 product)
    const void *A_lo = nullptr, *B_lo = nullptr;
    int a_mn = 0, b_mn = 0, lda = 0, ldb = 0;
    float *C = nullptr;            // fp32 output (may be nullptr if only Cb is wanted)
    void *Cb = nullptr;            // optional bf16 copy of the output
    int ldc = 0, ldcb = 0;
    const float *bias = nullptr;
    int M = 0, N = 0, K = 0;
    float beta = 0.f;
    int relu = 0;
    // optional: add sum(C^2) of the finished product to *sumsq (global-norm clip without a second pass);
    // *sumsq_done is set to 1 when the chosen kernel did it (persistent store epilogue), else 0
    double *sumsq = nullptr;
    int *sumsq_done = nullptr;
    // optional softmax statistics of the output rows (decoder-last forward feeding K4): per (256-column tile t, row r)
    // row_stats[t * M + r] = (max_j x_rj log2(e), sum_j 2^(x_rj log2(e) - max)) over the tile's columns, x including the bias.
    // *row_stats_tiles = number of column tiles when the chosen kernel wrote them (persistent store epilogue), else 0
    void *row_stats = nullptr;
    int *row_stats_tiles = nullptr;
};
bool gemm_tc_supported(const gemm_tc_args_t &g);
int launch_gemm_tc(const gemm_tc_args_t &g, cudaStream_t stream);
int launch_to_bf16(const float *x, void *y, size_t n, cudaStream_t st);
// x = hi + lo (two bf16 arrays): operands of the three-product fp32-tolerance GEMM ("dtype" 2)
int launch_split_bf16(const float *x, void *hi, void *lo, size_t n, cudaStream_t st);
// SWIZZLE_128B tiled map over a row-major bf16 matrix [outer x inner] (leading dimension ld elements); map = CUtensorMap*
int make_tma_map_bf16(void *map, const void *base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
                      uint32_t box_outer);
int make_tma_map_f32(void *map, const void *base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
                     uint32_t box_outer);

// ---- K1 sparse encoder (sparse_encoder.cu)
struct senc_args_t {
    const int64_t *indptr = nullptr;
    const int32_t *indices = nullptr;
    const float *vals = nullptr;
    const int32_t *cells = nullptr;   // [B] cell id per minibatch row (nullptr: identity)
    int B = 0, D = 0;
    int angular = 0;                  // 0: nb (Linear), 1: vmf (Angular + row L2 of log1p)
    const float *W1T = nullptr;       // [D x h1]
    int h1 = 0;
    const float *inv_t = nullptr;     // [D]
    const float *bias_eff = nullptr;  // [h1]
    const float *col_scale = nullptr; // [h1] (angular)
    int relu = 0;
    float *H1 = nullptr;              // [B x h1]
    const float *WnuT = nullptr;      // [D x H] (nb)
    const float *bnu = nullptr;       // [H]
    int H = 0;
    const float *wdepth = nullptr;    // [D]
    const float *bdepth = nullptr;    // [1]
    float *Hnu = nullptr;             // [B x H]
    float *dpre = nullptr;            // [B]
    float *rownorm = nullptr;         // [B] (angular)
};
int launch_senc_fwd(const senc_args_t &a, cudaStream_t st);
int launch_senc_bias(const float *W1T, const float *x_mean, const float *inv_t, const float *col_scale,
                     float *bias_eff, int D, int h1, int angular, cudaStream_t st);
// the same from the bf16 effective weight [D x h1] of the tensor-core path (bias_eff pre-initialised, -= the mat-vec)
bool senc_bias_bf16_supported(int h1);
int launch_senc_bias_bf16(const void *W1b, const float *x_mean, const float *inv_t, float *bias_eff, int D, int h1,
                          cudaStream_t st);

struct batch_csc_args_t {
    const int64_t *indptr = nullptr;
    const int32_t *indices = nullptr;
    const float *vals = nullptr;
    const int32_t *cells = nullptr;
    const float *rownorm = nullptr;   // vmf: entries are log1p(x) / ||log1p(x_i)||
    int B = 0, D = 0;
    int *counts = nullptr, *gene_ptr = nullptr, *cursor = nullptr;   // [D], [D+1], [D]
    int *ent_row = nullptr;
    float *ent_l = nullptr, *ent_x = nullptr;
};
int launch_batch_csc(const batch_csc_args_t &c, cudaStream_t st);

struct senc_bwd_args_t {
    const int *gene_ptr = nullptr, *ent_row = nullptr;
    const float *ent_l = nullptr, *ent_x = nullptr;
    int D = 0, h1 = 0, H = 0;
    int angular = 0, accumulate = 1;
    const float *delta = nullptr;     // [B x h1] dL/d(first-layer pre-activation)
    const float *Delta = nullptr;     // [h1] column sums of delta
    const float *W1T = nullptr, *col_scale = nullptr;
    const float *inv_t = nullptr, *x_mean = nullptr, *ln_x_sd = nullptr;
    float *dW1T = nullptr;            // [D x h1] gradient w.r.t. the EFFECTIVE weight
    float *dx_mean = nullptr, *dln_x_sd = nullptr;
    const float *dHnu = nullptr;      // [B x H]
    const float *ddpre = nullptr;     // [B]
    float *dWnuT = nullptr;           // [D x H]
    float *dwdepth = nullptr;         // [D]
    bool aux_done = false;            // tensor-core path: dWnuT / dwdepth were already issued (side stream, model.cu)
};
int launch_senc_bwd(const senc_bwd_args_t &a, cudaStream_t st);

// ---- K1 tensor-core path (senc_tc.cu): per-minibatch tiled-COO "panel" + densify-in-smem tcgen05 GEMMs
// Panel: the minibatch's nonzeros bucketed by (tile of 64 batch rows, chunk of 64 genes); bucket (T, k) is
// the run [bucket_ptr[T * (nK + 1) + k], bucket_ptr[T * (nK + 1) + k + 1]) of `ent` / `xraw`.
// ent = row_in_tile << 26 | gene_in_chunk << 20 | bf16(log1p(x) [/ ||log1p x_i||] / t_gene).
struct panel_args_t {
    const int64_t *indptr = nullptr;
    const int32_t *indices = nullptr;
    const float *vals = nullptr;
    const int32_t *cells = nullptr;
    int B = 0, D = 0, angular = 0;
    const int *unsorted = nullptr;                     // device flag: 0 = every CSR row has ascending gene ids
    int num_sms = 0;
    const float *inv_t = nullptr;
    const float *WnuT = nullptr, *bnu = nullptr;       // nb: nu encoder + depth on raw counts (nullable)
    int H = 0;
    const float *wdepth = nullptr, *bdepth = nullptr;
    float *Hnu = nullptr, *dpre = nullptr, *rownorm = nullptr;
    int *rowptr = nullptr;                             // [B + 1] workspace
    int *bucket_ptr = nullptr;                         // [nT * (nK + 1)]
    uint32_t *ent = nullptr;                           // [nnz of the minibatch]
    uint32_t *ent_lo = nullptr;                        // optional ("dtype" 2): same coordinates, value = what bf16 rounding of ent's value lost
    float *xraw = nullptr;                             // [nnz of the minibatch] raw counts, same order
};
inline int panel_tiles(int B) { return (B + 63) / 64; }
inline int panel_chunks(int D) { return (D + 63) / 64; }
int launch_panel_build(const panel_args_t &p, cudaStream_t st);

struct senc_tc_args_t {
    const int *bucket_ptr = nullptr;
    const uint32_t *ent = nullptr;
    int B = 0, D = 0, h1 = 0;
    const void *dense_bf16 = nullptr;   // fwd: effective W1^T [D x h1] bf16 ; bwd: delta [B x h1] bf16
    // fp32-tolerance mode ("dtype" 2): residual copies of both operands.  The kernel then walks its K axis three times —
    // (ent, dense_lo), (ent_lo, dense), (ent, dense) — into the same TMEM accumulator: A B ~ A_hi B_lo + A_lo B_hi + A_hi B_hi
    const uint32_t *ent_lo = nullptr;
    const void *dense_lo = nullptr;
    // forward epilogue: H1 = acc + bias_eff (ReLU optional), fp32 + bf16 copies
    const float *bias_eff = nullptr;
    int relu = 0;
    float *H1 = nullptr;
    void *H1_bf16 = nullptr;
    // backward epilogue (see senc_bwd_args_t for the formulas)
    const float *W1T = nullptr, *col_scale = nullptr, *Delta = nullptr, *inv_t = nullptr, *x_mean = nullptr;
    int angular = 0, accumulate = 1;
    float *dW1T = nullptr;
    float *S = nullptr;                 // [D x h1] raw product A^T delta (may alias dW1T when accumulate = 0)
    double *sumsq = nullptr;            // row pass: += sum(dW1T^2) (accumulate = 0 only)
    int tile0 = 0, ntiles = 0;          // backward + row pass: restrict to 128-gene tiles [tile0, tile0 + ntiles) (0 = all)
};
bool senc_tc_supported(int h1);
int launch_senc_tc_fwd(const senc_tc_args_t &a, cudaStream_t st);
int launch_senc_tc_bwd(const senc_tc_args_t &a, cudaStream_t st);
// row pass after launch_senc_tc_bwd: rank-1 correction of dW1, d x_mean, d ln_x_sd (formulas in senc_tc.cu)
int launch_senc_row_pass(const senc_tc_args_t &a, const float *ln_x_sd, float *dx_mean, float *dln_x_sd, cudaStream_t st);
// dWnu / dw_depth from the panel's raw counts
int launch_panel_aux_bwd(const int *bucket_ptr, const uint32_t *ent, const float *xraw, int B, int D, int H,
                         const float *dHnu, const float *ddpre, float *dWnuT, float *dwdepth, cudaStream_t st);
int launch_w1eff_bf16(const float *W1T, const float *col_scale, void *out, int D, int h1, cudaStream_t st);

// ---- K4 NB likelihood (nb_likelihood.cu)
struct nb_lik_args_t {
    const float *G = nullptr;         // [B x D] last decoder layer output WITHOUT bias (tuned kernel: the complete logits)
    const float *abias = nullptr;     // [D] mu_decoding.bias + covar_decoding.bias + mu_bias (tuned kernel: nullptr, it is in G)
    const void *rowmeta = nullptr;    // tuned kernel: float4[B] (max log2(e), 1 / sum exp2, depth, znu) from launch_nb_row_meta,
                                      // or a workspace the launcher fills itself when tile_stats is given
    const void *tile_stats = nullptr; // tuned kernel: the decoder-last GEMM's [n_tiles x B] softmax partials (gemm_tc_args_t::row_stats)
    int n_tiles = 0;
    const float *Wcd = nullptr;       // [D x C] covar_decoding.weight
    const float *cov = nullptr;       // [n_cells x C] covariates (nullptr: ones)
    const int64_t *indptr = nullptr;
    const int32_t *indices = nullptr;
    const float *vals = nullptr;
    const int32_t *cells = nullptr;
    const float *dpre = nullptr;      // [B] depth pre-activation
    const float *znu = nullptr;       // [B x Rnu]
    const float *Wnd = nullptr;       // [D x Rnu] nu_decoding.weight
    const float *ubias = nullptr;     // [D] nu_decoding.bias - nu_bias
    int B = 0, D = 0, C = 0, Rnu = 0;
    float inv_B = 0.f;
    float *dlogits = nullptr;         // [B x D]  (nullptr: forward only unless dlogits_bf16)
    void *dlogits_bf16 = nullptr;     // [B x D] bf16 instead of fp32 (tensor-core mode)
    void *dlogits_lo = nullptr;       // tuned kernel, with dlogits_bf16 ("dtype" 2): the bf16 residual, gradient = bf16 + lo
    float *ddpre = nullptr;           // [B]
    float *dznu = nullptr;            // [B x Rnu]
    float *col_du = nullptr;          // [D]       += sum_i du_ij
    float *col_duz = nullptr;         // [D x Rnu] += sum_i du_ij znu_ir
    double *loss_acc = nullptr;       // += sum_ij nllik
    float *recon_mu = nullptr, *recon_nu = nullptr;   // optional [B x D] (tests)
    float *row_lse = nullptr;         // optional [B]
    int allow_fast = 1;               // use the tuned Rnu = 1 / all-ones-covariate kernel when eligible
    const int *unsorted = nullptr;    // device flag (mmv_csr::d_unsorted); nullptr = treat rows as unsorted
    int64_t *slice_bounds = nullptr;  // workspace [B x 9] for the tuned kernel (per-row CSR offsets of the gene slices)
};
bool nb_likelihood_fast_eligible(const nb_lik_args_t &a);
// per-row metadata of the tuned kernel from the GEMM epilogue's [n_tiles x B] partials (tile_stats) or, without them, from
// one pass over the complete logits G
int launch_nb_row_meta(const float *G, const void *tile_stats, int n_tiles, const float *dpre, const float *znu, int B, int D,
                       void *meta, cudaStream_t stream);
int launch_nb_likelihood_fast(const nb_lik_args_t &a, int num_sms, cudaStream_t stream);
int launch_nb_likelihood(const nb_lik_args_t &a, int num_sms, cudaStream_t stream);

// ---- K5 vMF likelihood (vmf_likelihood.cu)
struct vmf_lik_args_t {
    const float *G = nullptr;         // [B x D] last decoder layer output WITHOUT bias
    const float *blast = nullptr;     // [D] decoding.bias
    const float *Wcd = nullptr, *bcd = nullptr;   // covar_decoding_ [D x C], [D]
    const float *cov = nullptr;
    const int64_t *indptr = nullptr;
    const int32_t *indices = nullptr;
    const float *vals = nullptr;
    const int32_t *cells = nullptr;
    const float *kappa = nullptr;     // device scalar
    int B = 0, D = 0, C = 0;
    float inv_B = 0.f, eps_obs = 0.f;
    float *dpre = nullptr;            // [B x D] (nullptr: forward only)
    void *dpre_bf16 = nullptr;        // tuned kernel only: the gradient as bf16 instead (feeds the tensor-core GEMMs)
    const int *unsorted = nullptr;    // device flag: 0 = every CSR row has ascending gene ids
    int64_t *slice_bounds = nullptr;  // [B x 9] workspace of the tuned kernel
    int allow_fast = 1;
    float *col_dq = nullptr;          // [D]
    float *col_dqc = nullptr;         // [D x C]
    double *t_acc = nullptr;          // += sum_i <yobs_i, recon_i>
    float *recon = nullptr;           // optional [B x D]
    float *row_t = nullptr;           // optional [B]
};
int launch_vmf_likelihood(const vmf_lik_args_t &a, int num_sms, cudaStream_t stream);
bool vmf_likelihood_fast_eligible(const vmf_lik_args_t &a);
int launch_vmf_likelihood_fast(const vmf_lik_args_t &a, int num_sms, cudaStream_t stream);

struct vmf_scalar_args_t {
    const float *kappa = nullptr, *ln_kappa = nullptr;
    const double *t_acc = nullptr;
    double *llik_sum = nullptr;
    float *dln_kappa = nullptr;
    int B = 0;                        // cells of THIS rank
    float inv_B_global = 0.f, inv_ranks = 1.f;
    float df = 0.f, eta = 0.f, kmin = 0.f, kmax = 0.f;
    float c_eta_nu_log2 = 0.f, c_fasterlgamma = 0.f, c_half_log_2pi = 0.f, c_half_D_fasterlog_2pi = 0.f;
};
int launch_vmf_kappa(const float *ln_kappa, float kmin, float kmax, float *kappa, cudaStream_t st);
int launch_vmf_scalar(const vmf_scalar_args_t &s, cudaStream_t st);

// ---- K3 + helpers (elementwise.cu)
int launch_gene_prep(const float *ln_x_sd, float eps, float *inv_t, int D, cudaStream_t st);
int launch_vec_combine(const float *a, const float *b, const float *c, float sb, float sc, float *out, int n,
                       cudaStream_t st);
int launch_nb_gene_prologue(const float *ln_x_sd, float eps, float *inv_t, const float *blast, const float *bcd,
                            const float *mu_bias, const float *Wcd, int C, const float *bnd, const float *nu_bias, float *abias,
                            float *ubias, int D, cudaStream_t st);
int launch_nb_gene_bias(const float *blast, const float *bcd, const float *mu_bias, const float *Wcd, int C, const float *bnd,
                        const float *nu_bias, float *abias, float *ubias, int D, cudaStream_t st);

struct reparam_args_t {
    const float *head = nullptr;      // [B x 2R] mean_raw | lnvar_raw
    const float *Wce = nullptr, *bce = nullptr;   // covar_encoding [R x C], [R] (nullable)
    const float *cov = nullptr;
    const int32_t *cells = nullptr;
    const float *eps = nullptr;       // [B x R] injected noise (nullptr: z = mean)
    float *eps_gen = nullptr;         // forward: draw the noise here (normal_draw(gen_key, gen_step, gen_off + idx)), store it to
    uint64_t gen_key = 0, gen_step = 0, gen_off = 0;   // eps_gen[idx] (== eps) and use it
    int B = 0, R = 0, C = 0;
    float *mean = nullptr, *lnvar = nullptr, *z = nullptr;   // [B x R]
    void *z_bf16 = nullptr;           // optional bf16 copy of z (feeds the decoder's first tensor-core GEMM)
    double *kl_acc = nullptr;
    // backward
    const float *dz = nullptr;
    float *dhead = nullptr;
    void *dhead_bf16 = nullptr;       // optional bf16 copy of dhead
    float kl_weight = 1.f, inv_B = 0.f;
};
int launch_reparam_fwd(const reparam_args_t &a, cudaStream_t st);
int launch_reparam_bwd(const reparam_args_t &a, cudaStream_t st);

// nu heads fused with their reparameterisation (elementwise.cu): head = Hnu W^T + b, W [2 Rnu x H]
struct nu_head_args_t {
    const float *Hnu = nullptr, *W = nullptr, *b = nullptr;
    const float *eps = nullptr;
    float *eps_gen = nullptr;         // as in reparam_args_t
    uint64_t gen_key = 0, gen_step = 0, gen_off = 0;
    int B = 0, H = 0, Rnu = 0;
    float *head = nullptr, *mean = nullptr, *lnvar = nullptr, *z = nullptr;
    double *kl_acc = nullptr;
    // backward
    const float *dz = nullptr, *ddpre = nullptr;
    float kl_weight = 1.f, inv_B = 0.f;
    float *dHnu = nullptr;            // [B x H] written
    float *dW = nullptr, *db = nullptr, *dbnu = nullptr, *ddepth_b = nullptr;   // += (gradient buffers)
};
bool nu_head_fused_supported(int H, int Rnu);
int launch_nu_head(const nu_head_args_t &a, bool bwd, cudaStream_t st);
// the mu reparam (forward or backward) and the nu heads in ONE launch (they are independent and adjacent in the step)
int launch_latent(const reparam_args_t &a, const nu_head_args_t &n, bool bwd, cudaStream_t st);

// ---- true vMF latent (vmf_latent.cu; SURVEY 8 f4, opt-in, no reference counterpart): same head / outputs as
// the Gaussian block above — mean = unit direction, lnvar = -log kappa broadcast, z on the sphere, kl_acc += KL
struct vmf_latent_args_t {
    reparam_args_t r;
    uint64_t seed = 0, draw = 0;      // Philox key: any (seed, draw, row) is reproducible on its own
    float kappa_max = 1e4f;
    double lgamma_half_d = 0.;        // lgamma(R / 2)
    float *w = nullptr;               // [B] sampled cosine <z, mu>: written forward, read backward
    int *attempts = nullptr;          // [B] proposals used (nullable)
    double *kl_aux = nullptr;         // [B x 2] A_d(kappa) and dKL/dkappa: written forward, read backward
};
int launch_vmf_latent_fwd(const vmf_latent_args_t &a, cudaStream_t st);
int launch_vmf_latent_bwd(const vmf_latent_args_t &a, cudaStream_t st);

struct colsum_args_t {
    const float *X = nullptr;
    const void *Xb = nullptr;         // bf16 input instead of X
    const void *Xb2 = nullptr;        // with Xb (bf16x8 kernel only): a second bf16 array added element-wise (hi + lo pair)
    int M = 0, N = 0, ldx = 0;
    float *out_sum = nullptr, *out_sum2 = nullptr, *out_sum3 = nullptr;   // [N] each, +=
    float *out_cov = nullptr;         // [N x C] += sum_i X[i, col] * cov[cell_i, c]
    const float *cov = nullptr;
    const int32_t *cells = nullptr;
    int C = 0;
    int rows_per_cta = 64;
    float scale2 = 1.f;               // out_sum2 += scale2 * sum (generic kernel only)
};
int launch_colsum(colsum_args_t a, cudaStream_t st);
// up to kColsumMaxSeg independent fp32 column sums per launch (generic kernel per segment; Xb segments are not accepted)
constexpr int kColsumMaxSeg = 12;
struct colsum_multi_args_t {
    colsum_args_t seg[kColsumMaxSeg];
    int first[kColsumMaxSeg] = {};    // first block of each segment
    int gx[kColsumMaxSeg] = {};       // blocks along the columns
    int n = 0;
};
int launch_colsum_multi(const colsum_args_t *segs, int n, cudaStream_t st);
int launch_relu_bwd(float *dX, const float *Y, size_t n, cudaStream_t st);

// ---- Angular (angular.cu)
int launch_angular_scale(const float *W, float *scale, int units, int inputs, size_t us, size_t is, cudaStream_t st);
int launch_angular_materialize(const float *W, const float *scale, float *Wp, int units, int inputs, cudaStream_t st);
int launch_angular_bwd(const float *W, const float *dWp, const float *scale, float *dot, float *dW,
                       int units, int inputs, size_t us, size_t is, cudaStream_t st);

// ---- K6 (optim.cu)
int launch_sumsq(const float *g, size_t n, double *out, int num_sms, cudaStream_t st);
// the same over several ranges in one launch (four per launch); *out += the total
// pre / npre: arrays of already-squared partial sums added as they are (fixed order)
int launch_sumsq_ranges(const float *const *g, const size_t *n, int count, double *out, int num_sms, cudaStream_t st,
                        const float *const *pre = nullptr, const int *npre = nullptr, int pre_count = 0);
struct adam_args_t {
    float *param = nullptr, *grad = nullptr, *m = nullptr, *v = nullptr;
    void *param_bf16 = nullptr;       // optional bf16 mirror refreshed in the same pass
    void *param_lo = nullptr;         // optional bf16 residual param - float(bf16(param)) ("dtype" 2), same pass
    size_t n = 0;
    const double *sumsq = nullptr;
    float *norm_out = nullptr;
    float max_norm = 1.f, weight_decay = 1e-4f, beta1 = 0.9f, beta2 = 0.999f, eps = 1e-8f;
    // 1 - beta rounded from DOUBLE like libtorch's Scalar(1 - beta) (1.f - 0.999f is off by 1.3e-5)
    float one_minus_beta1 = (float)(1.0 - 0.9), one_minus_beta2 = (float)(1.0 - 0.999);
    float bc2_sqrt = 1.f, neg_step = 0.f;
    int write_grad = 1;               // 0: leave .grad unscaled; the caller applies *coef_out when it is read
    float *coef_out = nullptr;        // clip coefficient min(1, max_norm / (||g|| + 1e-6))
};
int launch_adam(const adam_args_t &a, int num_sms, cudaStream_t st);
int launch_scale_by(float *x, size_t n, const float *coef, cudaStream_t st);

} // namespace mmv
