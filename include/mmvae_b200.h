/* mmvae_b200.h — C ABI of the B200-native mm-vae minibatch ELBO step.
 *
 * The reference (YPARK/mm-vae) has no FFI boundary: its seams are duck-typed
 * C++ template concepts inside one translation unit.  This
 * header is the boundary a maintainer binds instead of libtorch for the hot
 * path; each entry cites the reference code it replaces (paths relative to
 * the reference root).  Conventions: every function returns 0 (EXIT_SUCCESS)
 * or nonzero and never throws; the message is available from mmv_last_error();
 * handles are opaque and owned by the library; host pointers are borrowed for
 * the duration of the call; a ctx is not thread-safe, different ctxs are
 * independent; all device work of a ctx is issued on one CUDA stream and is
 * asynchronous unless stated.  There is no CPU fallback: without a CUDA device
 * mmv_ctx_create fails.
 */
#ifndef MMVAE_B200_H_
#define MMVAE_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define MMV_API __attribute__((visibility("default")))
#else
#define MMV_API
#endif

#define MMV_MAX_LAYERS 8
#define MMV_NAME_LEN 96

typedef struct mmv_ctx mmv_ctx;
typedef struct mmv_csr mmv_csr;
typedef struct mmv_model mmv_model;

/* ---- device context ------------------------------------------------------
 * replaces: training_options_t::device + model->to(device), mmvae_alg.hh:17,237 */
/* Number of CUDA devices this process can see (the multi-GPU mains shard the cells over them; the reference is
 * single device: training_options_t::device, mmvae_alg.hh:16-17). */
MMV_API int mmv_device_count(int *out);
MMV_API int mmv_ctx_create(int device, mmv_ctx **out);
MMV_API int mmv_ctx_destroy(mmv_ctx *ctx);
MMV_API int mmv_ctx_sync(mmv_ctx *ctx);
MMV_API const char *mmv_last_error(void);
/* CUDA-event timer on the ctx stream (bench/roofline measurement). */
MMV_API int mmv_timer_start(mmv_ctx *ctx);
MMV_API int mmv_timer_stop(mmv_ctx *ctx, float *elapsed_ms);   /* synchronises */
MMV_API int mmv_ctx_num_sms(mmv_ctx *ctx, int *out);
/* Number of kernels this library has launched in this process (bench accounting). */
MMV_API int mmv_launch_count(int64_t *out);

/* ---- resident sparse data: cells x genes CSR (+ dense covariates) ---------
 * replaces: mtx_data_block_t::read/torch_tensor/clear, mmvae_io.hh:81-88,208-256
 * (sparse text -> dense B x D per batch) and `.to(device)`, mmvae_alg.hh:271-272.
 * indptr[n_cells+1], indices[nnz] (gene ids, unique within a cell), values[nnz].
 * covar: n_cells x c_dim row-major, or NULL for the all-ones single covariate the
 * reference generates when --covar is absent (nb_vae_main.cc:68-73).           */
MMV_API int mmv_csr_upload(mmv_ctx *ctx, int64_t n_cells, int64_t n_genes, const int64_t *indptr,
                   const int32_t *indices, const float *values, const float *covar, int64_t c_dim,
                   mmv_csr **out);
/* Reusable slot for streaming minibatches from pinned host memory (end-to-end
 * path): create once with capacities, then mmv_csr_update copies a new
 * minibatch asynchronously on the ctx stream. */
MMV_API int mmv_csr_create(mmv_ctx *ctx, int64_t max_cells, int64_t max_nnz, int64_t n_genes, int64_t c_dim,
                   mmv_csr **out);
MMV_API int mmv_csr_update(mmv_csr *csr, int64_t n_cells, const int64_t *indptr, const int32_t *indices,
                   const float *values, const float *covar);
/* Zero-copy: adopt CSR arrays that already live on the ctx's device (e.g. a shard
 * generated or parsed there).  The arrays are BORROWED until mmv_csr_destroy.
 * max_row_nnz = longest cell (sizes the per-step transpose workspace). */
MMV_API int mmv_csr_wrap_device(mmv_ctx *ctx, int64_t n_cells, int64_t n_genes, int64_t nnz, int64_t max_row_nnz,
                        const int64_t *d_indptr, const int32_t *d_indices, const float *d_values,
                        const float *d_covar, int64_t c_dim, mmv_csr **out);
MMV_API int mmv_csr_destroy(mmv_csr *csr);
/* Pinned host allocation helpers for the streaming path. */
MMV_API int mmv_host_alloc(void **out, int64_t bytes);
MMV_API int mmv_host_free(void *p);

/* ---- models ---------------------------------------------------------------
 * replaces: nbvae_tImpl ctor nb.hh:299-401 (args nb.hh:214-221) and
 *           vmf_vae_tImpl ctor vmf.hh:307-389 (args vmf.hh:200-207).
 * Parameters are initialised like libtorch (kaiming-uniform(a=sqrt 5) weights,
 * U(-1/sqrt(fan_in), +) biases; x_mean 0, ln_x_sd 1, mu_bias/nu_bias 0,
 * ln_kappa log(kappa_min)) from `seed`.                                        */
typedef struct {
    int64_t x_dim, c_dim;
    int32_t n_enc;
    int64_t enc_dims[MMV_MAX_LAYERS];   /* --mean_encoding */
    int32_t n_dec;
    int64_t dec_dims[MMV_MAX_LAYERS];   /* --mean_decoding */
    int64_t mu_r_dim;                   /* --mean_latent */
    int64_t nu_h_dim;                   /* --overdisp_encoding */
    int64_t nu_r_dim;                   /* --overdisp_latent */
    int32_t do_relu;
} mmv_nb_dims;

typedef struct {
    int64_t x_dim, c_dim;
    int64_t z_dim;                      /* --latent */
    int32_t n_enc;
    int64_t enc_dims[MMV_MAX_LAYERS];   /* --encoding */
    int32_t n_dec;
    int64_t dec_dims[MMV_MAX_LAYERS];   /* --decoding */
    float kappa_min, kappa_max;
    int32_t do_relu;
} mmv_vmf_dims;

MMV_API int mmv_nb_model_create(mmv_ctx *ctx, const mmv_nb_dims *dims, uint64_t seed, mmv_model **out);
MMV_API int mmv_vmf_model_create(mmv_ctx *ctx, const mmv_vmf_dims *dims, uint64_t seed, mmv_model **out);
MMV_API int mmv_model_destroy(mmv_model *m);

/* Options.  key/value (doubles):
 *   "lr" (1e-3) "weight_decay" (1e-4) "grad_clip" (1.0)      mmvae_alg.hh:19-23,236
 *   "train_all_layers" (0): 0 = reference behaviour — the encoder/decoder
 *        Sequentials are never register_module'd (nb.hh:335-379, vmf.hh:341-385)
 *        so only the registered tensors are clipped/updated and the others'
 *        grads accumulate; 1 = optimise every tensor.
 *   "frozen_grads" (1): compute the gradients of the non-optimised tensors
 *        (what the reference's autograd does); 0 skips that dead work.
 *   "max_batch" : capacity hint (cells per step), grows on demand.
 *   "dtype" (0): 0 = fp32 FFMA GEMMs + gather first layer (1e-4 parity mode); 1 = bf16 operands on the tcgen05 tensor
 *        cores, fp32 accumulation (1e-2); 2 = the fp32 tolerance ON the tensor cores: every GEMM operand as a bf16
 *        pair hi + lo, three products per GEMM into one accumulator (1e-4; no reference counterpart).
 *   "debug_outputs" (0): keep recon_mu / recon_nu / recon for mmv_get_output.
 *   "comm_bf16" (1): data-parallel runs in bf16 mode ("dtype" 1) send the two D x h weight gradients over
 *        NVLink as bf16 (per-rank partial sums rounded once, summed by ncclAllReduce(bf16)); 0 or fp32 mode:
 *        every gradient is reduced in fp32.  No reference counterpart (the reference is single device).
 *   "side_stream" (1; the environment's MMV_SIDE overrides the default): work of a backward pass that only the
 *        optimiser reads (per-gene sums of dlogits, the hidden layers' weight-gradient GEMMs, bias-gradient column
 *        sums, dW_nu / dw_depth) and the first layer's effective bias run on a second, lowest-priority stream of
 *        the context next to the step's dependency chain; mmv_step joins it before the optimiser and before it
 *        returns, so nothing is in flight behind the caller's back.  0 = everything on the context's one stream.
 *        Same results either way (the sums that cross streams are atomic adds into distinct gradient tensors).
 *   "nb_fast" (1), "profile" (0): the tuned K4 / K5 kernels when eligible; per-kernel-group CUDA-event times
 *        (mmv_profile_read).
 *   "vmf_latent" (0): vmf models only, z_dim >= 3.  NOT in the reference (its vmf_vae reparameterises a
 *        Gaussian, vmf.hh:394-404, and uses kl_loss_normal, vmf.hh:410-414); this is the latent the north star
 *        names.  1 = the stacked head is read as a von Mises-Fisher posterior instead: mean direction
 *        mu = normalise(mean head + covar_enc(c)), concentration kappa = min(1 + softplus(mean_r of the second
 *        head), "vmf_kappa_max" (1e4)); z ~ vMF(mu, kappa) by Wood's rejection sampler (Philox keyed on the model
 *        seed, the draw counter "vmf_latent_draw" and the row; the cell's eps[1:] gives the tangent direction) and
 *        a Householder reflection; kl_mu = sum_i KL(vMF(mu_i, kappa_i) || uniform sphere); gradients flow
 *        through the reflection and, for kappa, the implicit-reparameterisation derivative of the sampled cosine.
 *        mmv_encode / the recorder then return mu (unit vectors) and -log kappa. */
MMV_API int mmv_model_set_option(mmv_model *m, const char *key, double value);

/* Parameters by the reference's named_parameters() keys ("x_mean",
 * "mu_encoding_1.weight", "covar_decoding.bias", ...; recorder file names
 * nb.hh:599-615 use the same keys), in libtorch layout ([out x in] weights). */
MMV_API int mmv_model_num_params(mmv_model *m, int *out);
MMV_API int mmv_model_param_info(mmv_model *m, int i, char name[MMV_NAME_LEN], int64_t *rows, int64_t *cols,
                         int *registered);
MMV_API int mmv_model_set_param(mmv_model *m, const char *name, const float *host, int64_t n);
MMV_API int mmv_model_get_param(mmv_model *m, const char *name, float *host, int64_t n);
MMV_API int mmv_model_get_grad(mmv_model *m, const char *name, float *host, int64_t n);

/* ---- the step ---------------------------------------------------------------
 * replaces: model->forward + loss_fun + zero_grad + backward + clip_grad_norm_ +
 * adam.step, mmvae_alg.hh:274-311, and the bootstrap index_select :292-301.
 *   cells[B]     : cell (CSR row) ids of the minibatch, mmvae_alg.hh:264-266
 *   boot_idx[B]  : NULL, or positions into cells[] (sampling with replacement):
 *                  the step runs on cells[boot_idx[r]]
 *   eps          : NULL -> z = mean (eval mode, nb.hh:469-471); otherwise injected
 *                  N(0,1) noise, nb: [B x mu_r_dim] then [B x nu_r_dim]; vmf: [B x z_dim]
 *                  (host pointer). MMV_EPS_PHILOX in flags draws it on device instead.
 *   kl_weight    : max(kl_max * exp(-kl_discount * epoch), kl_min), nb_vae_main.cc:27-32
 *   flags        : MMV_FWD | MMV_BWD | MMV_OPTIM
 *   out          : loss terms, written when the call returns (synchronises) unless NULL. */
enum { MMV_FWD = 1, MMV_BWD = 2, MMV_OPTIM = 4, MMV_EPS_PHILOX = 8, MMV_TERMS_DEFERRED = 16 };

typedef struct {
    float loss;       /* nb.hh:539-548 / vmf.hh:439: (recon + w*KL...) / B        */
    float recon;      /* nb: sum nllik (nb.hh:511-531); vmf: sum_i llik_i (vmf.hh:435-437) */
    float kl_mu;      /* nb: KL(mu_mean, mu_lnvar); vmf: KL(mean, lnvar)         */
    float kl_nu;      /* nb: KL(nu_mean, nu_lnvar); vmf: 0                       */
    float grad_norm;  /* total norm before clipping (MMV_OPTIM only)            */
} mmv_loss_terms;

MMV_API int mmv_step(mmv_model *m, mmv_csr *csr, const int64_t *cells, int64_t B, const int64_t *boot_idx,
             const float *eps, float kl_weight, int flags, mmv_loss_terms *out);

/* MMV_TERMS_DEFERRED (with out != NULL): the call does not wait for the device.  The step's sums are copied to
 * pinned memory behind an event and `out` receives the terms of the PREVIOUS deferred step (loss = NaN if none):
 * the epoch loss of mmvae_alg.hh:281,324 is a sum, so the host loop can stay one step ahead of the device instead
 * of paying the `.item()` round trip per batch.  mmv_step_collect waits for and returns the last pending terms. */
MMV_API int mmv_step_collect(mmv_model *m, mmv_loss_terms *out);

/* Per-kernel-group device times (CUDA events on the ctx stream), enabled with
 * mmv_model_set_option(m, "profile", 1).  mmv_profile_read synchronises, returns the
 * time and launch-group count accumulated for `name` since the last reset
 * ("" resets everything).  Names: see DESIGN.md / bench.py. */
MMV_API int mmv_profile_read(mmv_model *m, const char *name, float *total_ms, int *count);

/* Forward outputs of the last step by name (tests / recorder): nb: "mu_mean",
 * "mu_lnvar", "nu_mean", "nu_lnvar", "recon_depth" and, with debug_outputs,
 * "recon_mu", "recon_nu"; vmf: "mean", "lnvar", "kappa", "recon". */
MMV_API int mmv_get_output(mmv_model *m, const char *name, float *host, int64_t n);

/* Recorder path: eval-mode encoder WITHOUT covariates.
 * replaces: encode_mu(x) nb.hh:419-431,627-628 ; encode(x) vmf.hh:267-281,517 */
MMV_API int mmv_encode(mmv_model *m, mmv_csr *csr, const int64_t *cells, int64_t B, float *mean, float *lnvar);

/* ---- kernel-level entry for the unit tests --------------------------------------
 * C[M x N] = op(A) op(B) (+ bias) (+ beta C) (ReLU) with host fp32 operands, on the
 * bf16 tcgen05 path (use_tc = 1; operands rounded to bf16 on the device) or the fp32
 * FFMA path (use_tc = 0).  a_mn / b_mn: operand stored [K x M] / [K x N] (MN-major)
 * instead of [M x K] / [N x K] (K-major).  C is in/out. */
MMV_API int mmv_test_gemm(mmv_ctx *ctx, int use_tc, int M, int N, int K, int a_mn, int b_mn, const float *A,
                  const float *B, const float *bias, float beta, int relu, float *C);

/* ---- the same tensor-core GEMM for callers that cannot allocate device memory ------
 * No counterpart in the reference: added (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  C[M x N] = op(A) op(B) (+ bias) (+ beta C) (ReLU) on the
 * bf16 tcgen05 path, as mmv_test_gemm(use_tc = 1), but d_A / d_B / d_bias / d_C are DEVICE pointers owned by the
 * caller and the bf16 operand copies go to a static __device__ array of MMV_NOALLOC_SCRATCH_BYTES inside the library
 * (held by every GPU that loads it): the call makes no cudaMalloc / cudaFree / allocating library call and copies
 * nothing from or to the host.  Limit: round_up(2 M K, 256) + 2 N K <= MMV_NOALLOC_SCRATCH_BYTES, else an error.
 * Leading dimensions (K, or M / N for MN-major operands) must be multiples of 8, the pointers 16-byte aligned.
 * The launch is asynchronous on the context's stream (mmv_ctx_sync waits for it).  All calls on one GPU share the
 * scratch and it keeps the last call's contents (every call overwrites what it reads; nothing must start at zero):
 * calls that overlap on the GPU — two contexts or streams on one device — race and must be serialised by the caller.
 * The first call on a device also does the one-time set-up (scratch address, kernel attributes). */
#define MMV_NOALLOC_SCRATCH_BYTES (4 << 20)
MMV_API int mmv_gemm_bf16_noalloc(mmv_ctx *ctx, int M, int N, int K, int a_mn, int b_mn, const float *d_A,
                  const float *d_B, const float *d_bias, float beta, int relu, float *d_C);

/* ---- data-parallel gradient exchange (no reference counterpart) ----
 * One rank per GPU: separate processes (bench.py under torchrun) or the host threads of one process
 * (nb_vae / vmf_vae --ngpu G), each with its own mmv_ctx.  Rank 0 creates the id, the host program hands
 * the 128 bytes to every rank, every rank calls mmv_comm_init (concurrently: the call is collective).  After it, mmv_step sum-reduces the
 * flat gradient buffer (and the loss terms) over ranks with ncclAllReduce before
 * clip + Adam; each rank must scale by its share (the step divides by the GLOBAL
 * batch = B * nranks). */
MMV_API int mmv_comm_unique_id(unsigned char id[128]);
MMV_API int mmv_comm_init(mmv_ctx *ctx, int nranks, int rank, const unsigned char id[128]);
MMV_API int mmv_comm_destroy(mmv_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* MMVAE_B200_H_ */
