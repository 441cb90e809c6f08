/* TEST DOUBLE, not a CPU fallback and never shipped: a recording stand-in for libmmvae_b200.so that the CPU tests
 * put in front of the real library (LD_LIBRARY_PATH, the mains use RUNPATH) to check the HOST logic of
 * mm-vae_b200/host without a GPU — which C-ABI calls train_vae_model makes, in which order, with which flags,
 * KL weights and bootstrap indices (mmvae_alg.hh:254-334 of the reference).  It computes nothing: every call is
 * written as one line to $MMV_STUB_LOG and the loss terms are a fixed function of the call count.
 * Built by tests/test_host_schedule_cpu.py into a temporary directory. */
#include "mmvae_b200.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

struct mmv_ctx { int device, nranks, rank; };
struct mmv_csr { int64_t n_cells, n_genes, nnz, c_dim; int has_cov; mmv_ctx *ctx; };
struct mmv_model { int is_vmf; int64_t R; int pending; float pending_loss; int64_t evals; };

static FILE *logf_(void)
{
    static FILE *f = NULL;
    if (!f) {
        const char *p = getenv("MMV_STUB_LOG");
        f = p ? fopen(p, "w") : stderr;
    }
    return f;
}
#define LOG(...) do { fprintf(logf_(), __VA_ARGS__); fflush(logf_()); } while (0)

const char *mmv_last_error(void) { return "stub"; }
int mmv_device_count(int *out) { *out = 8; return 0; }
int mmv_comm_unique_id(unsigned char id[128]) { for (int i = 0; i < 128; ++i) id[i] = (unsigned char)(i * 7 + 1); LOG("comm_unique_id\n"); return 0; }
int mmv_comm_init(mmv_ctx *c, int nranks, int rank, const unsigned char id[128])
{
    int ok = 1;
    for (int i = 0; i < 128; ++i) ok = ok && id[i] == (unsigned char)(i * 7 + 1);
    c->nranks = nranks; c->rank = rank;
    LOG("comm_init device=%d nranks=%d rank=%d id_ok=%d\n", c->device, nranks, rank, ok);
    return 0;
}
int mmv_ctx_create(int device, mmv_ctx **out) { *out = calloc(1, sizeof(mmv_ctx)); (*out)->device = device; LOG("ctx_create device=%d\n", device); return 0; }
int mmv_ctx_destroy(mmv_ctx *c) { free(c); return 0; }
int mmv_ctx_sync(mmv_ctx *c) { (void)c; LOG("ctx_sync\n"); return 0; }

int mmv_csr_upload(mmv_ctx *ctx, int64_t n_cells, int64_t n_genes, const int64_t *indptr, const int32_t *indices,
                   const float *values, const float *covar, int64_t c_dim, mmv_csr **out)
{
    (void)indices; (void)values;
    mmv_csr *s = calloc(1, sizeof(mmv_csr));
    s->n_cells = n_cells; s->n_genes = n_genes; s->nnz = indptr[n_cells]; s->c_dim = c_dim; s->has_cov = covar != NULL; s->ctx = ctx;
    LOG("csr_upload dev=%d cells=%ld genes=%ld nnz=%ld covar=%d c_dim=%ld\n", ctx->device, (long)n_cells, (long)n_genes, (long)s->nnz, s->has_cov, (long)c_dim);
    *out = s;
    return 0;
}
int mmv_csr_destroy(mmv_csr *s) { free(s); return 0; }

int mmv_nb_model_create(mmv_ctx *ctx, const mmv_nb_dims *d, uint64_t seed, mmv_model **out)
{
    (void)ctx;
    mmv_model *m = calloc(1, sizeof(mmv_model));
    m->R = d->mu_r_dim;
    LOG("nb_model x=%ld c=%ld n_enc=%d n_dec=%d mu_r=%ld nu_h=%ld nu_r=%ld relu=%d seed=%llu\n", (long)d->x_dim, (long)d->c_dim,
        d->n_enc, d->n_dec, (long)d->mu_r_dim, (long)d->nu_h_dim, (long)d->nu_r_dim, d->do_relu, (unsigned long long)seed);
    *out = m;
    return 0;
}
int mmv_vmf_model_create(mmv_ctx *ctx, const mmv_vmf_dims *d, uint64_t seed, mmv_model **out)
{
    (void)ctx;
    mmv_model *m = calloc(1, sizeof(mmv_model));
    m->is_vmf = 1; m->R = d->z_dim;
    LOG("vmf_model x=%ld c=%ld n_enc=%d n_dec=%d z=%ld kmin=%g kmax=%g relu=%d seed=%llu\n", (long)d->x_dim, (long)d->c_dim, d->n_enc,
        d->n_dec, (long)d->z_dim, d->kappa_min, d->kappa_max, d->do_relu, (unsigned long long)seed);
    *out = m;
    return 0;
}
int mmv_model_destroy(mmv_model *m) { free(m); return 0; }
int mmv_model_set_option(mmv_model *m, const char *key, double value) { (void)m; LOG("set_option %s %g\n", key, value); return 0; }
int mmv_model_num_params(mmv_model *m, int *out) { (void)m; *out = 0; return 0; }
int mmv_model_param_info(mmv_model *m, int i, char name[MMV_NAME_LEN], int64_t *rows, int64_t *cols, int *registered)
{ (void)m; (void)i; name[0] = 0; *rows = *cols = 0; *registered = 0; return 1; }
int mmv_model_set_param(mmv_model *m, const char *name, const float *host, int64_t n) { (void)m; (void)host; LOG("set_param %s n=%ld\n", name, (long)n); return 0; }
int mmv_model_get_param(mmv_model *m, const char *name, float *host, int64_t n) { (void)m; (void)name; (void)host; (void)n; return 1; }

static void fill(mmv_loss_terms *t, float loss) { memset(t, 0, sizeof(*t)); t->loss = loss; }

int mmv_step(mmv_model *m, mmv_csr *csr, const int64_t *cells, int64_t B, const int64_t *boot_idx, const float *eps,
             float kl_weight, int flags, mmv_loss_terms *out)
{
    int64_t bmin = B, bmax = -1;
    if (boot_idx)
        for (int64_t r = 0; r < B; ++r) { if (boot_idx[r] < bmin) bmin = boot_idx[r]; if (boot_idx[r] > bmax) bmax = boot_idx[r]; }
    LOG("step dev=%d flags=%d B=%ld first=%ld last=%ld boot=%d bmin=%ld bmax=%ld eps=%d kl=%.9g out=%d ncells=%ld\n", csr->ctx->device, flags, (long)B,
        (long)cells[0], (long)cells[B - 1], boot_idx != NULL, (long)bmin, (long)bmax, eps != NULL, kl_weight, out != NULL,
        (long)csr->n_cells);
    if (!out) return 0;
    const float loss = 1000.f - (float)m->evals;       /* the k-th evaluation reports 1000 - k */
    m->evals += 1;
    if (flags & MMV_TERMS_DEFERRED) {
        if (m->pending) fill(out, m->pending_loss); else { fill(out, NAN); }
        m->pending = 1; m->pending_loss = loss;
    } else fill(out, loss);
    return 0;
}
int mmv_step_collect(mmv_model *m, mmv_loss_terms *out)
{
    LOG("collect pending=%d\n", m->pending);
    if (!m->pending) return 1;
    fill(out, m->pending_loss);
    m->pending = 0;
    return 0;
}
int mmv_encode(mmv_model *m, mmv_csr *csr, const int64_t *cells, int64_t B, float *mean, float *lnvar)
{
    LOG("encode dev=%d B=%ld first=%ld last=%ld\n", csr->ctx->device, (long)B, (long)cells[0], (long)cells[B - 1]);
    for (int64_t i = 0; i < B * m->R; ++i) { mean[i] = (float)cells[i / m->R]; lnvar[i] = -1.f; }
    return 0;
}
