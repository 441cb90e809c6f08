// TEST INFRASTRUCTURE (not shipped, never on the product path): compiles the scalar math header of
// the vMF latent, mm-vae_b200/csrc/vmf_latent_math.cuh, for the HOST so that tests/test_vmf_latent_cpu.py
// can compare exactly the code the GPU kernels run with scipy / mpmath / numpy restatements
// (oracle/vmf_latent_oracle.py).  Parity unpinned: the reference has no vMF latent.
#include "../mm-vae_b200/csrc/vmf_latent_math.cuh"

using namespace mmv::vmfl;

static const double kGL[128] = {
#include "../mm-vae_b200/csrc/gl64.inc"
};

extern "C" {

void
hc_philox(const uint32_t *c, const uint32_t *k, uint32_t *out)
{
    const u4_t r = philox4x32_10(u4_t{ c[0], c[1], c[2], c[3] }, k[0], k[1]);
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}

void
hc_log_bessel_ratio(double nu, double x, double *log_i, double *ratio)
{
    log_bessel_ratio(nu, x, *log_i, *ratio);
}

void
hc_kl(int d, double kappa, double *kl, double *dkl, double *A)
{
    kl_vmf_uniform(d, kappa, lgamma(0.5 * d), *kl, *dkl, *A);
}

double
hc_dwdk(int d, double kappa, double w)
{
    double log_i, A;
    log_bessel_ratio(0.5 * d - 1., kappa, log_i, A);
    const dwdk_t q = dwdk_setup(d, kappa, A, w);
    double s = 0.;
    for (int i = 0; i < 64; ++i) s += dwdk_node(q, kGL[i], kGL[64 + i]);
    return s;
}

// chi-square sums of one proposal, chunks added in order (the kernel adds them lane-strided + shuffles)
void
hc_chi2(int d, uint64_t seed, uint64_t draw, uint32_t row, uint32_t attempt, float *X, float *Y)
{
    const pkey_t k = make_key(seed, draw);
    float sx = 0.f, sy = 0.f;
    for (int q = 0; q < chi2_chunks(d); ++q) {
        float x, y;
        chi2_chunk(k, row, attempt, (uint32_t)q, d - 1, x, y);
        sx += x; sy += y;
    }
    *X = sx; *Y = sy;
}

// the whole rejection loop for one row; returns the number of proposals used
int
hc_wood(int d, double kappa, uint64_t seed, uint64_t draw, uint32_t row, float *w)
{
    const pkey_t k = make_key(seed, draw);
    const wood_t W = wood_setup(d, kappa);
    for (int attempt = 0; attempt < kMaxAttempts; ++attempt) {
        float X, Y;
        hc_chi2(d, seed, draw, row, (uint32_t)attempt, &X, &Y);
        if (wood_try(W, X, Y, accept_uniform(k, row, (uint32_t)attempt), *w)) return attempt + 1;
    }
    *w = (float)W.x0;
    return kMaxAttempts + 1;
}

} // extern "C"
