// Scalar math of the true von Mises-Fisher latent (named by the north
// star, NO reference code: vmf.hh:394-404 reparameterises a Gaussian) as host/device inline
// functions, so the same source runs inside vmf_latent.cu on the GPU and inside the host checker
// (oracle/vmf_latent_hostcheck.cu) that the CPU tests compare with scipy / mpmath.
//
//   * counter-based random numbers: Philox4x32-10 (Salmon et al., SC'11) keyed on (seed, draw) with the
//     counter (chunk, attempt, row, draw): any row is reproducible on its own;
//   * Wood's rejection sampler for the cosine w = <z, mu> (Wood 1994, "Simulation of the von Mises
//     Fisher distribution", Alg. VM*), Beta((d-1)/2, (d-1)/2) proposals built from chi-square sums
//     of Philox normals so a warp can draw one proposal cooperatively;
//   * log I_nu(kappa) and A = I_{nu+1} / I_nu by Miller's backward recurrence started 40 orders up
//     from the uniform (Debye) expansion; KL(vMF(mu, kappa) || U(S^{d-1})) and its kappa-derivative;
//   * d w / d kappa by implicit reparameterisation (Figurnov et al. 2018): -dF/dkappa / f(w) with the
//     CDF derivative as a 64-point Gauss-Legendre integral over the side of w that has no cancellation.
#pragma once

#include <cmath>
#include <cstdint>

#ifdef __CUDACC__
#define MMV_HD __host__ __device__ __forceinline__
#else
#define MMV_HD inline
#endif

namespace mmv {
namespace vmfl {

struct u4_t {
    uint32_t x, y, z, w;
};

MMV_HD u4_t
philox4x32_10(u4_t c, uint32_t k0, uint32_t k1)
{
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c.x, p1 = (uint64_t)0xCD9E8D57u * c.z;
        const u4_t n = { (uint32_t)(p1 >> 32) ^ c.y ^ k0, (uint32_t)p1, (uint32_t)(p0 >> 32) ^ c.w ^ k1, (uint32_t)p0 };
        c = n;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return c;
}

// 23 random bits + 1/2 -> strictly inside (0, 1), exactly representable
MMV_HD float
u01(uint32_t x)
{
    return ((float)(x >> 9) + 0.5f) * (1.f / 8388608.f);
}

MMV_HD void
box_muller(uint32_t a, uint32_t b, float &n0, float &n1)
{
    const float r = sqrtf(-2.f * logf(u01(a)));
    float s, c;
    sincosf(6.28318530717958648f * u01(b), &s, &c);
    n0 = r * c;
    n1 = r * s;
}

constexpr uint32_t kAcceptChunk = 0xFFFFFFFFu;   // counter word 0 of the acceptance uniform
constexpr uint32_t kKeyTag = 0x766d666cu;

struct pkey_t {
    uint32_t k0, k1, draw;
};

MMV_HD pkey_t
make_key(uint64_t seed, uint64_t draw)
{
    pkey_t k;
    k.k0 = (uint32_t)seed;
    k.k1 = (uint32_t)(seed >> 32) ^ (uint32_t)(draw >> 32) ^ kKeyTag;
    k.draw = (uint32_t)draw;
    return k;
}

// chunk q of one proposal: four normals; the first two feed X = chi^2 at indices 2q, 2q + 1, the
// last two feed Y at the same indices (indices >= d - 1 are masked)
MMV_HD void
chi2_chunk(const pkey_t &k, uint32_t row, uint32_t attempt, uint32_t q, int m1, float &X, float &Y)
{
    const u4_t c = { q, attempt, row, k.draw };
    const u4_t r = philox4x32_10(c, k.k0, k.k1);
    float g0, g1, g2, g3;
    box_muller(r.x, r.y, g0, g1);
    box_muller(r.z, r.w, g2, g3);
    const bool v0 = (int)(2 * q) < m1, v1 = (int)(2 * q + 1) < m1;
    X = (v0 ? g0 * g0 : 0.f) + (v1 ? g1 * g1 : 0.f);
    Y = (v0 ? g2 * g2 : 0.f) + (v1 ? g3 * g3 : 0.f);
}

MMV_HD int
chi2_chunks(int d)
{
    return (d - 1 + 1) / 2;   // ceil((d - 1) / 2) Philox calls per proposal
}

MMV_HD float
accept_uniform(const pkey_t &k, uint32_t row, uint32_t attempt)
{
    const u4_t c = { kAcceptChunk, attempt, row, k.draw };
    return u01(philox4x32_10(c, k.k0, k.k1).x);
}

// Wood (1994): b = (-2 kappa + sqrt(4 kappa^2 + (d-1)^2)) / (d-1), x0 = (1-b)/(1+b),
// c = kappa x0 + (d-1) log(1 - x0^2); proposal w = (1 - (1+b) Z) / (1 - (1-b) Z), Z ~ Beta;
// accept when kappa w + (d-1) log(1 - x0 w) - c >= log u.
struct wood_t {
    double b, x0, c, kappa, m1;
};

MMV_HD wood_t
wood_setup(int d, double kappa)
{
    wood_t W;
    W.m1 = (double)(d - 1);
    W.kappa = kappa;
    W.b = W.m1 / (2. * kappa + sqrt(4. * kappa * kappa + W.m1 * W.m1));   // cancellation-free form
    W.x0 = (1. - W.b) / (1. + W.b);
    W.c = kappa * W.x0 + W.m1 * log(1. - W.x0 * W.x0);
    return W;
}

MMV_HD bool
wood_try(const wood_t &W, float X, float Y, float u, float &w_out)
{
    const double Z = (double)X / ((double)X + (double)Y);
    const double w = (1. - (1. + W.b) * Z) / (1. - (1. - W.b) * Z);
    w_out = (float)w;
    return W.kappa * w + W.m1 * log(1. - W.x0 * w) - W.c >= log((double)u);
}

constexpr int kMaxAttempts = 64;      // acceptance is > 0.65 for every (d, kappa): 64 misses never happen
constexpr float kWClamp = 0.9999999f;

// ---------------------------------------------------------------------------------------------
// modified Bessel function of the first kind, log I_nu(x) and r = I_{nu+1}(x) / I_nu(x), x > 0
// ---------------------------------------------------------------------------------------------
MMV_HD double
debye_log_bessel(double N, double x)
{
    const double z = x / N, s = sqrt(1. + z * z), t = 1. / s, t2 = t * t;
    const double eta = s + log(z / (1. + s));
    const double u1 = t * (3. - 5. * t2) / 24.;
    const double u2 = t2 * (81. + t2 * (-462. + 385. * t2)) / 1152.;
    const double u3 = t * t2 * (30375. + t2 * (-369603. + t2 * (765765. - 425425. * t2))) / 414720.;
    const double iN = 1. / N;
    return N * eta - 0.5 * log(6.283185307179586477 * N) - 0.5 * log(s) + log1p(iN * (u1 + iN * (u2 + iN * u3)));
}

MMV_HD void
log_bessel_ratio(double nu, double x, double &log_i, double &ratio, double *ratio_next = nullptr)
{
    constexpr int M = 40;
    const double N = nu + (double)M;
    double lg = debye_log_bessel(N, x);
    double r = exp(debye_log_bessel(N + 1., x) - lg);      // I_{N+1} / I_N
    double r_up = r;
    for (int k = 0; k < M; ++k) {                           // r_{o-1} = 1 / (2 o / x + r_o);  I_{o-1} = I_o / r_{o-1}
        r_up = r;
        r = 1. / (2. * (N - (double)k) / x + r);
        lg -= log(r);
    }
    log_i = lg;
    ratio = r;
    if (ratio_next) *ratio_next = r_up;                     // I_{nu+2} / I_{nu+1}
}

// KL(vMF(mu, kappa) || uniform on S^{d-1}) = kappa A + nu log kappa - log I_nu(kappa) - nu log 2 - lgamma(d/2),
// nu = d/2 - 1, A = I_{d/2} / I_nu;  d KL / d kappa = kappa A'(kappa), A' = 1 - A^2 - (d-1) A / kappa, evaluated as
// A (r_{nu+1} - A) + A / kappa (from I_nu - I_{nu+2} = 2 (nu+1) I_{nu+1} / kappa): terms of size 1 / kappa cancel
// instead of terms of size 1
MMV_HD void
kl_vmf_uniform(int d, double kappa, double lgamma_half_d, double &kl, double &dkl, double &A)
{
    const double nu = 0.5 * (double)d - 1.;
    double log_i, r1;
    log_bessel_ratio(nu, kappa, log_i, A, &r1);
    kl = kappa * A + nu * log(kappa) - log_i - nu * 0.693147180559945309 - lgamma_half_d;
    dkl = A * (kappa * (r1 - A) + 1.);
}

// ---------------------------------------------------------------------------------------------
// implicit reparameterisation of the cosine: f(t) ~ exp(kappa t) (1 - t^2)^a on [-1, 1], a = (d-3)/2;
// dF/dkappa (w) = int_{-1}^{w} (t - A) f(t) dt = -int_{w}^{1} (t - A) f(t) dt, so
//   dw/dkappa = -int_{-1}^{w} (t - A) e^{g(t) - g(w)} dt = +int_{w}^{1} (t - A) e^{g(t) - g(w)} dt,
// g(t) = kappa t + a log(1 - t^2).  Below the mode the first form has no cancellation, above it the second.
// ---------------------------------------------------------------------------------------------
struct dwdk_t {
    double lo, h, sgn, w, kappa, a, A;
};

MMV_HD dwdk_t
dwdk_setup(int d, double kappa, double A, double w)
{
    dwdk_t q;
    q.a = 0.5 * (double)(d - 3);
    q.kappa = kappa;
    q.A = A;
    q.w = w;
    const double tstar = kappa / (q.a + sqrt(q.a * q.a + kappa * kappa));            // mode of f
    const double sig = q.a > 0. ? (1. - tstar * tstar) / sqrt(2. * q.a * (1. + tstar * tstar)) : 0.;
    const double span = 12. * sig + 40. / kappa;
    double hi;
    if (w <= tstar) {
        q.lo = fmax(-1., w - span);
        hi = w;
        q.sgn = -1.;
    } else {
        q.lo = w;
        hi = fmin(1., w + span);
        q.sgn = 1.;
    }
    q.h = 0.5 * (hi - q.lo);
    return q;
}

// contribution of one Gauss-Legendre node (x in [-1, 1], weight wt); the caller sums 64 of them
MMV_HD double
dwdk_node(const dwdk_t &q, double x, double wt)
{
    const double t = q.lo + q.h * (x + 1.);
    const double lg = q.kappa * (t - q.w) + q.a * (log1p(-t * t) - log1p(-q.w * q.w));
    return q.sgn * q.h * wt * (t - q.A) * exp(lg);
}

} // namespace vmfl
} // namespace mmv
