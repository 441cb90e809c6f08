// True von Mises-Fisher latent (named by the north star — "Wood rejection
// sampler with per-warp retry, Householder reflection, log-Bessel KL, warp shuffles for the row
// reductions").  NO reference code exists for it (vmf.hh:394-404 reparameterises a Gaussian and
// kl_loss_normal vmf.hh:410-414 is the Gaussian KL), so it is opt-in (model option "vmf_latent") and the
// default path stays reference-exact; parity is unpinned, the checker is oracle/vmf_latent_oracle.py.
//
// It takes the place of reparam_fwd/bwd (elementwise.cu) on the same stacked head [B x 2R]:
//   mu_i    = m_i / ||m_i||,  m_i = head[i, :R] (+ covar_enc(c_i), vmf.hh:261)
//   kappa_i = min(1 + softplus(mean_r head[i, R + r]), kappa_max)
//   w_i     ~ Wood(R, kappa_i) on Philox(seed, draw; row i)               — one warp per cell, retry per warp
//   v_i     = eps[i, 1:] / ||eps[i, 1:]||                                 — uniform tangent direction
//   z_i     = H(mu_i) (w_i, sqrt(1 - w_i^2) v_i),  H = I - 2 u u^T / u^T u,  u = e1 - mu_i
//   KL_i    = KL(vMF(mu_i, kappa_i) || U(S^{R-1}))                         — double precision, log-Bessel
// backward: dL/dz -> dhead through the reflection, the normalisation and — for kappa — the implicit
// reparameterisation gradient dw/dkappa (64-point Gauss-Legendre, two nodes per lane) plus kappa A'(kappa)
// from the KL.  One warp owns a row: every row reduction is a shuffle butterfly; B x R is tiny next to
// B x D, so the kernel is latency-, not bandwidth-bound (4096 x 128: 2 MB in, 2 MB out).
#include "common.cuh"
#include "kernels.hh"
#include "vmf_latent_math.cuh"

namespace mmv {

namespace {

using namespace vmfl;

constexpr int kWarpsPerCta = 4;

__constant__ double kGL[128] = {
#include "gl64.inc"
};

__device__ __forceinline__ double
warp_sum_f64(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

struct row_t {
    const float *hm, *hk;      // mean_raw, kappa_raw of this row
    int64_t cell;
    float inv_n, kpre, kappa;
    bool in_range;             // 1 + softplus(kpre) <= kappa_max: the clamp passes the gradient
};

__device__ __forceinline__ float
mean_raw(const reparam_args_t &a, const row_t &r, int k)
{
    float v = r.hm[k];
    if (a.Wce) {   // + covar_enc(c)  (vmf.hh:261)
        float hc = a.bce[k];
        for (int c = 0; c < a.C; ++c) hc = fmaf(a.cov ? a.cov[r.cell * a.C + c] : 1.f, a.Wce[k * a.C + c], hc);
        v += hc;
    }
    return v;
}

__device__ __forceinline__ row_t
row_setup(const vmf_latent_args_t &A, int i, int lane)
{
    const reparam_args_t &a = A.r;
    const int R = a.R;
    row_t r;
    r.hm = a.head + (size_t)i * 2 * R;
    r.hk = r.hm + R;
    r.cell = a.cells ? (int64_t)a.cells[i] : (int64_t)i;
    float n2 = 0.f, ks = 0.f;
    for (int k = lane; k < R; k += 32) {
        const float v = mean_raw(a, r, k);
        n2 = fmaf(v, v, n2);
        ks += r.hk[k];
    }
    n2 = warp_sum(n2);
    ks = warp_sum(ks);
    r.inv_n = 1.f / fmaxf(sqrtf(n2), 1e-12f);
    r.kpre = ks / (float)R;
    const float kraw = 1.f + softplus_f(r.kpre);
    r.in_range = kraw <= A.kappa_max;
    r.kappa = fminf(kraw, A.kappa_max);
    return r;
}

__global__ void __launch_bounds__(kWarpsPerCta * 32)
vmf_latent_fwd_kernel(vmf_latent_args_t A)
{
    const reparam_args_t &a = A.r;
    const int lane = threadIdx.x & 31, i = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5), R = a.R;
    if (i >= a.B) return;                      // warp-uniform
    const row_t r = row_setup(A, i, lane);
    float *mu = a.mean + (size_t)i * R, *z = a.z + (size_t)i * R;
    const float lnv = -logf(r.kappa);          // the recorder's second table: log(1 / kappa)
    for (int k = lane; k < R; k += 32) {
        mu[k] = mean_raw(a, r, k) * r.inv_n;
        a.lnvar[(size_t)i * R + k] = lnv;
    }
    if (lane == 0) {
        double kl, dkl, Ab;
        kl_vmf_uniform(R, (double)r.kappa, A.lgamma_half_d, kl, dkl, Ab);
        if (a.kl_acc) atomicAdd(a.kl_acc, kl);
        if (A.kl_aux) { A.kl_aux[2 * i] = Ab; A.kl_aux[2 * i + 1] = dkl; }   // the backward pass skips the recurrence
    }
    if (!a.eps) {                              // eval mode: z = mu (as reparameterize() returns the mean, vmf.hh:401)
        for (int k = lane; k < R; k += 32) z[k] = mu[k];
        if (lane == 0) {
            if (A.w) A.w[i] = 1.f;
            if (A.attempts) A.attempts[i] = 0;
        }
        return;
    }
    // ---- Wood's rejection sampler, the whole warp retries together
    const wood_t W = wood_setup(R, (double)r.kappa);
    const pkey_t key = make_key(A.seed, A.draw);
    const int nchunks = chi2_chunks(R);
    float w = (float)W.x0;
    int att = kMaxAttempts + 1;
    for (int t = 0; t < kMaxAttempts; ++t) {
        float X = 0.f, Y = 0.f;
        for (int q = lane; q < nchunks; q += 32) {
            float x, y;
            chi2_chunk(key, (uint32_t)i, (uint32_t)t, (uint32_t)q, R - 1, x, y);
            X += x;
            Y += y;
        }
        X = __shfl_sync(0xffffffffu, warp_sum(X), 0);
        Y = __shfl_sync(0xffffffffu, warp_sum(Y), 0);
        float wp;
        const bool ok = wood_try(W, X, Y, accept_uniform(key, (uint32_t)i, (uint32_t)t), wp);
        if (ok) {                              // warp-uniform: every lane holds the same X, Y, u
            w = wp;
            att = t + 1;
            break;
        }
    }
    w = fminf(fmaxf(w, -kWClamp), kWClamp);
    // ---- tangent direction, reflection e1 -> mu
    const float *e = a.eps + (size_t)i * R;
    float e2 = 0.f;
    for (int k = lane; k < R; k += 32)
        if (k >= 1) e2 = fmaf(e[k], e[k], e2);
    e2 = warp_sum(e2);
    const float inv_e = rsqrtf(fmaxf(e2, 1e-30f)), s = sqrtf(fmaxf(1.f - w * w, 1e-12f));
    float un2 = 0.f, dot = 0.f;
    for (int k = lane; k < R; k += 32) {
        const float zp = k == 0 ? w : s * e[k] * inv_e;
        const float u = (k == 0 ? 1.f : 0.f) - mu[k];
        un2 = fmaf(u, u, un2);
        dot = fmaf(u, zp, dot);
    }
    un2 = warp_sum(un2);
    dot = warp_sum(dot);
    const float coef = un2 > 1e-12f ? 2.f * dot / un2 : 0.f;
    for (int k = lane; k < R; k += 32) {
        const float zp = k == 0 ? w : s * e[k] * inv_e;
        const float u = (k == 0 ? 1.f : 0.f) - mu[k];
        z[k] = fmaf(-coef, u, zp);
    }
    if (lane == 0) {
        if (A.w) A.w[i] = w;
        if (A.attempts) A.attempts[i] = att;
    }
}

__global__ void __launch_bounds__(kWarpsPerCta * 32)
vmf_latent_bwd_kernel(vmf_latent_args_t A)
{
    const reparam_args_t &a = A.r;
    const int lane = threadIdx.x & 31, i = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5), R = a.R;
    if (i >= a.B) return;
    const row_t r = row_setup(A, i, lane);
    const float *mu = a.mean + (size_t)i * R, *gz = a.dz + (size_t)i * R;
    float *dm = a.dhead + (size_t)i * 2 * R, *dk = dm + R;
    double kl, dkl, Ab;
    if (A.kl_aux) { Ab = A.kl_aux[2 * i]; dkl = A.kl_aux[2 * i + 1]; }   // saved by the forward pass
    else kl_vmf_uniform(R, (double)r.kappa, A.lgamma_half_d, kl, dkl, Ab);
    double dkappa = (double)(a.kl_weight * a.inv_B) * dkl;
    if (!a.eps) {                              // z = mu
        float pm = 0.f;
        for (int k = lane; k < R; k += 32) pm = fmaf(gz[k], mu[k], pm);
        pm = warp_sum(pm);
        for (int k = lane; k < R; k += 32) dm[k] = (gz[k] - pm * mu[k]) * r.inv_n;
    } else {
        const float w = A.w[i];
        const float *e = a.eps + (size_t)i * R;
        float e2 = 0.f;
        for (int k = lane; k < R; k += 32)
            if (k >= 1) e2 = fmaf(e[k], e[k], e2);
        e2 = warp_sum(e2);
        const float inv_e = rsqrtf(fmaxf(e2, 1e-30f)), s = sqrtf(fmaxf(1.f - w * w, 1e-12f));
        float un2 = 0.f, c = 0.f, ug = 0.f;
        for (int k = lane; k < R; k += 32) {
            const float zp = k == 0 ? w : s * e[k] * inv_e;
            const float u = (k == 0 ? 1.f : 0.f) - mu[k];
            un2 = fmaf(u, u, un2);
            c = fmaf(u, zp, c);
            ug = fmaf(u, gz[k], ug);
        }
        un2 = warp_sum(un2);
        c = warp_sum(c);
        ug = warp_sum(ug);
        const bool refl = un2 > 1e-12f;
        const float iu = refl ? 1.f / un2 : 0.f;
        // dL/dz' = H dL/dz ; dL/dw = gz'_0 - (w / s) <gz'_{1:}, v>
        float g0 = 0.f, gv = 0.f, pm = 0.f;
        for (int k = lane; k < R; k += 32) {
            const float u = (k == 0 ? 1.f : 0.f) - mu[k];
            const float gzp = fmaf(-2.f * ug * iu, u, gz[k]);
            if (k == 0) g0 = gzp;
            else gv = fmaf(gzp, e[k] * inv_e, gv);
            // dL/dmu = -dL/du,  dL/du = -2 (c gz + ug z') / u.u + 4 c ug u / (u.u)^2
            const float zp = k == 0 ? w : s * e[k] * inv_e;
            const float dmu = 2.f * iu * (c * gz[k] + ug * zp) - 4.f * c * ug * iu * iu * u;
            pm = fmaf(dmu, mu[k], pm);
        }
        g0 = __shfl_sync(0xffffffffu, g0, 0);
        gv = warp_sum(gv);
        pm = warp_sum(pm);
        for (int k = lane; k < R; k += 32) {
            const float u = (k == 0 ? 1.f : 0.f) - mu[k];
            const float zp = k == 0 ? w : s * e[k] * inv_e;
            const float dmu = 2.f * iu * (c * gz[k] + ug * zp) - 4.f * c * ug * iu * iu * u;
            dm[k] = (dmu - pm * mu[k]) * r.inv_n;          // through mu = m / ||m||
        }
        const float gw = g0 - (w / s) * gv;
        // implicit reparameterisation: dw/dkappa, two Gauss-Legendre nodes per lane
        const dwdk_t q = dwdk_setup(R, (double)r.kappa, Ab, (double)w);
        const double part = dwdk_node(q, kGL[lane], kGL[64 + lane]) + dwdk_node(q, kGL[32 + lane], kGL[96 + lane]);
        dkappa += (double)gw * warp_sum_f64(part);
    }
    // kappa = min(1 + softplus(kpre), kappa_max), kpre = mean_r head[i, R + r]
    const float sig = r.kpre > 20.f ? 1.f : sigmoid_f(r.kpre);
    const float dkpre = r.in_range ? (float)dkappa * sig / (float)R : 0.f;
    for (int k = lane; k < R; k += 32) dk[k] = dkpre;
}

} // namespace

int
launch_vmf_latent_fwd(const vmf_latent_args_t &A, cudaStream_t st)
{
    MMV_CHECK(A.r.R >= 3, "vmf latent: latent dimension %d < 3", A.r.R);
    vmf_latent_fwd_kernel<<<ceil_div(A.r.B, kWarpsPerCta), kWarpsPerCta * 32, 0, st>>>(A); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

int
launch_vmf_latent_bwd(const vmf_latent_args_t &A, cudaStream_t st)
{
    MMV_CHECK(A.r.eps == nullptr || A.w != nullptr, "vmf latent backward: the forward pass did not save w");
    vmf_latent_bwd_kernel<<<ceil_div(A.r.B, kWarpsPerCta), kWarpsPerCta * 32, 0, st>>>(A); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

} // namespace mmv
