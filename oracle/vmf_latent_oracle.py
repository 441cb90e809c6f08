"""TEST INFRASTRUCTURE ONLY (never imported by the product path).

CPU restatement of the true von Mises-Fisher latent: Wood's rejection
sampler on the Philox stream the kernel uses, Householder reflection, KL(vMF || uniform) with
log-Bessel terms, and the implicit-reparameterisation gradient of the sampled cosine.

PARITY UNPINNED: the reference has NO code for any of this (`vmf.hh:394-404` reparameterises a
Gaussian).  What anchors it instead is the published algorithms — Wood (1994) for the
sampler, Davidson et al. (2018, "Hyperspherical VAE") for the reflection and the KL, Figurnov et al.
(2018) for implicit reparameterisation — evaluated here with independent numerics: numpy's Philox
restated below, scipy.special.ive for the Bessel functions, scipy.integrate.quad for the CDF
derivative, torch float64 autograd for every other derivative.
"""
import numpy as np
import scipy.integrate
import scipy.special
import torch

KEY_TAG = 0x766D666C
ACCEPT_CHUNK = 0xFFFFFFFF
MAX_ATTEMPTS = 64
W_CLAMP = np.float32(0.9999999)


def philox4x32_10(c, k0, k1):
    """c: uint32 array [..., 4]; k0, k1: uint32 (scalars or arrays broadcastable to c[..., 0])."""
    c = [np.asarray(c[..., i], dtype=np.uint64) for i in range(4)]
    k0 = np.asarray(k0, dtype=np.uint64) * np.ones_like(c[0])
    k1 = np.asarray(k1, dtype=np.uint64) * np.ones_like(c[0])
    M = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0 = np.uint64(0xD2511F53) * c[0]
        p1 = np.uint64(0xCD9E8D57) * c[2]
        c = [(p1 >> np.uint64(32)) ^ c[1] ^ k0, p1 & M, (p0 >> np.uint64(32)) ^ c[3] ^ k1, p0 & M]
        k0 = (k0 + np.uint64(0x9E3779B9)) & M
        k1 = (k1 + np.uint64(0xBB67AE85)) & M
    return np.stack([v.astype(np.uint32) for v in c], axis=-1)


def make_key(seed, draw):
    return (seed & 0xFFFFFFFF, ((seed >> 32) ^ (draw >> 32) ^ KEY_TAG) & 0xFFFFFFFF, draw & 0xFFFFFFFF)


def u01(x):
    return ((x >> np.uint32(9)).astype(np.float32) + np.float32(0.5)) * np.float32(1.0 / 8388608.0)


def box_muller(a, b):
    r = np.sqrt(np.float32(-2.0) * np.log(u01(a)))
    th = np.float32(6.28318530717958648) * u01(b)
    return r * np.cos(th), r * np.sin(th)


def chi2_sums(d, seed, draw, rows, attempt):
    """X, Y (float32) of one Beta proposal for every row in `rows`."""
    k0, k1, dr = make_key(seed, draw)
    nq = d // 2
    rows = np.asarray(rows, dtype=np.uint32)
    q = np.arange(nq, dtype=np.uint32)
    c = np.zeros((len(rows), nq, 4), dtype=np.uint32)
    c[..., 0] = q[None, :]
    c[..., 1] = attempt
    c[..., 2] = rows[:, None]
    c[..., 3] = dr
    r = philox4x32_10(c, k0, k1)
    g0, g1 = box_muller(r[..., 0], r[..., 1])
    g2, g3 = box_muller(r[..., 2], r[..., 3])
    v0 = (2 * q < d - 1)[None, :]
    v1 = (2 * q + 1 < d - 1)[None, :]
    X = np.where(v0, g0 * g0, 0).astype(np.float32) + np.where(v1, g1 * g1, 0).astype(np.float32)
    Y = np.where(v0, g2 * g2, 0).astype(np.float32) + np.where(v1, g3 * g3, 0).astype(np.float32)
    return X.sum(axis=1, dtype=np.float32), Y.sum(axis=1, dtype=np.float32)


def accept_uniform(seed, draw, rows, attempt):
    k0, k1, dr = make_key(seed, draw)
    rows = np.asarray(rows, dtype=np.uint32)
    c = np.zeros((len(rows), 4), dtype=np.uint32)
    c[:, 0] = ACCEPT_CHUNK
    c[:, 1] = attempt
    c[:, 2] = rows
    c[:, 3] = dr
    return u01(philox4x32_10(c, k0, k1)[:, 0])


def wood_sample(d, kappa, seed, draw, rows=None):
    """Wood (1994) on the kernel's random stream. kappa: [B] float64. Returns w [B] float32, attempts [B],
    margin [B] (how far the accepting test was from flipping, for tests that compare decisions)."""
    kappa = np.asarray(kappa, dtype=np.float64)
    B = len(kappa)
    rows = np.arange(B) if rows is None else np.asarray(rows)
    m1 = float(d - 1)
    b = m1 / (2 * kappa + np.sqrt(4 * kappa * kappa + m1 * m1))
    x0 = (1 - b) / (1 + b)
    c = kappa * x0 + m1 * np.log(1 - x0 * x0)
    w = x0.astype(np.float32).copy()
    att = np.full(B, MAX_ATTEMPTS + 1)
    margin = np.full(B, np.inf)
    live = np.ones(B, dtype=bool)
    for a in range(MAX_ATTEMPTS):
        if not live.any():
            break
        idx = np.nonzero(live)[0]
        X, Y = chi2_sums(d, seed, draw, rows[idx], a)
        u = accept_uniform(seed, draw, rows[idx], a)
        Z = X.astype(np.float64) / (X.astype(np.float64) + Y.astype(np.float64))
        wp = (1 - (1 + b[idx]) * Z) / (1 - (1 - b[idx]) * Z)
        lhs = kappa[idx] * wp + m1 * np.log(1 - x0[idx] * wp) - c[idx] - np.log(u.astype(np.float64))
        margin[idx] = np.minimum(margin[idx], np.abs(lhs))
        ok = lhs >= 0
        w[idx[ok]] = wp[ok].astype(np.float32)
        att[idx[ok]] = a + 1
        live[idx[ok]] = False
    return w, att, margin


def bessel_ratio(nu, kappa):
    return scipy.special.ive(nu + 1, kappa) / scipy.special.ive(nu, kappa)


def kl_vmf_uniform(d, kappa):
    """KL(vMF(mu, kappa) || U(S^{d-1})) and d/dkappa, float64 (Davidson et al. 2018, eq. 14-15)."""
    kappa = np.asarray(kappa, dtype=np.float64)
    nu = d / 2 - 1
    A = bessel_ratio(nu, kappa)
    log_i = np.log(scipy.special.ive(nu, kappa)) + kappa
    log_c = nu * np.log(kappa) - (d / 2) * np.log(2 * np.pi) - log_i
    log_area = np.log(2) + (d / 2) * np.log(np.pi) - scipy.special.gammaln(d / 2)
    kl = kappa * A + log_c + log_area
    dkl = kappa * (1 - A * A - (d - 1) * A / kappa)
    return kl, dkl, A


def dw_dkappa(d, kappa, w):
    """Implicit reparameterisation gradient of the cosine: -(dF/dkappa)(w) / f(w), adaptive quadrature."""
    a = (d - 3) / 2
    A = float(bessel_ratio(d / 2 - 1, kappa))
    lw = np.log1p(-w * w)

    def integrand(t):
        return (t - A) * np.exp(kappa * (t - w) + a * (np.log1p(-t * t) - lw))

    tstar = kappa / (a + np.sqrt(a * a + kappa * kappa))
    if w <= tstar:
        v, _ = scipy.integrate.quad(integrand, -1, w, epsabs=0, epsrel=1e-11, limit=400,
                                    points=[max(-1 + 1e-12, w - 50 / kappa)] if w - 50 / kappa > -1 else None)
        return -v
    v, _ = scipy.integrate.quad(integrand, w, 1, epsabs=0, epsrel=1e-11, limit=400,
                                points=[min(1 - 1e-12, w + 50 / kappa)] if w + 50 / kappa < 1 else None)
    return v


def forward_backward(head, eps, w, dz, kl_weight, inv_B, kappa_max=1e4, covar_term=None):
    """The latent block in float64 torch with autograd.
    head [B, 2R] (mean_raw | kappa_raw), eps [B, R] normals (tangent noise: columns 1..R-1) or None (z = mu),
    w [B] the sampled cosines (given: the rejection loop is not differentiable; its kappa-derivative
    enters through dw/dkappa), dz [B, R] upstream gradient.
    Returns dict(mean, kappa, z, kl_sum, dhead)."""
    B, R2 = head.shape
    R = R2 // 2
    h = torch.tensor(head, dtype=torch.float64, requires_grad=True)
    m = h[:, :R]
    if covar_term is not None:
        m = m + torch.tensor(covar_term, dtype=torch.float64)
    mu = m / m.norm(dim=1, keepdim=True).clamp_min(1e-12)
    kpre = h[:, R:].mean(dim=1)
    kap_raw = 1 + torch.nn.functional.softplus(kpre)
    kappa = kap_raw.clamp(max=kappa_max)
    kl, dkl, _ = kl_vmf_uniform(R, kappa.detach().numpy())
    kl_t = torch.tensor(kl)
    # KL as a function of kappa with the analytic derivative attached (first-order surrogate)
    kl_sur = kl_t + torch.tensor(dkl) * (kappa - kappa.detach())
    out = {"mean": mu.detach().numpy(), "kappa": kappa.detach().numpy(), "kl_sum": float(kl_t.sum())}
    if eps is None:
        out["z"] = mu.detach().numpy()
        return out
    wv = np.clip(np.asarray(w, dtype=np.float64), -float(W_CLAMP), float(W_CLAMP))
    g = np.array([dw_dkappa(R, float(k), float(x)) for k, x in zip(kappa.detach().numpy(), wv)])
    w_t = torch.tensor(wv) + torch.tensor(g) * (kappa - kappa.detach())
    e = torch.tensor(eps, dtype=torch.float64)[:, 1:]
    v = e / e.norm(dim=1, keepdim=True).clamp_min(1e-15)
    s = (1 - w_t * w_t).clamp_min(1e-12).sqrt()
    zp = torch.cat([w_t[:, None], s[:, None] * v], dim=1)
    e1 = torch.zeros(R, dtype=torch.float64)
    e1[0] = 1
    u = e1[None, :] - mu
    un2 = (u * u).sum(dim=1, keepdim=True)
    refl = zp - 2 * (u * zp).sum(dim=1, keepdim=True) / un2.clamp_min(1e-30) * u
    z = torch.where(un2 > 1e-12, refl, zp)
    out["z"] = z.detach().numpy()
    out["dw_dkappa"] = g
    if dz is not None:
        loss = (z * torch.tensor(dz, dtype=torch.float64)).sum() + kl_weight * inv_B * kl_sur.sum()
        loss.backward()
        out["dhead"] = h.grad.numpy()
    return out
