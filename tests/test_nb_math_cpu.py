"""Property checks ("property" row) of the arithmetic K4 is built on, in float64 on the CPU:
the x = 0 shortcut, the product forms for small counts, the shifted Stirling series for large counts and the
hand-derived gradients of SURVEY App. A against scipy's gammaln / digamma and finite differences.
These are the formulas of mm-vae_b200/csrc/nb_likelihood_fast.cu (count_terms, stirling, pair_fast) restated in
numpy; the kernels themselves are checked against the reference on the GPU (tests/test_gpu_parity.py)."""
import numpy as np
import scipy.special as sp

EPS = 1e-4


def nll_full(x, m, nu):            # nb.hh:522-528
    return (sp.gammaln(nu) + sp.gammaln(x + 1) - sp.gammaln(nu + x) + x * (np.log(m + nu) - np.log(m))
            + nu * (np.log(m + nu) - np.log(nu)))


def stirling(z):                   # the kernel's series for z >= 8
    lz, r = np.log(z), 1 / z
    r2 = r * r
    lg = (z - 0.5) * lz - z + 0.9189385332046727 + r * (1 / 12 - r2 * (1 / 360 - r2 / 1260))
    dg = lz - 0.5 * r - r2 * (1 / 12 - r2 * (1 / 120 - r2 / 252))
    return lg, dg


def count_terms(nu, x):
    """lgamma(nu) - lgamma(nu + x) + lgamma(x + 1) and digamma(nu) - digamma(nu + x) the way the kernel branches."""
    xi = int(x)
    if xi == x and 1 <= xi <= 8:
        P, dP = nu, 1.0
        for k in range(1, xi):
            f = nu + k
            dP = dP * f + P
            P *= f
        return sp.gammaln(xi + 1) - np.log(P), -dP / P
    assert x > 8
    P, dP, z = 1.0, 0.0, nu
    if nu < 8:
        P, dP = nu, 1.0
        for k in range(1, 8):
            f = nu + k
            dP = dP * f + P
            P *= f
        z = nu + 8
    lg_n, dg_n = stirling(z)
    lg_nx, dg_nx = stirling(nu + x)
    lg_x1, _ = stirling(x + 1)
    return (lg_n - np.log(P)) - lg_nx + lg_x1, (dg_n - dP / P) - dg_nx


def test_zero_count_shortcut():
    rng = np.random.default_rng(0)
    m, nu = rng.uniform(1e-4, 50, 1000), rng.uniform(2e-4, 1e4, 1000)
    assert np.allclose(nll_full(0.0, m, nu), nu * np.log1p(m / nu), rtol=1e-8, atol=1e-12)   # the full form cancels log(m + nu) - log(nu)


def test_count_terms_match_gammaln_digamma():
    rng = np.random.default_rng(1)
    worst_l = worst_d = 0.0
    for nu in np.concatenate([[2e-4, 1e-3, 0.5, 1.0, 7.99, 8.0, 8.01, 1e4 + 1e-4], rng.uniform(2e-4, 30, 40)]):
        for x in [1, 2, 3, 5, 8, 9, 9.5, 12, 40, 333, 2999, 8.5, 100.25, 1e5]:
            lg, dg = count_terms(float(nu), float(x))
            ref_l = sp.gammaln(nu) + sp.gammaln(x + 1) - sp.gammaln(nu + x)
            ref_d = sp.digamma(nu) - sp.digamma(nu + x)
            worst_l = max(worst_l, abs(lg - ref_l) / max(1.0, abs(ref_l)))
            worst_d = max(worst_d, abs(dg - ref_d) / max(1.0, abs(ref_d)))
    assert worst_l < 1e-7 and worst_d < 1e-8, (worst_l, worst_d)


def test_hand_derived_gradients_vs_finite_differences():
    """g = dl/dm = (nu + x)/(m + nu) - x/m ; dl/dnu = psi(nu) - psi(nu + x) + log((m + nu)/nu) + (x - m)/(m + nu)."""
    rng = np.random.default_rng(2)
    for _ in range(200):
        x = float(rng.choice([0, 1, 2, 5, 8, 9, 30, 400]))
        m, nu = float(rng.uniform(1e-3, 20)), float(rng.uniform(1e-2, 50))
        g = (nu + x) / (m + nu) - x / m
        dnu = sp.digamma(nu) - sp.digamma(nu + x) + np.log((m + nu) / nu) + (x - m) / (m + nu)
        hm, hn = 1e-6 * m, 1e-6 * nu
        fd_m = (nll_full(x, m + hm, nu) - nll_full(x, m - hm, nu)) / (2 * hm)
        fd_n = (nll_full(x, m, nu + hn) - nll_full(x, m, nu - hn)) / (2 * hn)
        assert abs(g - fd_m) <= 1e-6 * max(1.0, abs(fd_m)) and abs(dnu - fd_n) <= 1e-5 * max(1e-3, abs(fd_n)), (x, m, nu)
        if x >= 1:      # the kernel's split: dense x = 0 value + the count's correction
            _, dd = count_terms(nu, x)
            dnu0 = np.log1p(m / nu) - m / (m + nu)
            assert abs((dnu0 + dd + x / (m + nu)) - dnu) <= 1e-9 * max(1.0, abs(dnu))


def test_softmax_backward_row_identity():
    """dL/da_k = d p_k (g_k - s), s = sum_j g_j p_j (App. A) against a finite difference through softmax."""
    rng = np.random.default_rng(3)
    D, d, nu = 50, 3.7, rng.uniform(0.1, 5, 50)
    a = rng.normal(size=D)
    x = rng.poisson(0.3, size=D).astype(float)

    def loss(a):
        p = np.exp(a - a.max())
        p /= p.sum()
        return nll_full(x, p * d + EPS, nu + EPS).sum()

    p = np.exp(a - a.max())
    p /= p.sum()
    m = p * d + EPS
    g = (nu + EPS + x) / (m + nu + EPS) - x / m
    grad = d * p * (g - (g * p).sum())
    assert abs(grad.sum()) < 1e-12 * np.abs(grad).sum() + 1e-15          # rows of the softmax gradient sum to zero
    for k in range(0, D, 7):
        e = np.zeros(D)
        e[k] = 1e-6
        fd = (loss(a + e) - loss(a - e)) / 2e-6
        assert abs(grad[k] - fd) <= 1e-6 * max(1.0, abs(fd))


def test_plain_row_interval_test_is_sound():
    """The sweep without softplus threshold / clamp / mask is taken when zn*[wlo, whi] + [blo, bhi] (log2 domain) lies in
    (-13, 28.5): inside it softplus(u) is in [1e-4, 1e4] and u < 20, so clamp and threshold are inactive."""
    l2e = 1.4426950408889634
    for u2 in (-13.0, 28.5):
        u = u2 / l2e
        spv = np.log1p(np.exp(u))
        assert u < 20 and 1e-4 <= spv <= 1e4


def test_vmf_data_term_gradient_vs_finite_differences():
    """K5 (vmf.hh:283-290,419-440): recon = normalise(q), q = exp(pre) + cv; t = <yobs, recon>; per-cell loss -kappa t / B.
    Hand-derived: dL/dq = -kappa (yobs - recon t) / (B ||q||), dL/dpre = dL/dq exp(pre)."""
    rng = np.random.default_rng(4)
    D, B, kappa = 40, 7.0, 2.3
    pre, cv = rng.normal(size=D), rng.uniform(0, 0.3, size=D)
    x = rng.poisson(0.4, size=D).astype(float)
    yobs = np.log1p(x) + 1e-2 / D
    yobs /= np.linalg.norm(yobs)

    def loss(pre):
        q = np.exp(pre) + cv
        return -kappa * (yobs @ (q / np.linalg.norm(q))) / B

    q = np.exp(pre) + cv
    qn = np.linalg.norm(q)
    t = yobs @ (q / qn)
    dq = -kappa * (yobs - (q / qn) * t) / (B * qn)
    dpre = dq * np.exp(pre)
    assert abs(dq @ q) < 1e-12                      # the gradient is tangent to the sphere
    for k in range(0, D, 5):
        e = np.zeros(D)
        e[k] = 1e-6
        fd = (loss(pre + e) - loss(pre - e)) / 2e-6
        assert abs(dpre[k] - fd) <= 1e-7 * max(1e-3, abs(fd))
