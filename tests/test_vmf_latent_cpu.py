"""vMF latent, CPU side: the scalar math header the kernels run
(mm-vae_b200/csrc/vmf_latent_math.cuh, compiled for the host by oracle/vmf_latent_hostcheck.cc) against
independent numerics — Philox known answers (Random123), scipy Bessel functions, adaptive quadrature —
and the oracle's Wood sampler against the distribution it must produce.  Parity unpinned: no reference code."""
import ctypes
import os
import subprocess

import numpy as np
import pytest
import scipy.special

from oracle import vmf_latent_oracle as vo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "oracle", "_build", "libvmf_latent_hostcheck.so")


@pytest.fixture(scope="module")
def hc():
    if not os.path.exists(SO):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "hostcheck"], check=True)
    L = ctypes.CDLL(SO)
    L.hc_dwdk.restype = ctypes.c_double
    L.hc_dwdk.argtypes = [ctypes.c_int, ctypes.c_double, ctypes.c_double]
    L.hc_log_bessel_ratio.argtypes = [ctypes.c_double, ctypes.c_double, ctypes.POINTER(ctypes.c_double),
                                      ctypes.POINTER(ctypes.c_double)]
    L.hc_kl.argtypes = [ctypes.c_int, ctypes.c_double] + [ctypes.POINTER(ctypes.c_double)] * 3
    L.hc_wood.argtypes = [ctypes.c_int, ctypes.c_double, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint32,
                          ctypes.POINTER(ctypes.c_float)]
    L.hc_chi2.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint32,
                          ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_float)]
    return L


KAT = [((0, 0, 0, 0), (0, 0), (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
       ((0xFFFFFFFF,) * 4, (0xFFFFFFFF,) * 2, (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
       ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0),
        (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1))]


@pytest.mark.parametrize("c,k,want", KAT)
def test_philox_known_answers(hc, c, k, want):
    out = (ctypes.c_uint32 * 4)()
    hc.hc_philox((ctypes.c_uint32 * 4)(*c), (ctypes.c_uint32 * 2)(*k), out)
    assert tuple(out) == want
    got = vo.philox4x32_10(np.array([c], dtype=np.uint32), k[0], k[1])[0]
    assert tuple(int(v) for v in got) == want


def test_log_bessel_and_ratio_vs_scipy(hc):
    li, r = ctypes.c_double(), ctypes.c_double()
    for nu in (0.5, 1.0, 1.5, 4.0, 7.5, 15.0, 31.0, 63.0, 63.5, 127.0, 511.0):
        for x in (1.0, 1.5, 3.0, 10.0, 50.0, 100.0, 1e3, 1e4):
            ive = scipy.special.ive(nu, x)
            if ive == 0 or not np.isfinite(ive):
                continue
            hc.hc_log_bessel_ratio(nu, x, ctypes.byref(li), ctypes.byref(r))
            ref = np.log(ive) + x
            assert abs(li.value - ref) <= 2e-9 * max(1.0, abs(ref)), (nu, x, li.value, ref)
            refr = scipy.special.ive(nu + 1, x) / ive
            assert abs(r.value - refr) <= 1e-9 * refr, (nu, x, r.value, refr)


def test_kl_vs_oracle(hc):
    kl, dkl, A = ctypes.c_double(), ctypes.c_double(), ctypes.c_double()
    for d in (3, 4, 10, 32, 128, 256):
        for kappa in (1.0, 2.5, 20.0, 300.0, 5000.0):
            hc.hc_kl(d, kappa, ctypes.byref(kl), ctypes.byref(dkl), ctypes.byref(A))
            rk, rd, rA = vo.kl_vmf_uniform(d, kappa)
            if not np.isfinite(rk):
                continue
            assert abs(kl.value - rk) <= 1e-8 * max(1.0, abs(rk)), (d, kappa, kl.value, rk)
            # kappa (1 - A^2) - (d - 1) A: two terms of size d - 1 that cancel at large kappa
            assert abs(dkl.value - rd) <= 1e-7 * abs(rd) + 1e-8 * (d - 1), (d, kappa, dkl.value, rd)
            # derivative consistent with a central difference of the value
            h = 1e-4 * kappa
            fd = (vo.kl_vmf_uniform(d, kappa + h)[0] - vo.kl_vmf_uniform(d, kappa - h)[0]) / (2 * h)
            assert abs(dkl.value - fd) <= 1e-5 * abs(fd) + 1e-7 * (d - 1), (d, kappa, dkl.value, fd)
    # KL -> 0 as kappa -> 0 and grows with kappa
    hc.hc_kl(16, 1e-3, ctypes.byref(kl), ctypes.byref(dkl), ctypes.byref(A))
    assert abs(kl.value) < 1e-6


def test_implicit_gradient_vs_quadrature(hc):
    rng = np.random.default_rng(0)
    cases = [(3, 1.0, 0.2), (3, 50.0, 0.97), (8, 5.0, 0.3), (8, 5.0, -0.5), (128, 100.0, 0.85), (128, 1000.0, 0.99),
             (4, 3.0, 0.999), (4, 3.0, -0.999), (128, 1.0, 0.0), (5, 2.0, 0.84), (5, 2.0, 0.86), (3, 1e4, 0.9999)]
    for d in (3, 4, 5, 10, 32, 128):
        for kappa in (1.0, 3.0, 30.0, 300.0, 3000.0):
            for _ in range(2):
                w = rng.uniform(-1, 1) if kappa < 5 else 1 - rng.exponential((d - 1) / (2 * kappa))
                cases.append((d, kappa, float(np.clip(w, -0.9999, 0.99999))))
    for d, kappa, w in cases:
        got = hc.hc_dwdk(d, kappa, w)
        ref = vo.dw_dkappa(d, kappa, w)
        assert abs(got - ref) <= 2e-6 * abs(ref) + 1e-14, (d, kappa, w, got, ref)


def test_implicit_gradient_is_the_quantile_derivative():
    """dw/dkappa at fixed CDF level == finite difference of the quantile function."""
    import scipy.integrate
    import scipy.optimize
    d, kappa, w = 10, 6.0, 0.35
    a = (d - 3) / 2

    def cdf(x, k):
        f = lambda t: np.exp(k * (t - 1)) * (1 - t * t) ** a
        return scipy.integrate.quad(f, -1, x, epsrel=1e-13)[0] / scipy.integrate.quad(f, -1, 1, epsrel=1e-13)[0]

    u = cdf(w, kappa)
    h = 1e-4
    q = [scipy.optimize.brentq(lambda x: cdf(x, k) - u, -0.999, 0.999, xtol=1e-14) for k in (kappa - h, kappa + h)]
    fd = (q[1] - q[0]) / (2 * h)
    assert abs(vo.dw_dkappa(d, kappa, w) - fd) < 1e-6 * abs(fd)


def test_wood_stream_host_equals_oracle(hc):
    """Same Philox stream, same decisions: the header's sampler (host build) and the numpy restatement."""
    for d, kappa in ((3, 1.0), (10, 5.0), (32, 40.0), (128, 200.0), (128, 1.0), (33, 1e3)):
        B = 300
        w_o, att_o, margin = vo.wood_sample(d, np.full(B, kappa), seed=1234567890123, draw=7)
        w = ctypes.c_float()
        same = 0
        for i in range(B):
            att = hc.hc_wood(d, kappa, 1234567890123, 7, i, ctypes.byref(w))
            if margin[i] < 1e-4:      # a decision within rounding of flipping: either outcome is fine
                same += 1
                continue
            assert att == att_o[i], (d, kappa, i, att, att_o[i])
            assert abs(w.value - w_o[i]) <= 2e-5, (d, kappa, i, w.value, w_o[i])
            same += 1
        assert same == B
        assert att_o.max() <= vo.MAX_ATTEMPTS


def test_wood_distribution_moments():
    """E[w] = A_d(kappa) = I_{d/2} / I_{d/2-1} and Var[w] = 1 - A^2 - (d-1) A / kappa."""
    for d, kappa in ((3, 2.0), (8, 10.0), (32, 30.0), (128, 150.0)):
        B = 20000
        w, att, _ = vo.wood_sample(d, np.full(B, kappa), seed=99, draw=3)
        A = float(vo.bessel_ratio(d / 2 - 1, kappa))
        var = 1 - A * A - (d - 1) * A / kappa
        assert abs(w.mean() - A) < 5 * np.sqrt(var / B), (d, kappa, w.mean(), A)
        assert abs(w.var() - var) < 0.06 * var, (d, kappa, w.var(), var)
        assert att.mean() < 1.6          # Wood's envelope accepts > 65 %
