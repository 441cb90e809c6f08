"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's ELBO step.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module.  The product path (mm-vae_b200/) never
does; it fails loudly if its CUDA library is missing.

Every function restates reference code (file:line cited, paths relative to the
reference root) with the same ATen ops in the same order, on torch CPU, so
that the fp32 arithmetic is the reference's own.  Parity pin: the golden
vectors in tests/golden/*.npz were produced by oracle/_ref/ref_model_harness,
which compiles the reference's headers verbatim; tests/test_oracle_golden.py
checks this restatement against them.

North-star items with no reference code (Wood sampler, Householder
reflection, vMF/Bessel KL): parity unpinned — not restated here.
"""
from __future__ import annotations

import math
import re
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# --------------------------------------------------------------------------
# fastapprox scalars  (include/utils/fastlog.h:75-85, fastgamma.h:58-60)
# --------------------------------------------------------------------------
def fasterlog(x: float) -> float:
    """fastlog.h:75-85: float(bits(x)) * 8.2629582881927490e-8f - 87.989971088f"""
    bits = np.array([x], dtype=np.float32).view(np.uint32)[0]
    y = np.float32(bits)  # uint32 -> float conversion (rounds to nearest)
    y = np.float32(y * np.float32(8.2629582881927490e-8))
    return float(np.float32(y - np.float32(87.989971088)))


def fasterlgamma(x: float) -> float:
    """fastgamma.h:58-60 (all float arithmetic)."""
    f = np.float32
    x = f(x)
    t = f(f(-0.0810614667) - x)
    t = f(t - f(fasterlog(float(x))))
    t = f(t + f(f(f(0.5) + x) * f(fasterlog(float(f(f(1.0) + x))))))
    return float(t)


# --------------------------------------------------------------------------
# lbessel  (include/operators.hh:13-101)
# --------------------------------------------------------------------------
class _LBessel(torch.autograd.Function):
    @staticmethod
    def forward(ctx, kappa: Tensor, df: float) -> Tensor:
        # operators.hh:66-80; eta is stored as float (operators.hh:69)
        nu = float(df)
        eta = float(np.float32((nu + 0.5) / (2.0 * (nu + 1.0))))
        k = kappa.detach()
        stuff1 = nu * torch.log(k) + eta * k - (eta + nu) * math.log(2.0) - fasterlgamma(nu + 1)
        stuff2 = k - 0.5 * torch.log(k) - 0.5 * math.log(2.0 * math.pi)
        res = torch.le(k, nu).type_as(k).mul(stuff1) + torch.gt(k, nu).type_as(k).mul(stuff2)
        ctx.save_for_backward(kappa)
        ctx.df = np.float32(nu)  # backward reads `float df = df_.toDouble()` (operators.hh:22)
        return res

    @staticmethod
    def backward(ctx, grad):  # noqa: D401
        # operators.hh:28-37: Baricz-bound midpoint; the incoming grad is IGNORED.
        (x,) = ctx.saved_tensors
        df = float(ctx.df)
        lb = torch.sqrt(x * x * df / (df + 1.0) + df * df)
        ub = torch.sqrt(x * x + df * df)
        return 0.5 * (lb + ub) / x, None


def lbessel(kappa: Tensor, df: float) -> Tensor:
    return _LBessel.apply(kappa, df)


# --------------------------------------------------------------------------
# layer discovery from reference parameter names
# --------------------------------------------------------------------------
def _seq_layers(P: Dict[str, Tensor], stem: str, final: Optional[str]) -> List[str]:
    """Names of the Linear/Angular layers of a reference Sequential, in order.

    nb.hh:331-349 / :366-379, vmf.hh:336-355 / :372-385: hidden layers are
    '<stem>_<l>' (1-based); the closing layer is '<final>' (always for the
    decoders; for the encoders only when no hidden layer was asked for; the
    vmf encoder's closing layer is pushed unnamed -> key '0').
    """
    names = []
    l = 1
    while f"{stem}_{l}.weight" in P:
        names.append(f"{stem}_{l}")
        l += 1
    if final is not None and f"{final}.weight" in P:
        names.append(final)
    return names


def _linear(P, name, h):
    b = P.get(name + ".bias")
    return F.linear(h, P[name + ".weight"], b)


def _angular(P, name, h):
    # modules/angular.hh:34-42
    ww = F.normalize(F.relu(P[name + ".weight"]) + 1e-4, p=2, dim=1)
    return F.linear(h, ww, None)


def kl_loss(mean: Tensor, lnvar: Tensor) -> Tensor:
    # nb.hh:533-537, vmf.hh:410-414
    return -0.5 * torch.sum(1 + lnvar - mean.pow(2) - lnvar.exp())


def reparameterize(mu: Tensor, lnvar: Tensor, eps: Optional[Tensor]) -> Tensor:
    # nb.hh:462-472, vmf.hh:394-404 (eps injected instead of randn_like)
    if eps is None:
        return mu
    sig = lnvar.div(2.0).exp()
    return mu + eps.mul(sig)


# --------------------------------------------------------------------------
# NB-VAE  (include/models/nb.hh)
# --------------------------------------------------------------------------
def nb_encode_mu(P, x, c=None, relu=False):
    """nb.hh:403-431 (c=None is the recorder overload :419-431)."""
    eps = 1e-4
    x_sd = F.softplus(P["ln_x_sd"])
    h = torch.div(x.log1p() - P["x_mean"], x_sd + eps)
    for name in _seq_layers(P, "mu_encoding", "mu_encoding"):
        h = _linear(P, name, h)
        if relu:
            h = F.relu(h)
    lnvar = torch.clamp(_linear(P, "mu_representation_logvariance", h), -4.0, 4.0)
    mean = _linear(P, "mu_representation_mean", h)
    if c is not None:
        mean = mean + _linear(P, "covar_encoding", c)
    return mean, lnvar


def nb_forward(P, x, c, eps_mu, eps_nu, relu=False):
    """nb.hh:474-508.  eps_* = None reproduces eval mode (z = mean)."""
    mu_mean, mu_lnvar = nb_encode_mu(P, x, c, relu)
    z = reparameterize(mu_mean, mu_lnvar, eps_mu)
    # decode_mu nb.hh:433-442
    h = z
    dec = _seq_layers(P, "mu_decoding", "mu_decoding")
    for i, name in enumerate(dec):
        h = _linear(P, name, h)
        if relu and i + 1 < len(dec):
            h = F.relu(h)
    hc = _linear(P, "covar_decoding", c)
    recon_mu = torch.exp(torch.log_softmax(h + hc + P["mu_bias"], 1))
    # encode_nu nb.hh:444-451 (raw counts)
    hn = _linear(P, "nu_encoding", x)
    nu_lnvar = torch.clamp(_linear(P, "nu_representation_logvariance", hn), -4.0, 4.0)
    nu_mean = _linear(P, "nu_representation_mean", hn)
    zn = reparameterize(nu_mean, nu_lnvar, eps_nu)
    # decode_nu nb.hh:453-460
    recon_nu = torch.clamp(F.softplus(_linear(P, "nu_decoding", zn) - P["nu_bias"]), 1e-4, 1e4)
    # depth nb.hh:498
    depth = F.softplus(_linear(P, "depth", x))
    return dict(recon_mu=recon_mu, recon_nu=recon_nu, recon_depth=depth,
                mu_mean=mu_mean, mu_lnvar=mu_lnvar, nu_mean=nu_mean, nu_lnvar=nu_lnvar)


def nb_nllik(x, y):
    """nb.hh:511-531."""
    eps = 1e-4
    nu = y["recon_nu"] + eps
    mu = y["recon_mu"] * y["recon_depth"] + eps
    lg = torch.lgamma(nu) + torch.lgamma(x + 1.0)
    lg = lg - torch.lgamma(nu + x)
    denom = torch.log(mu + nu)
    pr = x.mul(denom - torch.log(mu))
    pr = pr + nu.mul(denom - torch.log(nu))
    return torch.sum(lg + pr)


def nb_loss(x, y, kl_weight=1.0):
    """nb.hh:539-548 -> dict(loss, nllik, kl_mu, kl_nu)."""
    nllik = nb_nllik(x, y)
    n = float(x.size(0))
    kl_mu = kl_loss(y["mu_mean"], y["mu_lnvar"])
    kl_nu = kl_loss(y["nu_mean"], y["nu_lnvar"])
    tot = (nllik + kl_mu * kl_weight + kl_nu * kl_weight) / n
    return dict(loss=tot, nllik=nllik, kl_mu=kl_mu, kl_nu=kl_nu)


# --------------------------------------------------------------------------
# vMF-VAE  (include/models/vmf.hh)
# --------------------------------------------------------------------------
def vmf_encode(P, x, c=None, relu=False):
    """vmf.hh:250-281."""
    eps = float(np.float32(1e-2) / np.float32(x.size(1)))
    xn = F.normalize(x.log1p(), p=2, dim=1)
    h = torch.div(torch.sub(xn, P["x_mean"]), F.softplus(P["ln_x_sd"]) + eps)
    names = _seq_layers(P, "encoding", None)
    if not names:
        names = ["0"]  # vmf.hh:350 pushes the closing Angular unnamed
    for name in names:
        h = _angular(P, name, h)
        if relu:
            h = F.relu(h)
    lnvar = torch.clamp(_linear(P, "representation_logvariance", h), -4.0, 4.0)
    mean = _linear(P, "representation_mean", h)
    if c is not None:
        mean = mean + _linear(P, "covar_encoding", c)
    return mean, lnvar


def vmf_forward(P, x, c, eps_z, kap_min=0.1, kap_max=10.0, relu=False):
    """vmf.hh:292-304."""
    mean, lnvar = vmf_encode(P, x, c, relu)
    z = reparameterize(mean, lnvar, eps_z)
    h = z
    dec = _seq_layers(P, "decoding", "decoding")
    for i, name in enumerate(dec):
        h = _linear(P, name, h)
        if relu and i + 1 < len(dec):
            h = F.relu(h)
    h = torch.exp(h)                                    # vmf.hh:287
    hc = _linear(P, "covar_decoding_", c)               # vmf.hh:288
    recon = F.normalize(h + hc, p=2, dim=1)
    kappa = torch.clamp(torch.exp(P["ln_kappa"]), kap_min, kap_max)  # vmf.hh:301
    return dict(recon=recon, mean=mean, lnvar=lnvar, kappa=kappa)


def vmf_loss(x, y, kl_weight=1.0):
    """vmf.hh:419-440 -> dict(loss, kl, llik_sum)."""
    D = x.size(1)
    eps = float(np.float32(1e-2) / np.float32(D))
    yobs = F.normalize(F.relu(x).log1p() + eps, p=2, dim=1)
    n = float(yobs.size(0))
    dd = float(D)
    df = float(np.float32(max(0.5 * dd - 1.0, 0.0)))
    kappa = y["kappa"]
    kl = kl_loss(y["mean"], y["lnvar"])
    llik = torch.sum(yobs * y["recon"], 1) * kappa
    llik = llik + (df * torch.log(kappa) - lbessel(kappa, df))
    llik = llik - 0.5 * dd * fasterlog(2.0 * math.pi)
    tot = kl / n * kl_weight - llik.sum() / n
    return dict(loss=tot, kl=kl, llik_sum=llik.sum())


# --------------------------------------------------------------------------
# optimiser  (mmvae_alg.hh:234-236,306-310; libtorch Adam / clip_grad_norm_)
# --------------------------------------------------------------------------
def clip_grad_norm_(grads: List[Tensor], max_norm: float) -> float:
    """libtorch clip_grad_norm_: scale = max_norm/(total+1e-6), applied if < 1."""
    total = torch.linalg.vector_norm(torch.stack([torch.linalg.vector_norm(g) for g in grads]))
    coef = max_norm / (float(total) + 1e-6)
    if coef < 1.0:
        for g in grads:
            g.mul_(coef)
    return float(total)


class Adam:
    """libtorch torch::optim::Adam with weight_decay as L2 (not AdamW)."""

    def __init__(self, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-4):
        self.lr, self.b1, self.b2, self.eps, self.wd = lr, betas[0], betas[1], eps, weight_decay
        self.state: Dict[str, Tuple[int, Tensor, Tensor]] = {}

    def step(self, params: Dict[str, Tensor], grads: Dict[str, Tensor]):
        for k, g in grads.items():
            p = params[k]
            t, m, v = self.state.get(k, (0, torch.zeros_like(p), torch.zeros_like(p)))
            t += 1
            g = g + self.wd * p
            m = m * self.b1 + (1 - self.b1) * g
            v = v * self.b2 + (1 - self.b2) * g * g
            bc1 = 1 - self.b1 ** t
            bc2 = 1 - self.b2 ** t
            denom = v.sqrt() / math.sqrt(bc2) + self.eps
            # addcdiv_(exp_avg, denom, -step_size): self + (value * t1) / t2 (ATen CPU order)
            params[k] = p + ((-(self.lr / bc1)) * m) / denom
            self.state[k] = (t, m, v)


def run_step(model: str, P: Dict[str, Tensor], x, c, eps, kl_weight, **kw):
    """One forward+loss+backward (mmvae_alg.hh:303-307). Returns (terms, outs, grads)."""
    Pg = {k: v.detach().clone().requires_grad_(True) for k, v in P.items()}
    if model == "nb":
        y = nb_forward(Pg, x, c, eps[0], eps[1], relu=kw.get("relu", False))
        terms = nb_loss(x, y, kl_weight)
    else:
        y = vmf_forward(Pg, x, c, eps[0], kw.get("kap_min", 0.1), kw.get("kap_max", 10.0),
                        relu=kw.get("relu", False))
        terms = vmf_loss(x, y, kl_weight)
    terms["loss"].sum().backward()
    grads = {k: (v.grad.detach() if v.grad is not None else torch.zeros_like(v)) for k, v in Pg.items()}
    return ({k: v.detach() for k, v in terms.items()}, {k: v.detach() for k, v in y.items()}, grads)
