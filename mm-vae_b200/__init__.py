"""mm-vae_b200 — B200-native minibatch ELBO step of YPARK/mm-vae.

This Python module is a thin ctypes binding over the C ABI declared in
include/mmvae_b200.h (libmmvae_b200.so, built in-tree by `make -C mm-vae_b200`
or __graft_entry__.build()).  It exists for the tests and bench.py; the
reference-facing host code is C++ (mm-vae_b200/host/, mirrors mmvae.hh /
mmvae_alg.hh / mmvae_io.hh / models/{nb,vmf}.hh).

There is NO CPU fallback: importing works anywhere (so the symbol table can be
checked), but every compute entry needs a B200 and raises otherwise, and a
missing library raises at import time.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmmvae_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "mmvae_b200.h")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `make -C {_HERE}` (or __graft_entry__.build()). "
        "mm-vae_b200 has no CPU or PyTorch fallback.")

lib = C.CDLL(LIB_PATH)

MMV_MAX_LAYERS = 8
MMV_NAME_LEN = 96
FWD, BWD, OPTIM, EPS_PHILOX, TERMS_DEFERRED = 1, 2, 4, 8, 16


class NbDims(C.Structure):
    _fields_ = [("x_dim", C.c_int64), ("c_dim", C.c_int64),
                ("n_enc", C.c_int32), ("enc_dims", C.c_int64 * MMV_MAX_LAYERS),
                ("n_dec", C.c_int32), ("dec_dims", C.c_int64 * MMV_MAX_LAYERS),
                ("mu_r_dim", C.c_int64), ("nu_h_dim", C.c_int64), ("nu_r_dim", C.c_int64),
                ("do_relu", C.c_int32)]


class VmfDims(C.Structure):
    _fields_ = [("x_dim", C.c_int64), ("c_dim", C.c_int64), ("z_dim", C.c_int64),
                ("n_enc", C.c_int32), ("enc_dims", C.c_int64 * MMV_MAX_LAYERS),
                ("n_dec", C.c_int32), ("dec_dims", C.c_int64 * MMV_MAX_LAYERS),
                ("kappa_min", C.c_float), ("kappa_max", C.c_float), ("do_relu", C.c_int32)]


class LossTerms(C.Structure):
    _fields_ = [("loss", C.c_float), ("recon", C.c_float), ("kl_mu", C.c_float),
                ("kl_nu", C.c_float), ("grad_norm", C.c_float)]

    def as_dict(self):
        return {k: float(getattr(self, k)) for k, _ in self._fields_}


_vp, _i64, _f32p = C.c_void_p, C.c_int64, C.POINTER(C.c_float)
_PROTOS = {
    "mmv_device_count": [C.POINTER(C.c_int)],
    "mmv_ctx_create": [C.c_int, C.POINTER(_vp)],
    "mmv_ctx_destroy": [_vp],
    "mmv_ctx_sync": [_vp],
    "mmv_timer_start": [_vp],
    "mmv_timer_stop": [_vp, C.POINTER(C.c_float)],
    "mmv_ctx_num_sms": [_vp, C.POINTER(C.c_int)],
    "mmv_launch_count": [C.POINTER(_i64)],
    "mmv_csr_upload": [_vp, _i64, _i64, _vp, _vp, _vp, _vp, _i64, C.POINTER(_vp)],
    "mmv_csr_create": [_vp, _i64, _i64, _i64, _i64, C.POINTER(_vp)],
    "mmv_csr_update": [_vp, _i64, _vp, _vp, _vp, _vp],
    "mmv_csr_wrap_device": [_vp, _i64, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _i64, C.POINTER(_vp)],
    "mmv_csr_destroy": [_vp],
    "mmv_host_alloc": [C.POINTER(_vp), _i64],
    "mmv_host_free": [_vp],
    "mmv_nb_model_create": [_vp, C.POINTER(NbDims), C.c_uint64, C.POINTER(_vp)],
    "mmv_vmf_model_create": [_vp, C.POINTER(VmfDims), C.c_uint64, C.POINTER(_vp)],
    "mmv_model_destroy": [_vp],
    "mmv_model_set_option": [_vp, C.c_char_p, C.c_double],
    "mmv_model_num_params": [_vp, C.POINTER(C.c_int)],
    "mmv_model_param_info": [_vp, C.c_int, C.c_char_p, C.POINTER(_i64), C.POINTER(_i64), C.POINTER(C.c_int)],
    "mmv_model_set_param": [_vp, C.c_char_p, _vp, _i64],
    "mmv_model_get_param": [_vp, C.c_char_p, _vp, _i64],
    "mmv_model_get_grad": [_vp, C.c_char_p, _vp, _i64],
    "mmv_step": [_vp, _vp, _vp, _i64, _vp, _vp, C.c_float, C.c_int, C.POINTER(LossTerms)],
    "mmv_step_collect": [_vp, C.POINTER(LossTerms)],
    "mmv_profile_read": [_vp, C.c_char_p, C.POINTER(C.c_float), C.POINTER(C.c_int)],
    "mmv_get_output": [_vp, C.c_char_p, _vp, _i64],
    "mmv_encode": [_vp, _vp, _vp, _i64, _vp, _vp],
    "mmv_test_gemm": [_vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, C.c_float, C.c_int, _vp],
    "mmv_gemm_bf16_noalloc": [_vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, C.c_float, C.c_int, _vp],
    "mmv_comm_unique_id": [_vp],
    "mmv_comm_init": [_vp, C.c_int, C.c_int, _vp],
    "mmv_comm_destroy": [_vp],
}
for _name, _args in _PROTOS.items():
    _fn = getattr(lib, _name)
    _fn.argtypes = _args
    _fn.restype = C.c_int
lib.mmv_last_error.restype = C.c_char_p
lib.mmv_last_error.argtypes = []
EXPORTED = sorted(list(_PROTOS) + ["mmv_last_error"])


class MMVError(RuntimeError):
    pass


def _chk(rc: int):
    if rc != 0:
        raise MMVError(lib.mmv_last_error().decode(errors="replace"))


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _arr(a, dtype):
    return None if a is None else np.ascontiguousarray(a, dtype=dtype)


class Context:
    """mmv_ctx: one device + one stream (replaces training_options_t::device)."""

    def __init__(self, device: int = 0):
        self.h = C.c_void_p()
        _chk(lib.mmv_ctx_create(device, C.byref(self.h)))

    def sync(self):
        _chk(lib.mmv_ctx_sync(self.h))

    def timer_start(self):
        _chk(lib.mmv_timer_start(self.h))

    def timer_stop(self) -> float:
        ms = C.c_float()
        _chk(lib.mmv_timer_stop(self.h, C.byref(ms)))
        return float(ms.value)

    @property
    def num_sms(self) -> int:
        n = C.c_int()
        _chk(lib.mmv_ctx_num_sms(self.h, C.byref(n)))
        return n.value

    def comm_init(self, nranks: int, rank: int, uid: bytes):
        buf = (C.c_ubyte * 128).from_buffer_copy(uid)
        _chk(lib.mmv_comm_init(self.h, nranks, rank, buf))

    def close(self):
        if self.h:
            lib.mmv_ctx_destroy(self.h)
            self.h = C.c_void_p()


def test_gemm(ctx, A, B, a_mn=False, b_mn=False, bias=None, beta=0.0, relu=False, C_in=None, use_tc=True):
    """Unit-test entry: A is [M,K] (or [K,M] if a_mn), B is [N,K] (or [K,N] if b_mn); returns C [M,N]."""
    A = _arr(A, np.float32)
    B = _arr(B, np.float32)
    M, K = (A.shape[1], A.shape[0]) if a_mn else A.shape
    N = B.shape[1] if b_mn else B.shape[0]
    out = np.zeros((M, N), np.float32) if C_in is None else np.array(C_in, dtype=np.float32, order="C")
    bias = _arr(bias, np.float32)
    _chk(lib.mmv_test_gemm(ctx.h, int(use_tc), M, N, K, int(a_mn), int(b_mn), _ptr(A), _ptr(B), _ptr(bias),
                           float(beta), int(relu), _ptr(out)))
    return out


NOALLOC_SCRATCH_BYTES = 4 << 20


def gemm_bf16_noalloc(ctx, M, N, K, d_A, d_B, d_C, a_mn=False, b_mn=False, d_bias=0, beta=0.0, relu=False):
    """mmv_gemm_bf16_noalloc: raw DEVICE pointers (e.g. torch tensor .data_ptr()); asynchronous on ctx's stream."""
    _chk(lib.mmv_gemm_bf16_noalloc(ctx.h, M, N, K, int(a_mn), int(b_mn), C.c_void_p(d_A), C.c_void_p(d_B),
                                   C.c_void_p(d_bias) if d_bias else None, float(beta), int(relu), C.c_void_p(d_C)))


def device_count() -> int:
    n = C.c_int()
    _chk(lib.mmv_device_count(C.byref(n)))
    return int(n.value)


def launch_count() -> int:
    n = _i64()
    _chk(lib.mmv_launch_count(C.byref(n)))
    return int(n.value)


def comm_unique_id() -> bytes:
    buf = (C.c_ubyte * 128)()
    _chk(lib.mmv_comm_unique_id(buf))
    return bytes(buf)


class CSR:
    """Device-resident cells x genes CSR (+ optional dense covariates)."""

    def __init__(self, ctx: Context, handle):
        self.ctx, self.h = ctx, handle

    @classmethod
    def upload(cls, ctx, indptr, indices, values, n_genes, covar=None, c_dim=1):
        indptr = _arr(indptr, np.int64)
        indices = _arr(indices, np.int32)
        values = _arr(values, np.float32)
        covar = _arr(covar, np.float32)
        if covar is not None:
            c_dim = covar.shape[1]
        h = C.c_void_p()
        _chk(lib.mmv_csr_upload(ctx.h, len(indptr) - 1, n_genes, _ptr(indptr), _ptr(indices), _ptr(values),
                                _ptr(covar), c_dim, C.byref(h)))
        return cls(ctx, h)

    @classmethod
    def from_dense(cls, ctx, x: np.ndarray, covar=None):
        """Rows of `x` = cells.  Helper for tests: dense -> CSR -> upload."""
        indptr, indices, values = dense_to_csr(x)
        return cls.upload(ctx, indptr, indices, values, x.shape[1], covar)

    @classmethod
    def create(cls, ctx, max_cells, max_nnz, n_genes, c_dim=0):
        h = C.c_void_p()
        _chk(lib.mmv_csr_create(ctx.h, max_cells, max_nnz, n_genes, c_dim, C.byref(h)))
        return cls(ctx, h)

    @classmethod
    def wrap_device(cls, ctx, n_cells, n_genes, nnz, max_row_nnz, d_indptr, d_indices, d_values, d_covar=0, c_dim=1):
        """Adopt device arrays (raw pointers, e.g. torch tensor .data_ptr()); caller keeps them alive."""
        h = C.c_void_p()
        _chk(lib.mmv_csr_wrap_device(ctx.h, n_cells, n_genes, nnz, max_row_nnz, C.c_void_p(d_indptr),
                                     C.c_void_p(d_indices), C.c_void_p(d_values),
                                     C.c_void_p(d_covar) if d_covar else None, c_dim, C.byref(h)))
        return cls(ctx, h)

    def update(self, n_cells, indptr, indices, values, covar=None):
        """Asynchronous H2D of a minibatch; the arrays must stay alive (pinned) until consumed."""
        _chk(lib.mmv_csr_update(self.h, n_cells, _ptr(indptr), _ptr(indices), _ptr(values), _ptr(covar)))

    def close(self):
        if self.h:
            lib.mmv_csr_destroy(self.h)
            self.h = C.c_void_p()


def dense_to_csr(x: np.ndarray) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    x = np.asarray(x)
    mask = x != 0
    indptr = np.concatenate([[0], np.cumsum(mask.sum(1))]).astype(np.int64)
    rows, cols = np.nonzero(mask)
    return indptr, cols.astype(np.int32), x[rows, cols].astype(np.float32)


def pinned_empty(shape, dtype) -> np.ndarray:
    """numpy view over cudaMallocHost memory (never freed: bench-lifetime buffers)."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    p = C.c_void_p()
    _chk(lib.mmv_host_alloc(C.byref(p), n))
    buf = (C.c_char * max(n, 1)).from_address(p.value)
    return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)


class Model:
    """mmv_model: nbvae_t (nb.hh) or vmf_vae_t (vmf.hh) behind the C ABI."""

    def __init__(self, ctx: Context, handle, kind: str):
        self.ctx, self.h, self.kind = ctx, handle, kind
        self._info = None

    @classmethod
    def nb(cls, ctx, x_dim, c_dim=1, enc=(), dec=(), mu_r=2, nu_h=1, nu_r=1, relu=False, seed=1):
        d = NbDims()
        d.x_dim, d.c_dim = x_dim, c_dim
        d.n_enc, d.n_dec = len(enc), len(dec)
        for i, v in enumerate(enc):
            d.enc_dims[i] = v
        for i, v in enumerate(dec):
            d.dec_dims[i] = v
        d.mu_r_dim, d.nu_h_dim, d.nu_r_dim, d.do_relu = mu_r, nu_h, nu_r, int(relu)
        h = C.c_void_p()
        _chk(lib.mmv_nb_model_create(ctx.h, C.byref(d), seed, C.byref(h)))
        return cls(ctx, h, "nb")

    @classmethod
    def vmf(cls, ctx, x_dim, c_dim=1, z_dim=2, enc=(), dec=(), kappa_min=0.1, kappa_max=10.0, relu=False, seed=1):
        d = VmfDims()
        d.x_dim, d.c_dim, d.z_dim = x_dim, c_dim, z_dim
        d.n_enc, d.n_dec = len(enc), len(dec)
        for i, v in enumerate(enc):
            d.enc_dims[i] = v
        for i, v in enumerate(dec):
            d.dec_dims[i] = v
        d.kappa_min, d.kappa_max, d.do_relu = kappa_min, kappa_max, int(relu)
        h = C.c_void_p()
        _chk(lib.mmv_vmf_model_create(ctx.h, C.byref(d), seed, C.byref(h)))
        return cls(ctx, h, "vmf")

    def set_option(self, key: str, value: float):
        _chk(lib.mmv_model_set_option(self.h, key.encode(), float(value)))

    def param_info(self) -> Dict[str, Tuple[int, int, bool]]:
        if self._info is None:
            n = C.c_int()
            _chk(lib.mmv_model_num_params(self.h, C.byref(n)))
            out = {}
            for i in range(n.value):
                name = C.create_string_buffer(MMV_NAME_LEN)
                r, c, reg = C.c_int64(), C.c_int64(), C.c_int()
                _chk(lib.mmv_model_param_info(self.h, i, name, C.byref(r), C.byref(c), C.byref(reg)))
                out[name.value.decode()] = (r.value, c.value, bool(reg.value))
            self._info = out
        return self._info

    def set_param(self, name: str, value: np.ndarray):
        v = _arr(value, np.float32)
        _chk(lib.mmv_model_set_param(self.h, name.encode(), _ptr(v), v.size))

    def set_params(self, params: Dict[str, np.ndarray]):
        for k, v in params.items():
            self.set_param(k, v)

    def _get(self, fn, name):
        r, c, _ = self.param_info()[name]
        out = np.empty(r * c, dtype=np.float32)
        _chk(fn(self.h, name.encode(), _ptr(out), out.size))
        return out

    def get_param(self, name: str) -> np.ndarray:
        return self._get(lib.mmv_model_get_param, name)

    def get_grad(self, name: str) -> np.ndarray:
        return self._get(lib.mmv_model_get_grad, name)

    def step(self, csr: CSR, cells, eps=None, kl_weight=1.0, flags=FWD | BWD | OPTIM, boot_idx=None,
             want_terms=True) -> Optional[dict]:
        cells = _arr(cells, np.int64)
        boot = _arr(boot_idx, np.int64)
        if isinstance(eps, (list, tuple)):
            eps = np.concatenate([np.asarray(e, dtype=np.float32).ravel() for e in eps])
        eps = _arr(eps, np.float32)
        terms = LossTerms()
        _chk(lib.mmv_step(self.h, csr.h, _ptr(cells), len(cells), _ptr(boot), _ptr(eps), kl_weight, flags,
                          C.byref(terms) if want_terms else None))
        return terms.as_dict() if want_terms else None

    def collect(self) -> dict:
        """Terms of the last step issued with TERMS_DEFERRED (a step call with that flag returns the PREVIOUS one's)."""
        terms = LossTerms()
        _chk(lib.mmv_step_collect(self.h, C.byref(terms)))
        return terms.as_dict()

    def profile_read(self, name: str = "") -> Tuple[float, int]:
        ms, cnt = C.c_float(), C.c_int()
        _chk(lib.mmv_profile_read(self.h, name.encode(), C.byref(ms), C.byref(cnt)))
        return float(ms.value), int(cnt.value)

    def output(self, name: str, n: int) -> np.ndarray:
        out = np.empty(n, dtype=np.float32)
        _chk(lib.mmv_get_output(self.h, name.encode(), _ptr(out), n))
        return out

    def encode(self, csr: CSR, cells, latent: int):
        cells = _arr(cells, np.int64)
        mean = np.empty((len(cells), latent), dtype=np.float32)
        lnvar = np.empty_like(mean)
        _chk(lib.mmv_encode(self.h, csr.h, _ptr(cells), len(cells), _ptr(mean), _ptr(lnvar)))
        return mean, lnvar

    def close(self):
        if self.h:
            lib.mmv_model_destroy(self.h)
            self.h = C.c_void_p()
