"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports
every symbol include/*.h declares, and refuses to compute without a GPU."""
import importlib
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def mmv():
    return importlib.import_module("mm-vae_b200")


def _declared_symbols():
    syms = set()
    inc = os.path.join(ROOT, "include")
    for fn in os.listdir(inc):
        if fn.endswith(".h"):
            text = open(os.path.join(inc, fn)).read()
            syms |= set(re.findall(r"MMV_API[^;()]*?\b(mmv_[a-z0-9_]+)\s*\(", text))
    return syms


def test_library_exports_every_declared_symbol(mmv):
    declared = _declared_symbols()
    assert len(declared) >= 25
    for s in sorted(declared):
        assert hasattr(mmv.lib, s), f"{s} declared in include/ but not exported"
    assert declared == set(mmv.EXPORTED), declared ^ set(mmv.EXPORTED)


def test_struct_layouts_match_header(mmv):
    # sizes implied by include/mmvae_b200.h (LP64): catches binding drift
    assert mmv.C.sizeof(mmv.NbDims) == 8 + 8 + 8 + 64 + 8 + 64 + 8 + 8 + 8 + 8
    assert mmv.C.sizeof(mmv.VmfDims) == 8 + 8 + 8 + 8 + 64 + 8 + 64 + 4 + 4 + 8
    assert mmv.C.sizeof(mmv.LossTerms) == 20


def test_no_cpu_fallback(mmv):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(mmv.MMVError, match="no CPU fallback|CUDA"):
        mmv.Context(0)


def test_dense_to_csr_roundtrip(mmv):
    rng = np.random.default_rng(0)
    x = (rng.random((7, 13)) < 0.3) * rng.integers(1, 9, (7, 13))
    indptr, idx, val = mmv.dense_to_csr(x.astype(np.float32))
    y = np.zeros_like(x, dtype=np.float32)
    for i in range(7):
        y[i, idx[indptr[i]:indptr[i + 1]]] = val[indptr[i]:indptr[i + 1]]
    assert np.array_equal(x.astype(np.float32), y)
