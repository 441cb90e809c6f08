"""Shared helpers for the tests (golden loading; no product code here)."""
import json
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

MODEL_CASES = ["nb_small", "nb_deep", "nb_nohidden", "nb_trainall", "nb_wide",
               "vmf_small", "vmf_deep", "vmf_nohidden", "vmf_trainall", "vmf_bigkappa",
               "nb_relu", "vmf_relu", "vmf_relu_trainall"]


class Golden:
    """One tests/golden/<case>.npz produced by the reference harness."""

    def __init__(self, name):
        z = np.load(os.path.join(GOLDEN, name + ".npz"))
        self.a = {k: z[k] for k in z.files}
        self.cfg = json.loads(bytes(self.a["config"]).decode())
        self.name = name

    def group(self, prefix):
        """{'name': array} of all arrays called '<prefix>__name'."""
        p = prefix + "__"
        return {k[len(p):]: v for k, v in self.a.items() if k.startswith(p)}

    def params(self, tag="init"):
        return self.group(tag + "__param")

    def eps(self, s):
        out, k = [], 0
        while f"s{s}__eps{k}" in self.a:
            out.append(self.a[f"s{s}__eps{k}"])
            k += 1
        return out

    def batch(self, s):
        r = self.a["ridx"][s]
        return self.a["x"][r], self.a["c"][r]


def rel_err(a, b):
    """Relative Frobenius error ||a-b|| / max(||b||, tiny)."""
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def to_t(d):
    return {k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in d.items()}
