"""The algorithmic work model behind bench.py's roofline numbers (SURVEY 8d contract figures) — CPU."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_bench():
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_contract_figures_cfg3():
    b = load_bench()
    w = dict(b.WORKLOAD, genes=20_000)
    B, D, nnz = 4096, 20_000, 0.05 * 20_000 * 4096
    k32 = b.kernel_work(w, B, bf16=False)
    # 8 * 81.9 M + 32.8 M ~ 688 MB ; K1 fwd ~ 90.9 MB ; fwd+bwd dense 126 MFLOP/cell
    assert abs(k32["nb_likelihood"][1] - 688e6) < 1e6
    assert abs(k32["senc_fwd"][1] - (8 * nnz + 8 * (B + 1) + 4 * D * 1024 + 4 * D * 4 + 4 * B * 1026)) < 1
    kb = b.kernel_work(w, B, bf16=True)
    # tensor-core mode: fp32 logits in, bf16 gradient out for BOTH likelihood kernels
    assert kb["nb_likelihood"][1] == 6.0 * B * D + 8 * nnz + 4.0 * D * 4 + 16 * B
    assert kb["vmf_likelihood"][1] == 6.0 * B * D + 8 * nnz + 4.0 * D * 1 + 16 * B
    dense = 3 * kb["gemm_dec_last_fwd"][1] + kb["gemm_other"][1]
    assert abs(dense / B - 126e6) < 1.5e6
    assert kb["senc_fwd"] == ("tensor", 2.0 * B * D * 1024)
