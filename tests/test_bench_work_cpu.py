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


def test_roofline_kernel_ignores_side_stream_groups():
    """Groups that run on the library's side stream report a span (eligible -> done, starved by the chain's kernels), not
    time added to the step: they carry no roofline fraction and cannot be the kernel `roofline` is quoted on — in the
    cfg-5 line of profiles/r02_step13 the dlogits column sums (0.63 ms span) would otherwise have displaced K5 / Adam."""
    b = load_bench()
    w = dict(b.WORKLOAD)
    B = 4096
    work = b.kernel_work(w, B, True, 1)
    k1c = b.k1_contract_bytes(w, B, True)
    pk = {"hbm": 6500.0, "tc_sustained": 1364.0, "src": "test"}
    n_prof, ms_prof = 10, 24.6
    prof = {"nb_likelihood": (4.9, 10), "colsum_dlogits": (6.3, 10), "gemm_other": (5.1, 120), "senc_fwd": (2.44, 10),
            "colsum": (0.67, 20), "panel_aux_bwd": (0.64, 10), "clip_adam": (3.0, 10)}
    kernels, dom = b.summarise_kernels(prof, n_prof, ms_prof, work, k1c, pk, True, True)
    assert dom == "nb_likelihood"
    for name in ("colsum_dlogits", "colsum", "panel_aux_bwd"):
        assert kernels[name]["stream"] == "side" and "frac" not in kernels[name] and "bound" not in kernels[name]
    assert "frac" in kernels["gemm_other"] and kernels["gemm_other"]["stream"].startswith("step")
    assert abs(kernels["nb_likelihood"]["frac"] - work["nb_likelihood"][1] / 0.49e-3 / 1e9 / 6500.0) < 1e-3
    assert "frac_hbm_contract" in kernels["senc_fwd"] and kernels["senc_fwd"]["bound"] == "tensor"
    # one stream (MMV_SIDE=0): the same groups are ordinary members of the chain again
    kernels, dom = b.summarise_kernels(prof, n_prof, ms_prof, work, k1c, pk, True, False)
    assert dom == "colsum_dlogits" and "stream" not in kernels["colsum_dlogits"] and "frac" in kernels["colsum_dlogits"]
