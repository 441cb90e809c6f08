"""Unit tests of the two GEMM kernels (K2) through the C ABI: the bf16 tcgen05 /
TMEM / TMA kernel against a float64 product of the bf16-ROUNDED operands (so only
accumulation order differs: tolerance 1e-5 relative), and the fp32 FFMA kernel."""
import importlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mmv():
    return importlib.import_module("mm-vae_b200")


@pytest.fixture(scope="module")
def ctx(mmv):
    c = mmv.Context(0)
    yield c
    c.close()


def bf16_round(x):
    return torch.from_numpy(x).to(torch.bfloat16).to(torch.float64).numpy()


SHAPES = [(128, 128, 64), (128, 128, 256), (256, 384, 1024), (200, 136, 72), (1000, 264, 520),
          (4096, 1024, 256), (72, 3000, 1024), (2048, 1024, 4096)]


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
def test_gemm_tc(mmv, ctx, M, N, K, a_mn, b_mn):
    if (a_mn and M % 8) or (b_mn and N % 8) or (not a_mn and K % 8) or (not b_mn and K % 8):
        pytest.skip("leading dimension not a multiple of 8")
    rng = np.random.default_rng(M * 7 + N * 3 + K)
    A = rng.normal(size=(M, K)).astype(np.float32)
    B = rng.normal(size=(N, K)).astype(np.float32)
    bias = rng.normal(size=N).astype(np.float32)
    C0 = rng.normal(size=(M, N)).astype(np.float32)
    ref = bf16_round(A) @ bf16_round(B).T + bias[None, :] + 0.5 * C0
    got = mmv.test_gemm(ctx, A.T.copy() if a_mn else A, B.T.copy() if b_mn else B, a_mn, b_mn, bias=bias,
                        beta=0.5, C_in=C0)
    err = np.abs(got - ref).max() / np.abs(ref).max()
    assert err < 1e-5, err
    got = mmv.test_gemm(ctx, A.T.copy() if a_mn else A, B.T.copy() if b_mn else B, a_mn, b_mn, relu=True)
    ref = np.maximum(bf16_round(A) @ bf16_round(B).T, 0)
    assert np.abs(got - ref).max() / np.abs(ref).max() < 1e-5


@pytest.mark.parametrize("M,N,K", [(37, 53, 19), (300, 130, 257)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True)])
def test_gemm_f32(mmv, ctx, M, N, K, a_mn, b_mn):
    rng = np.random.default_rng(5)
    A = rng.normal(size=(M, K)).astype(np.float32)
    B = rng.normal(size=(N, K)).astype(np.float32)
    ref = A.astype(np.float64) @ B.astype(np.float64).T
    got = mmv.test_gemm(ctx, A.T.copy() if a_mn else A, B.T.copy() if b_mn else B, a_mn, b_mn, use_tc=False)
    assert np.abs(got - ref).max() / np.abs(ref).max() < 1e-5


@pytest.mark.parametrize("M,N,K", [(1000, 20008, 200), (1152, 17000, 136)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("beta", [0.0, 1.0])
def test_gemm_tc_persistent(mmv, ctx, M, N, K, a_mn, b_mn, beta):
    """Many-tile products take the persistent 128 x 256 kernel (double-buffered TMEM accumulator, TMA
    store / reduce-add epilogue overlapped with the next tile): edge tiles in M and N, every layout."""
    if (a_mn and M % 8) or (b_mn and N % 8) or (not a_mn and K % 8) or (not b_mn and K % 8):
        pytest.skip("leading dimension not a multiple of 8")
    rng = np.random.default_rng(M + N + K)
    A = rng.normal(size=(M, K)).astype(np.float32)
    B = rng.normal(size=(N, K)).astype(np.float32)
    bias = rng.normal(size=N).astype(np.float32)
    C0 = rng.normal(size=(M, N)).astype(np.float32)
    ref = bf16_round(A) @ bf16_round(B).T + bias[None, :] + beta * C0
    got = mmv.test_gemm(ctx, A.T.copy() if a_mn else A, B.T.copy() if b_mn else B, a_mn, b_mn, bias=bias,
                        beta=beta, C_in=C0)
    err = np.abs(got - ref).max() / np.abs(ref).max()
    assert err < 1e-5, err
