"""tcgen05 TF32 slab GEMM self-test: descriptors, swizzled operand staging (K-major and MN-major), TMEM epilogue,
cluster (DSMEM) and deferred split-K against NumPy on tf32-truncated operands."""
import numpy as np
import pytest
import torch

from dqn_mgsc_zoo_b200 import _lib

pytestmark = pytest.mark.gpu


def tf32_trunc(x):
    """The tensor core reads fp32 operands as TF32 by ignoring the low 13 mantissa bits."""
    return (np.ascontiguousarray(x, np.float32).view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)


@pytest.mark.parametrize("M,N,K,a_mn,b_mn,bn,splits", [
    (128, 64, 32, 0, 0, 64, 1), (128, 64, 256, 0, 0, 64, 1), (300, 64, 576, 0, 0, 64, 3),
    (200, 32, 256, 0, 0, 32, 1), (256, 128, 512, 0, 0, 128, 4), (300, 100, 1000, 0, 0, 64, 4),
    (128, 64, 128, 1, 0, 64, 1), (128, 64, 128, 0, 1, 64, 1), (260, 64, 200, 1, 1, 64, 1),
    (132, 32, 256, 0, 1, 32, 1), (512, 96, 32, 1, 1, 128, 1), (3140, 64, 96, 1, 1, 64, 3),
    (260, 32, 12800, 1, 1, 32, 50), (64, 512, 3136, 0, 1, 128, 25),
])
def test_tf32_gemm(M, N, K, a_mn, b_mn, bn, splits):
    rng = np.random.default_rng(M + N + K)
    a = rng.normal(size=(M, K)).astype(np.float32)
    b = rng.normal(size=(N, K)).astype(np.float32)
    want = tf32_trunc(a).astype(np.float64) @ tf32_trunc(b).astype(np.float64).T
    ta = torch.from_numpy(np.ascontiguousarray(a.T if a_mn else a)).cuda()
    tb = torch.from_numpy(np.ascontiguousarray(b.T if b_mn else b)).cuda()
    tc = torch.full((M, N), float("nan"), device="cuda")
    _lib.call("dqnb_selftest_tf32_gemm", ta.data_ptr(), tb.data_ptr(), tc.data_ptr(), M, N, K, a_mn, b_mn, bn,
              splits, 0)
    got = tc.cpu().numpy().astype(np.float64)
    err = np.abs(got - want).max() / np.abs(want).max()
    exact = a.astype(np.float64) @ b.astype(np.float64).T
    err_exact = np.abs(got - exact).max() / np.abs(exact).max()
    print("tf32 gemm M%d N%d K%d: err vs truncated operands %.2e, vs exact %.2e" % (M, N, K, err, err_exact))
    assert err < 2e-5, err
    assert err_exact < 2e-3, err_exact
