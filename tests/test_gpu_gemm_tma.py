"""-m gpu: the persistent TMA-fed GEMM (gpr_b200/csrc/dgemm_tma.cu, option gemm_tma; off by default, kept as the measured
A/B alternative of profiles/r02_gemm_tma_probe.md) must give the cp.async kernel's results: products on ragged / lower /
beta shapes, the Cholesky's in-place call shape with ~1900 tiles (the case that exposed the setmaxnreg fault), and a whole
Learn against the oracle."""
import numpy as np
import pytest

from tests.synth import KERNELS, make_problem, rel_err

pytestmark = pytest.mark.gpu


def _gemm(h, M, N, K, alpha, beta, lower, A, B, Cm):
    h._ck(h.lib.gprb200_gemm_dev(h.h, 0, M, N, K, alpha, A.data_ptr(), A.stride(0), B.data_ptr(), B.stride(0), beta,
                                 Cm.data_ptr(), Cm.stride(0), lower))


def test_tma_gemm_matches_cp_async_kernel():
    import torch
    import gpr_b200
    h = gpr_b200.Handle(0)
    h.set_stream(torch.cuda.current_stream().cuda_stream)
    h.set_option("gemm_tma_min_tiles", 1)
    h.set_option("gemm_tma_min_k", 1)
    torch.manual_seed(1)
    for (M, N, K, alpha, beta, lower) in [(128, 128, 16, 1.0, 0.0, 0), (300, 200, 50, -1.0, 1.0, 0), (1000, 1000, 129, -1.0, 1.0, 1),
                                          (4097, 130, 1000, 2.0, 0.5, 0), (2050, 2050, 1024, -1.0, 1.0, 1)]:
        lda, ldc = (K + 1) // 2 * 2 + 2, (N + 1) // 2 * 2 + 4
        A = torch.randn(M, lda, dtype=torch.float64, device="cuda")[:, :K]
        B = A[:N] if (lower and M == N) else torch.randn(N, lda, dtype=torch.float64, device="cuda")[:, :K]
        C0 = torch.randn(M, ldc, dtype=torch.float64, device="cuda")
        out = {}
        for tma in (0, 1):
            h.set_option("gemm_tma", tma)
            Cf = C0.clone()
            n0 = h.launch_count()
            _gemm(h, M, N, K, alpha, beta, lower, A, B, Cf[:, :N])
            assert h.launch_count() == n0 + 1
            out[tma] = torch.tril(Cf[:, :N]) if lower else Cf[:, :N].clone()
        ref = alpha * (A @ B.T) + beta * C0[:, :N]
        ref = torch.tril(ref) if lower else ref
        scale = float(ref.abs().max())
        assert float((out[1] - ref).abs().max()) <= 1e-13 * scale
        assert float((out[1] - out[0]).abs().max()) <= 1e-13 * scale
    # the trailing update of the blocked Cholesky in place: operands and C inside one N x N buffer, 1830 tiles
    Nn, W = 8192, 256
    buf0 = torch.randn(Nn, Nn, dtype=torch.float64, device="cuda")
    r0 = 2 * W
    keep = {}
    for tma in (0, 1):
        h.set_option("gemm_tma", tma)
        buf = buf0.clone()
        _gemm(h, Nn - r0, Nn - r0, W, -1.0, 1.0, 1, buf[r0:, :W], buf[r0:, :W], buf[r0:, r0:])
        torch.cuda.synchronize()
        keep[tma] = torch.tril(buf[r0:, r0:]).clone()
        del buf
    assert float((keep[1] - keep[0]).abs().max()) <= 1e-11
    h.close()


def test_learn_with_tma_updates_matches_oracle():
    import gpr_b200
    from oracle import gpr_oracle as orc
    X, Y = make_problem(2600, 4, 2, seed=31)
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(KERNELS["sum3"]).lower()
    h.set_kernel(ops, params)
    h.set_option("gemm_tma", 1)
    h.set_option("gemm_tma_min_tiles", 1)
    h.set_option("nbo", 256)
    h.set_train(X, Y, 1.0)
    h.learn(0)
    o = orc.OracleGP(KERNELS["sum3"], X, Y, 1.0)
    o.learn()
    assert rel_err(h.alpha(), o.alpha) <= 1e-9
    h.close()
