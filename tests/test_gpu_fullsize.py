"""-m gpu: the BASELINE.json configurations at their FULL sizes, checked through size-independent properties (the CPU
oracle's O(N^3) inverse cannot be rerun per commit at these sizes):

  C3  N = 32768, Sum(Sum(Gaussian, Periodic), White), M = 3   residual of (K + s^2 I) alpha = y on sampled rows,
                                                              linearity of alpha in y, Predict(x_i) = y_i - s^2 alpha_i
  C2  N = 4096, d = 30, M = 786432 (Y = 25.8 GB on the device) the same identity over all 786432 columns and column
                                                              independence against single-column solves
  C4  N = 16384, RationalQuadratic                            gradient of the log-likelihood against five-point
                                                              differences of its (unclamped) value
Inputs are seeded; tolerances are fp64 with cond(K + s^2 I) <= N + 1 (sigma_n = 1, unit scales)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

C3_KERNEL = "SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))"


class _DeviceView:
    """zero-copy torch view of a device buffer owned by the library (CUDA array interface)"""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f8", "data": (int(ptr), False), "version": 2}


def _oracle_rows(ks, X, rows, chunk=32):
    """k(x_r, x_j) for the sampled rows r and every training point j, evaluated on the CPU by the oracle."""
    from oracle import gpr_oracle as orc
    k = orc.parse_kernel(ks)
    return np.concatenate([orc.kernel_matrix(k, X[rows[i:i + chunk]], X) for i in range(0, len(rows), chunk)], axis=0)


def _handle(ks, X, Y, sigma):
    import gpr_b200
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(ks).lower()
    h.set_kernel(ops, params)
    h.set_train(X, Y, sigma)
    return h, ops, np.array(params, dtype=float)


def test_c3_learn_predict_properties_at_n_32768():
    rng = np.random.default_rng(32768)
    N, d, sigma = 32768, 8, 1.0
    X = rng.uniform(-1, 1, (N, d))
    y1 = np.sin(3 * X[:, 0]) + 0.5 * np.cos(2 * X[:, 1] + X[:, 2]) + 0.1 * rng.standard_normal(N)
    y2 = np.tanh(X @ rng.standard_normal(d)) + 0.1 * rng.standard_normal(N)
    Y = np.stack([y1, y2, y1 + 2.0 * y2], axis=1)
    h, _, _ = _handle(C3_KERNEL, X, Y, sigma)
    h.learn(0)
    A = h.alpha()
    assert np.isfinite(A).all()
    # linearity of the solve in the labels
    assert np.abs(A[:, 2] - (A[:, 0] + 2.0 * A[:, 1])).max() <= 1e-10 * np.abs(A).max()
    # defining equation on sampled rows. The kernel rows come from the CPU ORACLE (oracle.gpr_oracle.kernel_matrix, the
    # numpy restatement of include/Kernel.h pinned to oracle/_ref), not from the device functor that built K: a kernel
    # value that is wrong the same way in the Gram build and in kernel_vector would otherwise cancel out.
    rows = rng.choice(N, 192, replace=False)
    Kr = _oracle_rows(C3_KERNEL, X, rows)                           # 192 x N, no noise term
    resid = Kr @ A + sigma * sigma * A[rows] - Y[rows]
    assert np.abs(resid).max() <= 1e-10 * np.abs(Y).max()
    # and the device's own kernel rows agree with the oracle's element-wise (the Gram parity bar, 1e-13)
    assert np.abs(h.kernel_vector(X[rows[:16]]) - Kr[:16]).max() <= 1e-13 * np.abs(Kr).max()
    # Predict at a training point: K_i alpha = y_i - s^2 alpha_i; the credible interval is real and non-negative
    mean, ci = h.predict(X[rows], want_ci=True)
    assert np.abs(mean - (Y[rows] - sigma * sigma * A[rows])).max() <= 1e-9 * np.abs(Y).max()
    assert np.isfinite(ci).all() and (ci >= 0).all()
    # log det from the factor is finite and below the Hadamard bound sum log(K_ii + s^2)
    kdiag = Kr[np.arange(len(rows)), rows].mean() + sigma * sigma
    assert np.isfinite(h.logdet()) and h.logdet() <= N * np.log(kdiag) + 1e-6
    h.close()


def test_c2_all_786432_columns_satisfy_the_training_point_identity():
    import torch
    import gpr_b200
    N, d, M, sigma = 4096, 30, 786432, 1.0
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev)
    g.manual_seed(4096)
    X = torch.randn(N, d, dtype=torch.float64, device=dev, generator=g)
    W = torch.randn(d, M, dtype=torch.float64, device=dev, generator=g) / d ** 0.5
    Y = torch.tanh(X @ W)
    Y += torch.randn(N, M, dtype=torch.float64, device=dev, generator=g)
    del W
    ks = "GaussianKernel(8.21583836257749,1,)"          # sigma = 1.5 sqrt(30)
    h = gpr_b200.Handle(0)
    h.set_stream(torch.cuda.current_stream(dev).cuda_stream)
    ops, params = gpr_b200.KernelFactory.GetKernel(ks).lower()
    h.set_kernel(ops, params)
    h.set_train_dev(X.data_ptr(), N, d, Y.data_ptr(), M, sigma)
    h.learn(0)
    h.synchronize()
    ap, ld = h.alpha_dev()
    alpha = torch.as_tensor(_DeviceView(ap, (N, ld)), device=dev)[:, :M]    # the library's own alpha, no copy
    rows = [0, 1777, 4095]
    mean = torch.empty(1, M, dtype=torch.float64, device=dev)
    scale = float(Y.abs().max())
    for i in rows:
        h.predict_dev(X[i:i + 1].data_ptr(), 1, mean.data_ptr(), M)
        torch.cuda.synchronize()
        err = (mean[0] - (Y[i] - sigma * sigma * alpha[i])).abs().max().item()
        assert err <= 1e-9 * scale
    # column independence: three columns solved on their own give the same regression vectors
    cols = [0, 393216, 786431]
    Xh = X.cpu().numpy()
    Yc = Y[:, cols].cpu().numpy()
    Afull = alpha[:, cols].cpu().numpy()
    del alpha
    h.close()
    del Y, mean
    torch.cuda.empty_cache()
    h1 = gpr_b200.Handle(0)
    h1.set_kernel(ops, params)
    h1.set_train(Xh, Yc, sigma)
    h1.learn(0)
    A1 = h1.alpha()
    assert np.abs(A1 - Afull).max() <= 1e-10 * np.abs(A1).max()
    h1.close()


def test_c4_loglik_gradient_is_the_derivative_of_the_value_at_n_16384():
    import gpr_b200
    rng = np.random.default_rng(16384)
    N, d = 16384, 8
    X = rng.uniform(-1, 1, (N, d))
    Y = (np.sin(3 * X[:, 0]) + 0.5 * X[:, 1] ** 2 + 0.3 * rng.standard_normal(N))[:, None]
    h, ops, p0 = _handle("RationalQuadraticKernel(1,1.5,2,)", X, Y, 1.0)
    TRUE = gpr_b200.capi.LOGLIK_TRUE
    v, g, _ = h.loglik(TRUE, True, True, False)
    assert np.isfinite(v).all() and np.isfinite(g).all()
    # alpha of this iteration against oracle-evaluated kernel rows (independent of the device functor)
    A = h.alpha()
    rows = rng.choice(N, 128, replace=False)
    Kr = _oracle_rows("RationalQuadraticKernel(1,1.5,2,)", X, rows)
    assert np.abs(Kr @ A + A[rows] - Y[rows]).max() <= 1e-10 * np.abs(Y).max()
    # value = -1/2 y^T alpha - 1/2 log det - N/2 log 2 pi with the quadratic form recomputed on the host
    assert abs(v[0] - (-0.5 * float(Y[:, 0] @ A[:, 0]) - 0.5 * h.logdet() - 0.5 * N * np.log(2 * np.pi))) <= 1e-8 * abs(v[0])

    def value(p):
        h.set_kernel(ops, p)
        return h.loglik(TRUE, True, False, False)[0][0]

    fd = np.zeros(len(p0))
    for i in range(len(p0)):
        e = np.zeros(len(p0))
        e[i] = 1e-3 * max(1.0, abs(p0[i]))
        fd[i] = (-value(p0 + 2 * e) + 8 * value(p0 + e) - 8 * value(p0 - e) + value(p0 - 2 * e)) / (12 * e[i])
    assert np.abs(g - fd).max() <= 1e-6 * np.abs(fd).max()
    # the reference's clamped value (include/Likelihood.h:180-188) stays finite although det leaves the double range
    h.set_kernel(ops, p0)
    vc = h.loglik(gpr_b200.capi.LOGLIK_REFERENCE_CLAMP, True, False, False)[0][0]
    assert np.isfinite(vc)
    h.close()
