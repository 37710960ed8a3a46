"""-m gpu: SparseGaussianLogLikelihood on the device (csrc/sparse.cu: gprb200_sparse_loglik through the C-ABI) against
the golden vectors of the reference's own class, the CPU oracle on seeded inputs, and - at n = 262144 where the
reference's n x n matrices (550 GB each, P + 1 of them) cannot exist - the property that the gradient is the derivative
of the value."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# (value, gradient) relative tolerances; the loose ones are the reference's own conditioning, see tests/test_oracle_sparselik.py
TOL = {"sparselik_gauss": (1e-8, 1e-8), "sparselik_sum": (1e-8, 1e-8), "sparselik_prod_white": (1e-8, 1e-8),
       "sparselik_rq_ragged": (1e-7, 1e-6), "sparselik_clamp": (1e-7, 1e-4)}
CASES = sorted(TOL)


def load(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    return {k: (z[k].item() if z[k].ndim == 0 else z[k]) for k in z.files}


def rel(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(np.asarray(b)).max(), 1e-300)


def build(ks, X, Y, Xm, sigma, jitter):
    import gpr_b200
    gp = gpr_b200.SparseGaussianProcess(gpr_b200.KernelFactory.GetKernel(ks), jitter)
    gp.SetSigma(sigma)
    for x, y in zip(X, Y):
        gp.AddSample(x, y)
    for x in Xm:
        gp.AddInducingSample(x, np.zeros(Y.shape[1]))
    return gp


@pytest.mark.parametrize("name", CASES)
def test_sparse_likelihood_matches_reference_golden(name):
    import gpr_b200
    g = load(name)
    tv, tg = TOL[name]
    lik = gpr_b200.SparseGaussianLogLikelihood()
    M = g["Y"].shape[1]
    for k in range(M):                       # the reference's likelihood is single-output: one problem per column
        gp = build(str(g["kernel"]), g["X"], g["Y"][:, k:k + 1], g["Xm"], float(g["sigma"]), float(g["jitter"]))
        assert rel(lik(gp), g["value"][k:k + 1]) <= tv
        v, gr = lik.GetValueAndParameterDerivatives(gp)
        assert rel(v, g["value_vd"][k:k + 1]) <= tv
        assert rel(gr, g["grad"][k]) <= tg
        assert rel(lik.GetParameterDerivatives(gp), g["grad"][k]) <= tg
    # all columns at once: per-column values and the Jacobian rows
    gp = build(str(g["kernel"]), g["X"], g["Y"], g["Xm"], float(g["sigma"]), float(g["jitter"]))
    v, J = lik.GetValueAndJacobian(gp)
    assert rel(v, g["value_vd"]) <= tv and rel(J, g["grad"]) <= tg
    if M > 1:
        with pytest.raises(gpr_b200.GprError):
            lik.GetParameterDerivatives(gp)  # include/SparseLikelihood.h:264


def test_sparse_likelihood_against_oracle_true_mode_and_errors():
    import gpr_b200
    from oracle.gpr_oracle import SparseLikOracle
    rng = np.random.default_rng(11)
    n, m, d, M = 900, 140, 3, 3
    ks = "SumKernel(ProductKernel(GaussianKernel(1.5,1.1,),PeriodicKernel(1,1.2,1.5,)),RationalQuadraticKernel(0.6,1.2,1.8,))"
    X = rng.uniform(-1.5, 1.5, (n, d))
    Y = np.tanh(X @ rng.standard_normal((d, M))) + 0.05 * rng.standard_normal((n, M))
    Xm = X[rng.choice(n, m, replace=False)] + 0.02 * rng.standard_normal((m, d))
    o = SparseLikOracle(ks, X, Y, Xm, 0.3, 0.05)
    gp = build(ks, X, Y, Xm, 0.3, 0.05)
    for mode, clamp in ((gpr_b200.capi.LOGLIK_REFERENCE_CLAMP, True), (gpr_b200.capi.LOGLIK_TRUE, False)):
        lik = gpr_b200.SparseGaussianLogLikelihood(mode)
        v, J = lik.GetValueAndJacobian(gp)
        assert rel(v, o.value(clamp=clamp, nan_to_min=True)) <= 1e-8
        assert rel(J, o.jacobian()) <= 1e-7
    # parameters changed on the kernel object are picked up (the optimiser loop does exactly this)
    k = gp.GetKernel()
    p = np.array(k.GetParameters(), dtype=float) * 1.05
    k.SetParameters(list(p))
    o2 = SparseLikOracle(k.ToString(), X, Y, Xm, 0.3, 0.05)
    assert rel(gpr_b200.SparseGaussianLogLikelihood()(gp), o2.value()) <= 1e-8
    with pytest.raises(gpr_b200.GprError):
        gpr_b200.SparseGaussianLogLikelihood()(gpr_b200.GaussianProcess(gpr_b200.GaussianKernel(1.0, 1.0)))
    empty = gpr_b200.SparseGaussianProcess(gpr_b200.GaussianKernel(1.0, 1.0), 0.1)
    empty.AddSample(np.zeros(2), np.zeros(1))
    with pytest.raises(gpr_b200.GprError):
        gpr_b200.SparseGaussianLogLikelihood()(empty)            # no inducing samples (:154)


def test_gradient_is_the_derivative_of_the_value_at_n_262144():
    import gpr_b200
    rng = np.random.default_rng(21)
    n, m, d = 262144, 1024, 3
    X = rng.uniform(-1, 1, (n, d))
    Y = np.sin(3 * X[:, :1]) + 0.5 * rng.standard_normal((n, 1))
    Xm = rng.uniform(-1, 1, (m, d))
    h = gpr_b200.Handle(0)
    kern = gpr_b200.KernelFactory.GetKernel("SumKernel(GaussianKernel(0.4,1.0,),RationalQuadraticKernel(0.5,0.8,2.0,),)")
    ops, p0 = kern.lower()
    p0 = np.array(p0, dtype=float)
    h.set_kernel(ops, p0)
    h.set_train(X, Y, 0.5)
    h._ck(h.lib.gprb200_sparse_set_inducing(h.h, Xm.ctypes.data, m, d, 0.1))   # cond(S) ~ 1e7: FD noise stays below 1e-6
    P = len(p0)

    def value(p):
        h.set_kernel(ops, p)
        v = np.empty(1)
        h._ck(h.lib.gprb200_sparse_loglik(h.h, gpr_b200.capi.LOGLIK_TRUE, v.ctypes.data, None, None))
        return v[0]

    v, g = np.empty(1), np.empty(P)
    h.set_kernel(ops, p0)
    h._ck(h.lib.gprb200_sparse_loglik(h.h, gpr_b200.capi.LOGLIK_TRUE, v.ctypes.data, g.ctypes.data, None))
    assert np.isfinite(v).all() and np.isfinite(g).all()
    fd = np.zeros(P)
    for i in range(P):
        e = np.zeros(P)
        e[i] = 1e-3 * max(1.0, abs(p0[i]))
        # five-point stencil: the length-scale direction has a large third derivative
        fd[i] = (-value(p0 + 2 * e) + 8 * value(p0 + e) - 8 * value(p0 - e) + value(p0 - 2 * e)) / (12 * e[i])
    assert rel(g, fd) <= 1e-5
    h.close()
