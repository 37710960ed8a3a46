"""-m gpu: the BASELINE.json configurations C1 and C2 with their ACTUAL kernels, noise levels and shapes, against the
reference's own code (oracle/_ref: lib/GaussianProcess.cpp compiled unmodified; LU inverse + C * Y, :642-672).

  C1  N = 1000, d = 1, x_i = 4 * 2 pi i / N, y = sin x, GaussianKernel(0.5, 1), sigma_n = 0.01, T = 1000
      (the scaled-up tests/GaussianProcessTest.cpp:35-76 of SURVEY 8(d)). cond(K + s^2 I) is ~1e6-1e7 here, orders
      worse than the sigma_n = 1 problems elsewhere in the suite; it is computed and printed. alpha is held to 1e-9
      against a long-double-refined solution; the reference's own explicit-inverse alpha is 3.9e-9 from it.
  C2  N = 4096, d = 30, GaussianKernel(1.5 sqrt(30), 1), sigma_n = 1, 32 of the 786432 output columns,
      Predict with T = 1 (the GEMV path of apps/GaussianProcessPredict.cpp:185-193) and T = 256 (batched GEMM).
  L0  GaussianLikelihood::operator() (include/Likelihood.h:106-139) against the `lik_m1` goldens.

oracle/_ref travels to the GPU box as a built .so (git-ignored, not gpurun-ignored); /root/reference is not read."""
import math
import os

import numpy as np
import pytest

from tests.synth import rel_err

pytestmark = pytest.mark.gpu
EPS = np.finfo(np.float64).eps


def _ref():
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref/libgpr_ref.so not built")
    return ref


def _gpu(ks, X, Y, sigma):
    import gpr_b200
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(ks).lower()
    h.set_kernel(ops, params)
    h.set_train(X, Y, sigma)
    h.learn(0)
    return h


def test_c1_exact_configuration_against_the_reference():
    ref = _ref()
    from oracle import gpr_oracle as orc
    N, T, sigma = 1000, 1000, 0.01
    ks = "GaussianKernel(0.5,1,)"
    X = (4.0 * 2.0 * math.pi * np.arange(N) / N)[:, None]
    Y = np.sin(X)
    Xs = (4.0 * 2.0 * math.pi * (np.arange(T) + 0.5) / T)[:, None]
    g = ref.RefGP(ks, sigma).add_samples(X, Y)
    g.initialize()
    a_ref = g.alpha()
    mean_ref = np.stack([g.predict(x) for x in Xs])
    ci_ref = np.array([g.ci(x) for x in Xs])
    K = orc.kernel_matrix(orc.parse_kernel(ks), X, X) + sigma * sigma * np.eye(N)
    ev = np.linalg.eigvalsh(K)
    cond = float(ev[-1] / ev[0])
    print("C1: cond(K + s^2 I) = %.3e" % cond)
    h = _gpu(ks, X, Y, sigma)
    a = h.alpha()
    mean, ci = h.predict(Xs, want_ci=True)
    one = np.stack([h.predict(x[None, :])[0] for x in Xs[::37]])
    h.close()
    # alpha. "Truth" = a double Cholesky solve of the same K refined four times with long-double residuals. Measured
    # here and on the B200: the reference's explicit LU inverse times Y (lib/GaussianProcess.cpp:661) is 3.9e-9 from
    # it at this conditioning (its residual K alpha - y is 4e-11), a factor-and-solve is 4e-12 from it. So the 1e-9
    # bar is held against the truth, and the distance to the reference is bounded by the reference's OWN error.
    import scipy.linalg as sl
    cf = sl.cho_factor(K, lower=True)
    at = sl.cho_solve(cf, Y).astype(np.longdouble)
    Kl, Yl = K.astype(np.longdouble), Y.astype(np.longdouble)
    for _ in range(4):
        at = at + sl.cho_solve(cf, (Yl - Kl @ at).astype(np.float64)).astype(np.longdouble)
    a_true = at.astype(np.float64)
    e_ours, e_ref, e_alpha = rel_err(a, a_true), rel_err(a_ref, a_true), rel_err(a, a_ref)
    print("C1: alpha vs refined truth: ours %.3e, reference %.3e; ours vs reference %.3e (cond * eps = %.3e)"
          % (e_ours, e_ref, e_alpha, cond * EPS))
    assert e_ours <= 1e-9
    assert e_alpha <= max(1e-9, 2.0 * e_ref)
    # the residual of the defining equation is at rounding level for OUR alpha (does not depend on cond)
    assert np.abs(K @ a - Y).max() <= 1e-11 * max(1.0, np.abs(a).max())
    assert rel_err(mean, mean_ref) <= 1e-9
    assert rel_err(one, mean_ref[::37]) <= 1e-9
    assert np.abs(mean[:, 0] - np.sin(Xs[:, 0])).max() <= 5e-3          # sanity only (not parity): it IS a sin regression; 1.1e-3 at the interval edge
    assert np.max(np.abs(ci - ci_ref)) <= 1e-8 * max(1.0, ci_ref.max())


def test_c2_shape_32_columns_predict_t1_and_t256_against_the_reference():
    ref = _ref()
    rng = np.random.default_rng(4096)
    N, d, M, T, sigma = 4096, 30, 32, 256, 1.0
    ks = "GaussianKernel(%.17g,1,)" % (1.5 * math.sqrt(30.0))
    X = rng.standard_normal((N, d))
    W = rng.standard_normal((d, M)) / math.sqrt(d)
    Y = np.tanh(X @ W) + rng.standard_normal((N, M))
    Xs = rng.standard_normal((T, d))
    ref.set_threads(os.cpu_count() or 1)
    g = ref.RefGP(ks, sigma).add_samples(X, Y)
    g.initialize()                     # OpenMP Gram + dgetrf/dgetri + C * Y at N = 4096: a few seconds
    a_ref = g.alpha()
    mean_ref = np.stack([g.predict(x) for x in Xs])
    ci_ref = np.array([g.ci(x) for x in Xs[:16]])
    h = _gpu(ks, X, Y, sigma)
    assert rel_err(h.alpha(), a_ref) <= 1e-9
    mean = h.predict(Xs)                                        # T = 256: DMMA GEMM
    assert rel_err(mean, mean_ref) <= 1e-9
    one = np.stack([h.predict(x[None, :])[0] for x in Xs[:24]])  # T = 1: gemv_t_kernel
    assert rel_err(one, mean_ref[:24]) <= 1e-9
    ci = h.credible_interval(Xs[:16])
    assert np.max(np.abs(ci - ci_ref)) <= 1e-8 * max(1.0, ci_ref.max())
    h.close()


GOLD_L0 = ["gaussexp", "gaussian", "general13", "periodic", "prod", "ragged130", "rq", "sum3", "white"]


@pytest.mark.parametrize("name", GOLD_L0)
def test_l0_gaussian_likelihood_matches_reference_golden(name):
    """exp(-1/2 y^T C y) det^-1/2 (2 pi)^-N/2 against the value the reference's own class produced
    (tests/golden/make_golden.py:56). Values range from 1e-10 down to 1e-60, so the comparison is relative."""
    import gpr_b200
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz"))
    gp = gpr_b200.GaussianProcess(gpr_b200.KernelFactory.GetKernel(str(z["kernel"])))
    gp.SetSigma(float(z["sigma"]))
    for x, y in zip(z["X"], z["Y"][:, :1]):
        gp.AddSample(x, y)
    v = gpr_b200.GaussianLikelihood()(gp)
    assert rel_err(v, z["lik_m1"]) <= 1e-8
