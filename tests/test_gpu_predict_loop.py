"""-m gpu: the loop body of apps/GaussianProcessPredict.cpp:185-193 - Predict + GetCredibleInterval for ONE test point
per call. From the third single-point call on the library answers the interval through inv(L) (a row-parallel
triangular matrix-vector product instead of a forward-substitution chain); both routes must agree with each other,
with the batched call, and with the oracle (lib/GaussianProcess.cpp:85-114) to the interval tolerance 1e-8."""
import numpy as np
import pytest

from tests.synth import KERNELS, make_problem

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,kname", [(700, "sum3"), (1333, "rq"), (2048, "general13")])
def test_single_point_interval_routes_agree_with_oracle(N, kname):
    import gpr_b200
    from oracle import gpr_oracle as orc
    X, Y = make_problem(N, 4, 2, seed=21)
    rng = np.random.default_rng(2)
    Xs = rng.uniform(-1, 1, (12, 4))
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(KERNELS[kname]).lower()
    h.set_kernel(ops, params)
    h.set_train(X, Y, 1.0)
    h.learn(0)
    batch_mean, batch_ci = h.predict(Xs, want_ci=True)               # T = 12: triangular solves with 12 right-hand sides
    loop = [h.predict(x[None, :], want_ci=True) for x in Xs]         # calls 1, 2: substitution chain; 3..12: inv(L) route
    ci_loop = np.array([c[0] for _, c in loop])
    mean_loop = np.stack([m[0] for m, _ in loop])
    o = orc.OracleGP(KERNELS[kname], X, Y, 1.0)
    o.learn()
    ci_ref = np.array([o.credible_interval(x) for x in Xs])
    scale = max(1.0, ci_ref.max())
    assert np.max(np.abs(ci_loop - ci_ref)) <= 1e-8 * scale
    assert np.max(np.abs(batch_ci - ci_ref)) <= 1e-8 * scale
    assert np.max(np.abs(ci_loop - batch_ci)) <= 1e-10 * scale
    assert np.max(np.abs(mean_loop - batch_mean)) <= 1e-10 * np.max(np.abs(batch_mean))
    # posterior covariance of two different points through the same route (operator(), :85-99)
    cov = np.array([h.posterior_cov(Xs[i:i + 1], Xs[i + 1:i + 2])[0] for i in range(6)])
    cov_ref = np.array([o.cov(Xs[i], Xs[i + 1]) for i in range(6)])
    assert np.max(np.abs(cov - cov_ref)) <= 1e-8 * max(1.0, np.max(np.abs(cov_ref)))
    # new kernel parameters invalidate inv(L) together with the factor
    h.set_kernel_params(np.array(params) * 1.1)
    h.learn(0)
    o2 = orc.OracleGP(orc.set_params(orc.parse_kernel(KERNELS[kname]), list(np.array(params) * 1.1)), X, Y, 1.0)
    o2.learn()
    for x in Xs[:5]:
        assert abs(h.credible_interval(x[None, :])[0] - o2.credible_interval(x)) <= 1e-8 * scale
    h.set_option("ci_winv", 0)
    off = np.array([h.credible_interval(x[None, :])[0] for x in Xs[:5]])
    h.set_option("ci_winv", 2)
    on = np.array([h.credible_interval(x[None, :])[0] for x in Xs[:5]])
    assert np.max(np.abs(on - off)) <= 1e-10 * scale
    h.close()
