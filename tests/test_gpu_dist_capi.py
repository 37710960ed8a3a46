"""-m gpu: libgprb200_dist.so (include/gprb200_dist.h; the block-cyclic Learn with NCCL from C++) on one B200 - world 1 runs
the same ownership / stream / event code with the exchanges skipped - against the oracle, the single-GPU library and the
torch.distributed form of the schedule. World sizes 2..8 (1 x g and Pr x Pc grids): tools/dist_capi_check.py under torchrun,
records in profiles/."""
import numpy as np
import pytest

from tests.synth import KERNELS, make_problem, rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,nb,M,kname", [(700, 128, 3, "sum3"), (1500, 256, 1, "rq"), (2048, 1024, 2, "general13"),
                                          (130, 128, 1, "gaussian"), (3000, 512, 1, "sum3")])
def test_dist_capi_learn_matches_oracle_and_single_gpu(N, nb, M, kname):
    import gpr_b200
    from gpr_b200 import capi_dist
    from oracle import gpr_oracle as orc
    X, Y = make_problem(N, 4, M, seed=3)
    ops_, params = gpr_b200.KernelFactory.GetKernel(KERNELS[kname]).lower()
    d = capi_dist.DistHandle(0)
    d.set_kernel(ops_, params)
    alpha, logdet = d.learn(X, Y, 1.0, nb=nb)
    assert d.counters()["launches"] > 0 and d.counters()["collectives"] == 0
    # the flop the run issued is what the host-only plan (tests/test_dist_plan.py) predicts for this rank
    assert d.counters()["gemm_flops"] == capi_dist.plan(N, nb, 1, 1, 0)["flops"]
    o = orc.OracleGP(KERNELS[kname], X, Y, 1.0)
    o.learn()
    assert rel_err(alpha, o.alpha) <= 1e-9      # BASELINE.json tolerance on alpha
    ld = float(np.linalg.slogdet(o.kernel_matrix() + np.eye(N))[1])
    assert abs(logdet - ld) <= 1e-10 * abs(ld)
    h = gpr_b200.Handle(0)
    h.set_kernel(ops_, params)
    h.set_train(X, Y, 1.0)
    h.learn(0)
    assert rel_err(alpha, h.alpha()) <= 1e-11
    # a second right-hand side against the kept factor, and a second learn on the same handle
    Y2 = np.cos(Y) + 0.5
    a2 = d.solve(Y2)
    h.set_train(X, Y2, 1.0)
    h.learn(0)
    assert rel_err(a2, h.alpha()) <= 1e-11
    alpha3, _ = d.learn(X[: N // 2], Y[: N // 2], 0.5, nb=nb)
    h.set_train(X[: N // 2], Y[: N // 2], 0.5)
    h.learn(0)
    assert rel_err(alpha3, h.alpha()) <= 1e-11
    h.close()
    d.close()


def test_dist_capi_reports_info_of_a_non_spd_matrix():
    import gpr_b200
    from gpr_b200 import capi_dist
    X, Y = make_problem(600, 3, 1, seed=5)
    X[300] = X[7]                               # a duplicated sample and no noise: a zero pivot at row 301
    ops_, params = gpr_b200.KernelFactory.GetKernel(KERNELS["gaussian"]).lower()
    d = capi_dist.DistHandle(0)
    d.set_kernel(ops_, params)
    with pytest.raises(gpr_b200.GprError) as e:
        d.learn(X, Y, 0.0, nb=128)
    assert e.value.code > 0 and "positive definite" in str(e.value)
    with pytest.raises(gpr_b200.GprError):
        d.solve(Y)
    d.close()


def test_dist_capi_argument_errors():
    import gpr_b200
    from gpr_b200 import capi_dist
    X, Y = make_problem(300, 3, 1, seed=5)
    d = capi_dist.DistHandle(0)
    with pytest.raises(gpr_b200.GprError, match="kernel first"):
        d.learn(X, Y, 1.0)
    ops_, params = gpr_b200.KernelFactory.GetKernel(KERNELS["gaussian"]).lower()
    d.set_kernel(ops_, params)
    with pytest.raises(gpr_b200.GprError, match="multiple of 128"):
        d.learn(X, Y, 1.0, nb=200)
    d.close()


def test_python_mirror_set_distributed_matches_the_plain_model():
    """gpr_b200.GaussianProcess.SetDistributed (mirror of the C++ host class): Initialize on the (one-rank) grid, Predict
    locally, posterior covariance / credible interval through the solve on the grid."""
    import gpr_b200
    X, Y = make_problem(900, 3, 2, seed=21)
    gps = []
    for dist_on in (True, False):
        gp = gpr_b200.GaussianProcess(gpr_b200.KernelFactory.GetKernel(KERNELS["sum3"]))
        gp.SetSigma(0.6)
        for x, y in zip(X, Y):
            gp.AddSample(x, y)
        if dist_on:
            gp.SetDistributed(0, 1)
        gp.Initialize()
        gps.append(gp)
    xs, ys = X[5] + 0.03, X[9] - 0.02
    assert rel_err(gps[0].Predict(xs), gps[1].Predict(xs)) <= 1e-11
    assert abs(gps[0](xs, ys) - gps[1](xs, ys)) <= 1e-11
    assert abs(gps[0].GetCredibleInterval(xs) - gps[1].GetCredibleInterval(xs)) <= 1e-10
