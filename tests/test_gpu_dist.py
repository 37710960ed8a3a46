"""GPU tests of gpr_b200.dist on one B200 (world_size 1: same code path, DeviceOps through the C-ABI, no process
group) against the oracle and against the single-GPU library; the N>1 schedule itself is covered on CPU by
tests/test_dist_gloo.py and on 2+ GPUs by tools/dist_check.py (torchrun)."""
import numpy as np
import pytest

from tests.synth import KERNELS, make_problem, rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,nb,M,kname", [(700, 128, 3, "sum3"), (1500, 256, 1, "rq"), (2048, 1024, 2, "general13"),
                                          (130, 128, 1, "gaussian")])
def test_block_cyclic_device_ops_match_oracle(N, nb, M, kname):
    import gpr_b200
    from gpr_b200 import dist as gd
    from oracle import gpr_oracle as orc
    X, Y = make_problem(N, 4, M, seed=3)
    ops_, params = gpr_b200.KernelFactory.GetKernel(KERNELS[kname]).lower()
    ops = gd.DeviceOps(0, ops_, params, X, 1.0)
    gp = gd.BlockCyclicGP(ops, N, nb=nb)
    alpha, logdet, info = gp.learn(Y)
    assert info == 0
    o = orc.OracleGP(KERNELS[kname], X, Y, 1.0)
    o.learn()
    assert rel_err(alpha.cpu().numpy(), o.alpha) <= 1e-9  # BASELINE.json tolerance on alpha
    ld = float(np.linalg.slogdet(o.kernel_matrix() + np.eye(N))[1])
    assert abs(logdet - ld) <= 1e-10 * abs(ld)
    # and bit-for-bit-close to the single-GPU path of the library
    h = gpr_b200.Handle(0)
    h.set_kernel(ops_, params)
    h.set_train(X, Y, 1.0)
    h.learn(0)
    assert rel_err(alpha.cpu().numpy(), h.alpha()) <= 1e-11
    h.close()
    ops.close()


def test_sharded_outputs_single_rank_is_the_plain_model():
    import gpr_b200
    from gpr_b200 import dist as gd
    X, Y = make_problem(500, 5, 37, seed=9)
    ops_, params = gpr_b200.KernelFactory.GetKernel(KERNELS["gaussian"]).lower()
    h = gpr_b200.Handle(0)
    h.set_kernel(ops_, params)
    g = gd.ShardedOutputsGP(h)
    g.set_train(X, Y, 1.0)
    g.learn()
    Xs = X[:7] + 0.01
    assert g.shard == (0, 37)
    assert np.array_equal(g.predict(Xs, gather=True), h.predict(Xs))
    h.close()


@pytest.mark.parametrize("N,block,M,kname", [(700, 128, 2, "rq"), (1500, 512, 1, "sum3"), (1111, 256, 3, "general13")])
def test_sharded_likelihood_device_ops_match_oracle_and_single_gpu(N, block, M, kname):
    """ShardedLikelihood on one rank (same code path as g ranks, no process group) against the oracle's
    GaussianLogLikelihood gradient (include/Likelihood.h:231-285) and against gprb200_loglik."""
    import gpr_b200
    from gpr_b200 import dist as gd
    from oracle import gpr_oracle as orc
    X, Y = make_problem(N, 4, M, seed=8)
    ops_, params = gpr_b200.KernelFactory.GetKernel(KERNELS[kname]).lower()
    lo = gd.DeviceLikOps(0, ops_, params, X, Y, 1.0)
    sl = gd.ShardedLikelihood(lo, block=block)
    p = np.array(params) * 1.03
    v, g = sl.value_and_gradient(p, mode=gpr_b200.capi.LOGLIK_TRUE)
    k = orc.set_params(orc.parse_kernel(KERNELS[kname]), list(p))
    o = orc.OracleGP(k, X, Y, 1.0)
    o.learn()
    assert rel_err(g, o.loglik_grad()) <= 1e-8          # BASELINE.json tolerance on the likelihood gradient
    assert rel_err(v, o.loglik_value(clamp=False)) <= 1e-8
    h = gpr_b200.Handle(0)
    h.set_kernel(ops_, list(p))
    h.set_train(X, Y, 1.0)
    v1, g1, _ = h.loglik(gpr_b200.capi.LOGLIK_TRUE, True, True, False)
    assert rel_err(g, g1) <= 1e-10 and rel_err(v, v1) <= 1e-13
    h.close()
    lo.close()


def test_sharded_credible_interval_single_rank_is_the_plain_call():
    import gpr_b200
    from gpr_b200 import dist as gd
    X, Y = make_problem(600, 5, 4, seed=10)
    ops_, params = gpr_b200.KernelFactory.GetKernel(KERNELS["sum3"]).lower()
    h = gpr_b200.Handle(0)
    h.set_kernel(ops_, params)
    g = gd.ShardedOutputsGP(h)
    g.set_train(X, Y, 1.0)
    g.learn()
    Xs = X[:9] + 0.02
    assert np.array_equal(g.credible_interval(Xs), h.credible_interval(Xs))
    from oracle import gpr_oracle as orc
    o = orc.OracleGP(KERNELS["sum3"], X, Y, 1.0)
    o.learn()
    assert np.max(np.abs(g.credible_interval(Xs) - np.array([o.credible_interval(x) for x in Xs]))) <= 1e-8
    h.close()
