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
