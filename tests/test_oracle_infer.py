"""CPU: the restated update rules of GaussianProcessInference (oracle/gpr_oracle.infer_optimize*, and the product's
Python mirror gpr_b200.inference with an injected likelihood) against the reference's own optimiser
(oracle/_ref: include/GaussianProcessInference.h compiled unmodified) and its golden outputs.

The reference takes pinv(g g^T) with an ABSOLUTE cut-off (include/Prior.h:38-57). g g^T has rank one, so the
result is dominated by how the SVD rounds the zero singular values: eigen_lite's one-sided Jacobi, LAPACK and real
Eigen all differ at O(1) after a few iterations. The rules are therefore pinned PER STEP with the reference's own
pinv injected; whole trajectories are pinned only against the same-pinv reference run."""
import os

import numpy as np
import pytest

from oracle import gpr_oracle as orc

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["infer_opt1_gaussexp", "infer_opt1_rq", "infer_opt2_periodic", "infer_opt2_sum"]


def load(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    return {k: (z[k].item() if z[k].ndim == 0 else z[k]) for k in z.files}


def ref_or_skip():
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    return ref


@pytest.mark.parametrize("name", CASES)
def test_restated_rules_reproduce_every_step_of_the_reference_optimiser(name):
    """One step from every point of the reference's trajectory (bit-exact starting point, the reference's own
    likelihood and pinv) lands on the reference's next point."""
    ref = ref_or_skip()
    g = load(name)
    ks, X, Y, sigma = str(g["kernel"]), g["X"], g["Y"], float(g["sigma"])
    traj = g["traj"]
    P = traj.shape[1]
    mask = None if len(g["mask"]) == 0 else [bool(m) for m in g["mask"]]
    gp = ref.RefGP(ks, sigma).add_samples(X, Y)
    worst = 0.0
    for k in range(len(traj) - 1):
        if int(g["which"]) == 1:
            nxt = orc.infer_optimize(lambda p: gp.set_params(p).loglik_value_grad(P), traj[k], float(g["step"]), 1,
                                     pinv=ref.pinv)
        else:
            nxt = orc.infer_optimize2(lambda p: gp.set_params(p).loglik_value_jac(P), traj[k], float(g["step"]), 1, mask,
                                      pinv=ref.pinv)
        worst = max(worst, np.abs(np.array(nxt) - traj[k + 1]).max())
    assert worst <= 1e-12, worst


class _Kernel:
    def __init__(self, p):
        self.p = list(p)

    def GetParameters(self):
        return list(self.p)

    def SetParameters(self, p):
        self.p = list(p)


class _GP:
    def __init__(self, p):
        self.k = _Kernel(p)

    def GetKernel(self):
        return self.k


@pytest.mark.parametrize("name", CASES)
def test_python_mirror_applies_the_same_rules(name):
    """gpr_b200.GaussianProcessInference, one iteration from every point of the reference's trajectory with the
    reference's likelihood injected; the mirror's own pinv (one-sided Jacobi, the algorithm the reference is compiled
    against here) reproduces the reference's step, rounding noise included."""
    ref = ref_or_skip()
    from gpr_b200.inference import GaussianProcessInference
    g = load(name)
    ks, X, Y, sigma = str(g["kernel"]), g["X"], g["Y"], float(g["sigma"])
    traj = g["traj"]
    P = traj.shape[1]
    rgp = ref.RefGP(ks, sigma).add_samples(X, Y)

    class Lh:
        def GetValueAndParameterDerivatives(self, gp):
            return rgp.set_params(gp.k.p).loglik_value_grad(P)

        def GetValueAndJacobian(self, gp):
            return rgp.set_params(gp.k.p).loglik_value_jac(P)

    worst = 0.0
    for k in range(len(traj) - 1):
        gpi = GaussianProcessInference(Lh(), _GP(traj[k]), float(g["step"]), 1)   # its own Jacobi-SVD pinv
        if len(g["mask"]):
            gpi.SetParametersToOptimize([bool(m) for m in g["mask"]])
        (gpi.Optimize if int(g["which"]) == 1 else gpi.Optimize2)()
        worst = max(worst, np.abs(np.array(gpi.GetParameters()) - traj[k + 1]).max())
        assert len(gpi.trace) == 1
    assert worst <= 1e-12, worst


def test_optimize2_stops_when_the_likelihood_stops_changing():
    from gpr_b200.inference import GaussianProcessInference
    calls = []

    class Lh:
        def GetValueAndJacobian(self, gp):
            calls.append(1)
            return np.array([-3.0]), np.array([[0.5, -0.25]])

    gpi = GaussianProcessInference(Lh(), _GP([1.0, 2.0]), 0.1, 50)
    gpi.Optimize2()
    assert len(calls) == 2                       # iteration 1 sees the value of iteration 0 and breaks (:169-174)
    p = orc.infer_optimize2(lambda q: (np.array([-3.0]), np.array([[0.5, -0.25]])), [1.0, 2.0], 0.1, 50)
    assert np.allclose(p, gpi.GetParameters(), rtol=0, atol=1e-15)


def test_pinv_matches_reference_on_well_conditioned_matrices():
    ref = ref_or_skip()
    from gpr_b200.inference import pinv
    rng = np.random.default_rng(0)
    for shape in ((3, 3), (5, 2), (2, 4)):
        A = rng.standard_normal(shape)
        assert np.abs(pinv(A) - ref.pinv(A)).max() <= 1e-12
        assert np.abs(orc.pinv_abs(A) - ref.pinv(A)).max() <= 1e-12
    for P in (2, 3, 5):                                   # the optimiser's case: rank one, noise-dominated
        gvec = rng.standard_normal(P) * 4
        A = np.outer(gvec, gvec)
        assert np.array_equal(pinv(A), ref.pinv(A))
    Z = np.zeros((3, 3))
    assert np.array_equal(pinv(Z), Z) and np.array_equal(ref.pinv(Z), Z)      # everything <= eps is dropped
