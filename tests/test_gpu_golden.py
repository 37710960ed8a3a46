"""-m gpu: the CUDA path against the committed golden vectors (outputs of the reference's own code,
tests/golden/make_golden.py). Does not need /root/reference or oracle/_ref at run time."""
import glob
import os

import numpy as np
import pytest

from tests.synth import rel_err

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(os.path.basename(f)[:-4] for f in glob.glob(os.path.join(GOLD, "*.npz")) if "clamp" not in f and "pca_" not in f and "infer_" not in f and "sparse" not in f)


def load(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    return {k: (z[k].item() if z[k].ndim == 0 else z[k]) for k in z.files}


def handle(g, cols=None):
    import gpr_b200
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(str(g["kernel"])).lower()
    h.set_kernel(ops, params)
    Y = g["Y"] if cols is None else g["Y"][:, :cols]
    h.set_train(g["X"], Y, float(g["sigma"]))
    return h


@pytest.mark.parametrize("name", CASES)
def test_cuda_matches_reference_golden(name):
    g = load(name)
    h = handle(g)
    assert rel_err(h.gram(False), g["K"]) <= 1e-13
    assert rel_err(h.dgram(), g["dK"]) <= 1e-13
    h.learn(0)
    assert rel_err(h.alpha(), g["alpha"]) <= 1e-9
    assert rel_err(h.core(), g["core"]) <= 1e-8
    mean, ci = h.predict(g["Xs"], want_ci=True)
    assert rel_err(mean, g["mean"]) <= 1e-9
    assert np.max(np.abs(ci - g["ci"])) <= 1e-8 * max(1.0, np.max(g["ci"]))
    one = np.stack([h.predict(x[None, :])[0] for x in g["Xs"]])  # T = 1: the GEMV path the app uses
    assert rel_err(one, g["mean"]) <= 1e-9
    cov = h.posterior_cov(g["Xs"], g["Xs"][::-1].copy())
    assert np.max(np.abs(cov - g["cov"])) <= 1e-8 * max(1.0, np.max(np.abs(g["cov"])))
    assert rel_err(h.kernel_vector(g["Xs"]), g["kvec"]) <= 1e-13
    m, D = h.predict_derivative(g["Xs"][2])
    assert rel_err(m, g["dmean"]) <= 1e-9 and rel_err(D, g["D"]) <= 1e-9
    _, grad, _ = h.loglik(0, False, True, False)
    assert rel_err(grad, g["grad"]) <= 1e-8
    h1 = handle(g, cols=1)
    v, grad1, _ = h1.loglik(0, True, True, False)
    assert rel_err(v, g["ll_value_m1"]) <= 1e-8
    assert rel_err(grad1, g["ll_grad_m1"]) <= 1e-8
    _, _, J = h1.loglik(0, True, False, True)
    assert rel_err(J, g["ll_jac_m1"]) <= 1e-8


@pytest.mark.parametrize("tag", ["clamp_over", "clamp_under"])
def test_cuda_reproduces_determinant_clamp(tag):
    g = load(tag)
    h = handle(g)
    v, _, _ = h.loglik(0, True, False, False)
    assert rel_err(v, g["ll_value"]) <= 1e-8
    vt, _, _ = h.loglik(1, True, False, False)
    assert abs(vt[0] - g["ll_value"][0]) > 100


def test_load_path_uses_stored_core_without_refactorising():
    g = load("sum3")
    import gpr_b200
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(str(g["kernel"])).lower()
    h.set_kernel(ops, params)
    h.set_state(g["X"], g["Y"], g["alpha"], g["core"], float(g["sigma"]))
    n0 = h.launch_count()
    mean, ci = h.predict(g["Xs"], want_ci=True)
    assert rel_err(mean, g["mean"]) <= 1e-12           # alpha is the reference's own
    assert np.max(np.abs(ci - g["ci"])) <= 1e-10
    assert h.phase_ms()["potrf"] == 0.0                # lib/GaussianProcess.cpp:267: Load does not re-learn
    assert h.launch_count() > n0
