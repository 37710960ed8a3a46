"""CPU tests (-m "not gpu"): pin the numpy oracle (oracle/gpr_oracle.py).

  1. against the committed golden vectors, generated from oracle/_ref = the reference's OWN sources
     compiled here (tests/golden/make_golden.py);
  2. against oracle/_ref directly on fresh seeds, when the built library is present;
  3. the reference's own test programs (tests/*.cpp compiled unmodified, oracle/Makefile: ref_tests)
     must pass when present.
"""
import glob
import os
import subprocess

import numpy as np
import pytest

from oracle import gpr_oracle as orc
from oracle import ref
from tests.synth import KERNELS, make_problem, rel_err

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(os.path.basename(f)[:-4] for f in glob.glob(os.path.join(GOLD, "*.npz")) if "clamp" not in f and "pca_" not in f and "infer_" not in f and "sparse" not in f)


def load(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    return {k: (z[k].item() if z[k].ndim == 0 else z[k]) for k in z.files}


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_golden(name):
    g = load(name)
    o = orc.OracleGP(str(g["kernel"]), g["X"], g["Y"], float(g["sigma"]))
    o.learn()
    assert rel_err(o.kernel_matrix(), g["K"]) <= 2e-15
    assert rel_err(o.derivative_kernel_matrix(), g["dK"]) <= 5e-15
    assert rel_err(o.alpha, g["alpha"]) <= 1e-12
    assert rel_err(o.C, g["core"]) <= 1e-12
    assert rel_err(o.predict_batch(g["Xs"]), g["mean"]) <= 1e-12
    assert rel_err([o.credible_interval(x) for x in g["Xs"]], g["ci"]) <= 1e-11
    assert rel_err([o.cov(a, b) for a, b in zip(g["Xs"], g["Xs"][::-1])], g["cov"]) <= 1e-11
    m, D = o.predict_derivative(g["Xs"][2])
    assert rel_err(m, g["dmean"]) <= 1e-12 and rel_err(D, g["D"]) <= 1e-12
    assert rel_err(o.loglik_grad(), g["grad"]) <= 1e-11
    o1 = orc.OracleGP(str(g["kernel"]), g["X"], g["Y"][:, :1], float(g["sigma"]))
    assert rel_err(o1.loglik_value(), g["ll_value_m1"]) <= 1e-12
    assert rel_err(o1.loglik_jacobian(), g["ll_jac_m1"]) <= 1e-11
    assert rel_err(o1.loglik_grad(), g["ll_grad_m1"]) <= 1e-11
    assert orc.get_params(o.k) == pytest.approx(list(g["params"]), abs=0)


@pytest.mark.parametrize("tag", ["clamp_over", "clamp_under"])
def test_oracle_determinant_clamp(tag):
    g = load(tag)
    o = orc.OracleGP(str(g["kernel"]), g["X"], g["Y"], float(g["sigma"]))
    assert rel_err(o.loglik_value(clamp=True), g["ll_value"]) <= 1e-12
    assert abs(o.loglik_value(clamp=False)[0] - g["ll_value"][0]) > 100  # the clamp really is active


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("kname", list(KERNELS))
def test_oracle_matches_reference_library(kname):
    X, Y = make_problem(90, 3, 2, seed=11, dup=2)
    ks = KERNELS[kname]
    g = ref.RefGP(ks, 0.9).add_samples(X, Y).initialize()
    o = orc.OracleGP(ks, X, Y, 0.9)
    o.learn()
    P = orc.num_params(o.k)
    assert rel_err(o.kernel_matrix(), g.kernel_matrix()) <= 2e-15
    assert rel_err(o.derivative_kernel_matrix(), g.dkernel_matrix(P)) <= 5e-15
    assert rel_err(o.alpha, g.alpha()) <= 1e-12
    x = np.array([0.3, -0.2, 0.7])
    assert rel_err(o.predict(x), g.predict(x)) <= 1e-12
    assert abs(o.cov(x, -x) - g.cov(x, -x)) <= 1e-12
    assert rel_err(o.loglik_grad(), g.loglik_grad(P)) <= 1e-11
    # LAPACKUtils wrappers (include/LAPACKUtils.h:85-111; tests/InversionMethodsTest.cpp:147-175)
    assert rel_err(orc.lu_invert(o.Kn), ref.lu_invert(o.Kn)) <= 1e-13
    assert rel_err(orc.chol_invert(o.Kn), ref.chol_invert(o.Kn)) <= 1e-13


REF_TESTS = sorted(glob.glob(os.path.join(os.path.dirname(GOLD), "..", "oracle", "_ref", "tests", "*")))


@pytest.mark.skipif(not REF_TESTS, reason="reference test programs not built (make -C oracle ref_tests)")
@pytest.mark.parametrize("exe", REF_TESTS, ids=[os.path.basename(e) for e in REF_TESTS])
def test_reference_own_test_programs_pass(exe, tmp_path):
    name = os.path.basename(exe)
    # several programs seed their measurement noise with time(0) (e.g. tests/ScopeTest.cpp:66), so their error
    # thresholds are statistical: allow a re-draw after a full second
    import time
    for attempt in range(4):
        p = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True, timeout=600)
        out = p.stdout + p.stderr
        if p.returncode == 0 or name in ("KernelDerivativeTest", "InversionMethodsTest", "GaussianProcessTest"):
            break
        time.sleep(1.1)
    if name == "PosteriorProcessTest":
        # Test 2 multiplies arrays of different lengths (tests/PosteriorProcessTest.cpp:151), which only "works" with
        # real Eigen in release mode; Test 1 (credible interval == 2 sqrt(gp(x,x))) is the one that pins this path.
        assert "Test 1: sinus regression...  [passed]" in out
        return
    assert p.returncode == 0, out[-2000:]
    assert "failed" not in out.lower(), out[-2000:]
