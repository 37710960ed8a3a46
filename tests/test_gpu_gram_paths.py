"""-m gpu: the three value-Gram paths (row-group term form, per-tile term form, per-entry interpreter) and the two
derivative-contraction paths agree, including the automatic fall-backs: d too large to keep a tile's points resident in
shared memory, and a kernel tree whose sum-of-products expansion exceeds the term limits."""
import numpy as np
import pytest

from tests.synth import KERNELS, make_problem, rel_err

pytestmark = pytest.mark.gpu


def _gram(ks, X, terms, lower=None):
    import gpr_b200
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(ks).lower()
    h.set_kernel(ops, params)
    h.set_train(X, np.zeros((len(X), 1)), 0.3)
    h.set_option("gram_terms", terms)
    K = h.gram(add_noise=True)
    h.close()
    return K


@pytest.mark.parametrize("kname", ["sum3", "prod", "general13", "rq", "white"])
def test_value_paths_agree(kname):
    X, _ = make_problem(333, 5, 1, seed=4, dup=2)
    K2, K1, K0 = (_gram(KERNELS[kname], X, t) for t in (2, 1, 0))
    assert np.array_equal(K2, K2.T) and np.array_equal(K1, K1.T)
    assert rel_err(K2, K0) <= 1e-14 and rel_err(K1, K0) <= 1e-14


def test_large_d_falls_back_and_matches_the_oracle():
    from oracle import gpr_oracle
    rng = np.random.default_rng(8)
    N, d = 150, 300                                 # 3 * d * 64 doubles = 460 KB: not resident, chunked kernel
    X = rng.uniform(-1, 1, (N, d)) / np.sqrt(d)
    ks = "SumKernel(GaussianKernel(1.2,1,),PeriodicKernel(1,2,1.5,),)"
    K = _gram(ks, X, 2)
    Kref = gpr_oracle.OracleGP(ks, X, np.zeros((N, 1)), 0.3).kernel_matrix() + 0.09 * np.eye(N)
    assert rel_err(K, Kref) <= 1e-13 and np.array_equal(K, K.T)


def test_term_expansion_overflow_uses_the_interpreter():
    from oracle import gpr_oracle
    # (G + G)(G + G)(G + G)(G + G): 16 terms of 4 leaves > 12 terms -> interpreter
    s = "SumKernel(GaussianKernel(1.1,0.9,),GaussianKernel(2.0,0.7,))"
    ks = "ProductKernel(ProductKernel(%s,%s),ProductKernel(%s,%s))" % (s, s, s, s)
    X, _ = make_problem(140, 3, 1, seed=9, dup=0)
    K = _gram(ks, X, 2)
    Kref = gpr_oracle.OracleGP(ks, X, np.zeros((len(X), 1)), 0.3).kernel_matrix() + 0.09 * np.eye(len(X))
    assert rel_err(K, Kref) <= 1e-13


@pytest.mark.parametrize("kname", ["sum3", "general13", "rq"])
def test_contraction_paths_agree(kname):
    import gpr_b200
    X, Y = make_problem(500, 4, 2, seed=6, dup=2)
    out = []
    for terms in (2, 0):
        h = gpr_b200.Handle(0)
        ops, params = gpr_b200.KernelFactory.GetKernel(KERNELS[kname]).lower()
        h.set_kernel(ops, params)
        h.set_train(X, Y, 0.8)
        h.set_option("gram_terms", terms)
        v, g, _ = h.loglik(gpr_b200.capi.LOGLIK_TRUE, True, True, False)
        out.append((v, g))
        h.close()
    assert rel_err(out[0][0], out[1][0]) <= 1e-12
    assert np.abs(out[0][1] - out[1][1]).max() <= 1e-11 * max(1.0, np.abs(out[1][1]).max())
