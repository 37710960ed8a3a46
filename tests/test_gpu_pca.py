"""-m gpu: PCA on the device (csrc/pca.cu through the C-ABI) against the golden vectors of the reference's own
PCA<double> (tests/golden/pca_*.npz) and against the CPU oracle on seeded inputs; size-independent properties at the
BASELINE field size D = 3 * 64^3. Singular vectors are sign-aligned before comparing (defined up to sign)."""
import os

import numpy as np
import pytest

from tests.synth import make_pca_data, make_problem, rel_err

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    return {k: (z[k].item() if z[k].ndim == 0 else z[k]) for k in z.files}


def signs(U, Uref):
    s = np.sign(np.sum(U * Uref, axis=0))
    s[s == 0] = 1.0
    return s


@pytest.mark.parametrize("name", ["pca_small", "pca_odd"])
def test_fit_matches_reference_golden(name):
    from gpr_b200.pca import PCA
    g = load(name)
    nF = int(g["nF"])
    n = g["X"].shape[1]
    p = PCA(g["X"], nF)
    assert rel_err(p.GetMean(), g["mean"]) <= 1e-14
    sigma, U = p.GetEigenvalues(), p.GetMatrixU()
    live = n - 1                                            # mean-free data: the last mode is numerically null
    assert np.abs(sigma[:live] / g["sigma"][:live] - 1).max() <= 1e-9   # RELATIVE, down to sigma_k / sigma_0 = 3e-5
    assert sigma[live] <= 1e-6 * sigma[0]
    assert np.abs(np.abs(np.sum(U * g["U"], axis=0))[:live] - 1).max() <= 1e-9
    sg = signs(U, g["U"])
    assert rel_err(p.DimensionalityReductionFast(g["Xt"]) * sg[:nF, None], g["F_fast"]) <= 1e-9
    assert rel_err((p.DimensionalityReduction(g["Xt"]) * sg[:, None])[:live], g["F_all"][:live]) <= 1e-7
    assert rel_err(p.GetReconstruction(g["W"] * sg[:nF, None]), g["R"]) <= 1e-9
    assert p.stat("sweeps") > 0
    p.close()


@pytest.mark.parametrize("name", ["pca_small", "pca_odd"])
def test_stored_basis_matches_reference_golden(name, tmp_path):
    """PCA(path, nFeatures): Mean/Sigma/U.bin written by the reference's layout -> identical projections."""
    import gpr_b200
    from gpr_b200.pca import PCA
    g = load(name)
    nF = int(g["nF"])
    prefix = str(tmp_path / "model-")
    gpr_b200.WriteMatrix(g["mean"][:, None], prefix + "Mean.bin")
    gpr_b200.WriteMatrix(g["sigma"][:, None], prefix + "Sigma.bin")
    gpr_b200.WriteMatrix(g["U"], prefix + "U.bin")
    assert open(prefix + "Mean.bin", "rb").read() == g["mean_bin"].tobytes()   # byte-identical to the reference's file
    p = PCA(prefix, nF)
    assert rel_err(p.GetReconstruction(g["W"]), g["R"]) <= 1e-14
    assert rel_err(p.DimensionalityReductionFast(g["Xt"]), g["F_fast"]) <= 1e-12
    live = g["X"].shape[1] - 1
    assert rel_err(p.DimensionalityReduction(g["Xt"])[:live], g["F_all"][:live]) <= 1e-9
    assert rel_err(p.GetReconstruction(g["W"][:, 0]), g["R"][:, :1]) <= 1e-14      # a single frame (the app's call)
    p.WriteMatricesToFile(str(tmp_path / "copy-"))
    assert open(str(tmp_path / "copy-U.bin"), "rb").read() == open(prefix + "U.bin", "rb").read()
    p.close()


@pytest.mark.parametrize("D,n,nF,T", [(4097, 33, 7, 11), (20000, 128, 40, 9), (1500, 2, 1, 1)])
def test_fit_project_reconstruct_against_oracle(D, n, nF, T):
    from gpr_b200.pca import PCA
    from oracle.gpr_oracle import PCAOracle
    X = make_pca_data(D, n, rank=min(n - 1, 14), decay=1e-3, seed=D)
    Xt = make_pca_data(D, T, rank=min(n - 1, 14), decay=1e-3, seed=D + 1)
    o = PCAOracle(X, nF)
    p = PCA(X, nF)
    sigma, U = p.GetEigenvalues(), p.GetMatrixU()
    keep = o.sigma > 1e-5 * o.sigma[0]
    assert np.abs(sigma[keep] / o.sigma[keep] - 1).max() <= 1e-9
    dots = np.abs(np.sum(U * o.U, axis=0))
    well = keep.copy()
    well[1:] &= (o.sigma[:-1] - o.sigma[1:])[: n - 1] > 1e-3 * o.sigma[1:]      # separated from the previous mode
    well[:-1] &= (o.sigma[:-1] - o.sigma[1:]) > 1e-3 * o.sigma[:-1]             # and from the next
    assert np.abs(dots[well] - 1).max() <= 1e-8
    G = U[:, keep].T @ U[:, keep]
    assert np.abs(G - np.eye(G.shape[0])).max() <= 1e-9                          # orthonormal columns
    if well[:nF].all():
        sg = signs(U, o.U)
        assert rel_err(p.DimensionalityReductionFast(Xt) * sg[:nF, None], o.project(Xt, fast=True)) <= 1e-8
        W = np.random.default_rng(1).standard_normal((nF, T))
        assert rel_err(p.GetReconstruction(W * sg[:nF, None]), o.reconstruct(W)) <= 1e-9
    p.close()


def test_one_pass_is_less_accurate_on_small_modes():
    """Documents why the fit runs two Gram/Jacobi passes: one pass squares the condition number."""
    from gpr_b200.pca import PCA
    from oracle.gpr_oracle import PCAOracle
    X = make_pca_data(6000, 32, rank=12, decay=1e-5, noise=0.0, seed=3)
    o = PCAOracle(X, 4)
    keep = o.sigma > 1e-6 * o.sigma[0]
    e1 = np.abs(PCA(X, 4, passes=1).GetEigenvalues()[keep] / o.sigma[keep] - 1).max()
    e2 = np.abs(PCA(X, 4, passes=2).GetEigenvalues()[keep] / o.sigma[keep] - 1).max()
    assert e2 <= 1e-8 and e2 <= e1


def test_field_size_properties():
    """D = 3 * 64^3 (the DVF of BASELINE config C2): orthonormality, round trip, linearity."""
    from gpr_b200.pca import PCA
    D, n, nF = 3 * 64 ** 3, 48, 16
    X = make_pca_data(D, n, rank=16, decay=1e-2, noise=0.0, seed=2)
    p = PCA(X, nF)
    back = p.GetReconstruction(p.DimensionalityReductionFast(X[:, :5]))
    assert rel_err(back, X[:, :5]) <= 1e-9                                        # rank-16 data, 16 modes
    W1, W2 = np.random.default_rng(0).standard_normal((2, nF, 3))
    mean = p.GetMean()[:, None]
    lin = (p.GetReconstruction(W1) - mean) + 2.0 * (p.GetReconstruction(W2) - mean)
    assert rel_err(p.GetReconstruction(W1 + 2.0 * W2) - mean, lin) <= 1e-12
    F = p.DimensionalityReductionFast(p.GetReconstruction(W1))
    assert rel_err(F, W1) <= 1e-9                                                 # basis^T basisInverse = I
    p.close()


def test_predict_reconstruct_is_predict_then_reconstruct():
    import gpr_b200
    from gpr_b200.pca import PCA
    D, n, nF, N, d = 30000, 40, 8, 300, 4
    Xf = make_pca_data(D, n, rank=10, decay=1e-2, seed=7)
    h = gpr_b200.Handle(0)
    p = PCA(Xf, nF, handle=h)
    X, Y = make_problem(N, d, nF, seed=3)
    ops, params = gpr_b200.KernelFactory.GetKernel("GaussianKernel(1.3,1.1,)").lower()
    h.set_kernel(ops, params)
    h.set_train(X, Y, 0.3)
    h.learn(0)
    Xs, _ = make_problem(11, d, 1, seed=4)
    fields, w = p.PredictReconstruct(h, Xs, want_weights=True)
    mean = h.predict(Xs)
    assert rel_err(w, mean) <= 1e-14
    assert rel_err(fields, p.GetReconstruction(mean.T)) <= 1e-14
    one = p.PredictReconstruct(h, Xs[0])
    assert rel_err(one, fields[:, :1]) <= 1e-14
    p.close()
    h.close()


def test_error_behaviour():
    import gpr_b200
    from gpr_b200.pca import PCA
    X = make_pca_data(500, 10, rank=4, seed=1)
    with pytest.raises(gpr_b200.GprError):
        PCA(X, 11)                                           # more features than samples
    with pytest.raises(gpr_b200.GprError):
        PCA(X.T.copy(), 2)                                   # samples x features: D < n is rejected
    p = PCA(X, 3)
    with pytest.raises(gpr_b200.GprError):
        p.GetReconstruction(np.zeros((4, 1)))                # wrong number of weights
    with pytest.raises(gpr_b200.GprError):
        p.DimensionalityReduction(np.zeros((499, 1)))        # wrong field size
    Xn = X.copy()
    Xn[3, 2] = np.nan
    with pytest.raises(gpr_b200.GprError):
        PCA(Xn, 3)
    p.close()
