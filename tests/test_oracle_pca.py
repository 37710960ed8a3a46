"""CPU: the numpy restatement of include/PCA.h (oracle/gpr_oracle.PCAOracle) against the golden vectors produced by
the reference's own PCA<double> (tests/golden/make_golden_pca.py), and against oracle/_ref live where it is built.
Singular vectors are defined up to sign: modes are sign-aligned before comparing."""
import os

import numpy as np
import pytest

from oracle.gpr_oracle import PCAOracle
from tests.synth import make_pca_data, rel_err

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["pca_small", "pca_odd"]


def load(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    return {k: (z[k].item() if z[k].ndim == 0 else z[k]) for k in z.files}


def signs(U, Uref):
    s = np.sign(np.sum(U * Uref, axis=0))
    s[s == 0] = 1.0
    return s


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_golden(name):
    g = load(name)
    nF = int(g["nF"])
    o = PCAOracle(g["X"], nF)
    n = g["X"].shape[1]
    assert rel_err(o.mean, g["mean"]) <= 1e-14
    assert np.abs(o.sigma - g["sigma"]).max() <= 1e-13 * g["sigma"][0]
    sg = signs(o.U, g["U"])
    live = n - 1                                   # mean-free data: the last mode is numerically null
    assert np.abs(np.abs(np.sum(o.U * g["U"], axis=0))[:live] - 1).max() <= 1e-9
    assert rel_err(o.project(g["Xt"], fast=True) * sg[:nF, None], g["F_fast"]) <= 1e-10
    assert rel_err((o.project(g["Xt"]) * sg[:, None])[:live], g["F_all"][:live]) <= 1e-8
    assert rel_err(o.reconstruct(g["W"] * sg[:nF, None]), g["R"]) <= 1e-12
    # from the stored matrices (the file constructor, include/PCA.h:51-65): no sign freedom left
    o2 = PCAOracle(mean=g["mean"], sigma=g["sigma"], U=g["U"], n_features=nF)
    assert rel_err(o2.reconstruct(g["W"]), g["R"]) <= 1e-14
    assert rel_err(o2.project(g["Xt"], fast=True), g["F_fast"]) <= 1e-12


def test_oracle_matches_live_reference():
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    X = make_pca_data(900, 20, rank=9, decay=1e-2, seed=5)
    r = ref.RefPCA(X, 4)
    o = PCAOracle(X, 4)
    m, s, U = r.get()
    assert rel_err(o.mean, m) <= 1e-14 and np.abs(o.sigma - s).max() <= 1e-13 * s[0]
    sg = signs(o.U, U)
    Xt = make_pca_data(900, 3, rank=9, decay=1e-2, seed=6)
    assert rel_err(o.project(Xt, fast=True) * sg[:4, None], r.project(Xt, fast=True)) <= 1e-10
    W = np.arange(12.0).reshape(4, 3) / 7
    assert rel_err(o.reconstruct(W * sg[:4, None]), r.reconstruct(W)) <= 1e-12


def test_projection_reconstruction_round_trip():
    X = make_pca_data(640, 16, rank=6, decay=1e-1, noise=0.0, seed=9)
    o = PCAOracle(X, 6)
    back = o.reconstruct(o.project(X, fast=True))
    assert rel_err(back, X) <= 1e-10                 # rank-6 data is reproduced by 6 modes
    ev = o.explained_variance()
    assert np.all(np.diff(ev) >= -1e-15) and abs(ev[-1] - 1.0) <= 1e-15
