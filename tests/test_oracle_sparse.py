"""CPU: oracle/gpr_oracle.SparseOracle (numpy restatement of include/SparseGaussianProcess.h, regression part) against
the golden vectors of the reference's own SparseGaussianProcess<double> (tests/golden/make_golden_sparse.py)."""
import os

import numpy as np
import pytest

from oracle.gpr_oracle import SparseOracle
from tests.synth import rel_err

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["sparse_gauss", "sparse_sum", "sparse_rq_ragged"]


def load(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    return {k: (z[k].item() if z[k].ndim == 0 else z[k]) for k in z.files}


@pytest.mark.parametrize("name", CASES)
def test_sparse_oracle_matches_reference_golden(name):
    g = load(name)
    o = SparseOracle(str(g["kernel"]), g["X"], g["Y"], g["Xm"], float(g["sigma"]), float(g["jitter"]))
    Xs = g["Xs"]
    mean = np.stack([o.predict(x) for x in Xs])
    var = np.array([o.cov(x, x) for x in Xs])
    cov = np.array([o.cov(Xs[i], Xs[(i + 1) % len(Xs)]) for i in range(len(Xs))])
    assert rel_err(mean, g["mean"]) <= 1e-9
    assert np.abs(var - g["var"]).max() <= 1e-8 * max(1.0, np.abs(g["var"]).max())
    assert np.abs(cov - g["cov"]).max() <= 1e-8 * max(1.0, np.abs(g["var"]).max())
