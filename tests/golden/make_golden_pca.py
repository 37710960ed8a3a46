"""Generate tests/golden/pca_*.npz from oracle/_ref: the reference's own PCA<double> (include/PCA.h) compiled
unmodified against third_party/eigen_lite (BDCSVD -> LAPACK dgesdd). Run where /root/reference exists:

    make -C oracle && python tests/golden/make_golden_pca.py
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402
from tests.synth import make_pca_data  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
CASES = [  # name, D, n, nFeatures, rank, decay, seed
    ("pca_small", 1200, 24, 6, 10, 1e-3, 11),
    ("pca_odd", 777, 17, 5, 8, 1e-2, 12),      # odd n (Jacobi pads), D not a multiple of anything
]


def main():
    for name, D, n, nF, rank, decay, seed in CASES:
        X = make_pca_data(D, n, rank=rank, decay=decay, seed=seed)
        Xt = make_pca_data(D, 5, rank=rank, decay=decay, seed=seed + 100)
        W = np.random.default_rng(seed).standard_normal((nF, 3))
        r = ref.RefPCA(X, nF)
        mean, sigma, U = r.get()
        F_fast = r.project(Xt, fast=True)
        F_all = r.project(Xt)
        R = r.reconstruct(W)
        with tempfile.TemporaryDirectory() as td:   # file constructor: PCA(path, nFeatures), include/PCA.h:51-65
            r.write(os.path.join(td, "m-"))
            r2 = ref.RefPCA(prefix=os.path.join(td, "m-"), n_features=nF)
            R2 = r2.reconstruct(W, D=D)
            mean_bin = open(os.path.join(td, "m-Mean.bin"), "rb").read()
        assert np.array_equal(R, R2)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), X=X, Xt=Xt, W=W, nF=nF, mean=mean, sigma=sigma, U=U,
                            F_fast=F_fast, F_all=F_all, R=R, mean_bin=np.frombuffer(mean_bin, dtype=np.uint8))
        print(name, "sigma", sigma[:4], "...", sigma[-2:])


if __name__ == "__main__":
    main()
