"""Generate tests/golden/sparse_*.npz from oracle/_ref: the reference's own SparseGaussianProcess<double>
(include/SparseGaussianProcess.h) compiled unmodified. Run where /root/reference exists:

    make -C oracle && python tests/golden/make_golden_sparse.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
CASES = [  # name, kernel, n, m, d, M, sigma, jitter
    ("sparse_gauss", "GaussianKernel(1.3,1.1,)", 300, 40, 2, 2, 0.2, 0.1),
    ("sparse_sum", "SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,),)", 260, 33, 3, 1, 0.3, 0.05),
    ("sparse_rq_ragged", "RationalQuadraticKernel(1.0,1.5,2.0,)", 391, 130, 2, 3, 0.15, 0.01),
]


def main():
    for name, ks, n, m, d, M, sigma, jitter in CASES:
        rng = np.random.default_rng(len(name) + n)
        X = rng.uniform(-2, 2, (n, d))
        Y = np.stack([np.sin(1.7 * X[:, 0] + 0.5 * k) + 0.4 * np.cos(X[:, -1] * (k + 1)) for k in range(M)], axis=1)
        Y += 0.05 * rng.standard_normal((n, M))
        idx = rng.choice(n, m, replace=False)
        Xm, Ym = X[idx] + 0.01 * rng.standard_normal((m, d)), Y[idx]
        Xs = rng.uniform(-2, 2, (7, d))
        g = ref.RefSparseGP(ks, sigma, jitter).add(X, Y).add(Xm, Ym, inducing=True)
        mean = np.stack([g.predict(x) for x in Xs])
        var = np.array([g.cov(x, x) for x in Xs])
        cov = np.array([g.cov(Xs[i], Xs[(i + 1) % len(Xs)]) for i in range(len(Xs))])
        np.savez_compressed(os.path.join(OUT, name + ".npz"), kernel=ks, X=X, Y=Y, Xm=Xm, Ym=Ym, Xs=Xs, sigma=sigma,
                            jitter=jitter, mean=mean, var=var, cov=cov)
        print(name, mean[0], var[:2], cov[:2])


if __name__ == "__main__":
    main()
