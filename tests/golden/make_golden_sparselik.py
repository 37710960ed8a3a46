"""Generate tests/golden/sparselik_*.npz from oracle/_ref: the reference's own SparseGaussianLogLikelihood<double>
(include/SparseLikelihood.h) compiled unmodified. The reference's likelihood is defined for ONE output column
(include/SparseLikelihood.h:170, :264), so every column of Y is run as its own single-output problem. Run where
/root/reference exists:

    make -C oracle && python tests/golden/make_golden_sparselik.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
CASES = [  # name, kernel, P, n, m, d, M, sigma, jitter
    ("sparselik_gauss", "GaussianKernel(1.3,1.1,)", 2, 300, 40, 2, 2, 0.2, 0.1),
    ("sparselik_sum", "SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,),)", 5, 260, 33, 3, 1, 0.3, 0.05),
    ("sparselik_rq_ragged", "RationalQuadraticKernel(1.0,1.5,2.0,)", 3, 391, 130, 2, 3, 0.15, 0.01),
    ("sparselik_prod_white", "SumKernel(ProductKernel(GaussianKernel(1.5,1,),PeriodicKernel(1,1.5,2,)),WhiteKernel(0.3,))",
     6, 200, 64, 2, 1, 0.4, 0.02),
    # sigma^(2n) leaves the long-double range: the complexity term is clamped (:177-185, :351-359)
    ("sparselik_clamp", "GaussianKernel(1.0,1.0,)", 2, 2600, 30, 1, 1, 0.1, 0.1),
]


def main():
    for name, ks, P, n, m, d, M, sigma, jitter in CASES:
        rng = np.random.default_rng(len(name) + n)
        X = rng.uniform(-2, 2, (n, d))
        Y = np.stack([np.sin(1.7 * X[:, 0] + 0.5 * k) + 0.4 * np.cos(X[:, -1] * (k + 1)) for k in range(M)], axis=1)
        Y += 0.05 * rng.standard_normal((n, M))
        idx = rng.choice(n, m, replace=False)
        Xm, Ym = X[idx] + 0.01 * rng.standard_normal((m, d)), Y[idx]
        value_op, value_vd, grad, grad_only = [], [], [], []
        for k in range(M):
            g = ref.RefSparseGP(ks, sigma, jitter).add(X, Y[:, k:k + 1]).add(Xm, Ym[:, k:k + 1], inducing=True)
            value_op.append(g.loglik(0, P)[0][0])
            grad_only.append(g.loglik(1, P)[1])
            v, gr, _ = g.loglik(2, P)
            value_vd.append(v[0])
            grad.append(gr)
        value_op, value_vd, grad, grad_only = map(np.array, (value_op, value_vd, grad, grad_only))
        assert np.array_equal(grad, grad_only)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), kernel=ks, X=X, Y=Y, Xm=Xm, Ym=Ym, sigma=sigma, jitter=jitter,
                            value=value_op, value_vd=value_vd, grad=grad)
        print(name, value_op, grad[0])


if __name__ == "__main__":
    main()
