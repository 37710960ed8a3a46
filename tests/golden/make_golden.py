"""Generate tests/golden/*.npz from oracle/_ref (the reference's own code compiled here; see
oracle/Makefile). Run in the build container, where /root/reference exists:

    make -C oracle && python tests/golden/make_golden.py

The fixtures travel with the repository; nothing at test time needs /root/reference.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402
from oracle import gpr_oracle as orc  # noqa: E402
from tests.synth import KERNELS, make_problem  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))

CASES = [  # name, kernel, N, d, M, sigma, dup
    ("gaussian", "gaussian", 48, 3, 2, 0.8, 0),
    ("gaussexp", "gaussexp", 40, 2, 1, 1.0, 0),
    ("white", "white", 40, 2, 2, 0.5, 3),
    ("rq", "rq", 48, 3, 1, 1.0, 0),
    ("periodic", "periodic", 40, 1, 1, 0.7, 0),
    ("sum3", "sum3", 56, 2, 3, 1.0, 2),
    ("prod", "prod", 40, 2, 1, 0.9, 0),
    ("general13", "general13", 64, 2, 2, 1.0, 2),
    ("ragged130", "gaussian", 130, 5, 3, 1.0, 0),
]


def main():
    for name, kname, N, d, M, sigma, dup in CASES:
        ks = KERNELS[kname]
        X, Y = make_problem(N, d, M, seed=100 + len(name), dup=dup)
        Xs, _ = make_problem(9, d, 1, seed=200 + len(name))
        Xs[0] = X[1]  # a test point equal to a training point (WhiteKernel fires there too)
        g = ref.RefGP(ks, sigma).add_samples(X, Y).initialize()
        P = orc.num_params(orc.parse_kernel(ks))
        out = dict(kernel=ks, X=X, Y=Y, sigma=sigma, Xs=Xs,
                   K=g.kernel_matrix(), dK=g.dkernel_matrix(P), alpha=g.alpha(), core=g.core(),
                   mean=np.stack([g.predict(x) for x in Xs]),
                   ci=np.array([g.ci(x) for x in Xs]),
                   cov=np.array([g.cov(a, b) for a, b in zip(Xs, Xs[::-1])]),
                   kvec=np.stack([g.kernel_vector(x) for x in Xs]),
                   grad=g.loglik_grad(P))
        m, D = g.predict_derivative(Xs[2])
        out["dmean"], out["D"] = m, D
        # likelihood value / Jacobian: the reference defines them for one output column only
        g1 = ref.RefGP(ks, sigma).add_samples(X, Y[:, :1])
        v, J = g1.loglik_value_jac(P)
        out["ll_value_m1"], out["ll_jac_m1"] = v, J
        out["ll_grad_m1"] = g1.loglik_grad(P)
        out["lik_m1"] = g1.lik_value()
        tostr, params = ref.kernel_tostring(ks)
        out["tostring"], out["params"] = tostr, params
        np.savez_compressed(os.path.join(OUT, f"{name}.npz"), **out)
        print(name, "K", out["K"].shape, "dK", out["dK"].shape)
    # determinant clamp (include/Likelihood.h:180-188): det overflows / underflows a double
    for tag, ks, sigma in (("clamp_over", "GaussianKernel(0.05,3,)", 1.0), ("clamp_under", "GaussianKernel(0.05,0.01,)", 0.1)):
        X, Y = make_problem(900, 2, 1, seed=3)
        g = ref.RefGP(ks, sigma).add_samples(X, Y)
        np.savez_compressed(os.path.join(OUT, f"{tag}.npz"), kernel=ks, X=X, Y=Y, sigma=sigma, ll_value=g.loglik_value())
        print(tag, g.loglik_value())


if __name__ == "__main__":
    main()
