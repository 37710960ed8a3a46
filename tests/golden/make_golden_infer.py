"""Generate tests/golden/infer_*.npz from oracle/_ref: the reference's own GaussianProcessInference<double>
(include/GaussianProcessInference.h) driving its own GaussianLogLikelihood. Run where /root/reference exists:

    make -C oracle && python tests/golden/make_golden_infer.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def data(n, d, M, seed):
    rng = np.random.default_rng(seed)
    X = rng.uniform(-2, 2, (n, d))
    Y = np.stack([np.sin(1.5 * X[:, 0] + 0.4 * k) + 0.3 * np.cos(X[:, -1] * (k + 1)) for k in range(M)], axis=1)
    return X, Y + 0.05 * rng.standard_normal((n, M))


# M = 1 throughout: with M > 1 the reference assigns the M x M matrix Y^T C Y to a VectorType
# (include/Likelihood.h:175, :296), which is undefined in Eigen and rejected by eigen_lite.
CASES = [  # name, kernel, n, d, M, sigma, which, step, iterations, mask
    ("infer_opt1_gaussexp", "GaussianExpKernel(0.3,0.2,)", 60, 1, 1, 0.1, 1, 0.1, 25, None),
    ("infer_opt1_rq", "RationalQuadraticKernel(1.2,0.8,1.5,)", 50, 2, 1, 0.2, 1, 0.1, 15, None),
    ("infer_opt2_periodic", "PeriodicKernel(1.5,0.5,4,)", 48, 1, 1, 0.1, 2, 0.1, 20, [0, 1, 0]),
    ("infer_opt2_sum", "SumKernel(GaussianKernel(1.1,0.9,),PeriodicKernel(1.0,2.0,1.0,),)", 40, 2, 1, 0.3, 2, 0.05, 12, None),
]


def main():
    for name, ks, n, d, M, sigma, which, step, iters, mask in CASES:
        X, Y = data(n, d, M, len(name))
        g = ref.RefGP(ks, sigma).add_samples(X, Y)
        P = len(ref.kernel_tostring(ks)[1])
        p0 = np.array(ref.kernel_tostring(ks)[1])
        # the whole trajectory: GetParameters() after k iterations, k = 0 .. iters (each from a fresh run; the
        # reference is deterministic, so run k+1 continues run k bit for bit)
        traj = np.stack([p0] + [g_params(ks, sigma, X, Y, which, step, k, mask, P) for k in range(1, iters + 1)])
        final = g.infer(which, step, iters, P, mask)
        assert np.array_equal(final, traj[-1])
        np.savez_compressed(os.path.join(OUT, name + ".npz"), kernel=ks, X=X, Y=Y, sigma=sigma, which=which, step=step,
                            iterations=iters, mask=np.array([] if mask is None else mask, dtype=np.int32), params0=p0,
                            traj=traj, final=final)
        print(name, p0, "->", final)


def g_params(ks, sigma, X, Y, which, step, iters, mask, P):
    return ref.RefGP(ks, sigma).add_samples(X, Y).infer(which, step, iters, P, mask)


if __name__ == "__main__":
    main()
