"""Seeded synthetic inputs shared by the parity tests (SURVEY.md §8a conventions: X ~ U[-1,1]^d or
N(0,1)^d, Y smooth function + noise, sigma_n = 1 and O(1) scales so that cond(K + s^2 I) <= N + 1)."""
import numpy as np

SEED = 0x6770725F62323030 % (2 ** 63)

KERNELS = {
    "gaussian": "GaussianKernel(1.7,1.3,)",
    "gaussexp": "GaussianExpKernel(0.4,0.2,)",
    "white": "SumKernel(GaussianKernel(1.5,1,),WhiteKernel(0.3,))",
    "rq": "RationalQuadraticKernel(1.2,1.5,2,)",
    "periodic": "PeriodicKernel(1.1,2,0.9,)",
    "sum3": "SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))",
    "prod": "ProductKernel(GaussianKernel(2.5,1.2,),PeriodicKernel(0.9,1.5,1.3,))",
    "general13": ("SumKernel(SumKernel(SumKernel(GaussianKernel(2,1.1,),ProductKernel(GaussianKernel(3,0.8,),"
                  "PeriodicKernel(0.7,1.3,1.1,))),RationalQuadraticKernel(0.9,2.2,1.5,)),"
                  "SumKernel(GaussianKernel(0.8,0.5,),WhiteKernel(0.2,)))"),
}


def make_problem(N, d, M, seed=0, dist="uniform", dup=0):
    rng = np.random.default_rng(SEED + seed)
    X = rng.uniform(-1, 1, (N, d)) if dist == "uniform" else rng.standard_normal((N, d))
    if dup:  # exact duplicates exercise WhiteKernel's ||x-y|| == 0 test off the diagonal
        X[-dup:] = X[:dup]
    W = rng.standard_normal((d, M)) / np.sqrt(d)
    Y = np.tanh(X @ W) + 0.1 * rng.standard_normal((N, M))
    return X, Y


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    den = max(np.max(np.abs(b)), 1e-300)
    return float(np.max(np.abs(a - b)) / den)


def make_pca_data(D, n, rank=12, decay=1e-3, noise=1e-4, seed=0, offset=2.5):
    """D x n data matrix, one sample per column: `rank` smooth modes with geometrically decaying amplitudes
    (1 .. decay), white noise at `noise`, and a non-zero mean; the shape of a displacement-field training set
    (include/DataParser.h:536-613) at test size."""
    rng = np.random.default_rng(seed)
    grid = np.linspace(0.0, 1.0, D)
    modes = np.stack([np.sin((k + 1) * np.pi * grid + 0.3 * k) for k in range(rank)], axis=1)
    amp = np.geomspace(1.0, decay, rank)
    X = modes @ (rng.standard_normal((rank, n)) * amp[:, None]) + noise * rng.standard_normal((D, n))
    return X + offset * np.cos(3.0 * grid)[:, None]
