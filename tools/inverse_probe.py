"""A/B of the explicit inverse (solve.cu): structured right-looking sweeps (K = 128 per pass) vs TRTRI + LAUUM with
long-K GEMMs. Checks both against numpy at a ragged N, then times N = 8192 / 16384 / 32768."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import gpr_b200  # noqa: E402


def make(N, d=8, seed=0):
    rng = np.random.default_rng(seed)
    X = rng.uniform(-1, 1, (N, d))
    Y = np.sin(3 * X[:, :1]) + 0.1 * rng.standard_normal((N, 1))
    return X, Y


def run(N, long_k, check=False):
    X, Y = make(N)
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel("RationalQuadraticKernel(1,1.5,2,)").lower()
    h.set_kernel(ops, params)
    h.set_option("inverse_long_k", long_k)
    h.set_train(X, Y, 1.0)
    h.learn(gpr_b200.capi.WANT_CORE)
    torch.cuda.synchronize()
    ms = h.phase_ms()["inverse"]
    out = None
    if check:
        C = h.core()
        K = h.gram(True)
        out = (np.abs(C @ K - np.eye(N)).max(), np.abs(C - C.T).max())
    else:
        for _ in range(2):
            h.set_train(X, Y, 1.0)
            h.learn(gpr_b200.capi.WANT_CORE)
            ms = min(ms, h.phase_ms()["inverse"])
    h.close()
    return ms, out


if __name__ == "__main__":
    for N in (300, 1111, 3000):
        for lk in (0, 1):
            print("N", N, "long_k", lk, run(N, lk, check=True), flush=True)
    for N in (8192, 16384, 32768):
        for lk in (0, 1):
            ms, _ = run(N, lk)
            print("N", N, "long_k", lk, "inverse ms %.2f  TFLOP/s %.2f" % (ms, 2 * N ** 3 / 3 / ms * 1e-9), flush=True)
