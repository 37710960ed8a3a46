"""Once, for the record: the reference's OWN Learn path at the full BASELINE size
C3 (N = 32768: OpenMP Gram + dgetrf/dgetri + C * Y, lib/GaussianProcess.cpp:642-672 via oracle/_ref) on the GPU box's
host cores, against the GPU run on the same inputs: alpha, predictive mean, credible interval, log det - with the
seconds. Then the likelihood gradient at C4 (N = 16384, RationalQuadratic) from the numpy oracle's explicit LU inverse
and materialised dK stack (include/Likelihood.h:268-282; tr(A D_p) is evaluated as sum(A o D_p), D_p symmetric, instead
of P more N^3 products) against gprb200_loglik. Writes gpurun_out/r02_c3_reference_full.json.

TEST INFRASTRUCTURE: uses oracle/ as the checker. Not part of the product or of bench.py."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (C3 inputs and constants)
import gpr_b200  # noqa: E402
from oracle import gpr_oracle as orc  # noqa: E402
from oracle import ref  # noqa: E402


def rel(a, b):
    return float(np.max(np.abs(np.asarray(a) - np.asarray(b))) / max(np.max(np.abs(b)), 1e-300))


def avail_gb():
    try:
        import psutil
        return psutil.virtual_memory().available / 1e9
    except Exception:
        return 0.0


out = {"host_cores": os.cpu_count(), "host_ram_available_GB": avail_gb()}
N = int(sys.argv[1]) if len(sys.argv) > 1 else bench.C3_N
N4 = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
threads = os.cpu_count() or 1
os.environ["OMP_NUM_THREADS"] = str(threads)

# ---- C3 ---------------------------------------------------------------------------------------------------------
X, Y = bench.c3_inputs(N, seed=0)
rng = np.random.default_rng(99)
Xs = rng.uniform(-1, 1, (256, bench.C3_D))
h = gpr_b200.Handle(0)
ops, params = gpr_b200.KernelFactory.GetKernel(bench.C3_KERNEL).lower()
h.set_kernel(ops, params)
t0 = time.perf_counter()
h.set_train(X, Y, bench.C3_SIGMA)
h.learn(0)
a_gpu = h.alpha()
gpu_s = time.perf_counter() - t0
mean_gpu, ci_gpu = h.predict(Xs, want_ci=True)
ld_gpu = h.logdet()
h.close()
out["c3"] = {"N": N, "gpu_learn_s_host_to_host": gpu_s}

ref.set_threads(threads)
t0 = time.perf_counter()
g = ref.RefGP(bench.C3_KERNEL, bench.C3_SIGMA).add_samples(X, Y)
g.initialize()
ref_s = time.perf_counter() - t0
a_ref = g.alpha()
mean_ref = np.stack([g.predict(x) for x in Xs])
t0 = time.perf_counter()
ci_ref = np.array([g.ci(x) for x in Xs[:32]])
ci_s = time.perf_counter() - t0
out["c3"].update({"reference_learn_s": ref_s, "reference_threads": threads,
                  "reference_useful_tflops": bench.learn_flops(N, 1) / ref_s * 1e-12,
                  "speedup_host_to_host": ref_s / gpu_s,
                  "alpha_rel": rel(a_gpu, a_ref), "mean_rel": rel(mean_gpu, mean_ref),
                  "ci_abs_over_max": float(np.max(np.abs(ci_gpu[:32] - ci_ref)) / max(1.0, ci_ref.max())),
                  "reference_ci_s_per_point": ci_s / 32})
print(json.dumps(out), flush=True)
if avail_gb() > 45.0:
    t0 = time.perf_counter()
    K = g.kernel_matrix()
    K[np.diag_indices(N)] += bench.C3_SIGMA ** 2
    L = np.linalg.cholesky(K)
    ld_ref = 2.0 * float(np.sum(np.log(np.diag(L))))
    out["c3"].update({"logdet_gpu": ld_gpu, "logdet_host_cholesky_of_reference_K": ld_ref,
                      "logdet_rel": abs(ld_gpu - ld_ref) / abs(ld_ref), "logdet_s": time.perf_counter() - t0})
    del K, L
else:
    out["c3"]["logdet"] = "skipped: not enough host RAM for a second and third N x N matrix"
del g
print(json.dumps(out), flush=True)

# ---- C4 gradient ----------------------------------------------------------------------------------------------------
ks = "RationalQuadraticKernel(1,1.5,2,)"
rng = np.random.default_rng(bench.SEED + 4)
X4 = rng.uniform(-1, 1, (N4, bench.C3_D))
Y4 = np.sin(3.0 * X4[:, :1]) + 0.1 * rng.standard_normal((N4, 1))
h = gpr_b200.Handle(0)
ops, params = gpr_b200.KernelFactory.GetKernel(ks).lower()
h.set_kernel(ops, params)
h.set_train(X4, Y4, 1.0)
h.loglik(1, True, True, False)
t0 = time.perf_counter()
v_gpu, g_gpu, _ = h.loglik(1, True, True, False)
gpu4_s = time.perf_counter() - t0
h.close()
t0 = time.perf_counter()
k = orc.parse_kernel(ks)
Kn = orc.kernel_matrix(k, X4, X4) + np.eye(N4)
C = orc.lu_invert(Kn)                                   # dgetrf + dgetri, as include/LAPACKUtils.h:38-56
alpha = C @ Y4
A = alpha @ alpha.T - C
D = orc.kernel_derivative(k, X4, X4)                    # (P, N, N)
g_cpu = np.array([0.5 * float(np.sum(A * D[p])) for p in range(D.shape[0])])
sign, ld = np.linalg.slogdet(Kn)
v_cpu = -0.5 * float(Y4[:, 0] @ alpha[:, 0]) - 0.5 * ld - 0.5 * N4 * np.log(2 * np.pi)
cpu4_s = time.perf_counter() - t0
out["c4"] = {"N": N4, "kernel": ks, "gpu_iteration_s": gpu4_s, "oracle_s": cpu4_s, "grad_gpu": [float(x) for x in g_gpu],
             "grad_oracle": [float(x) for x in g_cpu], "grad_rel": rel(g_gpu, g_cpu), "value_rel": abs(v_gpu[0] - v_cpu) / abs(v_cpu),
             "note": "oracle: explicit LU inverse, materialised dK stack, tr(A D_p) as sum(A o D_p); true (unclamped) log det"}
print(json.dumps(out), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "r02_c3_reference_full.json"), "w"), indent=1)
