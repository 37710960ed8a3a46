"""C2 end to end at FULL size from pageable host vectors: N = 4096 samples, d = 30, M = 786432 output
columns - the motion model's label matrix is 25.8 GB, stored as the reference stores it (one allocation per sample,
std::vector<VectorType>); Learn = host label vectors -> Gram + Cholesky + solve of all columns -> host regression vectors
(25.8 GB back), through gprb200_learn_rows (upload / solve / download pipelined over column blocks). Checked on sampled
columns against the CPU oracle's kernel matrix. Writes gpurun_out/r02_c2_e2e_full.json.

TEST INFRASTRUCTURE: uses oracle/ as the checker."""
import json, os, sys, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import gpr_b200  # noqa: E402
from oracle import gpr_oracle as orc  # noqa: E402

N, d = 4096, 30
M = int(sys.argv[1]) if len(sys.argv) > 1 else 786432
import psutil
out = {"N": N, "d": d, "M": M, "label_gigabytes": 8.0 * N * M * 1e-9, "host_ram_available_GB_before": psutil.virtual_memory().available / 1e9}
rng = np.random.default_rng(5)
X = rng.standard_normal((N, d))
t0 = time.perf_counter()
base = rng.standard_normal(M)
ai, bi = rng.standard_normal(N), rng.standard_normal(N)
rows = [base * ai[i] + bi[i] for i in range(N)]          # one allocation per sample, pageable
out["generate_s"] = time.perf_counter() - t0
A = np.empty((N, M))
A[::512] = 0.0                                            # touch part of the destination; the rest faults in the warm-up
h = gpr_b200.Handle(0)
ops, params = gpr_b200.KernelFactory.GetKernel(bench.C2_KERNEL).lower()
h.set_kernel(ops, params)
xr = [x for x in X]
t0 = time.perf_counter()
h.learn_rows(xr, rows, bench.C2_SIGMA, A)                 # warm-up: pinned rings, device buffers, page faults of A
out["first_call_s"] = time.perf_counter() - t0
ts = []
for _ in range(2):
    t0 = time.perf_counter()
    h.learn_rows(xr, rows, bench.C2_SIGMA, A)
    ts.append(time.perf_counter() - t0)
out["learn_e2e_pipelined_s"] = min(ts)
out["learn_e2e_pipelined_s_all"] = ts
flops = bench.learn_flops(N, M)
out["useful_tflops_e2e"] = flops / min(ts) * 1e-12
out["host_bytes_moved_GB"] = 2 * 8.0 * N * M * 1e-9
out["host_GBs_each_way_if_serial"] = out["host_bytes_moved_GB"] / min(ts)
# device-resident solve of the same model for comparison (what `value` of the C2 bench object times)
cols = rng.choice(M, 24, replace=False)
k = orc.parse_kernel(bench.C2_KERNEL)
K = orc.kernel_matrix(k, X, X) + bench.C2_SIGMA ** 2 * np.eye(N)
Y = np.stack([rows[i][cols] for i in range(N)])
res = K @ A[:, cols] - Y
out["rel_residual_sampled_columns"] = float(np.max(np.abs(res)) / np.max(np.abs(Y)))
out["note"] = ("label vectors y_i = a_i * base + b_i (rank-2 Y: cheap to generate at 25.8 GB; the solve does not exploit it); "
               "residual |K alpha - y| / max|y| on 24 sampled columns with K from the CPU oracle")
h.close()
print(json.dumps(out), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "r02_c2_e2e_full.json"), "w"), indent=1)
