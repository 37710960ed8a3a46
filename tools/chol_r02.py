"""Round-2 probe of the diagonal-tile chain: correctness of the tile kernels against numpy (spd_inverse: factor +
inverse + log det of random SPD matrices, ragged sizes), then the Cholesky phase time at N = 4096 ... 32768 for
potrf_tile = 2 (round 1) and 3 (DMMA tile), with cuSOLVER's potrf (torch.linalg.cholesky; yardstick only, no product
path calls it) on the same box. Writes gpurun_out/chol_r02.json."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpr_b200  # noqa: E402

out = {"gpu": torch.cuda.get_device_name(0), "tile_check": [], "chol": [], "cusolver": []}
rng = np.random.default_rng(5)
h = gpr_b200.Handle(0)
for tile in (3, 2):
    h.set_option("potrf_tile", tile)
    for n in (1, 31, 32, 33, 100, 128, 129, 200, 257, 640, 1000):
        B = rng.standard_normal((n, n))
        A = B @ B.T / n + np.eye(n) * (0.05 if n > 1 else 1.0)
        inv, ld = h.spd_inverse(A)
        ref = np.linalg.inv(A)
        sgn, ldr = np.linalg.slogdet(A)
        out["tile_check"].append({"tile": tile, "n": n, "inv_rel": float(np.abs(inv - ref).max() / np.abs(ref).max()),
                                  "logdet_abs": float(abs(ld - ldr))})
    A = np.eye(200)   # a non-SPD matrix must report LAPACK-style INFO
    A[150, 150] = -1.0
    try:
        h.spd_inverse(A)
        out["tile_check"].append({"tile": tile, "non_spd": "NOT DETECTED"})
    except gpr_b200.GprError as e:
        out["tile_check"].append({"tile": tile, "non_spd_info": e.code})
h.close()
print(json.dumps(out["tile_check"]), flush=True)

KS = "SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))"
sizes = [int(a) for a in sys.argv[1:]] or [4096, 8192, 16384, 32768]
for N in sizes:
    X = torch.rand(N, 8, dtype=torch.float64, device="cuda") * 2 - 1
    Y = torch.randn(N, 1, dtype=torch.float64, device="cuda")
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(KS).lower()
    h.set_kernel(ops, params)
    h.set_train_dev(X.data_ptr(), N, 8, Y.data_ptr(), 1, 1.0)
    alphas = {}
    for tile in (2, 3, 2, 3):
        h.set_option("potrf_tile", tile)
        h.learn(0)
        t = []
        for _ in range(3):
            h.learn(0)
            t.append(h.phase_ms()["potrf"])
        alphas[tile] = h.alpha()
        rec = {"N": N, "tile": tile, "potrf_ms": round(min(t), 3), "tflops": round(N ** 3 / 3.0 / min(t) * 1e-9, 2)}
        out["chol"].append(rec)
        print(json.dumps(rec), flush=True)
    out["chol"].append({"N": N, "alpha_tile3_vs_tile2_rel": float(np.abs(alphas[3] - alphas[2]).max() / np.abs(alphas[2]).max())})
    K = torch.empty(N, (N + 15) // 16 * 16, dtype=torch.float64, device="cuda")   # cuSOLVER yardstick, same matrix
    h._ck(h.lib.gprb200_gram_dev(h.h, K.data_ptr(), K.stride(0), 1))
    h.synchronize()
    h.close()
    Kc = K[:, :N].contiguous()
    del K
    torch.linalg.cholesky(Kc[:1024, :1024])
    best = 1e30
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        L = torch.linalg.cholesky(Kc)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
        del L
    rec = {"N": N, "cusolver_potrf_ms": round(best, 3), "tflops": round(N ** 3 / 3.0 / best * 1e-9, 2)}
    out["cusolver"].append(rec)
    print(json.dumps(rec), flush=True)
    del Kc, X, Y
    torch.cuda.empty_cache()

os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/chol_r02.json", "w"), indent=1)
