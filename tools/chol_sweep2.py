import json, os, sys, torch
sys.path.insert(0, ".")
import gpr_b200
from tests.synth import KERNELS
for N in (8192, 16384, 24576, 32768):
    X = torch.rand(N, 8, dtype=torch.float64, device="cuda") * 2 - 1
    Y = torch.randn(N, 1, dtype=torch.float64, device="cuda")
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(KERNELS["gaussian"]).lower()
    h.set_kernel(ops, params)
    h.set_train_dev(X.data_ptr(), N, 8, Y.data_ptr(), 1, 1.0)
    for nbo in (256, 384, 512, 768, 1024):
        h.set_option("nbo", nbo)
        best = 1e9
        for _ in range(3):
            h.learn(0)
            best = min(best, h.phase_ms()["potrf"])
        print(json.dumps({"N": N, "nbo": nbo, "potrf_ms": round(best, 2), "TFs": round(N ** 3 / 3.0 / best * 1e-9, 2)}), flush=True)
    h.close()
