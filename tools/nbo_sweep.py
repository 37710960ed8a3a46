"""Outer-block (NBO) sweep of the single-GPU Cholesky: python tools/nbo_sweep.py N..."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpr_b200
KS = "SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))"
for N in [int(a) for a in sys.argv[1:]] or [8192, 12288, 16384, 24576, 32768]:
    X = torch.rand(N, 8, dtype=torch.float64, device="cuda") * 2 - 1
    Y = torch.randn(N, 1, dtype=torch.float64, device="cuda")
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(KS).lower()
    h.set_kernel(ops, params)
    h.set_train_dev(X.data_ptr(), N, 8, Y.data_ptr(), 1, 1.0)
    for nbo in (128, 256, 384, 512, 768, 1024):
        for split in (0, 1):
            h.set_option("nbo", nbo); h.set_option("panel_split", split)
            h.learn(0)
            t = []
            for _ in range(3):
                h.learn(0); t.append(h.phase_ms()["potrf"])
            print(json.dumps({"N": N, "nbo": nbo, "panel_split": split, "potrf_ms": round(min(t), 3),
                              "tflops": round(N ** 3 / 3.0 / min(t) * 1e-9, 2)}), flush=True)
    h.close()
