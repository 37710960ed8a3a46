"""A/B of Cholesky options on one B200: python tools/chol_ab.py N... ; each config = dict of set_option values."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpr_b200
KS = "SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))"
CONFIGS = [{"panel_split": 0, "gemm_small": 0}, {"panel_split": 0, "gemm_small": 1}, {"panel_split": 1, "gemm_small": 0},
           {"panel_split": 1, "gemm_small": 1}]
for N in [int(a) for a in sys.argv[1:]] or [2048, 4096, 8192, 16384, 32768]:
    X = torch.rand(N, 8, dtype=torch.float64, device="cuda") * 2 - 1
    Y = torch.randn(N, 1, dtype=torch.float64, device="cuda")
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(KS).lower()
    h.set_kernel(ops, params)
    h.set_train_dev(X.data_ptr(), N, 8, Y.data_ptr(), 1, 1.0)
    base = None
    for cfg in CONFIGS:
        for k, v in cfg.items():
            h.set_option(k, v)
        h.learn(0)
        t = []
        for _ in range(3):
            h.learn(0)
            t.append(h.phase_ms()["potrf"])
        a = h.alpha()
        if base is None:
            base = a
        print(json.dumps({"N": N, **cfg, "potrf_ms": round(min(t), 3), "tflops": round(N ** 3 / 3.0 / min(t) * 1e-9, 2),
                          "alpha_vs_first": float(np.abs(a - base).max() / np.abs(base).max())}), flush=True)
    h.close()
