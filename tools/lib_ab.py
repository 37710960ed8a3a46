"""A/B of whole library builds on one B200 (GPRB200_LIB selects the .so): min of 4 timed Cholesky factorisations per N."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpr_b200
KS = "SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))"
out = {"lib": os.environ.get("GPRB200_LIB", "default")}
for N in [int(a) for a in sys.argv[1:]] or [8192, 16384, 32768]:
    X = torch.rand(N, 8, dtype=torch.float64, device="cuda") * 2 - 1
    Y = torch.randn(N, 1, dtype=torch.float64, device="cuda")
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(KS).lower()
    h.set_kernel(ops, params)
    h.set_train_dev(X.data_ptr(), N, 8, Y.data_ptr(), 1, 1.0)
    h.learn(0)
    t = []
    for _ in range(4):
        h.learn(0)
        t.append(h.phase_ms()["potrf"])
    out[str(N)] = round(min(t), 2)
    h.close()
print(json.dumps(out), flush=True)
