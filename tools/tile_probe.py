import os, torch, gpr_b200
from tests.synth import KERNELS
N = 4096
X = torch.rand(N, 8, dtype=torch.float64, device="cuda") * 2 - 1
Y = torch.randn(N, 1, dtype=torch.float64, device="cuda")
h = gpr_b200.Handle(0)
ops, params = gpr_b200.KernelFactory.GetKernel(KERNELS["gaussian"]).lower()
h.set_kernel(ops, params)
h.set_train_dev(X.data_ptr(), N, 8, Y.data_ptr(), 1, 1.0)
for v in (1, 2):
    h.set_option("potrf_tile", v)
    h.learn(0)
    print(v, h.phase_ms()["potrf"])
