"""Small driver for ncu captures of the round-2 kernels: single-point credible intervals through inv(L) (trmv_lower_kernel)
at N = 16384 and the large-d Gram (gram_mma_transform_kernel + DMMA GEMM) at d = 238, N = 8192."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpr_b200
rng = np.random.default_rng(1)
N = 16384
X = rng.uniform(-1, 1, (N, 8)); Y = np.sin(3 * X[:, :1])
h = gpr_b200.Handle(0)
ops, params = gpr_b200.KernelFactory.GetKernel("GaussianKernel(2,1,)").lower()
h.set_kernel(ops, params); h.set_train(X, Y, 1.0); h.learn(0)
for i in range(8):
    h.credible_interval(X[i:i + 1] + 0.01)
h.close()
N, d = 8192, 238
X = torch.randn(N, d, dtype=torch.float64, device="cuda") * 3
Y = torch.zeros(N, 1, dtype=torch.float64, device="cuda")
K = torch.empty(N, N, dtype=torch.float64, device="cuda")
h = gpr_b200.Handle(0)
ops, params = gpr_b200.KernelFactory.GetKernel("GaussianKernel(35,30,)").lower()
h.set_kernel(ops, params); h.set_train_dev(X.data_ptr(), N, d, Y.data_ptr(), 1, 1.0)
for _ in range(2):
    h._ck(h.lib.gprb200_gram_dev(h.h, K.data_ptr(), N, 1))
h.synchronize(); h.close()
print("ok")
