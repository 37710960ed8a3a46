"""One warm-up + one timed Learn at N (default 4096) for ncu launch lists: python tools/chol_one.py N"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpr_b200
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
X = torch.rand(N, 8, dtype=torch.float64, device="cuda") * 2 - 1
Y = torch.randn(N, 1, dtype=torch.float64, device="cuda")
h = gpr_b200.Handle(0)
ops, params = gpr_b200.KernelFactory.GetKernel("SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))").lower()
h.set_kernel(ops, params)
h.set_train_dev(X.data_ptr(), N, 8, Y.data_ptr(), 1, 1.0)
h.learn(0)
n0 = h.launch_count()
h.learn(0)
print(N, h.launch_count() - n0, h.phase_ms())
h.close()
