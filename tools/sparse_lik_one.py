"""One sparse log-likelihood + gradient evaluation at n = 262144, m = 4096 (for an ncu launch list)."""
import sys
import numpy as np
sys.path.insert(0, ".")
import gpr_b200
n, m, d = 262144, 4096, 8
rng = np.random.default_rng(91)
X = rng.uniform(-1, 1, (n, d))
Y = np.sin(3.0 * X[:, :1]) + 0.1 * rng.standard_normal((n, 1))
Xm = rng.uniform(-1, 1, (m, d))
h = gpr_b200.Handle(0)
ops, params = gpr_b200.KernelFactory.GetKernel("SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))").lower()
h.set_kernel(ops, params)
h.set_train(X, Y, 1.0)
h._ck(h.lib.gprb200_sparse_set_inducing(h.h, Xm.ctypes.data, m, d, 1e-2))
v, g = np.empty(1), np.empty(len(params))
h._ck(h.lib.gprb200_sparse_loglik(h.h, 1, v.ctypes.data, g.ctypes.data, None))
print(v, g)
h.close()
