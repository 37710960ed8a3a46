"""Times the N = 32768 factorisation with whichever gpr_b200 package is in the CURRENT directory (A/B of library
builds on the same box: `cd ab_old && python ../tools/potrf_time.py`)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.getcwd())
import gpr_b200  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
rng = np.random.default_rng(0)
X = rng.uniform(-1, 1, (N, 8))
Y = np.sin(3 * X[:, :1])
h = gpr_b200.Handle(0)
ops, params = gpr_b200.KernelFactory.GetKernel(
    "SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))").lower()
h.set_kernel(ops, params)
h.set_train(X, Y, 1.0)
t = []
for _ in range(5):
    h.learn(0)
    torch.cuda.synchronize()
    t.append(h.phase_ms()["potrf"])
print(os.path.abspath(gpr_b200.__file__), "potrf ms", ["%.1f" % v for v in t], "min %.2f" % min(t))
