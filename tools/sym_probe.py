import sys
import numpy as np
sys.path.insert(0, ".")
import gpr_b200
from tests.synth import KERNELS, make_problem
X, Y = make_problem(300, 3, 1, seed=1, dup=3)
for kname in ("general13", "rq", "prod"):
    if kname not in KERNELS:
        continue
    for terms in (2, 1, 0):
        h = gpr_b200.Handle(0)
        ops, params = gpr_b200.KernelFactory.GetKernel(KERNELS[kname]).lower()
        h.set_kernel(ops, params)
        h.set_train(X, Y, 0.7)
        h.set_option("gram_terms", terms)
        K = h.gram(add_noise=False)
        bad = np.argwhere(K != K.T)
        print(kname, "terms", terms, "asym entries", len(bad), bad[:6].tolist(),
              [(K[i, j], K[j, i], K[i, j] - K[j, i]) for i, j in bad[:3]])
        h.close()
