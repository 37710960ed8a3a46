"""One Gram build per kernel family (for ncu): python tools/gram_one.py N [terms]"""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import gpr_b200  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
terms = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dev = torch.device("cuda", 0)
for ks in ("GaussianKernel(2,1,)", "SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))"):
    rng = np.random.default_rng(3)
    X = rng.uniform(-1, 1, (N, 8))
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(ks).lower()
    h.set_kernel(ops, params)
    h.set_train(X, rng.standard_normal((N, 1)), 1.0)
    h.set_option("gram_terms", terms)
    out = torch.empty((N, N), dtype=torch.float64, device=dev)
    for _ in range(2):
        h._ck(h.lib.gprb200_gram_block_dev(h.h, 0, N, 0, N, out.data_ptr(), N))
    torch.cuda.synchronize()
    print(ks, float(out[5, 7]))
    # fused derivative contraction (gram_dcontract_kernel) inside one log-likelihood + gradient evaluation
    v, g, _ = h.loglik(1, True, True, False)
    print("loglik", v, g)
    h.close()
