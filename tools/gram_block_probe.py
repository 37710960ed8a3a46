"""Tall block-column Gram (the block-cyclic layout's unit of work): rows x 1024 block at a row/column offset."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import gpr_b200
N = 131072
rng = np.random.default_rng(3)
X = rng.uniform(-1, 1, (N, 8))
h = gpr_b200.Handle(0)
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream(dev)
h.set_stream(stream.cuda_stream)
ops, params = gpr_b200.KernelFactory.GetKernel("SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))").lower()
h.set_kernel(ops, params)
h.set_train(X, None, 1.0)
for terms in (0, 2):
    h.set_option("gram_terms", terms)
    for r0 in (0, 65536, 122880):
        rows = N - r0
        out = torch.empty((rows, 1024), dtype=torch.float64, device=dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for it in range(3):
            if it == 1:
                e0.record(stream)
            h._ck(h.lib.gprb200_gram_block_dev(h.h, r0, rows, r0, 1024, out.data_ptr(), 1024))
        e1.record(stream)
        torch.cuda.synchronize()
        print(f"terms {terms} block {rows} x 1024 at {r0}: {e0.elapsed_time(e1) / 2:.3f} ms", flush=True)
h.close()
