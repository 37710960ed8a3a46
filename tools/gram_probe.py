"""Gram-build A/B: term-form value kernel (gram_terms=1) vs the per-entry interpreter (0), lower triangle as in Learn.
Device time from the library's own phase timers (CUDA events around the Gram inside gprb200_factorize is mixed with
potrf, so the Gram is timed through gprb200_gram_block_dev on a preallocated torch buffer)."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import gpr_b200  # noqa: E402

KERNELS = {
    "gauss": "GaussianKernel(2,1,)",
    "c3": "SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))",
    "rq": "RationalQuadraticKernel(1,1.5,2,)",
    "general13": None,
}


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
    dev = torch.device("cuda", 0)
    for name, ks in KERNELS.items():
        for d in (8, 30):
            rng = np.random.default_rng(3)
            X = rng.uniform(-1, 1, (N, d))
            Y = rng.standard_normal((N, 1))
            k = gpr_b200.GetGeneralKernel([1.0 + 0.05 * i for i in range(13)]) if ks is None else gpr_b200.KernelFactory.GetKernel(ks)
            ops, params = k.lower()
            h = gpr_b200.Handle(0)
            stream = torch.cuda.current_stream(dev)
            h.set_stream(stream.cuda_stream)
            h.set_kernel(ops, params)
            h.set_train(X, Y, 1.0)
            out = torch.empty((N, N), dtype=torch.float64, device=dev)
            res = {}
            for terms in (0, 1, 2):
                h.set_option("gram_terms", terms)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                for it in range(3):
                    if it == 1:
                        e0.record(stream)
                    h._ck(h.lib.gprb200_gram_block_dev(h.h, 0, N, 0, N, out.data_ptr(), N))
                e1.record(stream)
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / 2
                res[terms] = (ms, out[:2048, :2048].clone())
            diff = (res[0][1] - res[2][1]).abs().max().item()
            rel = ((res[0][1] - res[2][1]).abs() / res[0][1].abs().clamp_min(1e-300)).max().item()
            print(f"{name:10s} d={d:2d} N={N}: full N x N cross-form Gram  interp {res[0][0]:7.2f} ms  terms {res[1][0]:7.2f} ms  rows {res[2][0]:7.2f} ms "
                  f"({8.0 * N * N / res[2][0] * 1e-6:7.0f} GB/s written)  max|diff| {diff:.2e} max rel {rel:.2e}", flush=True)
            h.close()
            del out


if __name__ == "__main__":
    main()
