"""GEMM / Cholesky variant probe on one B200 (not the bench contract): times gprb200_gemm_dev with CUDA events."""
import json, sys
import torch
import gpr_b200
from tests.synth import KERNELS


def time_gemm(h, form, M, N, K, beta, lower, reps=3):
    A = torch.randn(M if form != 2 else K, K if form != 2 else M, dtype=torch.float64, device="cuda")
    B = torch.randn(N if form == 0 else K, K if form == 0 else N, dtype=torch.float64, device="cuda")
    Cm = torch.zeros(M, N, dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream()
    h.set_stream(st.cuda_stream)
    def run():
        h._ck(h.lib.gprb200_gemm_dev(h.h, form, M, N, K, 1.0, A.data_ptr(), A.shape[1], B.data_ptr(), B.shape[1],
                                     beta, Cm.data_ptr(), N, lower))
    run(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(reps):
        run()
    e1.record(st); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    flops = 2.0 * M * N * K * (0.5 * (1 + 1.0 / M) if lower else 1.0)
    return ms, flops / ms * 1e-9


def main():
    h = gpr_b200.Handle(0)
    shapes = [("syrk_lower K=1024", 0, 30720, 30720, 1024, 1.0, 1), 
              ("leftlook 16k x 1k x 16k", 0, 16384, 1024, 16384, 1.0, 0), 
              ("square 8192^3 NT", 0, 8192, 8192, 8192, 0.0, 0), ("square 8192^3 TN", 2, 8192, 8192, 8192, 0.0, 0)]
    import os
    variants = [int(v) for v in os.environ.get('VARIANTS', '0,1').split(',')]
    h.set_option('gemm_stages', int(os.environ.get('STAGES', '3')))
    for variant in variants:
        h.set_option("gemm_variant", variant)
        for name, form, M, N, K, beta, lower in shapes:
            ms, tf = time_gemm(h, form, M, N, K, beta, lower)
            print(json.dumps({"variant": variant, "shape": name, "ms": round(ms, 3), "TFLOPs": round(tf, 2)}), flush=True)
    # whole Cholesky
    N = 32768
    X = torch.rand(N, 8, dtype=torch.float64, device="cuda") * 2 - 1
    Y = torch.randn(N, 1, dtype=torch.float64, device="cuda")
    ops, params = gpr_b200.KernelFactory.GetKernel(KERNELS["gaussian"]).lower()
    h.set_kernel(ops, params)
    h.set_train_dev(X.data_ptr(), N, 8, Y.data_ptr(), 1, 1.0)
    for variant in variants:
        h.set_option("gemm_variant", variant)
        for nbo in (1024,):
            h.set_option("nbo", nbo)
            h.learn(0); h.learn(0)
            ms = h.phase_ms()
            print(json.dumps({"variant": variant, "nbo": nbo, "potrf_ms": ms["potrf"],
                              "potrf_TFs": N ** 3 / 3.0 / ms["potrf"] * 1e-9, "solve_ms": ms["solve"]}), flush=True)
    h.close()


if __name__ == "__main__":
    main()
