"""Ad-hoc per-phase performance probe (not the bench contract): prints phase times from the handle's
CUDA-event timers."""
import json, sys, time
import numpy as np
import torch
import gpr_b200
from tests.synth import KERNELS

def probe_learn(N, d, M, kstr, sigma=1.0):
    torch.manual_seed(0)
    X = (torch.rand(N, d, dtype=torch.float64, device="cuda") * 2 - 1)
    Y = torch.randn(N, M, dtype=torch.float64, device="cuda")
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(kstr).lower()
    h.set_kernel(ops, params)
    h.set_train_dev(X.data_ptr(), N, d, Y.data_ptr(), M, sigma)
    out = {}
    for it in range(2):
        h.learn(0)
        ms = h.phase_ms()
        out = {"N": N, "d": d, "M": M, "kernel": kstr.split("(")[0],
               "gram_ms": ms["gram"], "gram_GBs_lower": 4.0 * N * N / ms["gram"] * 1e-6,
               "potrf_ms": ms["potrf"], "potrf_TFs": N ** 3 / 3.0 / ms["potrf"] * 1e-9,
               "solve_ms": ms["solve"], "solve_TFs": 2.0 * N * N * M / ms["solve"] * 1e-9}
    print(json.dumps(out), flush=True)
    h.close()

def probe_predict(N, d, M):
    X = torch.randn(N, d, dtype=torch.float64, device="cuda")
    Y = torch.randn(N, M, dtype=torch.float64, device="cuda")
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel("GaussianKernel(8.2,1,)").lower()
    h.set_kernel(ops, params)
    h.set_train_dev(X.data_ptr(), N, d, Y.data_ptr(), M, 1.0)
    del Y
    t0 = time.time(); h.learn(0); h.synchronize(); t1 = time.time()
    ms = h.phase_ms()
    print(json.dumps({"N": N, "M": M, "learn_wall_s": t1 - t0, "solve_ms": ms["solve"],
                      "solve_TFs": 2.0 * N * N * M / ms["solve"] * 1e-9, "potrf_ms": ms["potrf"]}), flush=True)
    xs = torch.randn(1, d, dtype=torch.float64, device="cuda")
    mean = torch.empty(1, M, dtype=torch.float64, device="cuda")
    for it in range(3):
        h.predict_dev(xs.data_ptr(), 1, mean.data_ptr(), M)
    h.synchronize()
    best = 1e9
    for it in range(5):
        h.predict_dev(xs.data_ptr(), 1, mean.data_ptr(), M)
        best = min(best, h.phase_ms()["predict"])
    print(json.dumps({"predict_T1_ms": best, "GBs": 8.0 * N * M / best * 1e-6, "pred_per_s": 1000.0 / best}), flush=True)
    for T in (16, 256):
        xs = torch.randn(T, d, dtype=torch.float64, device="cuda")
        mean = torch.empty(T, M, dtype=torch.float64, device="cuda")
        h.predict_dev(xs.data_ptr(), T, mean.data_ptr(), M)
        h.predict_dev(xs.data_ptr(), T, mean.data_ptr(), M)
        t = h.phase_ms()["predict"]
        print(json.dumps({"predict_T": T, "ms": t, "TFs": 2.0 * T * N * M / t * 1e-9, "pred_per_s": T * 1000.0 / t}), flush=True)
    h.close()

if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "learn"):
        probe_learn(4096, 8, 1, KERNELS["gaussian"])
        probe_learn(8192, 8, 1, KERNELS["gaussian"])
        probe_learn(16384, 8, 64, KERNELS["rq"])
        probe_learn(32768, 8, 1, KERNELS["gaussian"])
        probe_learn(32768, 8, 64, KERNELS["sum3"])
    if which in ("all", "predict"):
        probe_predict(4096, 30, 786432)


def probe_chol_sweep(Ns=(16384, 32768)):
    for N in Ns:
        X = (torch.rand(N, 8, dtype=torch.float64, device="cuda") * 2 - 1)
        Y = torch.randn(N, 1, dtype=torch.float64, device="cuda")
        h = gpr_b200.Handle(0)
        ops, params = gpr_b200.KernelFactory.GetKernel(KERNELS["gaussian"]).lower()
        h.set_kernel(ops, params)
        h.set_train_dev(X.data_ptr(), N, 8, Y.data_ptr(), 1, 1.0)
        for nbo in (256, 512, 1024):
            for la in (0, 1):
                for st in (3, 4):
                    h.set_option("nbo", nbo); h.set_option("lookahead", la); h.set_option("gemm_stages", st)
                    best = 1e9
                    for it in range(2):
                        h.factorize()
                        best = min(best, h.phase_ms()["potrf"])
                    print(json.dumps({"N": N, "nbo": nbo, "lookahead": la, "stages": st, "potrf_ms": round(best, 2),
                                      "TFs": round(N ** 3 / 3.0 / best * 1e-9, 2)}), flush=True)
        h.close()


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "chol":
    probe_chol_sweep()
