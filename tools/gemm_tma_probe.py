"""A/B of the persistent TMA GEMM (csrc/dgemm_tma.cu, option gemm_tma) against the cp.async kernel (csrc/dgemm.cu) on one
B200: (1) correctness on ragged / lower / beta shapes against torch.matmul (yardstick only) and on the EXACT call shapes
of the blocked Cholesky (operands and C inside one N x N buffer, ld = N), (2) CUDA-event timings of the trailing-update
shapes, (3) the whole factorisation. python tools/gemm_tma_probe.py [check|time|all]"""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpr_b200
what = sys.argv[1] if len(sys.argv) > 1 else "all"
h = gpr_b200.Handle(0)
st = torch.cuda.current_stream()
h.set_stream(st.cuda_stream)
h.set_option("gemm_tma_min_tiles", 1)
h.set_option("gemm_tma_min_k", 1)
out = {"check": [], "inplace": [], "time": [], "chol": []}


def gemm(M, N, K, alpha, beta, lower, A, B, Cm):
    h._ck(h.lib.gprb200_gemm_dev(h.h, 0, M, N, K, alpha, A.data_ptr(), A.stride(0), B.data_ptr(), B.stride(0), beta,
                                 Cm.data_ptr(), Cm.stride(0), lower))


torch.manual_seed(3)
if what in ("check", "all"):
    for (M, N, K, alpha, beta, lower) in [(128, 128, 16, 1.0, 0.0, 0), (300, 200, 50, -1.0, 1.0, 0), (1000, 1000, 129, -1.0, 1.0, 1),
                                          (2050, 2050, 1024, -1.0, 1.0, 1), (4097, 130, 1000, 2.0, 0.5, 0), (5000, 5000, 256, -1.0, 1.0, 1)]:
        lda = (K + 1) // 2 * 2 + 2
        ldc = (N + 1) // 2 * 2 + 4
        A = torch.randn(M, lda, dtype=torch.float64, device="cuda")[:, :K]
        B = A[:N] if (lower and M == N) else torch.randn(N, lda, dtype=torch.float64, device="cuda")[:, :K]
        C0 = torch.randn(M, ldc, dtype=torch.float64, device="cuda")[:, :N]
        res = {}
        for tma in (0, 1):
            h.set_option("gemm_tma", tma)
            Cf = torch.empty(M, ldc, dtype=torch.float64, device="cuda")
            Cf[:, :N] = C0
            gemm(M, N, K, alpha, beta, lower, A, B, Cf[:, :N])
            torch.cuda.synchronize()
            ref = alpha * (A @ B.T) + beta * C0
            err = (Cf[:, :N] - ref).abs()
            res[tma] = float((torch.tril(err) if lower else err).max() / ref.abs().max())
        rec = {"M": M, "N": N, "K": K, "lower": lower, "rel_err_cp_async": res[0], "rel_err_tma": res[1]}
        out["check"].append(rec); print(json.dumps(rec), flush=True)
    # the Cholesky's own call shapes: one N x N buffer, panel = columns [k0, k0 + W) of rows r0.., C = trailing block
    for (Nn, W, k0, look) in [(4096, 256, 0, 256), (8192, 256, 0, 256), (8192, 256, 0, 0), (8192, 256, 2048, 256), (8192, 512, 1024, 512),
                              (16384, 512, 0, 512)]:
        ld = Nn
        buf0 = torch.randn(Nn, ld, dtype=torch.float64, device="cuda")
        r0 = k0 + W + look
        mb = Nn - r0
        res = {}
        keep = {}
        for tma in (0, 1):
            h.set_option("gemm_tma", tma)
            buf = buf0.clone()
            P = buf[r0:, k0:k0 + W]
            Cm = buf[r0:, r0:]
            gemm(mb, mb, W, -1.0, 1.0, 1, P, P, Cm)
            torch.cuda.synchronize()
            keep[tma] = torch.tril(Cm).clone()
            del buf
        ref = torch.tril(buf0[r0:, r0:] - buf0[r0:, k0:k0 + W] @ buf0[r0:, k0:k0 + W].T)
        d = (keep[1] - keep[0]).abs()
        rec = {"N": Nn, "W": W, "k0": k0, "mb": mb, "tma_vs_cp_async": float(d.max()),
 "tma_vs_torch": float((keep[1] - ref).abs().max()),
               "cp_async_vs_torch": float((keep[0] - ref).abs().max())}
        if rec["tma_vs_cp_async"] > 1e-9:
            bad = torch.nonzero(d > 1e-9)
            rec["nbad"] = int(bad.shape[0]); rec["first_bad"] = bad[0].tolist()
            rec["bad_tiles_sample"] = sorted({(int(a) // 128, int(b) // 128) for a, b in bad[:: max(1, bad.shape[0] // 2000)].tolist()})[:40]
        out["inplace"].append(rec); print(json.dumps(rec), flush=True)
        del buf0, keep, ref, d
        torch.cuda.empty_cache()

if what in ("time", "all"):
    def time_gemm(M, N, K, beta, lower, reps=3):
        A = torch.randn(M, K, dtype=torch.float64, device="cuda")
        B = A if (lower or M == N) else torch.randn(N, K, dtype=torch.float64, device="cuda")
        Cm = torch.zeros(M, N, dtype=torch.float64, device="cuda")
        gemm(M, N, K, -1.0, beta, lower, A, B, Cm); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(reps):
            gemm(M, N, K, -1.0, beta, lower, A, B, Cm)
        e1.record(st); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        return ms, 2.0 * M * N * K * (0.5 * (1 + 1.0 / M) if lower else 1.0) / ms * 1e-9
    for name, M, N, K, beta, lower in [("T^b(0) lower 30720^2 K=1024", 30720, 30720, 1024, 1.0, 1), ("T^b lower 16384^2 K=1024", 16384, 16384, 1024, 1.0, 1),
                                       ("T^b lower 8192^2 K=512", 8192, 8192, 512, 1.0, 1), ("T^b lower 4096^2 K=256", 4096, 4096, 256, 1.0, 1),
                                       ("T^a 30720 x 1024 K=1024", 30720, 1024, 1024, 1.0, 0), ("dist column 65536 x 1024 K=1024", 65536, 1024, 1024, 1.0, 0),
                                       ("square 8192^3", 8192, 8192, 8192, 0.0, 0)]:
        for tma in (0, 1):
            h.set_option("gemm_tma", tma)
            ms, tf = time_gemm(M, N, K, beta, lower)
            rec = {"shape": name, "tma": tma, "ms": round(ms, 3), "TFLOPs": round(tf, 2), "frac_of_37.1": round(tf / 37.1, 3)}
            out["time"].append(rec); print(json.dumps(rec), flush=True)
    h.set_option("gemm_tma_min_tiles", 64); h.set_option("gemm_tma_min_k", 256)
    KS = "SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))"
    for N in (8192, 16384, 32768):
        X = torch.rand(N, 8, dtype=torch.float64, device="cuda") * 2 - 1
        Y = torch.randn(N, 1, dtype=torch.float64, device="cuda")
        ops, params = gpr_b200.KernelFactory.GetKernel(KS).lower()
        h.set_kernel(ops, params)
        h.set_train_dev(X.data_ptr(), N, 8, Y.data_ptr(), 1, 1.0)
        al = {}
        for tma in (0, 1):
            h.set_option("gemm_tma", tma)
            h.learn(0)
            t = []
            for _ in range(2):
                h.learn(0); t.append(h.phase_ms()["potrf"])
            al[tma] = h.alpha()
            rec = {"N": N, "tma": tma, "potrf_ms": round(min(t), 2), "TFLOPs": round(N ** 3 / 3.0 / min(t) * 1e-9, 2)}
            out["chol"].append(rec); print(json.dumps(rec), flush=True)
        rec = {"N": N, "alpha_tma_vs_cp_async": float(abs(al[1] - al[0]).max() / abs(al[0]).max())}
        out["chol"].append(rec); print(json.dumps(rec), flush=True)
h.close()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/r02_gemm_tma_probe.json", "w"), indent=1)
