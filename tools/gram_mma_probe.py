"""Times the Gram build for a large input dimension: DMMA cross-term path (gram_mma.cu) vs the direct-difference
kernels, symmetric full K (gprb200_gram_dev) on one B200. python tools/gram_mma_probe.py"""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpr_b200
out = []
for ks, d, N, xs in (("GaussianKernel(35,30,)", 238, 32768, 3.0), ("GaussianKernel(35,30,)", 238, 16384, 3.0),
                     ("GaussianKernel(8.215838362577491,1,)", 30, 32768, 1.0), ("GaussianKernel(12,1,)", 64, 32768, 1.0)):
    X = torch.randn(N, d, dtype=torch.float64, device="cuda") * xs
    Y = torch.zeros(N, 1, dtype=torch.float64, device="cuda")
    ldk = (N + 15) // 16 * 16
    K = torch.empty(N, ldk, dtype=torch.float64, device="cuda")
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(ks).lower()
    h.set_kernel(ops, params)
    h.set_option("gram_mma_min_d", 16)
    h.set_train_dev(X.data_ptr(), N, d, Y.data_ptr(), 1, 1.0)
    rec = {"kernel": ks, "d": d, "N": N}
    keep = {}
    for mma in (1, 0):
        h.set_option("gram_mma", mma)
        t = []
        for _ in range(3):
            h._ck(h.lib.gprb200_gram_dev(h.h, K.data_ptr(), ldk, 1))
            t.append(h.phase_ms()["gram"])
        keep[mma] = K[:2048, :2048].clone()
        key = "dmma" if mma else "direct"
        rec[key + "_ms"] = round(min(t), 3)
        rec[key + "_GBs_written"] = round(8.0 * N * N / min(t) * 1e-6, 1)
        rec[key + "_pair_TFLOPs"] = round(N * (N + 1) / 2 * 2.0 * d / min(t) * 1e-9, 2)
    rec["max_abs_diff_block"] = float((keep[1] - keep[0]).abs().max())
    # Learn's lower-triangle build
    for mma in (1, 0):
        h.set_option("gram_mma", mma)
        h.factorize(); h.factorize()
        rec[("dmma" if mma else "direct") + "_learn_gram_ms"] = round(h.phase_ms()["gram"], 3)
    h.close()
    out.append(rec)
    print(json.dumps(rec), flush=True)
    del X, K
    torch.cuda.empty_cache()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/r02_gram_mma_probe.json", "w"), indent=1)
