import json, torch, gpr_b200
from tests.synth import KERNELS
for N in (8192, 16384, 32768, 49152):
    X = torch.rand(N, 8, dtype=torch.float64, device="cuda") * 2 - 1
    Y = torch.randn(N, 1, dtype=torch.float64, device="cuda")
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(KERNELS["gaussian"]).lower()
    h.set_kernel(ops, params)
    h.set_train_dev(X.data_ptr(), N, 8, Y.data_ptr(), 1, 1.0)
    ref = None
    for left in (0, 1):
        for nbo in (512, 1024, 2048):
            h.set_option("chol_left", left); h.set_option("nbo", nbo)
            h.learn(0); h.learn(0)
            a = h.alpha()
            if ref is None: ref = a
            ms = h.phase_ms()["potrf"]
            print(json.dumps({"N": N, "left": left, "nbo": nbo, "potrf_ms": round(ms, 2), "TFs": round(N**3/3/ms*1e-9, 2),
                              "alpha_diff": float(abs(a - ref).max() / abs(ref).max())}), flush=True)
    h.close()
