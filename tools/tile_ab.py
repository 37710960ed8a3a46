import json, torch, gpr_b200
from tests.synth import KERNELS
for N in (8192, 32768):
    X = torch.rand(N, 8, dtype=torch.float64, device="cuda") * 2 - 1
    Y = torch.randn(N, 1, dtype=torch.float64, device="cuda")
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(KERNELS["gaussian"]).lower()
    h.set_kernel(ops, params)
    h.set_train_dev(X.data_ptr(), N, 8, Y.data_ptr(), 1, 1.0)
    for la in (0, 1):
        for v in (1, 2, 1, 2):
            h.set_option("potrf_tile", v); h.set_option("lookahead", la)
            h.learn(0); h.learn(0)
            print(json.dumps({"N": N, "tile": v, "lookahead": la, "potrf_ms": round(h.phase_ms()["potrf"], 2)}), flush=True)
    h.close()
