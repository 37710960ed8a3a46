import json, os, torch, gpr_b200
from tests.synth import KERNELS
N = int(os.environ.get("N", 32768))
X = torch.rand(N, 8, dtype=torch.float64, device="cuda") * 2 - 1
Y = torch.randn(N, 1, dtype=torch.float64, device="cuda")
h = gpr_b200.Handle(0)
ops, params = gpr_b200.KernelFactory.GetKernel(KERNELS["gaussian"]).lower()
h.set_kernel(ops, params)
h.set_train_dev(X.data_ptr(), N, 8, Y.data_ptr(), 1, 1.0)
for nbo in (512, 768, 1024, 1536, 2048):
    for la in (0, 1):
        h.set_option("nbo", nbo); h.set_option("lookahead", la)
        h.learn(0); h.learn(0)
        ms = h.phase_ms()
        print(json.dumps({"nbo": nbo, "lookahead": la, "potrf_ms": round(ms["potrf"], 1), "TFs": round(N ** 3 / 3.0 / ms["potrf"] * 1e-9, 2)}), flush=True)
