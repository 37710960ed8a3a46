"""A/B of library builds (GPRB200_LIB) on the Gram kernels: lower-triangle Learn Gram and full symmetric K at N = 32768."""
import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpr_b200
out = {"lib": os.environ.get("GPRB200_LIB", "default")}
N = 32768
for name, ks in (("c3", "SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))"), ("gauss", "GaussianKernel(2,1,)")):
    X = torch.rand(N, 8, dtype=torch.float64, device="cuda") * 2 - 1
    Y = torch.randn(N, 1, dtype=torch.float64, device="cuda")
    K = torch.empty(N, N, dtype=torch.float64, device="cuda")
    h = gpr_b200.Handle(0)
    h.set_stream(torch.cuda.current_stream().cuda_stream)
    ops, params = gpr_b200.KernelFactory.GetKernel(ks).lower()
    h.set_kernel(ops, params)
    h.set_train_dev(X.data_ptr(), N, 8, Y.data_ptr(), 1, 1.0)
    def full():
        h._ck(h.lib.gprb200_gram_dev(h.h, K.data_ptr(), N, 1))
    for _ in range(2):
        full()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        full()
    e1.record(); torch.cuda.synchronize()
    out[name + "_full_K_ms"] = round(e0.elapsed_time(e1) / 5, 3)
    h.close()
    del K
print(json.dumps(out), flush=True)
