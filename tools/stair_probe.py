"""One GPU: the staircase GEMM (gprb200_gemm_stair_dev) against the plain call on the trailing-update shape of the
block-cyclic Cholesky (tall C, K = nb): cost of the extended kernel variant and of the tiles that exit at the stair."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpr_b200

h = gpr_b200.Handle(0)
h.set_stream(torch.cuda.current_stream().cuda_stream)
out = []
for (M, N, K) in [(65536, 8192, 1024), (32768, 16384, 1024), (65536, 1024, 1024)]:
    A = torch.randn(M, K, dtype=torch.float64, device="cuda")
    B = torch.randn(N, K, dtype=torch.float64, device="cuda")
    Cm = torch.zeros(M, N, dtype=torch.float64, device="cuda")
    zero = torch.zeros(N // 64 + 1, dtype=torch.int32, device="cuda")
    diag = (torch.arange(N // 64 + 1, dtype=torch.int32, device="cuda") * 64 * (M // N)).contiguous()   # half of the tiles dead
    def plain():
        h._ck(h.lib.gprb200_gemm_dev(h.h, 0, M, N, K, -1.0, A.data_ptr(), K, B.data_ptr(), K, 1.0, Cm.data_ptr(), N, 0))
    def stair(t):
        h._ck(h.lib.gprb200_gemm_stair_dev(h.h, M, N, K, -1.0, A.data_ptr(), K, B.data_ptr(), K, 1.0, Cm.data_ptr(), N, t.data_ptr(), 0))
    res = {"M": M, "N": N, "K": K}
    for name, fn, frac in [("plain", plain, 1.0), ("stair_all_live", lambda: stair(zero), 1.0), ("stair_half_dead", lambda: stair(diag), 0.5)]:
        for _ in range(2):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            fn()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        res[name] = {"ms": ms, "TFs_nominal": 2.0 * M * N * K * frac / ms * 1e-9}
    out.append(res)
    print(json.dumps(res), flush=True)
    del A, B, Cm
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/r02_stair_probe.json", "w"), indent=1)
