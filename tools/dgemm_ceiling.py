"""Measure the cuBLAS DGEMM ceiling (roofline denominator for the fp64 kernels) on this box.

cuBLAS is used ONLY as the yardstick; no product path calls it."""
import json, sys, time
import torch

def bench(n, reps=5, sustained_s=0.0):
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    c = torch.empty_like(a)
    for _ in range(3):
        torch.matmul(a, b, out=c)
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, b, out=c); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    out = {"n": n, "burst_tflops": 2.0 * n**3 / best * 1e-9}
    if sustained_s > 0:
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        t0 = time.time(); k = 0
        e0.record()
        while time.time() - t0 < sustained_s:
            for _ in range(4):
                torch.matmul(a, b, out=c); k += 1
            torch.cuda.synchronize()
        e1.record(); torch.cuda.synchronize()
        out["sustained_tflops"] = 2.0 * n**3 * k / e0.elapsed_time(e1) * 1e-9
    return out

if __name__ == "__main__":
    res = {"gpu": torch.cuda.get_device_name(0), "dgemm": [bench(4096), bench(8192, sustained_s=3.0), bench(16384, reps=3)]}
    # syrk-like and potrf from cusolver for context (yardstick only)
    n = 16384
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    spd = a @ a.T + n * torch.eye(n, dtype=torch.float64, device="cuda")
    torch.linalg.cholesky(spd[:1024, :1024])
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); L = torch.linalg.cholesky(spd); e1.record(); torch.cuda.synchronize()
    res["cusolver_potrf_16384_tflops"] = n**3 / 3.0 / e0.elapsed_time(e1) * 1e-9
    print(json.dumps(res))
    with open(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/dgemm_ceiling.json", "w") as f:
        json.dump(res, f)
