"""Efficiency of the Cholesky trailing update as a function of the trailing size: C (m x m, lower tiles) -= P P^T with
P m x 1024 (T^b), and the block-column update m x 1024 by K = 1024 (T^a). Live CUDA-event timing, warm."""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
import gpr_b200  # noqa: E402

h = gpr_b200.Handle(0)
st = torch.cuda.current_stream()
h.set_stream(st.cuda_stream)
K = 1024
N = 32768
A = torch.randn(N, N, device="cuda", dtype=torch.float64)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def run(M, Ncols, lower, reps=3):
    P = A[N - M:, :K]
    Cm = A[N - M:, N - Ncols:] if not lower else A[N - M:, N - M:]
    B = P if lower else A[N - Ncols:, :K]
    args = (h.h, 0, M, Cm.shape[1], K, ctypes.c_double(-1.0), P.data_ptr(), P.stride(0), B.data_ptr(), B.stride(0),
            ctypes.c_double(1.0), Cm.data_ptr(), Cm.stride(0), 1 if lower else 0)
    h._ck(h.lib.gprb200_gemm_dev(*args))
    e0.record(st)
    for _ in range(reps):
        h._ck(h.lib.gprb200_gemm_dev(*args))
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    fl = (M * (M + 1.0) * K) if lower else 2.0 * M * Cm.shape[1] * K
    return ms, fl / ms * 1e-9


for m in (31744, 24576, 16384, 12288, 8192, 6144, 4096, 3072, 2048, 1024):
    ms, tf = run(m, m, True)
    ms2, tf2 = run(m, 1024, False)
    print("m %6d  T^b %8.3f ms %6.2f TF (%.1f%%)   T^a %7.3f ms %6.2f TF (%.1f%%)" % (m, ms, tf, tf / 0.371, ms2, tf2, tf2 / 0.371), flush=True)
