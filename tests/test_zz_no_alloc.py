"""-m gpu: gprb200_gemv_t_noalloc_dev, the allocation-free Predict GEMV (scratch in a static __device__ array,
gpr_b200/csrc/predict.cu), called through the C-ABI and compared with the existing entry point gprb200_predict_dev
(T = 1) on two calls in a row. Collected last on purpose.

The entry point has no counterpart in the reference."""
import ctypes as C

import numpy as np
import pytest

from tests.synth import KERNELS, make_problem, rel_err

pytestmark = pytest.mark.gpu


def _model(N, d, M):
    import gpr_b200
    X, Y = make_problem(N, d, M, seed=77)
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(KERNELS["sum3"]).lower()
    h.set_kernel(ops, params)
    h.set_train(X, Y, 1.0)
    h.learn(0)
    return h


def test_noalloc_gemv_matches_predict_dev_twice_in_a_row():
    import torch
    N, d, M = 2000, 5, 301            # 8 row splits x 302 columns of partial sums -> the static scratch is used
    h = _model(N, d, M)
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(11)
    Xs = rng.uniform(-1, 1, (2, d))
    ap, ld = h.alpha_dev()            # the library's own regression vectors, N x ld on the device
    Kx = torch.from_numpy(h.kernel_vector(Xs)).to(dev)          # 2 x N kernel vectors (existing entry point)
    xs_dev = torch.from_numpy(Xs).to(dev)
    want = torch.empty(2, M, dtype=torch.float64, device=dev)
    got = torch.full((2, M), float("nan"), dtype=torch.float64, device=dev)
    for t in range(2):                # two calls in a row: the second one finds the first one's partials in the scratch
        h.predict_dev(xs_dev[t:t + 1].data_ptr(), 1, want[t:t + 1].data_ptr(), M)
        h._ck(h.lib.gprb200_gemv_t_noalloc_dev(h.h, C.c_void_p(Kx[t].data_ptr()), C.c_void_p(ap), ld, N, M,
                                               C.c_void_p(got[t].data_ptr())))
    h.synchronize()
    torch.cuda.synchronize()
    w, g = want.cpu().numpy(), got.cpu().numpy()
    assert np.isfinite(g).all()
    assert rel_err(g[0], w[0]) <= 1e-9 and rel_err(g[1], w[1]) <= 1e-9      # the suite's tolerance on the mean
    assert np.abs(g[0] - g[1]).max() > 0                                     # two different test points, two answers
    # and against the oracle-independent host product of the same operands
    A = h.alpha()
    assert rel_err(g, Kx.cpu().numpy() @ A) <= 1e-12
    h.close()


def test_noalloc_gemv_rejects_shapes_beyond_the_static_scratch():
    import torch
    import gpr_b200
    h = gpr_b200.Handle(0)
    buf = torch.zeros(16, dtype=torch.float64, device="cuda")
    p = C.c_void_p(buf.data_ptr())
    # N = 4096, M = 2^20: 5 row splits x 2^20 columns > 4 Mi doubles -> refused before anything is launched
    rc = h.lib.gprb200_gemv_t_noalloc_dev(h.h, p, p, 1 << 20, 4096, 1 << 20, p)
    assert rc == -2
    msg = h.lib.gprb200_last_error(h.h).decode()
    assert "static scratch" in msg and "4194304" in msg
    rc = h.lib.gprb200_gemv_t_noalloc_dev(h.h, p, p, 3, 8, 3, p)             # odd leading dimension
    assert rc == -2
    h.close()
