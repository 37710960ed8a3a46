"""-m gpu: the two GEMM forms the block-cyclic Cholesky adds to the C-ABI - gprb200_gemm_stair_dev (staircase C: tiles that
lie before their columns' first needed row are skipped, everything at or below the stair is the full product) and
gprb200_gemm_dev form 3 (NT with a lower-triangular B, k loop cut at the tile's last column) - against torch fp64."""
import pytest

pytestmark = pytest.mark.gpu


def _handle():
    import torch
    import gpr_b200
    h = gpr_b200.Handle(0)
    h.set_stream(torch.cuda.current_stream().cuda_stream)
    return h


@pytest.mark.parametrize("M,N,K,row0", [(1024, 512, 256, 0), (1500, 700, 128, 384), (4096, 2048, 1024, 1024), (130, 64, 40, 0)])
def test_stair_gemm_is_the_full_product_below_the_stair(M, N, K, row0):
    import torch
    h = _handle()
    torch.manual_seed(M + N)
    A = torch.randn(M, K, dtype=torch.float64, device="cuda")
    B = torch.randn(N, K, dtype=torch.float64, device="cuda")
    ldc = N + (N & 1) + 2
    C0 = torch.randn(M, ldc, dtype=torch.float64, device="cuda")
    units = (N + 63) // 64
    # a staircase in the caller's row numbering: unit u needs the rows from stair[u] on
    stair = (row0 + torch.arange(units, device="cuda") * max(1, (M // max(units, 1)) // 64) * 64).to(torch.int32).contiguous()
    Cf = C0.clone()
    n0 = h.launch_count()
    h._ck(h.lib.gprb200_gemm_stair_dev(h.h, M, N, K, -1.0, A.data_ptr(), K, B.data_ptr(), K, 1.0, Cf.data_ptr(), ldc,
                                       stair.data_ptr(), row0))
    assert h.launch_count() == n0 + 1
    ref = C0[:, :N] - A @ B.T
    rows = row0 + torch.arange(M, device="cuda")[:, None]
    need = rows >= stair.long().repeat_interleave(64)[:N][None, :]
    scale = float(ref.abs().max())
    assert float(((Cf[:, :N] - ref) * need).abs().max()) <= 1e-13 * scale
    # what lies above the stair is either untouched or the full product - never anything else
    above = ~need
    d_old, d_new = (Cf[:, :N] - C0[:, :N]).abs(), (Cf[:, :N] - ref).abs()
    assert bool(((torch.minimum(d_old, d_new) <= 1e-13 * scale) | ~above).all())
    # and at least some tiles were skipped when the stair leaves whole 128-row tiles out
    if M >= 1024:
        assert bool(((d_old == 0) & above).any())
    assert torch.equal(Cf[:, N:], C0[:, N:])
    h.close()


@pytest.mark.parametrize("M,N", [(2048, 1024), (1000, 384), (300, 130)])
def test_form3_lower_triangular_b(M, N):
    import torch
    h = _handle()
    torch.manual_seed(M)
    A = torch.randn(M, N, dtype=torch.float64, device="cuda")
    W = torch.tril(torch.randn(N, N, dtype=torch.float64, device="cuda"))
    Wd = W + torch.triu(torch.full_like(W, float("nan")), diagonal=129)   # far above the diagonal is never read
    X = torch.empty(M, N, dtype=torch.float64, device="cuda")
    h._ck(h.lib.gprb200_gemm_dev(h.h, 3, M, N, N, 1.0, A.data_ptr(), N, Wd.data_ptr(), N, 0.0, X.data_ptr(), N, 0))
    ref = A @ W.T
    assert float((X - ref).abs().max()) <= 1e-13 * float(ref.abs().max())
    h.close()
