"""-m gpu: large label / regression matrices through the pinned staging ring (csrc/hostio.cu): bit-exact round trips
against the plain cudaMemcpy path, ragged shapes that split pieces by rows and by columns, and the row-pointer form
(the reference's std::vector<VectorType> storage, lib/GaussianProcess.cpp:48-49)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _learned(X, Y, rows=False, stream=1):
    import gpr_b200
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.GaussianKernel(1.5, 1.0).lower()
    h.set_kernel(ops, params)
    h.set_option("stream_io", stream)
    if rows:
        h.set_train_rows([x for x in X], [y.copy() for y in Y], 0.5)     # every label vector its own allocation
    else:
        h.set_train(X, Y, 0.5)
    h.learn(0)
    A = h.alpha()
    h.close()
    return A


@pytest.mark.parametrize("N,M", [(64, 200003), (3, 9000001), (700, 16411)])
def test_streamed_transfers_are_bit_identical_to_the_plain_path(N, M):
    rng = np.random.default_rng(N + M)
    X = rng.uniform(-1, 1, (N, 2))
    Y = rng.standard_normal((N, M))
    plain = _learned(X, Y, stream=0)
    assert np.array_equal(_learned(X, Y, stream=1), plain)            # contiguous Y >= 64 MB: streamed both ways
    assert np.array_equal(_learned(X, Y, rows=True, stream=1), plain)  # one pointer per label vector


def test_row_pointer_form_on_a_small_problem_and_errors():
    import gpr_b200
    rng = np.random.default_rng(5)
    X = rng.uniform(-1, 1, (40, 3))
    Y = rng.standard_normal((40, 5))
    assert np.array_equal(_learned(X, Y, rows=True), _learned(X, Y, stream=0))
    h = gpr_b200.Handle(0)
    h.set_kernel(*gpr_b200.GaussianKernel(1.0, 1.0).lower())
    h.set_train_rows([x for x in X], None, 0.5)                        # samples only (no labels)
    assert h.kernel_vector(X[:2]).shape == (2, 40)
    h.close()


def test_pipelined_learn_rows_equals_the_three_step_path():
    """gprb200_learn_rows (upload / solve / download overlapped over column blocks) against set_train + learn +
    get_alpha: three blocks of 32768 columns at N = 2048, the last one ragged; Predict afterwards uses the resident alpha."""
    import gpr_b200
    rng = np.random.default_rng(77)
    N, M = 2048, 70001
    X = rng.uniform(-1, 1, (N, 3))
    Y = rng.standard_normal((N, M))
    ref = _learned(X, Y, stream=0)
    h = gpr_b200.Handle(0)
    h.set_kernel(*gpr_b200.GaussianKernel(1.5, 1.0).lower())
    A = np.empty((N, M))
    h.learn_rows([x for x in X], [y.copy() for y in Y], 0.5, A)
    # same factor, same algorithm; a narrow last block takes the right-looking instead of the left-looking triangular
    # solve (solve.cu), so the sums are ordered differently: agreement to rounding, not bit for bit
    scale = np.abs(ref).max()
    assert np.abs(A - ref).max() <= 1e-12 * scale
    assert np.array_equal(A[:, :65536], ref[:, :65536])         # full-width blocks take the same path: identical
    assert np.array_equal(h.alpha(), A)                         # the device copy is complete as well
    mean = h.predict(X[:3])
    assert np.abs(mean - (Y[:3] - 0.25 * ref[:3])).max() <= 1e-9 * np.abs(Y).max()
    h.learn_rows([x for x in X], [y.copy() for y in Y], 0.5, None)   # no host copy requested
    assert np.array_equal(h.alpha(), A)
    h.close()
