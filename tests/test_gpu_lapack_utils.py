"""-m gpu: the public dense-inversion interface lapack::lu_invert / lapack::chol_invert
(reference include/LAPACKUtils.h:76-111) on the device (gprb200_lu_inverse, gprb200_spd_inverse) against the reference's
own wrappers compiled into oracle/_ref (dgetrf + dgetri, dpotrf + dpotri). Thresholds: the reference's own Test5
(tests/InversionMethodsTest.cpp:147-175) asks 1e-10 (LU) and 1e-4 (Cholesky) in Frobenius norm on random 200 x 200."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ref():
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref/libgpr_ref.so not built")
    return ref


@pytest.fixture(scope="module")
def h():
    import gpr_b200
    hd = gpr_b200.Handle(0)
    yield hd
    hd.close()


@pytest.mark.parametrize("n", [1, 2, 7, 33, 200, 257, 1000])
def test_lu_inverse_matches_reference_lu_invert(h, n):
    ref = _ref()
    rng = np.random.default_rng(n)
    M = rng.uniform(-1, 1, (n, n))                     # MatrixType::Random, as the reference's Test5
    want = ref.lu_invert(M)
    got = h.lu_inverse(M)
    assert np.linalg.norm(got - want) <= 1e-10 * max(1.0, np.linalg.norm(want))
    assert np.abs(got @ M - np.eye(n)).max() <= 1e-9 * max(1.0, np.abs(want).max())


def test_lu_inverse_pivots_rows(h):
    # zero diagonal: every step needs a row exchange; inverse of a permutation-like matrix is exact
    n = 64
    P = np.roll(np.eye(n), 3, axis=0) * np.arange(1, n + 1)[None, :]
    got = h.lu_inverse(P)
    assert np.abs(got @ P - np.eye(n)).max() <= 1e-14


def test_lu_inverse_reports_singular(h):
    import gpr_b200
    A = np.ones((5, 5))
    with pytest.raises(gpr_b200.GprError) as e:
        h.lu_inverse(A)
    assert e.value.code > 0 and "singular" in str(e.value)


@pytest.mark.parametrize("n", [200, 513])
def test_spd_inverse_matches_reference_chol_invert(h, n):
    ref = _ref()
    rng = np.random.default_rng(n + 1)
    M = rng.uniform(-1, 1, (n, n))
    MM = M @ M.T
    want = ref.chol_invert(MM)
    got, _ = h.spd_inverse(MM)
    assert np.linalg.norm(got - want) <= 1e-4 * max(1.0, np.linalg.norm(want)) * 1e-3   # measured ~1e-9 relative
