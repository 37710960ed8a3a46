"""CPU ORACLE (numpy/scipy) for the GaussianProcess Learn/Predict hot path of agiger/GPR.

TEST INFRASTRUCTURE ONLY. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this file; nothing under gpr_b200/ does.

It restates the reference's algorithm function by function (citations are file:line under
/root/reference) and calls the SAME LAPACK routines the reference calls for its default inversion
path (dgetrf + dgetri, include/LAPACKUtils.h:38-56), through scipy.linalg.lapack.

Parity status: pinned against (a) oracle/_ref — the reference's own Kernel.h / KernelFactory.h /
lib/GaussianProcess.cpp / Likelihood.h compiled unmodified against an Eigen stand-in
(tests/test_oracle_vs_ref.py, run in the build container where /root/reference exists), (b) the
committed golden vectors generated from oracle/_ref (tests/golden/), and (c) the reference's own
deterministic known-answer tests (tests/test_reference_known_answers.py).
Third-party arithmetic not in /root/reference: Eigen 3 (README.md:63 names 3.2.4, unpinned by the
build) for GEMM/GEMV/determinant; system LAPACK (unpinned) for dgetrf/dgetri/dpotrf/dpotri.
"""
import math

import numpy as np
from scipy.linalg import lapack

LDBL_MIN = np.finfo(np.longdouble).tiny
LDBL_MAX = np.finfo(np.longdouble).max


# ------------------------------------------------------------------------------------------------
# kernels (include/Kernel.h). A kernel is a nested tuple: (name, params...) or (name, k1, k2).
# ------------------------------------------------------------------------------------------------
def parse_kernel(s):
    """KernelFactory::GetKernel (include/KernelFactory.h:83-178) for ToString()-style strings."""
    s = s.strip()

    def rec(pos):
        lp = s.index("(", pos)
        name = s[pos:lp].strip()
        pos = lp + 1
        if name in ("SumKernel", "ProductKernel"):
            k1, pos = rec(pos)
            while s[pos] in ", ":
                pos += 1
            k2, pos = rec(pos)
            while s[pos] in ", ":
                pos += 1
            assert s[pos] == ")"
            return (name, k1, k2), pos + 1
        rp = s.index(")", pos)
        vals = [float(t) for t in s[pos:rp].split(",") if t.strip()]
        return (name,) + tuple(vals), rp + 1

    return rec(0)[0]


def num_params(k):
    if k[0] in ("SumKernel", "ProductKernel"):
        return num_params(k[1]) + num_params(k[2])
    return len(k) - 1


def get_params(k):
    if k[0] in ("SumKernel", "ProductKernel"):
        return get_params(k[1]) + get_params(k[2])
    return list(k[1:])


def set_params(k, p):
    p = list(p)
    if k[0] in ("SumKernel", "ProductKernel"):
        n1 = num_params(k[1])
        return (k[0], set_params(k[1], p[:n1]), set_params(k[2], p[n1:]))
    return (k[0],) + tuple(p)


def _pairwise_diff(A, B):
    return A[:, None, :] - B[None, :, :]  # (na, nb, d)


def kernel_matrix(k, A, B):
    """k(a_i, b_j) for all pairs, vectorised restatement of the scalar functors."""
    name = k[0]
    if name == "SumKernel":  # include/Kernel.h:165-167
        return kernel_matrix(k[1], A, B) + kernel_matrix(k[2], A, B)
    if name == "ProductKernel":  # :314-316
        return kernel_matrix(k[1], A, B) * kernel_matrix(k[2], A, B)
    D = _pairwise_diff(A, B)
    if name == "PeriodicKernel":  # :912-920, params [scale, b, sigma]
        scale, b, sigma = k[1:]
        s = np.sum(np.sin(b * D) ** 2, axis=2)
        return scale * scale * np.exp(-0.5 * s / (sigma * sigma))
    r = np.sqrt(np.sum(D * D, axis=2))  # (x-y).norm()
    rr = r * r
    if name == "GaussianKernel":  # :465-468, params [sigma, scale]
        sigma, scale = k[1:]
        return scale * scale * np.exp(-0.5 * rr / (sigma * sigma))
    if name == "GaussianExpKernel":  # :580-585
        sigma, scale = k[1:]
        es, eg = math.exp(scale), math.exp(sigma)
        return es * es * np.exp(-0.5 * rr / (eg * eg))
    if name == "WhiteKernel":  # :695-702
        (scale,) = k[1:]
        return np.where(r == 0, scale * scale, 0.0)
    if name == "RationalQuadraticKernel":  # :794-797, params [scale, sigma, alpha]
        scale, sigma, alpha = k[1:]
        return scale * scale * np.power(1 + 0.5 * rr / (sigma * sigma * alpha), -alpha)
    raise ValueError(name)


def kernel_derivative(k, A, B):
    """Kernel::GetDerivative for all pairs -> array (P, na, nb)."""
    name = k[0]
    if name == "SumKernel":  # :169-178
        return np.concatenate([kernel_derivative(k[1], A, B), kernel_derivative(k[2], A, B)], axis=0)
    if name == "ProductKernel":  # :318-327
        d1 = kernel_derivative(k[1], A, B) * kernel_matrix(k[2], A, B)[None]
        d2 = kernel_derivative(k[2], A, B) * kernel_matrix(k[1], A, B)[None]
        return np.concatenate([d1, d2], axis=0)
    D = _pairwise_diff(A, B)
    if name == "PeriodicKernel":  # :922-948
        scale, b, sigma = k[1:]
        f1 = np.sum(np.sin(b * D) ** 2, axis=2)
        f2 = np.sum(2 * D * np.cos(b * D) * np.sin(b * D), axis=2)
        e = np.exp(-0.5 * f1 / (sigma * sigma))
        return np.stack([2 * scale * e, -0.5 * scale * scale * e * f2 / (sigma * sigma),
                         scale * scale * e * f1 / (sigma ** 3)])
    r = np.sqrt(np.sum(D * D, axis=2))
    rr = r * r
    if name == "GaussianKernel":  # :471-479
        sigma, scale = k[1:]
        f = np.exp(-0.5 * rr / (sigma * sigma))
        return np.stack([scale * scale * rr / (sigma ** 3) * f, 2 * scale * f])
    if name == "GaussianExpKernel":  # :588-598
        sigma, scale = k[1:]
        f1, f2 = math.exp(-2 * sigma), math.exp(2 * sigma)
        return np.stack([rr * np.exp(-0.5 * f1 * ((4 * sigma - 4 * scale) * f2 + rr)),
                         2 * np.exp(0.5 * f1 * (4 * f2 * scale - rr))])
    if name == "WhiteKernel":  # :704-713
        (scale,) = k[1:]
        return np.where(r == 0, 2 * scale, 0.0)[None]
    if name == "RationalQuadraticKernel":  # :799-808
        scale, sigma, alpha = k[1:]
        f = 0.5 * rr / (sigma * sigma * alpha) + 1
        return np.stack([2 * scale * np.power(f, -alpha),
                         scale * scale * rr * np.power(f, -alpha - 1) / sigma ** 3,
                         scale * scale * ((rr / (2 * sigma * sigma * f * alpha)) - np.log(f)) * np.power(f, -alpha)])
    raise ValueError(name)


# ------------------------------------------------------------------------------------------------
# dense algebra (include/LAPACKUtils.h)
# ------------------------------------------------------------------------------------------------
def lu_invert(K):
    """lapack::lu_invert (include/LAPACKUtils.h:38-56, :85-97): dgetrf + dgetri on the row-major buffer
    (LAPACK sees K^T; the inverse of the transpose is the transpose of the inverse, so the row-major
    result is inv(K))."""
    a = np.array(K.T, dtype=np.float64, order="F")  # row-major K viewed column-major == K^T
    lu, piv, info = lapack.dgetrf(a, overwrite_a=True)
    if info != 0:
        raise np.linalg.LinAlgError("lu_invert: error in inversion. Matrix is singular or contains an illegal value.")
    inv, info = lapack.dgetri(lu, piv, overwrite_lu=True)
    if info != 0:
        raise np.linalg.LinAlgError("lu_invert: error in inversion. Matrix is singular or contains an illegal value.")
    return np.ascontiguousarray(inv.T)


def chol_invert(K):
    """lapack::chol_invert (include/LAPACKUtils.h:59-73, :100-111): dpotrf('L') + dpotri('L') on the
    row-major buffer, then mirror the (row-major) upper triangle."""
    a = np.array(K.T, dtype=np.float64, order="F")
    c, info = lapack.dpotrf(a, lower=1, overwrite_a=True)
    inv, info = lapack.dpotri(c, lower=1, overwrite_c=True)
    rm = np.ascontiguousarray(inv.T)  # back to row-major indexing: A[i*N + j]
    iu = np.triu_indices(rm.shape[0])
    rm.T[iu] = rm[iu]  # A[j*N+i] = A[i*N+j] for j >= i
    return rm


# ------------------------------------------------------------------------------------------------
# GaussianProcess (lib/GaussianProcess.cpp)
# ------------------------------------------------------------------------------------------------
class OracleGP:
    def __init__(self, kernel, X, Y, sigma, inversion="lu"):
        self.k = parse_kernel(kernel) if isinstance(kernel, str) else kernel
        self.X = np.asarray(X, dtype=np.float64)
        self.Y = np.asarray(Y, dtype=np.float64)
        self.sigma = float(sigma)
        self.inversion = inversion
        self.C = None
        self.alpha = None

    # ComputeKernelMatrixInternal (:384-402): upper triangle evaluated, mirrored
    def kernel_matrix(self):
        K = kernel_matrix(self.k, self.X, self.X)
        K = np.triu(K) + np.triu(K, 1).T
        if not np.all(np.isfinite(K)):
            raise FloatingPointError(
                "GaussianProcess::ComputeKernelMatrixInternal: kernel matrix contains entries which are not finite.")
        return K

    # ComputeCoreMatrix (:498-510) with AddNoiseToKernelMatrix (:375-381) and InvertKernelMatrix (:531-618)
    def core_matrix(self):
        K = self.kernel_matrix()
        K[np.diag_indices_from(K)] += self.sigma * self.sigma
        self.Kn = K
        return lu_invert(K) if self.inversion == "lu" else chol_invert(K)

    # ComputeRegressionVectors (:642-672)
    def learn(self):
        self.C = self.core_matrix()
        self.alpha = self.C @ self.Y  # :661
        return self.alpha

    def kernel_vector(self, x):  # :684-693
        return kernel_matrix(self.k, np.atleast_2d(x), self.X)[0]

    def predict(self, x):  # :55-61
        return self.kernel_vector(x) @ self.alpha

    def predict_batch(self, Xs):
        return kernel_matrix(self.k, np.atleast_2d(Xs), self.X) @ self.alpha

    def predict_derivative(self, x):  # :65-81, :697-706
        x = np.asarray(x, dtype=np.float64).reshape(-1)
        Kx = self.kernel_vector(x)
        Xd = x[None, :] - self.X
        D = np.stack([-Xd.T @ (Kx * self.alpha[:, i]) for i in range(self.alpha.shape[1])], axis=1)
        return Kx @ self.alpha, D

    def cov(self, x, y):  # operator() :85-99
        Kx, Ky = self.kernel_vector(x), self.kernel_vector(y)
        kxy = kernel_matrix(self.k, np.atleast_2d(x), np.atleast_2d(y))[0, 0]
        return kxy - Kx @ self.C @ Ky

    def credible_interval(self, x):  # :103-114
        return 2.0 * math.sqrt(max(0.0, self.cov(x, x)))

    def derivative_kernel_matrix(self):  # :472-495 -> (P*N, N)
        D = kernel_derivative(self.k, self.X, self.X)
        D = np.triu(D) + np.transpose(np.triu(D, 1), (0, 2, 1))
        return D.reshape(-1, self.X.shape[0])

    # ---- GaussianLogLikelihood (include/Likelihood.h:166-344) ------------------------------------
    def _logdet(self):
        # ComputeCoreMatrixWithDeterminant (:513-528) takes the long-double determinant by LU; the sign
        # and log-magnitude from dgetrf pivots are the same quantity without the overflow.
        lu, piv, info = lapack.dgetrf(np.array(self.Kn.T, order="F"))
        d = np.diag(lu)
        sign = np.prod(np.sign(d)) * (-1.0) ** np.sum(piv != np.arange(len(piv)))
        return sign, np.sum(np.log(np.abs(d).astype(np.longdouble)))

    def _cp(self, clamp=True):
        sign, ld = self._logdet()
        if not clamp:
            return -0.5 * ld
        with np.errstate(over="ignore", under="ignore"):
            det = sign * np.exp(ld)          # long double
            det = np.longdouble(np.float64(det))  # GetCoreMatrix returns TScalarType (:77-79)
        if det <= LDBL_MIN:  # :180-188
            return -0.5 * np.log(LDBL_MIN)
        if det > LDBL_MAX:
            return -0.5 * np.log(LDBL_MAX)
        return -0.5 * np.log(det)

    def loglik_value(self, clamp=True):
        C = self.core_matrix()
        N = C.shape[0]
        df = -0.5 * np.einsum("nk,nm,mk->k", self.Y, C, self.Y)  # diagonal of -0.5 Y^T C Y (:175)
        cp = self._cp(clamp)
        ct = -N / 2.0 * math.log(2 * math.pi)  # :192
        value = (df.astype(np.longdouble) + (cp + np.longdouble(ct))).astype(np.float64)
        if np.isinf(value.sum()):
            raise FloatingPointError(
                "GaussianLogLikelihood::GetValueAndParameterDerivatives: likelihood is infinite.")
        self.C = C
        return value

    def loglik_grad(self):  # :204-229 / :268-282: C subtracted ONCE for all output columns
        C = self.core_matrix()
        alpha = C @ self.Y
        D = self.derivative_kernel_matrix()
        N = C.shape[0]
        P = D.shape[0] // N
        A = alpha @ alpha.T - C
        return np.array([0.5 * np.trace(A @ D[p * N:(p + 1) * N]) for p in range(P)])

    def loglik_jacobian(self):  # :287-344
        C = self.core_matrix()
        D = self.derivative_kernel_matrix()
        N = C.shape[0]
        P = D.shape[0] // N
        J = np.zeros((self.Y.shape[1], P))
        for i in range(self.Y.shape[1]):
            a = C @ self.Y[:, i:i + 1]
            A = a @ a.T - C
            for p in range(P):
                J[i, p] = 0.5 * np.trace(A @ D[p * N:(p + 1) * N])
        return J


# ---------------------------------------------------------------------------------------------------
# PCA: restatement of include/PCA.h. X is D x n with one sample per column.
# ---------------------------------------------------------------------------------------------------
class PCAOracle:
    """include/PCA.h:33-49 (fit), :92-108 (project), :110-115 (reconstruct), :117-124 (explained variance).
    numpy's svd is LAPACK gesdd, the same divide-and-conquer family as Eigen::BDCSVD; singular vectors are
    defined up to sign, so comparisons are made on sign-invariant quantities."""

    def __init__(self, X=None, n_features=0, mean=None, sigma=None, U=None):
        self.n_features = int(n_features)
        if X is not None:
            X = np.asarray(X, dtype=np.float64)
            self.mean = X.mean(axis=1)                                   # PCA.h:37  X.rowwise().mean()
            Xc = X - self.mean[:, None]                                  # :38
            U, s, _ = np.linalg.svd(Xc, full_matrices=False)             # :41  BDCSVD, ComputeThinU
            self.sigma = s / np.sqrt(X.shape[1])                         # :44
            self.U = U                                                   # :45
        else:
            self.mean, self.sigma, self.U = (np.asarray(a, dtype=np.float64) for a in (mean, sigma, U))
            self.mean, self.sigma = self.mean.ravel(), self.sigma.ravel()
        with np.errstate(divide="ignore"):
            self.basis = self.U / self.sigma[None, :]                    # :46  U * sigma.asDiagonal().inverse()
        self.basis_inverse = (self.U * self.sigma[None, :])[:, :self.n_features]   # :47-48

    def project(self, X, fast=False):
        Xc = np.asarray(X, dtype=np.float64) - self.mean[:, None]        # :93
        if fast:
            return self.basis[:, :self.n_features].T @ Xc                # :68, :102-106
        return self.basis.T @ Xc                                         # :94 (`nFeatures << rows` keeps every row)

    def reconstruct(self, W):
        return self.basis_inverse @ np.asarray(W, dtype=np.float64) + self.mean[:, None]   # :110-115

    def explained_variance(self):
        """The intent of :117-124 (cumulative sigma, normalised). The reference indexes the VECTOR m_sigma as
        (i, i), which reads out of bounds for i >= 1 in a column vector; not reproduced."""
        c = np.cumsum(self.sigma)
        return c / c[-1]


# ---------------------------------------------------------------------------------------------------
# GaussianProcessInference: restatement of include/GaussianProcessInference.h
# ---------------------------------------------------------------------------------------------------
def pinv_abs(m, epsilon=np.finfo(np.float64).eps):
    """include/Prior.h:38-57: SVD pseudo-inverse, singular values <= epsilon (ABSOLUTE) are dropped."""
    m = np.atleast_2d(np.asarray(m, dtype=np.float64))
    u, s, vt = np.linalg.svd(m, full_matrices=True)
    k = len(s)
    inv = np.array([0.0 if v <= epsilon else 1.0 / v for v in s])
    return (vt[:k].T * inv[None, :]) @ u[:, :k].T


def _mm(A, B):
    """Dense product with the plain i-k-j loop order of an unblocked row-major GEMM (what the reference's Eigen
    expression evaluates to for these tiny operands). The order matters: pinv(g g^T) g cancels catastrophically
    (entries of 1e14 producing an O(1) result), so the optimiser's step IS the rounding of this sum."""
    A, B = np.atleast_2d(np.asarray(A, dtype=np.float64)), np.asarray(B, dtype=np.float64)
    vec = B.ndim == 1
    B2 = B[:, None] if vec else B
    out = np.zeros((A.shape[0], B2.shape[1]))
    for i in range(A.shape[0]):
        for k in range(A.shape[1]):
            a = A[i, k]
            for j in range(B2.shape[1]):
                out[i, j] += a * B2[k, j]
    return out[:, 0] if vec else out


def _signed_log(v):
    return np.log(1 + v) if v >= 0 else -np.log(1 + abs(v))


def infer_optimize(value_grad, params, step, iterations, pinv=None):
    """GaussianProcessInference::Optimize (:84-156). value_grad(params) -> (value (M,), grad (P,)).
    pinv: the pseudo-inverse to use (default pinv_abs; tests inject the reference's own gpr::pinv, because the
    pseudo-inverse of the rank-1 matrix g g^T with an ABSOLUTE cut-off is dominated by the SVD's rounding noise and
    differs between SVD implementations)."""
    pinv = pinv or pinv_abs
    params = [float(p) for p in params]
    for _ in range(iterations):
        value, grad = value_grad(params)
        sign = -1.0 if value[0] > 0 else 1.0                                   # :95
        update = _mm(pinv(np.outer(grad, grad)), grad)                         # :97
        for p in range(len(params)):
            if update[p] == 0:                                                 # :118-126
                params[p] += step ** 3 * _signed_log(grad[p]) * sign
            else:                                                              # :127-136
                params[p] -= step * _signed_log(update[p] * value[0]) * sign
    return params


def infer_optimize2(value_jac, params, step, iterations, mask=None, pinv=None):
    """GaussianProcessInference::Optimize2 (:158-229). value_jac(params) -> (value (M,), J (M, P))."""
    pinv = pinv or pinv_abs
    params = [float(p) for p in params]
    mask = [True] * len(params) if mask is None else list(mask)
    previous = None
    for it in range(iterations):
        value, J = value_jac(params)
        value = np.array(value, dtype=np.float64)
        if it > 0 and np.linalg.norm(previous - value) == 0:                   # :169-174
            break
        value = np.where(value > 0, -value, value)                             # :176-179
        update = _mm(_mm(pinv(_mm(J.T, J)), J.T), value)                       # :182
        for p in range(len(params)):
            if mask[p]:                                                        # :195-203
                params[p] -= step * _signed_log(update[p])
        previous = value
    return params


# ---------------------------------------------------------------------------------------------------
# SparseGaussianProcess, regression part: restatement of include/SparseGaussianProcess.h
# ---------------------------------------------------------------------------------------------------
class SparseOracle:
    """PreComputeRegression (:274-312), Predict (:84-90), operator() (:92-106) with the reference's explicit
    inverses (LU, include/LAPACKUtils.h:85-97) and its order of products."""

    def __init__(self, kernel, X, Y, Xm, sigma, jitter):
        self.k = parse_kernel(kernel) if isinstance(kernel, str) else kernel
        self.X, self.Y, self.Xm = (np.asarray(a, dtype=np.float64) for a in (X, Y, Xm))
        K = kernel_matrix(self.k, self.Xm, self.Xm) + jitter * np.eye(len(self.Xm))    # :175-180
        self.Kinv = lu_invert(K)                                                        # :283
        Knm = kernel_matrix(self.k, self.X, self.Xm)                                    # :285-286 (n x m)
        is2 = 1.0 / (sigma * sigma)
        S = K + is2 * Knm.T @ Knm                                                       # :292
        self.Sigma = lu_invert(S)                                                       # :293
        self.alpha = self.Kinv @ (is2 * K @ self.Sigma @ Knm.T @ self.Y)                # :296
        self.R = self.Kinv @ (K @ self.Sigma @ K) @ self.Kinv                           # :299

    def kx(self, x):
        return kernel_matrix(self.k, np.atleast_2d(x), self.Xm)[0]

    def predict(self, x):
        return self.kx(x) @ self.alpha

    def cov(self, x, y):
        kx, ky = self.kx(x), self.kx(y)
        kxy = kernel_matrix(self.k, np.atleast_2d(x), np.atleast_2d(y))[0, 0]
        return kxy - kx @ self.Kinv @ ky + kx @ self.R @ ky


# ---------------------------------------------------------------------------------------------------
# SparseGaussianLogLikelihood: restatement of include/SparseLikelihood.h:111-551 with the
# reference's explicit matrices (n x n C_inv, the stacked n x n grad(C) blocks), so small cases only.
# ---------------------------------------------------------------------------------------------------
def _ld_logabsdet(A):
    """long-double determinant by LU (Eigen's PartialPivLU on cast<long double>()) as (sign, log|det|)."""
    lu, piv, info = lapack.dgetrf(np.array(A.T, order="F"))
    d = np.diag(lu)
    sign = np.prod(np.sign(d)) * (-1.0) ** np.sum(piv != np.arange(len(piv)))
    return sign, np.sum(np.log(np.abs(d).astype(np.longdouble)))


class SparseLikOracle:
    def __init__(self, kernel, X, Y, Xm, sigma, jitter):
        self.k = parse_kernel(kernel) if isinstance(kernel, str) else kernel
        self.X, self.Y, self.Xm = (np.asarray(a, dtype=np.float64) for a in (X, Y, Xm))
        self.sigma, self.jitter = float(sigma), float(jitter)

    def _core(self):
        """ComputeCoreMatrices (include/SparseGaussianProcess.h:323-350) + EfficientInversion (:132-138)."""
        n, m = len(self.X), len(self.Xm)
        K = kernel_matrix(self.k, self.Xm, self.Xm) + self.jitter * np.eye(m)
        K_inv = lu_invert(K)
        Knm = kernel_matrix(self.k, self.X, self.Xm)
        s2 = self.sigma * self.sigma
        A_inv = np.full(n, 1.0 / s2)
        Cm = lu_invert(K + (Knm.T * A_inv) @ Knm)                      # inv(B_inv + X' inv(A) X), B_inv = K
        C_inv = np.diag(A_inv) - (A_inv[:, None] * Knm) @ Cm @ (Knm.T * A_inv)
        return K, K_inv, Knm, s2, C_inv

    def _cp(self, K, K_inv, Knm, s2, nan_to_min):
        """EfficientDeterminant (:142-150) and the clamp (:177-185 for operator(), :351-359 otherwise)."""
        n = len(self.X)
        with np.errstate(over="ignore", under="ignore", invalid="ignore"):
            sb, lb = _ld_logabsdet(K_inv)
            det_B = sb * np.exp(lb)
            if np.isinf(det_B):
                det_B = LDBL_MAX
            ss, ls = _ld_logabsdet(K + Knm.T @ Knm / s2)
            prod = np.exp(np.longdouble(n) * np.log(np.longdouble(s2)))
            det = det_B * prod * (ss * np.exp(ls))
        if det <= LDBL_MIN or (nan_to_min and np.isnan(det)):
            return -0.5 * np.log(LDBL_MIN)
        if det > LDBL_MAX:
            return -0.5 * np.log(LDBL_MAX)
        return -0.5 * np.log(det)

    def true_logdet(self):
        """log det(sigma^2 I + Knm K^-1 Kmn) without the long-double range (what LOGLIK_TRUE uses)."""
        K, K_inv, Knm, s2, _ = self._core()
        return float(-_ld_logabsdet(K)[1] + len(self.X) * np.log(np.longdouble(s2)) + _ld_logabsdet(K + Knm.T @ Knm / s2)[1])

    def value(self, clamp=True, nan_to_min=False):
        K, K_inv, Knm, s2, C_inv = self._core()
        n = len(self.X)
        df = -0.5 * np.einsum("nk,nm,mk->k", self.Y, C_inv, self.Y)      # diagonal of -0.5 Y' C_inv Y (:170)
        cp = self._cp(K, K_inv, Knm, s2, nan_to_min) if clamp else np.longdouble(-0.5 * self.true_logdet())
        ct = -(n / 2.0) * math.log(2 * math.pi)                           # :198
        v = (df.astype(np.longdouble) + (cp + np.longdouble(ct))).astype(np.float64)
        if np.isnan(v.sum()):
            raise FloatingPointError("SparseLikelihood::GetValueAndParameterDerivative: likelihood value is not a number.")
        return v

    def _grad_blocks(self):
        K, K_inv, Knm, s2, C_inv = self._core()
        n, m = Knm.shape
        Kmm_d = kernel_derivative(self.k, self.Xm, self.Xm)               # ComputeDerivativeKernelMatrix (:389-397)
        Knm_d = kernel_derivative(self.k, self.X, self.Xm)                # ComputeDerivativeKernelVectorMatrix (:237-264)
        A = [Knm_d[p] @ K_inv @ Knm.T - Knm @ K_inv @ Kmm_d[p] @ K_inv @ Knm.T + Knm @ K_inv @ Knm_d[p].T
             for p in range(len(Kmm_d))]                                  # :255-259
        return C_inv, A

    def grad(self):
        """GetParameterDerivatives (:219-285); the reference throws unless M == 1 (:264)."""
        if self.Y.shape[1] != 1:
            raise ValueError("SparseLikelihood::GetParameterDerivatives: dimension missmatch in calculating derivative of data fit term")
        C_inv, A = self._grad_blocks()
        y = self.Y[:, 0]
        dt = np.array([0.5 * y @ C_inv @ Ap @ C_inv @ y for Ap in A])     # :261-266
        ct = np.array([-0.5 * np.trace(C_inv @ Ap) for Ap in A])          # :270-274
        return dt + ct

    def jacobian(self):
        """GetValueAndJacobian (:411-535): J(i, p) = 0.5 y_i' C_inv A_p C_inv y_i - 0.5 tr(C_inv A_p). The reference's own
        last line `df_jac.colwise() + cp_grad` (:530) adds a P-vector to the columns of an M x P matrix and its value
        line (:440) assigns an M x M product to a vector, so the call is only defined for M == P == 1; the evident
        intent is implemented: row i equals GetParameterDerivatives of the single-output problem on column i."""
        C_inv, A = self._grad_blocks()
        M, P = self.Y.shape[1], len(A)
        df = np.array([[0.5 * self.Y[:, i] @ C_inv @ A[p] @ C_inv @ self.Y[:, i] for p in range(P)] for i in range(M)])
        cp = np.array([-0.5 * np.trace(C_inv @ Ap) for Ap in A])
        return df + cp[None, :]
