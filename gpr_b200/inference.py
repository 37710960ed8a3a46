"""GaussianProcessInference of the reference (include/GaussianProcessInference.h:40-243) over the device-resident
likelihood: the samples, the factor and the inverse stay on the GPU across iterations, P kernel parameters go in and
M values + P (or M x P) derivatives come out per iteration. Same constructor arguments,
method names and update rules as the reference; the C++ twin is gpr_b200/cxx/include/GaussianProcessInference.h."""
import numpy as np


def _jacobi_svd(A):
    """One-sided Jacobi (Hestenes) SVD of a tiny matrix, A V = U S: the same algorithm and sweep order as the
    JacobiSVD the C++ host API is compiled against (third_party/eigen_lite), so both mirrors round the (numerically
    zero) trailing singular values of g g^T the same way. Returns U (m x n), S (n, descending), V (n x n)."""
    W = np.array(A, dtype=np.float64)
    m, n = W.shape
    V = np.eye(n)
    eps = np.finfo(np.float64).eps
    for _ in range(60):
        off = 0.0
        for p in range(n - 1):
            for q in range(p + 1, n):
                a = b = c = 0.0
                for i in range(m):
                    a += W[i, p] * W[i, p]
                    b += W[i, q] * W[i, q]
                    c += W[i, p] * W[i, q]
                if abs(c) <= eps * np.sqrt(a * b):
                    continue
                off = max(off, abs(c) / np.sqrt(a * b))
                zeta = (b - a) / (2 * c)
                t = (1.0 if zeta >= 0 else -1.0) / (abs(zeta) + np.sqrt(1 + zeta * zeta))
                cs = 1 / np.sqrt(1 + t * t)
                sn = cs * t
                for M in (W, V):
                    for i in range(M.shape[0]):
                        mp, mq = M[i, p], M[i, q]
                        M[i, p] = cs * mp - sn * mq
                        M[i, q] = sn * mp + cs * mq
        if off == 0:
            break
    norms = np.zeros(n)
    for j in range(n):
        sj = 0.0
        for i in range(m):
            sj += W[i, j] * W[i, j]
        norms[j] = np.sqrt(sj)
    order = sorted(range(n), key=lambda j: -norms[j])     # stable, descending
    S = norms[order]
    U = np.zeros((m, n))
    for k, j in enumerate(order):
        if S[k] > 0:
            U[:, k] = W[:, j] / S[k]
    return U, S, V[:, order]


def pinv(m, epsilon=np.finfo(np.float64).eps):
    """SVD pseudo-inverse with an ABSOLUTE cut-off on the singular values (include/Prior.h:38-57)."""
    m = np.atleast_2d(np.asarray(m, dtype=np.float64))
    if m.shape[0] < m.shape[1]:
        return pinv(m.T, epsilon).T
    U, S, V = _jacobi_svd(m)
    inv = np.array([0.0 if v <= epsilon else 1.0 / v for v in S])
    return _mm(V * inv[None, :], U.T)


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


def _slog(v):
    return np.log1p(v) if v >= 0 else -np.log1p(abs(v))


class GaussianProcessInference:
    def __init__(self, lh, gp, stepwidth, iterations, pinv=None):
        self._pinv = pinv or globals()["pinv"]   # additive: the pseudo-inverse is a customisation point
        self.m_Likelihood = lh
        self.m_GaussianProcess = gp
        self.m_StepWidth = float(stepwidth)
        self.m_StepWidth3 = float(stepwidth) ** 3
        self.m_NumberOfIterations = int(iterations)
        self.m_Parameters = list(gp.GetKernel().GetParameters())
        self.m_ParametersToOptimize = [True] * len(self.m_Parameters)
        self.trace = []   # additive: (parameters, likelihood) at the start of every iteration of the last call

    def GetParameters(self):
        return list(self.m_Parameters)

    def SetParametersToOptimize(self, v):
        for i in range(min(len(v), len(self.m_ParametersToOptimize))):
            self.m_ParametersToOptimize[i] = bool(v[i])

    def Optimize(self, output=False, exp_output=False):      # :84-156
        self.trace = []
        for _ in range(self.m_NumberOfIterations):
            try:
                self.m_GaussianProcess.GetKernel().SetParameters(self.m_Parameters)
                value, grad = self.m_Likelihood.GetValueAndParameterDerivatives(self.m_GaussianProcess)
            except Exception as e:                            # the reference prints "[failed] ..." and returns
                if output:
                    print("[failed]", e)
                return
            value, grad = np.asarray(value, dtype=np.float64), np.asarray(grad, dtype=np.float64)
            self.trace.append((list(self.m_Parameters), value.copy()))
            sign = -1.0 if value[0] > 0 else 1.0
            update = _mm(self._pinv(np.outer(grad, grad)), grad)
            for p in range(len(self.m_Parameters)):
                if update[p] == 0:                            # damped gradient step
                    self.m_Parameters[p] += self.m_StepWidth3 * _slog(grad[p]) * sign
                else:                                         # damped Gauss-Newton step
                    self.m_Parameters[p] -= self.m_StepWidth * _slog(update[p] * value[0]) * sign
            if output:
                shown = np.exp(self.m_Parameters) if exp_output else np.asarray(self.m_Parameters)
                print("Likelihood", value, "gradient", grad, "new parameters", shown)

    def Optimize2(self, output=False, exp_output=False):     # :158-229
        self.trace = []
        previous = None
        for it in range(self.m_NumberOfIterations):
            try:
                self.m_GaussianProcess.GetKernel().SetParameters(self.m_Parameters)
                value, J = self.m_Likelihood.GetValueAndJacobian(self.m_GaussianProcess)
            except Exception as e:
                if output:
                    print("[failed]", e)
                return
            value, J = np.array(value, dtype=np.float64), np.atleast_2d(np.asarray(J, dtype=np.float64))
            self.trace.append((list(self.m_Parameters), value.copy()))
            if it > 0 and np.linalg.norm(previous - value) == 0:
                break
            value = np.where(value > 0, -value, value)
            update = _mm(_mm(self._pinv(_mm(J.T, J)), J.T), value)
            for p in range(len(self.m_Parameters)):
                if self.m_ParametersToOptimize[p]:
                    self.m_Parameters[p] -= self.m_StepWidth * _slog(update[p])
            if output:
                shown = np.exp(self.m_Parameters) if exp_output else np.asarray(self.m_Parameters)
                print("Likelihood", value, "update", update, "new parameters", shown)
            previous = value
