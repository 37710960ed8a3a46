"""ctypes access to oracle/_ref/libgpr_ref.so: the reference's OWN code (lib/GaussianProcess.cpp,
include/Kernel.h, KernelFactory.h, Likelihood.h, LAPACKUtils.h compiled unmodified from /root/reference
against third_party/eigen_lite and scipy's OpenBLAS; recipe: oracle/Makefile).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU baseline legs.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(_HERE, "_ref", "libgpr_ref.so")
_lib = None


def available():
    return os.path.exists(REF_LIB)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(REF_LIB)
        _lib.ref_last_error.restype = C.c_char_p
        _lib.ref_gp_new.restype = C.c_void_p
        _lib.ref_gp_new.argtypes = [C.c_char_p, C.c_double]
    return _lib


class RefError(RuntimeError):
    pass


def _ck(rc):
    if rc != 0:
        raise RefError(lib().ref_last_error().decode())


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def set_threads(n):
    lib().ref_set_blas_threads(int(n))
    os.environ["OMP_NUM_THREADS"] = str(int(n))


def kernel_eval(ks, x, y):
    x, y = _f(x).ravel(), _f(y).ravel()
    out = C.c_double()
    _ck(lib().ref_kernel_eval(ks.encode(), _p(x), _p(y), x.size, C.byref(out)))
    return out.value


def kernel_deriv(ks, x, y):
    x, y = _f(x).ravel(), _f(y).ravel()
    out = np.zeros(64)
    n = C.c_int()
    _ck(lib().ref_kernel_deriv(ks.encode(), _p(x), _p(y), x.size, _p(out), C.byref(n)))
    return out[:n.value].copy()


def kernel_tostring(ks):
    buf = C.create_string_buffer(4096)
    params = np.zeros(64)
    n = C.c_int()
    _ck(lib().ref_kernel_tostring(ks.encode(), buf, 4096, _p(params), C.byref(n)))
    return buf.value.decode(), params[:n.value].copy()


def lu_invert(A):
    A = _f(A).copy()
    _ck(lib().ref_lu_invert(_p(A), A.shape[0]))
    return A


def chol_invert(A):
    A = _f(A).copy()
    _ck(lib().ref_chol_invert(_p(A), A.shape[0]))
    return A


class RefGP:
    """gpr::GaussianProcess<double> of the reference, driven through its public API."""

    def __init__(self, kernel_string, sigma=0.0):
        self.h = C.c_void_p(lib().ref_gp_new(kernel_string.encode(), float(sigma)))
        if not self.h:
            raise RefError(lib().ref_last_error().decode())
        self.N = self.d = self.M = 0

    def __del__(self):
        try:
            if self.h:
                lib().ref_gp_free(self.h)
                self.h = None
        except Exception:
            pass

    def add_samples(self, X, Y):
        X, Y = _f(X), _f(Y)
        self.N += X.shape[0]
        self.d, self.M = X.shape[1], Y.shape[1]
        _ck(lib().ref_gp_add_samples(self.h, _p(X), C.c_long(X.shape[0]), self.d, _p(Y), self.M))
        return self

    def set_sigma(self, s):
        _ck(lib().ref_gp_set_sigma(self.h, C.c_double(s)))

    def set_kernel(self, ks):
        _ck(lib().ref_gp_set_kernel(self.h, ks.encode()))

    def set_inversion(self, m):
        _ck(lib().ref_gp_set_inversion(self.h, int(m)))

    def set_efficient_storage(self, s):
        _ck(lib().ref_gp_set_efficient_storage(self.h, int(s)))

    def initialize(self):
        _ck(lib().ref_gp_initialize(self.h))
        return self

    def dims(self):
        N, d, M, cs = C.c_long(), C.c_int(), C.c_int(), C.c_long()
        _ck(lib().ref_gp_dims(self.h, C.byref(N), C.byref(d), C.byref(M), C.byref(cs)))
        self.N, self.d, self.M = N.value, d.value, M.value
        return N.value, d.value, M.value, cs.value

    def predict(self, x):
        x = _f(x).ravel()
        out = np.empty(self.M)
        _ck(lib().ref_gp_predict(self.h, _p(x), x.size, _p(out)))
        return out

    def predict_derivative(self, x):
        x = _f(x).ravel()
        mean, D = np.empty(self.M), np.empty((self.d, self.M))
        _ck(lib().ref_gp_predict_derivative(self.h, _p(x), x.size, _p(mean), _p(D)))
        return mean, D

    def cov(self, x, y):
        x, y = _f(x).ravel(), _f(y).ravel()
        out = C.c_double()
        _ck(lib().ref_gp_cov(self.h, _p(x), _p(y), x.size, C.byref(out)))
        return out.value

    def ci(self, x):
        x = _f(x).ravel()
        out = C.c_double()
        _ck(lib().ref_gp_ci(self.h, _p(x), x.size, C.byref(out)))
        return out.value

    def kernel_matrix(self):
        K = np.empty((self.N, self.N))
        _ck(lib().ref_gp_kernel_matrix(self.h, _p(K)))
        return K

    def dkernel_matrix(self, P):
        D = np.empty((P * self.N, self.N))
        _ck(lib().ref_gp_dkernel_matrix(self.h, _p(D)))
        return D

    def kernel_vector(self, x):
        x = _f(x).ravel()
        out = np.empty(self.N)
        _ck(lib().ref_gp_kernel_vector(self.h, _p(x), x.size, _p(out)))
        return out

    def alpha(self):
        A = np.empty((self.N, self.M))
        _ck(lib().ref_gp_alpha(self.h, _p(A)))
        return A

    def core(self):
        cs = self.dims()[3]
        Cm = np.empty((cs, cs))
        if cs:
            _ck(lib().ref_gp_core(self.h, _p(Cm)))
        return Cm

    def save(self, prefix):
        _ck(lib().ref_gp_save(self.h, prefix.encode()))

    def load(self, prefix):
        _ck(lib().ref_gp_load(self.h, prefix.encode()))
        self.dims()

    def equal(self, other):
        return bool(lib().ref_gp_equal(self.h, other.h))

    def kernel_string(self):
        buf = C.create_string_buffer(4096)
        _ck(lib().ref_gp_kernel_string(self.h, buf, 4096))
        return buf.value.decode()

    def loglik_value(self):
        v = np.empty(self.M)
        _ck(lib().ref_loglik_value(self.h, _p(v)))
        return v

    def loglik_grad(self, P):
        g = np.empty(P)
        _ck(lib().ref_loglik_grad(self.h, _p(g)))
        return g

    def loglik_value_grad(self, P):
        v, g = np.empty(self.M), np.empty(P)
        _ck(lib().ref_loglik_value_grad(self.h, _p(v), _p(g)))
        return v, g

    def loglik_value_jac(self, P):
        v, J = np.empty(self.M), np.empty((self.M, P))
        _ck(lib().ref_loglik_value_jac(self.h, _p(v), _p(J)))
        return v, J

    def set_params(self, p):
        p = _f(p).ravel()
        _ck(lib().ref_gp_set_params(self.h, _p(p), p.size))
        return self

    def infer(self, which, step, iterations, P, mask=None):
        """GaussianProcessInference::Optimize (which=1) / Optimize2 (which=2) of the reference; returns GetParameters()."""
        out = np.empty(P)
        n = C.c_int()
        m = None if mask is None else np.ascontiguousarray(mask, dtype=np.int32)
        lib().ref_infer.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        _ck(lib().ref_infer(self.h, int(which), float(step), int(iterations), None if m is None else _p(m),
                            0 if m is None else len(m), _p(out), C.byref(n)))
        return out[:n.value]

    def lik_value(self):
        v = np.empty(self.M)
        _ck(lib().ref_lik_value(self.h, _p(v)))
        return v


class RefSparseGP:
    """The reference's SparseGaussianProcess<double> (include/SparseGaussianProcess.h) through oracle/_ref."""

    def __init__(self, kernel_string, sigma, jitter):
        L = lib()
        L.ref_sgp_new.restype = C.c_void_p
        L.ref_sgp_new.argtypes = [C.c_char_p, C.c_double, C.c_double]
        L.ref_sgp_free.argtypes = [C.c_void_p]
        L.ref_sgp_add.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_int, C.c_void_p, C.c_int, C.c_int]
        L.ref_sgp_predict.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.ref_sgp_cov.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        self.h = L.ref_sgp_new(kernel_string.encode(), float(sigma), float(jitter))
        if not self.h:
            raise RefError(L.ref_last_error().decode())
        self.M = None

    def __del__(self):
        try:
            if self.h:
                lib().ref_sgp_free(self.h)
                self.h = None
        except Exception:
            pass

    def add(self, X, Y, inducing=False):
        X, Y = _f(X), _f(Y)
        self.M = Y.shape[1]
        _ck(lib().ref_sgp_add(self.h, _p(X), X.shape[0], X.shape[1], _p(Y), Y.shape[1], int(inducing)))
        return self

    def predict(self, x):
        x = _f(x).ravel()
        out = np.empty(self.M)
        _ck(lib().ref_sgp_predict(self.h, _p(x), x.size, _p(out)))
        return out

    def cov(self, x, y):
        x, y = _f(x).ravel(), _f(y).ravel()
        out = C.c_double()
        _ck(lib().ref_sgp_cov(self.h, _p(x), _p(y), x.size, C.byref(out)))
        return out.value

    def set_params(self, params):
        p = _f(params).ravel()
        lib().ref_sgp_set_params.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        _ck(lib().ref_sgp_set_params(self.h, _p(p), p.size))

    def loglik(self, what, P):
        """SparseGaussianLogLikelihood (include/SparseLikelihood.h): what = 0 operator(), 1 GetParameterDerivatives,
        2 GetValueAndParameterDerivatives, 3 GetValueAndJacobian. Returns (value[M], grad[P], jac[M x P]) with None for
        what the call does not produce."""
        lib().ref_sgp_loglik.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        value, grad, jac = np.empty(self.M), np.empty(P), np.empty((self.M, P))
        _ck(lib().ref_sgp_loglik(self.h, int(what), _p(value), _p(grad), _p(jac)))
        return (value if what in (0, 2, 3) else None, grad if what in (1, 2) else None, jac if what == 3 else None)


def pinv(A):
    """gpr::pinv (include/Prior.h:38-57)."""
    A = _f(np.atleast_2d(A))
    out = np.empty((A.shape[1], A.shape[0]))
    _ck(lib().ref_pinv(_p(A), A.shape[0], A.shape[1], _p(out)))
    return out


class RefPCA:
    """The reference's PCA<double> (include/PCA.h) through oracle/_ref. X is D x n, one sample per column."""

    def __init__(self, X=None, n_features=0, prefix=None):
        L = lib()
        L.ref_pca_fit.restype = C.c_void_p
        L.ref_pca_fit.argtypes = [C.c_void_p, C.c_long, C.c_long, C.c_int]
        L.ref_pca_load.restype = C.c_void_p
        L.ref_pca_load.argtypes = [C.c_char_p, C.c_int]
        L.ref_pca_free.argtypes = [C.c_void_p]
        L.ref_pca_get.argtypes = [C.c_void_p] * 4
        L.ref_pca_write.argtypes = [C.c_void_p, C.c_char_p]
        L.ref_pca_project.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_long, C.c_int, C.c_void_p, C.c_void_p]
        L.ref_pca_reconstruct.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_long, C.c_void_p]
        self.n_features = int(n_features)
        if prefix is not None:
            self.p = L.ref_pca_load(str(prefix).encode(), self.n_features)
            if not self.p:
                raise RefError(L.ref_last_error().decode())
            self.D = self.n = None
        else:
            X = _f(X)
            self.D, self.n = X.shape
            self.p = L.ref_pca_fit(_p(X), self.D, self.n, self.n_features)
            if not self.p:
                raise RefError(L.ref_last_error().decode())

    def close(self):
        if self.p:
            lib().ref_pca_free(self.p)
            self.p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def get(self):
        mean, sigma, U = np.empty(self.D), np.empty(self.n), np.empty((self.D, self.n))
        _ck(lib().ref_pca_get(self.p, _p(mean), _p(sigma), _p(U)))
        return mean, sigma, U

    def write(self, prefix):
        _ck(lib().ref_pca_write(self.p, str(prefix).encode()))

    def project(self, X, fast=False):
        X = _f(X)
        D, ncols = X.shape
        out = np.empty((max(self.n or 0, self.n_features, 1) if not fast else self.n_features, ncols))
        if not fast and self.n is None:
            raise RefError("project(fast=False) needs the number of modes; use a fitted object")
        rows = C.c_long()
        _ck(lib().ref_pca_project(self.p, _p(X), D, ncols, int(bool(fast)), _p(out), C.byref(rows)))
        return out[:rows.value]

    def reconstruct(self, W, D=None):
        W = _f(W)
        nF, ncols = W.shape
        out = np.empty((D or self.D, ncols))
        _ck(lib().ref_pca_reconstruct(self.p, _p(W), nF, ncols, _p(out)))
        return out
