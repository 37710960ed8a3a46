"""ctypes binding of the C-ABI in include/gprb200.h (libgprb200.so, sm_100a CUDA).

There is deliberately NO fallback: if the shared library is missing or no B200 is present the
import of the library / creation of a handle raises.  Nothing under oracle/ is imported here.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GPRB200_LIB") or os.path.join(_HERE, "libgprb200.so")   # override: A/B builds (tools/)

# every symbol include/gprb200.h declares (tests/test_abi.py checks header <-> library <-> this table)
_PROTOS = {
    "gprb200_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int]),
    "gprb200_destroy": (C.c_int, [C.c_void_p]),
    "gprb200_last_error": (C.c_char_p, [C.c_void_p]),
    "gprb200_version": (C.c_char_p, []),
    "gprb200_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "gprb200_synchronize": (C.c_int, [C.c_void_p]),
    "gprb200_launch_count": (C.c_int64, [C.c_void_p]),
    "gprb200_phase_ms": (C.c_int, [C.c_void_p, C.c_void_p]),
    "gprb200_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "gprb200_get_stat": (C.c_int, [C.c_void_p, C.c_char_p, C.POINTER(C.c_double)]),
    "gprb200_set_kernel": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int]),
    "gprb200_set_kernel_params": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "gprb200_set_train": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int64, C.c_double]),
    "gprb200_set_train_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int64, C.c_double]),
    "gprb200_set_sigma": (C.c_int, [C.c_void_p, C.c_double]),
    "gprb200_set_state": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_int64,
                                    C.c_void_p, C.c_double]),
    "gprb200_gram": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "gprb200_gram_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int]),
    "gprb200_learn": (C.c_int, [C.c_void_p, C.c_uint]),
    "gprb200_factorize": (C.c_int, [C.c_void_p]),
    "gprb200_solve_alpha": (C.c_int, [C.c_void_p]),
    "gprb200_get_alpha": (C.c_int, [C.c_void_p, C.c_void_p]),
    "gprb200_get_alpha_dev": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]),
    "gprb200_get_factor_dev": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.POINTER(C.c_void_p)]),
    "gprb200_dgram_contract_rows_dev": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p]),
    "gprb200_get_core": (C.c_int, [C.c_void_p, C.c_void_p]),
    "gprb200_get_factor": (C.c_int, [C.c_void_p, C.c_void_p]),
    "gprb200_get_logdet": (C.c_int, [C.c_void_p, C.POINTER(C.c_double)]),
    "gprb200_spd_inverse": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.POINTER(C.c_double)]),
    "gprb200_lu_inverse": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "gprb200_gemm_dev": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_double, C.c_void_p, C.c_int64,
                                   C.c_void_p, C.c_int64, C.c_double, C.c_void_p, C.c_int64, C.c_int]),
    "gprb200_gemm_stair_dev": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_double, C.c_void_p, C.c_int64, C.c_void_p,
                                         C.c_int64, C.c_double, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]),
    "gprb200_gram_block_dev": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_int64]),
    "gprb200_gram_pair_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]),
    "gprb200_potrf_panel_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p,
                                          C.POINTER(C.c_int)]),
    "gprb200_potrf_panel_async_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_int64,
                                                C.c_void_p]),
    "gprb200_trsm_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                   C.c_int64, C.c_int]),
    "gprb200_trtri_cols_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64,
                                         C.c_void_p, C.c_int64]),
    "gprb200_logdet_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.POINTER(C.c_double)]),
    "gprb200_predict": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "gprb200_predict_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p]),
    "gprb200_gemv_t_noalloc_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]),
    "gprb200_posterior_cov": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "gprb200_predict_derivative": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gprb200_kernel_vector": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "gprb200_dgram": (C.c_int, [C.c_void_p, C.c_void_p]),
    "gprb200_loglik": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gprb200_quadform": (C.c_int, [C.c_void_p, C.c_void_p]),
    "gprb200_learn_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int64, C.c_double, C.c_void_p]),
    "gprb200_set_train_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int64, C.c_double]),
    "gprb200_sparse_set_inducing": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_double]),
    "gprb200_sparse_set_jitter": (C.c_int, [C.c_void_p, C.c_double]),
    "gprb200_sparse_learn": (C.c_int, [C.c_void_p]),
    "gprb200_sparse_get_alpha": (C.c_int, [C.c_void_p, C.c_void_p]),
    "gprb200_sparse_loglik": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gprb200_sparse_dgram_cross": (C.c_int, [C.c_void_p, C.c_void_p]),
    "gprb200_sparse_predict": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "gprb200_sparse_cov": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "gprb200_pca_create": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "gprb200_pca_destroy": (C.c_int, [C.c_void_p]),
    "gprb200_pca_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "gprb200_pca_fit": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64]),
    "gprb200_pca_fit_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64]),
    "gprb200_pca_set": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64]),
    "gprb200_pca_get": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gprb200_pca_dims": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "gprb200_pca_get_stat": (C.c_int, [C.c_void_p, C.c_char_p, C.POINTER(C.c_double)]),
    "gprb200_pca_project": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]),
    "gprb200_pca_project_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_int64]),
    "gprb200_pca_reconstruct": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]),
    "gprb200_pca_reconstruct_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int64, C.c_int64, C.c_void_p,
                                              C.c_int64]),
    "gprb200_predict_reconstruct": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "gprb200_predict_reconstruct_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                                  C.c_void_p, C.c_int64]),
}

WANT_CORE = 1
LOGLIK_REFERENCE_CLAMP = 0
LOGLIK_TRUE = 1
PHASES = ("gram", "potrf", "solve", "inverse", "predict", "dgram", "loglik", "reserved")

_lib = None


def load_library():
    """dlopen libgprb200.so and attach prototypes. Raises if the CUDA library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `make -C gpr_b200/csrc` (or __graft_entry__.build()); "
            "gpr_b200 has no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _PROTOS.items():
        if os.environ.get("GPRB200_LIB") and not hasattr(lib, name):
            continue  # an older A/B build (tools/lib_ab.py) may lack the newest entry points
        fn = getattr(lib, name)  # AttributeError here means header and library diverged
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class GprError(RuntimeError):
    """Carries the C-ABI return code; str(e) is the reference-style message."""

    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Handle:
    """Thin RAII wrapper over gprb200_handle; host-pointer entry points take/return numpy arrays."""

    def __init__(self, device=0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.gprb200_create(C.byref(h), int(device))
        if rc != 0:
            raise GprError(rc, self.lib.gprb200_last_error(None).decode())
        self.h = h
        self.N = self.M = self.d = 0
        self.P = 0

    def close(self):
        if getattr(self, "h", None):
            self.lib.gprb200_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise GprError(rc, self.lib.gprb200_last_error(self.h).decode())

    # -- model ---------------------------------------------------------------------------------
    def set_kernel(self, ops, params):
        ops = np.ascontiguousarray(ops, dtype=np.int32)
        params = _f64(params)
        self._ck(self.lib.gprb200_set_kernel(self.h, _ptr(ops), len(ops), _ptr(params), len(params)))
        self.P = len(params)

    def set_kernel_params(self, params):
        params = _f64(params)
        self._ck(self.lib.gprb200_set_kernel_params(self.h, _ptr(params), len(params)))

    def set_train(self, X, Y, sigma):
        X = _f64(X)
        Y = _f64(Y) if Y is not None else None
        self.N, self.d = X.shape
        self.M = Y.shape[1] if Y is not None else 0
        self._ck(self.lib.gprb200_set_train(self.h, _ptr(X), self.N, self.d, _ptr(Y), self.M, float(sigma)))

    def set_train_rows(self, Xrows, Yrows, sigma):
        """one numpy vector per sample / label (the reference's vector-of-vectors storage); labels are streamed"""
        Xrows = [_f64(x).ravel() for x in Xrows]
        Yrows = [_f64(y).ravel() for y in Yrows] if Yrows is not None else None
        self.N, self.d = len(Xrows), Xrows[0].size
        self.M = Yrows[0].size if Yrows else 0
        xp = (C.c_void_p * self.N)(*[x.ctypes.data for x in Xrows])
        yp = (C.c_void_p * self.N)(*[y.ctypes.data for y in Yrows]) if Yrows else None
        self._ck(self.lib.gprb200_set_train_rows(self.h, xp, self.N, self.d, yp, self.M, float(sigma)))

    def learn_rows(self, Xrows, Yrows, sigma, alpha_out=None):
        """pipelined AddSample x N + Initialize (+ host copy of alpha): upload, solve and download in column blocks"""
        Xrows = [_f64(x).ravel() for x in Xrows]
        Yrows = [_f64(y).ravel() for y in Yrows]
        self.N, self.d, self.M = len(Xrows), Xrows[0].size, Yrows[0].size
        xp = (C.c_void_p * self.N)(*[x.ctypes.data for x in Xrows])
        yp = (C.c_void_p * self.N)(*[y.ctypes.data for y in Yrows])
        self._ck(self.lib.gprb200_learn_rows(self.h, xp, self.N, self.d, yp, self.M, float(sigma), _ptr(alpha_out)))
        return alpha_out

    def set_train_dev(self, dX_ptr, N, d, dY_ptr, M, sigma):
        self.N, self.d, self.M = int(N), int(d), int(M)
        self._ck(self.lib.gprb200_set_train_dev(self.h, C.c_void_p(dX_ptr), N, d, C.c_void_p(dY_ptr), M, float(sigma)))

    def set_sigma(self, sigma):
        self._ck(self.lib.gprb200_set_sigma(self.h, float(sigma)))

    def set_state(self, X, Y, alpha, core, sigma):
        X = _f64(X)
        alpha = _f64(alpha)
        Y = _f64(Y) if Y is not None else None
        core = _f64(core) if core is not None and core.size else None
        self.N, self.d = X.shape
        self.M = alpha.shape[1]
        self._ck(self.lib.gprb200_set_state(self.h, _ptr(X), self.N, self.d, _ptr(Y), _ptr(alpha), self.M, _ptr(core),
                                            float(sigma)))

    def set_stream(self, stream_ptr):
        self._ck(self.lib.gprb200_set_stream(self.h, C.c_void_p(stream_ptr)))

    def set_option(self, name, value):
        self._ck(self.lib.gprb200_set_option(self.h, name.encode(), int(value)))

    def get_stat(self, name):
        v = C.c_double()
        self._ck(self.lib.gprb200_get_stat(self.h, name.encode(), C.byref(v)))
        return v.value

    def synchronize(self):
        self._ck(self.lib.gprb200_synchronize(self.h))

    def launch_count(self):
        return int(self.lib.gprb200_launch_count(self.h))

    def phase_ms(self):
        ms = np.zeros(8)
        self._ck(self.lib.gprb200_phase_ms(self.h, _ptr(ms)))
        return dict(zip(PHASES, ms.tolist()))

    # -- learn ---------------------------------------------------------------------------------
    def gram(self, add_noise=False):
        K = np.empty((self.N, self.N))
        self._ck(self.lib.gprb200_gram(self.h, _ptr(K), int(add_noise)))
        return K

    def learn(self, flags=0):
        self._ck(self.lib.gprb200_learn(self.h, flags))

    def factorize(self):
        self._ck(self.lib.gprb200_factorize(self.h))

    def solve_alpha(self):
        self._ck(self.lib.gprb200_solve_alpha(self.h))

    def alpha(self):
        A = np.empty((self.N, self.M))
        self._ck(self.lib.gprb200_get_alpha(self.h, _ptr(A)))
        return A

    def alpha_dev(self):
        p = C.c_void_p()
        ld = C.c_int64()
        self._ck(self.lib.gprb200_get_alpha_dev(self.h, C.byref(p), C.byref(ld)))
        return p.value, ld.value

    def factor_dev(self):
        """(device pointer of L, leading dimension, device pointer of the inverted 128 x 128 diagonal blocks)"""
        p, q = C.c_void_p(), C.c_void_p()
        ld = C.c_int64()
        self._ck(self.lib.gprb200_get_factor_dev(self.h, C.byref(p), C.byref(ld), C.byref(q)))
        return p.value, ld.value, q.value

    def dgram_contract_rows_dev(self, row0, nrows, ncols, dS_ptr, lds):
        out = np.empty(self.P)
        self._ck(self.lib.gprb200_dgram_contract_rows_dev(self.h, int(row0), int(nrows), int(ncols), C.c_void_p(dS_ptr), int(lds),
                                                          _ptr(out)))
        return out

    def core(self):
        Cm = np.empty((self.N, self.N))
        self._ck(self.lib.gprb200_get_core(self.h, _ptr(Cm)))
        return Cm

    def factor(self):
        L = np.empty((self.N, self.N))
        self._ck(self.lib.gprb200_get_factor(self.h, _ptr(L)))
        return L

    def spd_inverse(self, A):
        """InvertKernelMatrix for an arbitrary SPD host matrix; returns (inverse, logdet)."""
        A = _f64(A)
        n = A.shape[0]
        out = np.empty((n, n))
        v = C.c_double()
        self._ck(self.lib.gprb200_spd_inverse(self.h, _ptr(A), n, _ptr(out), C.byref(v)))
        return out, v.value

    def lu_inverse(self, A):
        """lapack::lu_invert for a general square host matrix."""
        A = _f64(A)
        n = A.shape[0]
        out = np.empty((n, n))
        self._ck(self.lib.gprb200_lu_inverse(self.h, _ptr(A), n, _ptr(out)))
        return out

    def logdet(self):
        v = C.c_double()
        self._ck(self.lib.gprb200_get_logdet(self.h, C.byref(v)))
        return v.value

    # -- predict -------------------------------------------------------------------------------
    def predict(self, Xs, want_ci=False):
        Xs = _f64(np.atleast_2d(Xs))
        T = Xs.shape[0]
        mean = np.empty((T, self.M))
        ci = np.empty(T) if want_ci else None
        self._ck(self.lib.gprb200_predict(self.h, _ptr(Xs), T, _ptr(mean), _ptr(ci)))
        return (mean, ci) if want_ci else mean

    def credible_interval(self, Xs):
        Xs = _f64(np.atleast_2d(Xs))
        ci = np.empty(Xs.shape[0])
        self._ck(self.lib.gprb200_predict(self.h, _ptr(Xs), Xs.shape[0], None, _ptr(ci)))
        return ci

    def predict_dev(self, dXs_ptr, T, dmean_ptr, ld_mean, dci_ptr=None):
        self._ck(self.lib.gprb200_predict_dev(self.h, C.c_void_p(dXs_ptr), T, C.c_void_p(dmean_ptr), ld_mean,
                                              C.c_void_p(dci_ptr) if dci_ptr else None))

    def posterior_cov(self, Xa, Xb):
        Xa = _f64(np.atleast_2d(Xa))
        Xb = _f64(np.atleast_2d(Xb))
        out = np.empty(Xa.shape[0])
        self._ck(self.lib.gprb200_posterior_cov(self.h, _ptr(Xa), _ptr(Xb), Xa.shape[0], _ptr(out)))
        return out

    def predict_derivative(self, x):
        x = _f64(x).reshape(-1)
        mean = np.empty(self.M)
        D = np.empty((self.d, self.M))
        self._ck(self.lib.gprb200_predict_derivative(self.h, _ptr(x), _ptr(mean), _ptr(D)))
        return mean, D

    def kernel_vector(self, Xs):
        Xs = _f64(np.atleast_2d(Xs))
        out = np.empty((Xs.shape[0], self.N))
        self._ck(self.lib.gprb200_kernel_vector(self.h, _ptr(Xs), Xs.shape[0], _ptr(out)))
        return out

    # -- likelihood ----------------------------------------------------------------------------
    def dgram(self):
        D = np.empty((self.P * self.N, self.N))
        self._ck(self.lib.gprb200_dgram(self.h, _ptr(D)))
        return D

    def loglik(self, mode=LOGLIK_REFERENCE_CLAMP, want_value=True, want_grad=False, want_jac=False):
        value = np.empty(self.M) if want_value else None
        grad = np.empty(self.P) if want_grad else None
        jac = np.empty((self.M, self.P)) if want_jac else None
        self._ck(self.lib.gprb200_loglik(self.h, int(mode), _ptr(value), _ptr(grad), _ptr(jac)))
        return value, grad, jac

    def quadform(self):
        q = np.empty(self.M)
        self._ck(self.lib.gprb200_quadform(self.h, _ptr(q)))
        return q
