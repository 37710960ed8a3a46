"""PCA<T> of the reference (include/PCA.h) over the device implementation in csrc/pca.cu.

Same method names and argument meaning as the reference class: data matrices are D x n with ONE SAMPLE PER COLUMN
(the MatrixType DataParser builds, include/DataParser.h:536-613), `nFeatures` is the number of leading modes kept for
reconstruction. Files written by WriteMatricesToFile are the reference's Mean.bin / Sigma.bin / U.bin
(include/PCA.h:126-134) in MatrixIO's binary format (lib/MatrixIO.cpp:38-99).
"""
import ctypes as C

import numpy as np

from .capi import GprError, Handle, _f64, _ptr


class PCA:
    def __init__(self, X_or_path, nFeatures, handle=None, device=0, passes=None):
        self._own = handle is None
        self.handle = handle if handle is not None else Handle(device)
        self.lib = self.handle.lib
        self.p = C.c_void_p()
        self._ck(self.lib.gprb200_pca_create(self.handle.h, C.byref(self.p)))
        self.nFeatures = int(nFeatures)
        if passes is not None:
            self._ck(self.lib.gprb200_pca_set_option(self.p, b"passes", int(passes)))
        if isinstance(X_or_path, str):                                       # PCA(std::string path, int nFeatures)
            from .gp import ReadMatrix
            mean = _f64(ReadMatrix(X_or_path + "Mean.bin")).ravel()
            sigma = _f64(ReadMatrix(X_or_path + "Sigma.bin")).ravel()
            U = _f64(ReadMatrix(X_or_path + "U.bin"))
            self._ck(self.lib.gprb200_pca_set(self.p, _ptr(mean), _ptr(sigma), _ptr(U), U.shape[0], U.shape[1],
                                              self.nFeatures))
        else:                                                                # PCA(MatrixType& X, int nFeatures)
            X = _f64(X_or_path)
            if X.ndim != 2:
                raise GprError(-2, "PCA: need a D x n data matrix")
            self._ck(self.lib.gprb200_pca_fit(self.p, _ptr(X), X.shape[0], X.shape[1], self.nFeatures))
        D, n, nF = C.c_int64(), C.c_int64(), C.c_int64()
        self._ck(self.lib.gprb200_pca_dims(self.p, C.byref(D), C.byref(n), C.byref(nF)))
        self.D, self.n = D.value, n.value

    @classmethod
    def from_arrays(cls, mean, sigma, U, nFeatures, handle=None, device=0):
        self = cls.__new__(cls)
        self._own = handle is None
        self.handle = handle if handle is not None else Handle(device)
        self.lib = self.handle.lib
        self.p = C.c_void_p()
        self._ck(self.lib.gprb200_pca_create(self.handle.h, C.byref(self.p)))
        self.nFeatures = int(nFeatures)
        mean, sigma, U = _f64(mean).ravel(), _f64(sigma).ravel(), _f64(U)
        self._ck(self.lib.gprb200_pca_set(self.p, _ptr(mean), _ptr(sigma), _ptr(U), U.shape[0], U.shape[1], self.nFeatures))
        self.D, self.n = U.shape
        return self

    @classmethod
    def fit_dev(cls, dX_ptr, ldx, D, n, nFeatures, handle, passes=None):
        """Fit from a D x n matrix already in HBM (device pointer, leading dimension ldx)."""
        self = cls.__new__(cls)
        self._own = False
        self.handle = handle
        self.lib = handle.lib
        self.p = C.c_void_p()
        self._ck(self.lib.gprb200_pca_create(handle.h, C.byref(self.p)))
        self.nFeatures = int(nFeatures)
        if passes is not None:
            self._ck(self.lib.gprb200_pca_set_option(self.p, b"passes", int(passes)))
        self._ck(self.lib.gprb200_pca_fit_dev(self.p, C.c_void_p(dX_ptr), ldx, D, n, self.nFeatures))
        self.D, self.n = int(D), int(n)
        return self

    def reconstruct_dev(self, dW_ptr, ldw, transposed, T, dOut_ptr, ldo):
        self._ck(self.lib.gprb200_pca_reconstruct_dev(self.p, C.c_void_p(dW_ptr), ldw, int(transposed), self.nFeatures, T,
                                                      C.c_void_p(dOut_ptr), ldo))

    def project_dev(self, dX_ptr, ldx, T, rows_out, dF_ptr, ldf):
        self._ck(self.lib.gprb200_pca_project_dev(self.p, C.c_void_p(dX_ptr), ldx, T, rows_out, C.c_void_p(dF_ptr), ldf))

    def predict_reconstruct_dev(self, gp_handle, dXs_ptr, T, dFields_ptr, ldo, dW_ptr=None, ldw=0):
        self._ck(self.lib.gprb200_predict_reconstruct_dev(gp_handle.h, self.p, C.c_void_p(dXs_ptr), T,
                                                          C.c_void_p(dFields_ptr), ldo,
                                                          C.c_void_p(dW_ptr) if dW_ptr else None, ldw))

    def _ck(self, rc):
        if rc != 0:
            raise GprError(rc, self.lib.gprb200_last_error(self.handle.h).decode())

    def close(self):
        if getattr(self, "p", None) is not None and self.p:
            self.lib.gprb200_pca_destroy(self.p)
            self.p = None
        if self._own and self.handle is not None:
            self.handle.close()
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- the reference's accessors (include/PCA.h:71-90) ----
    def _get(self, mean=False, sigma=False, U=False):
        m = np.empty(self.D) if mean else None
        s = np.empty(self.n) if sigma else None
        u = np.empty((self.D, self.n)) if U else None
        self._ck(self.lib.gprb200_pca_get(self.p, _ptr(m), _ptr(s), _ptr(u)))
        return m, s, u

    def GetMean(self):
        return self._get(mean=True)[0]

    def GetEigenvalues(self):
        return self._get(sigma=True)[1]

    def GetMatrixU(self):
        return self._get(U=True)[2]

    def GetBasis(self, nFeatures=0):
        _, s, u = self._get(sigma=True, U=True)
        with np.errstate(divide="ignore", invalid="ignore"):
            basis = u / s[None, :]
        return basis[:, :nFeatures] if 0 < nFeatures < basis.shape[1] else basis

    def PrecomputeTranspose(self):
        """The compact transposed basis already lives on the device (csrc/pca.cu: dBasis)."""

    def DimensionalityReduction(self, X, nFeatures=0):
        """basis^T (X - mean): every mode (the reference's `nFeatures << rows` test is always true, PCA.h:95)."""
        rows = nFeatures if 0 < nFeatures <= self.n else self.n
        return self._project(X, rows)

    def DimensionalityReductionFast(self, X):
        return self._project(X, self.nFeatures)

    def _project(self, X, rows):
        X = _f64(X)
        if X.ndim == 1:
            X = X[:, None]
        if X.shape[0] != self.D:
            raise GprError(-2, "PCA::DimensionalityReduction: dimension mismatch")
        F = np.empty((rows, X.shape[1]))
        self._ck(self.lib.gprb200_pca_project(self.p, _ptr(X), X.shape[1], rows, _ptr(F)))
        return F

    def GetReconstruction(self, weights):
        W = _f64(weights)
        if W.ndim == 1:
            W = W[:, None]
        out = np.empty((self.D, W.shape[1]))
        self._ck(self.lib.gprb200_pca_reconstruct(self.p, _ptr(W), W.shape[0], W.shape[1], _ptr(out)))
        return out

    def GetExplainedVariance(self):
        """Cumulative sigma, normalised (the intent of include/PCA.h:117-124; the reference indexes the vector
        m_sigma as (i, i), which is out of bounds for i >= 1)."""
        c = np.cumsum(self.GetEigenvalues())
        return c / c[-1]

    def WriteMatricesToFile(self, path):
        from .gp import WriteMatrix
        m, s, u = self._get(mean=True, sigma=True, U=True)
        WriteMatrix(m[:, None], path + "Mean.bin")
        WriteMatrix(s[:, None], path + "Sigma.bin")
        WriteMatrix(u, path + "U.bin")

    # ---- additive: fused Predict -> GetReconstruction on the device ----
    def PredictReconstruct(self, gp_handle, Xs, want_weights=False):
        """fields (D x T) for T test points through the GP held by `gp_handle` (same Handle this PCA was built on)."""
        Xs = _f64(np.atleast_2d(Xs))
        T = Xs.shape[0]
        fields = np.empty((self.D, T))
        w = np.empty((T, self.nFeatures)) if want_weights else None
        self._ck(self.lib.gprb200_predict_reconstruct(gp_handle.h, self.p, _ptr(Xs), T, _ptr(fields), _ptr(w)))
        return (fields, w) if want_weights else fields

    def stat(self, name):
        v = C.c_double()
        self._ck(self.lib.gprb200_pca_get_stat(self.p, name.encode(), C.byref(v)))
        return v.value
