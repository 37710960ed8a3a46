"""SparseGaussianProcess of the reference (include/SparseGaussianProcess.h:30-420), regression part, over
csrc/sparse.cu: same method names (AddInducingSample, ClearInducingSamples, SetJitter, Predict, __call__ for
operator()(x, y)). Dense samples come from the GaussianProcess base class (AddSample)."""
import numpy as np

from .capi import GprError, _f64, _ptr, LOGLIK_REFERENCE_CLAMP
from .gp import GaussianProcess


class SparseGaussianProcess(GaussianProcess):
    def __init__(self, kernel, jitter=0.0, device=0):
        super().__init__(kernel, device=device)
        self.m_Jitter = float(jitter)
        self.m_InducingSampleVectors = []
        self.m_InducingLabelVectors = []
        self._sparse_dirty = True

    def AddInducingSample(self, x, y):
        x, y = _f64(x).ravel(), _f64(y).ravel()
        if not self.m_InducingSampleVectors and not self.m_SampleVectors:
            self.m_InputDimension = x.size
        if not self.m_InducingLabelVectors and not self.m_LabelVectors:
            self.m_OutputDimension = y.size
        self._check_in(x, "SparseGaussianProcess::AddInducingSample: ")
        self._check_out(y, "SparseGaussianProcess::AddInducingSample: ")
        self.m_InducingSampleVectors.append(x)
        self.m_InducingLabelVectors.append(y)
        self._sparse_dirty = True

    def ClearInducingSamples(self):
        self.m_InducingSampleVectors, self.m_InducingLabelVectors = [], []
        self._sparse_dirty = True

    def GetNumberOfInducingSamples(self):
        return len(self.m_InducingSampleVectors)

    def GetJitter(self):
        return self.m_Jitter

    def SetJitter(self, jitter):
        self.m_Jitter = float(jitter)
        self._sparse_dirty = True

    def _sync_sparse(self):
        """kernel, dense samples, sigma and inducing points resident on the device; nothing is factored"""
        self._sync_kernel()
        self._sync_train()
        self._h.set_sigma(self.m_Sigma)
        h = self._h
        if self._sparse_dirty:
            Xm = _f64(np.stack(self.m_InducingSampleVectors))
            h._ck(h.lib.gprb200_sparse_set_inducing(h.h, _ptr(Xm), Xm.shape[0], Xm.shape[1], self.m_Jitter))
            self._sparse_dirty = False

    def Initialize(self):
        if not self.m_InducingSampleVectors:
            raise GprError(-2, "SparseGaussianProcess::Initialize: no inducing samples defined during initialization")
        if not self.m_SampleVectors:
            raise GprError(-2, "SparseGaussianProcess::Initialize: no dense samples defined during initialization")
        self._sync_sparse()
        self._h._ck(self._h.lib.gprb200_sparse_learn(self._h.h))

    def RegressionVectors(self):
        self.Initialize()
        a = np.empty((len(self.m_InducingSampleVectors), self.m_OutputDimension))
        self._h._ck(self._h.lib.gprb200_sparse_get_alpha(self._h.h, _ptr(a)))
        return a

    def Predict(self, x):
        self.Initialize()
        x = _f64(x).ravel()
        self._check_in(x, "GaussianProcess::Predict: ")
        out = np.empty((1, self.m_OutputDimension))
        self._h._ck(self._h.lib.gprb200_sparse_predict(self._h.h, _ptr(x), 1, _ptr(out), None))
        return out[0]

    def PredictBatch(self, Xs, want_var=False):
        self.Initialize()
        Xs = _f64(np.atleast_2d(Xs))
        mean = np.empty((Xs.shape[0], self.m_OutputDimension))
        var = np.empty(Xs.shape[0]) if want_var else None
        self._h._ck(self._h.lib.gprb200_sparse_predict(self._h.h, _ptr(Xs), Xs.shape[0], _ptr(mean), _ptr(var)))
        return (mean, var) if want_var else mean

    def __call__(self, x, y):
        self.Initialize()
        x, y = _f64(x).ravel(), _f64(y).ravel()
        self._check_in(x, "SparseGaussianProcess::(): ")
        self._check_in(y, "SparseGaussianProcess::(): ")
        out = np.empty(1)
        self._h._ck(self._h.lib.gprb200_sparse_cov(self._h.h, _ptr(x), _ptr(y), 1, _ptr(out)))
        return float(out[0])


class SparseGaussianLogLikelihood:
    """include/SparseLikelihood.h:111-551 over gprb200_sparse_loglik. `mode` selects the reference's long-double range
    clamp of the complexity term (default) or the true -0.5 log det(sigma^2 I + Knm K^-1 Kmn)."""

    def __init__(self, mode=LOGLIK_REFERENCE_CLAMP):
        self.mode = mode

    def ToString(self):
        return "SparseGaussianLogLikelihood"

    def _run(self, gp, want_value, want_grad, want_jac):
        if not isinstance(gp, SparseGaussianProcess):
            raise GprError(-2, "SparseGaussianLogLikelihood: cannot cast to SparseGaussianProcess")
        if gp.GetNumberOfInducingSamples() == 0:
            raise GprError(-2, "SparseLikelihood::GetValueAndParameterDerivative: there are no inducing samples specified")
        gp._sync_sparse()
        h = gp._h
        M, P = gp.m_OutputDimension, gp.m_Kernel.GetNumberOfParameters()
        v = np.empty(M) if want_value else None
        g = np.empty(P) if want_grad else None
        j = np.empty((M, P)) if want_jac else None
        h._ck(h.lib.gprb200_sparse_loglik(h.h, int(self.mode), _ptr(v), _ptr(g), _ptr(j)))
        return v, g, j

    def __call__(self, gp):  # :152-217
        return self._run(gp, True, False, False)[0]

    def GetParameterDerivatives(self, gp):  # :219-285
        return self._run(gp, False, True, False)[1]

    def GetValueAndParameterDerivatives(self, gp):  # :287-409
        v, g, _ = self._run(gp, True, True, False)
        return v, g

    def GetValueAndJacobian(self, gp):  # :411-535
        v, _, j = self._run(gp, True, False, True)
        return v, j
