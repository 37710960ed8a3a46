"""Python mirror of the reference's GaussianProcess<T> / GaussianLogLikelihood<T> interface.

Same method names, argument meaning and error behaviour as /root/reference/include/GaussianProcess.h
and include/Likelihood.h (the reference throws std::string; here GprError / ValueError carry the same
message). All numerical work goes through the C-ABI (gpr_b200.capi.Handle -> libgprb200.so); there is no
numpy implementation of the path in this package.
"""
import os

import numpy as np

from .capi import Handle, GprError, WANT_CORE, LOGLIK_REFERENCE_CLAMP, LOGLIK_TRUE
from .kernels import Kernel, KernelFactory

FullPivotLU, JacobiSVD, BDCSVD, SelfAdjointEigenSolver = 0, 1, 2, 3  # include/GaussianProcess.h:66


# ---- MatrixIO (lib/MatrixIO.cpp:38-99): ASCII "rows cols\n" + raw row-major scalars -------------
def WriteMatrix(matrix, filename, dtype=np.float64):
    m = np.ascontiguousarray(matrix, dtype=dtype)
    if m.ndim == 1:
        m = m.reshape(-1, 1)
    with open(filename, "wb") as f:
        f.write(f"{m.shape[0]} {m.shape[1]}\n".encode())
        f.write(m.tobytes(order="C"))


def ReadMatrix(filename, dtype=np.float64):
    with open(filename, "rb") as f:
        header = f.readline()
        try:
            rows, cols = (int(t) for t in header.split()[:2])
        except Exception:
            raise ValueError(f"ReadMatrix: header is corrupt (filename {filename}).")
        data = np.frombuffer(f.read(rows * cols * np.dtype(dtype).itemsize), dtype=dtype)
    return data.reshape(rows, cols).astype(np.float64)


class GaussianProcess:
    def __init__(self, kernel, device=0, dtype=np.float64):
        if not isinstance(kernel, Kernel):
            raise TypeError("GaussianProcess needs a gpr_b200.kernels.Kernel")
        self.m_Kernel = kernel
        self.m_Sigma = 0.0
        self.m_SampleVectors = []
        self.m_LabelVectors = []
        self.m_Initialized = False
        self.m_InputDimension = 0
        self.m_OutputDimension = 0
        self.m_InvMethod = FullPivotLU
        self.m_EfficientStorage = False
        self.debug = False
        self.dtype = dtype  # float32 instantiation == cast at the boundary (include/LAPACKUtils.h:87,96)
        self._h = Handle(device)
        self._sent_kernel = None
        self._train_dirty = True
        self._alpha = None       # host mirror of m_RegressionVectors (lazy)
        self._core = None        # host mirror of m_CoreMatrix (lazy)
        self._core_size = 0      # what m_CoreMatrix.diagonalSize() would be
        self._loaded = False
        self._device = device
        self._dist = None        # capi_dist.DistHandle of this rank after SetDistributed

    # ---- helpers -----------------------------------------------------------------------------
    def _sync_kernel(self):
        ops, params = self.m_Kernel.lower()
        key = (tuple(ops), tuple(params))
        if key != self._sent_kernel:
            if self._sent_kernel is not None and self._sent_kernel[0] == key[0]:
                self._h.set_kernel_params(params)
            else:
                self._h.set_kernel(ops, params)
            self._sent_kernel = key

    def _sync_train(self):
        if self._train_dirty:
            X = np.stack(self.m_SampleVectors).astype(np.float64)
            Y = np.stack(self.m_LabelVectors).astype(np.float64)
            self._h.set_train(X, Y, self.m_Sigma)
            self._train_dirty = False

    def _check_in(self, x, prefix):
        if x.size != self.m_InputDimension:
            raise ValueError(f"{prefix}dimension of input vector ({x.size}) does not correspond to the input dimension "
                             f"({self.m_InputDimension}).")

    def _check_out(self, y, prefix):
        if y.size != self.m_OutputDimension:
            raise ValueError(f"{prefix}dimension of output vector ({y.size}) does not correspond to the output "
                             f"dimension ({self.m_OutputDimension}).")

    def _cast(self, a):
        return np.asarray(a, dtype=np.float64).astype(self.dtype).astype(np.float64) if self.dtype != np.float64 else a

    # ---- reference API -----------------------------------------------------------------------
    def AddSample(self, x, y):  # lib/GaussianProcess.cpp:36-51
        x = np.asarray(x, dtype=np.float64).reshape(-1)
        y = np.asarray(y, dtype=np.float64).reshape(-1)
        if not self.m_SampleVectors:
            self.m_InputDimension = x.size
        if not self.m_LabelVectors:
            self.m_OutputDimension = y.size
        self._check_in(x, "GaussianProcess::AddSample: ")
        self._check_out(y, "GaussianProcess::AddSample: ")
        self.m_SampleVectors.append(x.copy())
        self.m_LabelVectors.append(y.copy())
        self.m_Initialized = False
        self._train_dirty = True

    def SetDistributed(self, rank, world, unique_id=None, grid=None):
        """ONE model over several GPUs (mirror of gpr::GaussianProcess<T>::SetDistributed; not in the reference): call on
        every rank before Initialize(). unique_id: capi_dist.get_unique_id() of rank 0 (capi_dist.exchange_id() carries it
        over a torch.distributed group). Initialize() then factors on the grid; Predict runs locally, __call__ and
        GetCredibleInterval solve on the grid and are collective."""
        from . import capi_dist
        self._dist = capi_dist.DistHandle(self._device, rank, world, unique_id, grid)
        self.m_Initialized = False

    def _kernel_value(self, x, y):
        """k(x, y) on the device (the Python kernel classes carry parameters only): a one-sample model whose kernel vector
        at x is the value."""
        if getattr(self, "_pair_h", None) is None:
            self._pair_h = Handle(self._device)
        ops, params = self.m_Kernel.lower()
        self._pair_h.set_kernel(ops, params)
        self._pair_h.set_train(y[None, :], np.zeros((1, 1)), 0.0)
        return float(self._pair_h.kernel_vector(x[None, :])[0, 0])

    def _initialize_distributed(self):
        ops, params = self.m_Kernel.lower()
        X = np.stack(self.m_SampleVectors).astype(np.float64)
        Y = np.stack(self.m_LabelVectors).astype(np.float64)
        self._dist.set_kernel(ops, params)
        alpha, self._dist_logdet = self._dist.learn(X, Y, self.m_Sigma)
        self._h.set_kernel(ops, params)
        self._sent_kernel = (tuple(ops), tuple(params))
        self._h.set_state(X, Y, alpha, None, self.m_Sigma)      # alpha on this rank's GPU, like a loaded model
        self._train_dirty = False
        self._alpha = alpha
        self._core = None
        self._core_size = 0 if self.m_EfficientStorage else len(self.m_SampleVectors)
        self._loaded = False
        self.m_Initialized = True

    def Initialize(self):  # lib/GaussianProcess.cpp:118-130
        if self.m_Initialized:
            return
        if not self.m_SampleVectors:
            raise GprError(-2, "GaussianProcess::Initialize: no input samples defined during initialization")
        if not self.m_LabelVectors:
            raise GprError(-2, "GaussianProcess::Initialize: no ouput labels defined during initialization")
        if self._dist is not None:
            return self._initialize_distributed()
        self._sync_kernel()
        self._sync_train()
        self._h.set_sigma(self.m_Sigma)
        self._h.learn(0)
        self._alpha = None
        self._core = None
        self._core_size = 0 if self.m_EfficientStorage else len(self.m_SampleVectors)
        self._loaded = False
        self.m_Initialized = True

    def Predict(self, x):  # lib/GaussianProcess.cpp:55-61
        self.Initialize()
        x = np.asarray(x, dtype=np.float64).reshape(-1)
        self._check_in(x, "GaussianProcess::Predict: ")
        self._sync_kernel()
        return self._cast(self._h.predict(x[None, :])[0])

    def PredictBatch(self, Xs, want_ci=False):
        """Additive batch entry point (T test points per launch); not in the reference."""
        self.Initialize()
        Xs = np.atleast_2d(np.asarray(Xs, dtype=np.float64))
        self._sync_kernel()
        return self._h.predict(Xs, want_ci=want_ci)

    def PredictDerivative(self, x):  # lib/GaussianProcess.cpp:65-81 -> (mean, D[d x M])
        self.Initialize()
        x = np.asarray(x, dtype=np.float64).reshape(-1)
        self._check_in(x, "GaussianProcess::PredictDerivative: ")
        self._sync_kernel()
        mean, D = self._h.predict_derivative(x)
        return self._cast(mean), self._cast(D)

    def __call__(self, x, y):  # lib/GaussianProcess.cpp:85-99
        self.Initialize()
        x = np.asarray(x, dtype=np.float64).reshape(-1)
        y = np.asarray(y, dtype=np.float64).reshape(-1)
        self._check_in(x, "GaussianProcess::(): ")
        self._check_in(y, "GaussianProcess::(): ")
        self._sync_kernel()
        if self._core_size == 0:
            self._core_size = len(self.m_SampleVectors)  # the reference rebuilds m_CoreMatrix here (:95-97)
        if self._dist is not None:      # the factor lives on the grid: k(x, y) - Kx^T (K + s^2 I)^-1 Ky, collective
            Kx, Ky = self._h.kernel_vector(x[None, :])[0], self._h.kernel_vector(y[None, :])[0]
            S = self._dist.solve(Ky[:, None])[:, 0]
            return float(self._kernel_value(x, y) - np.dot(Kx.astype(np.longdouble), S.astype(np.longdouble)))
        return float(self._h.posterior_cov(x[None, :], y[None, :])[0])

    def GetCredibleInterval(self, x):  # lib/GaussianProcess.cpp:103-114
        self.Initialize()
        x = np.asarray(x, dtype=np.float64).reshape(-1)
        self._check_in(x, "GaussianProcess::GetCredibleIntervall: ")
        c = self(x, x)
        return 2.0 * np.sqrt(max(0.0, c))

    def GetKernel(self):
        return self.m_Kernel

    def SetKernel(self, k):
        self.m_Kernel = k
        self.m_Initialized = False

    def DebugOn(self):
        self.debug = True

    def GetNumberOfSamples(self):
        return len(self.m_SampleVectors)

    def GetSigma(self):
        return self.m_Sigma

    def GetSigmaSquared(self):
        return self.m_Sigma * self.m_Sigma

    def SetSigma(self, sigma):
        self.m_Sigma = float(sigma)
        self.m_Initialized = False

    def GetNumberOfInputDimensions(self):
        return self.m_InputDimension

    def SetInversionMethod(self, m):
        # All four reference methods (LU, two SVDs, eigen/Cholesky) are inverses of the same SPD matrix;
        # the B200 path always factors by Cholesky. The enum is kept for source compatibility.
        self.m_InvMethod = m

    def GetInversionMethod(self):
        return self.m_InvMethod

    def GetEfficientStorage(self):
        return self.m_EfficientStorage

    def SetEfficientStorage(self, s):
        self.m_EfficientStorage = bool(s)

    # host mirrors -------------------------------------------------------------------------------
    def GetRegressionVectors(self):
        if self._alpha is None:
            self._alpha = self._cast(self._h.alpha())
        return self._alpha

    def GetCoreMatrix(self):
        if self._core_size == 0:
            return np.zeros((0, 0))
        if self._core is None:
            self._sync_kernel()
            self._core = self._cast(self._h.core())
        return self._core

    # ---- Save / Load (lib/GaussianProcess.cpp:134-268) -------------------------------------------
    def Save(self, prefix):
        if not self.m_Initialized:
            raise GprError(-2, "GaussianProcess::Save: gaussian process is not initialized.")
        WriteMatrix(self.GetRegressionVectors(), prefix + "-RegressionVectors.txt", self.dtype)
        if self.m_EfficientStorage:
            self._core_size = 0
            self._core = None
        WriteMatrix(self.GetCoreMatrix(), prefix + "-CoreMatrix.txt", self.dtype)
        WriteMatrix(np.stack(self.m_SampleVectors).T, prefix + "-SampleVectors.txt", self.dtype)
        WriteMatrix(np.stack(self.m_LabelVectors).T, prefix + "-LabelVectors.txt", self.dtype)
        with open(prefix + "-ParameterFile.txt", "w") as f:
            # sigma is streamed before setprecision -> 6 significant digits (:174)
            f.write("%g %d %d %d %d " % (self.m_Sigma, self.m_InputDimension, self.m_OutputDimension,
                                         int(self.m_EfficientStorage), int(self.debug)))
            f.write(self.m_Kernel.ToString())

    def Load(self, prefix):
        files = {}
        for tag in ("RegressionVectors", "CoreMatrix", "SampleVectors", "LabelVectors"):
            fn = f"{prefix}-{tag}.txt"
            if not os.path.isfile(fn):
                raise GprError(-2, f"GaussianProcess::Load: {fn} does not exist or is a directory.")
            files[tag] = ReadMatrix(fn, self.dtype)
        pf = prefix + "-ParameterFile.txt"
        if not os.path.isfile(pf):
            raise GprError(-2, f"GaussianProcess::Load: {pf} does not exist or is a directory.")
        line = open(pf).readline()
        tok = line.split(None, 5)
        try:
            self.m_Sigma = float(tok[0])
            self.m_InputDimension = int(tok[1])
            self.m_OutputDimension = int(tok[2])
            self.m_EfficientStorage = bool(int(tok[3]))
            self.debug = bool(int(tok[4]))
        except Exception:
            raise GprError(-2, "GaussianProcess::Load: parameter file is corrupt")
        self.m_Kernel = KernelFactory.GetKernel(tok[5].split()[0] if len(tok) > 5 else "")
        X = files["SampleVectors"].T
        Y = files["LabelVectors"].T
        self.m_SampleVectors = [r.copy() for r in X]
        self.m_LabelVectors = [r.copy() for r in Y]
        alpha = files["RegressionVectors"]
        core = files["CoreMatrix"]
        self._sent_kernel = None
        self._sync_kernel()
        self._h.set_state(X, Y, alpha, core if core.size else None, self.m_Sigma)
        self._train_dirty = False
        self._alpha = alpha
        self._core = core if core.size else None
        self._core_size = core.shape[0]
        self._loaded = True
        self.m_Initialized = True  # :267 — no re-factorisation on Load

    def __eq__(self, b):  # lib/GaussianProcess.cpp:292-360
        if not isinstance(b, GaussianProcess):
            return False
        ra, rb = self.GetRegressionVectors(), b.GetRegressionVectors()
        if ra.shape != rb.shape or np.linalg.norm(ra - rb) > 0:
            return False
        if self._core_size != b._core_size:
            return False
        if len(self.m_SampleVectors) != len(b.m_SampleVectors):
            return False
        cast = lambda v: v.astype(self.dtype)
        if any(np.linalg.norm(cast(x) - cast(y)) > 0 for x, y in zip(self.m_SampleVectors, b.m_SampleVectors)):
            return False
        if len(self.m_LabelVectors) != len(b.m_LabelVectors):
            return False
        if any(np.linalg.norm(cast(x) - cast(y)) > 0 for x, y in zip(self.m_LabelVectors, b.m_LabelVectors)):
            return False
        if self.m_Kernel != b.m_Kernel:
            return False
        return (self.m_Sigma == b.m_Sigma and self.m_Initialized == b.m_Initialized
                and self.m_InputDimension == b.m_InputDimension and self.m_OutputDimension == b.m_OutputDimension
                and self.m_EfficientStorage == b.m_EfficientStorage and self.debug == b.debug)

    def __ne__(self, b):
        return not self.__eq__(b)

    __hash__ = None

    # ---- protected virtuals the Likelihood classes reach through friendship -----------------------
    def ComputeKernelMatrix(self):  # lib/GaussianProcess.cpp:363-372
        self._sync_kernel()
        self._sync_train()
        return self._h.gram(add_noise=False)

    def ComputeDerivativeKernelMatrix(self):  # :460-495
        self._sync_kernel()
        self._sync_train()
        return self._h.dgram()

    def ComputeKernelVector(self, x):  # :675-693
        if not self.m_Initialized:
            raise GprError(-2, "GaussianProcess::ComputeKernelVectorInternal: gaussian process is not initialized.")
        self._sync_kernel()
        return self._h.kernel_vector(np.asarray(x, dtype=np.float64).reshape(1, -1))[0]

    def ComputeLabelMatrix(self):  # :621-639
        if not self.m_LabelVectors:
            raise GprError(-2, "GaussianProcess::ComputeLabelMatrixInternal: no ouput labels defined.")
        return np.stack(self.m_LabelVectors)


class GaussianLogLikelihood:
    """include/Likelihood.h:152-356. `mode` selects the reference's clamped complexity term (default)
    or the true -0.5 logdet."""

    def __init__(self, mode=LOGLIK_REFERENCE_CLAMP):
        self.mode = mode

    def ToString(self):
        return "GaussianLogLikelihood"

    def _prep(self, gp):
        gp._sync_kernel()
        gp._sync_train()
        gp._h.set_sigma(gp.m_Sigma)

    def __call__(self, gp):  # :166-202
        self._prep(gp)
        return gp._h.loglik(self.mode, True, False, False)[0]

    def GetParameterDerivatives(self, gp):  # :204-229
        self._prep(gp)
        return gp._h.loglik(self.mode, False, True, False)[1]

    def GetValueAndParameterDerivatives(self, gp):  # :231-285
        self._prep(gp)
        v, g, _ = gp._h.loglik(self.mode, True, True, False)
        return v, g

    def GetValueAndJacobian(self, gp):  # :287-344
        self._prep(gp)
        v, _, j = gp._h.loglik(self.mode, True, False, True)
        return v, j


class GaussianLikelihood:
    """include/Likelihood.h:95-150: exp(df) * det^-1/2 * (2 pi)^-N/2 from the same device quantities."""

    def ToString(self):
        return "GaussianLikelihood"

    def __call__(self, gp):
        gp._sync_kernel()
        gp._sync_train()
        gp._h.set_sigma(gp.m_Sigma)
        gp._h.learn(0)
        q = gp._h.quadform()
        det = np.exp(np.longdouble(gp._h.logdet()))
        det = float(det)  # GetCoreMatrix returns TScalarType (:77-79)
        cp = 1.0 / np.sqrt(np.finfo(np.longdouble).tiny) if det <= 0 else 1.0 / np.sqrt(det)
        ct = 1.0 / np.power(2 * np.pi, len(gp.m_SampleVectors) / 2.0)
        return np.exp(-0.5 * q) * float(cp) * ct
