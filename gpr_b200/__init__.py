"""gpr_b200 — B200-native GaussianProcess Learn/Predict hot path behind the reference's API.

Layout: csrc/ (hand-written sm_100a CUDA + the C-ABI of include/gprb200.h), include/ (C++ host API
with the reference's class names), capi.py / kernels.py / gp.py (Python mirror used by tests and
bench.py). Importing the package loads libgprb200.so; there is no CPU fallback.
"""
from .capi import load_library, Handle, GprError, LIB_PATH  # noqa: F401

load_library()

from .kernels import (  # noqa: E402,F401
    Kernel, GaussianKernel, GaussianExpKernel, WhiteKernel, RationalQuadraticKernel, PeriodicKernel,
    SumKernel, ProductKernel, KernelFactory, GetGeneralKernel)
from .gp import (  # noqa: E402,F401
    GaussianProcess, GaussianLogLikelihood, GaussianLikelihood, ReadMatrix, WriteMatrix,
    FullPivotLU, JacobiSVD, BDCSVD, SelfAdjointEigenSolver)
from .inference import GaussianProcessInference  # noqa: E402,F401
from .pca import PCA  # noqa: E402,F401
from .sparse import SparseGaussianProcess, SparseGaussianLogLikelihood  # noqa: E402,F401
