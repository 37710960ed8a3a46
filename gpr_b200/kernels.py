"""Host-side kernel descriptors: the reference's Kernel / KernelFactory interface for Python callers.

Names, constructor argument order, parameter order, ToString() format and the zero-parameter
errors follow /root/reference/include/Kernel.h and KernelFactory.h; the objects only DESCRIBE a
kernel tree — evaluation happens on the GPU through the flat post-order program produced by
`lower()` (gprb200_set_kernel, include/gprb200.h).
"""
from . import capi  # noqa: F401  (fails loudly if the CUDA library is missing)

OP_GAUSSIAN, OP_GAUSSIAN_EXP, OP_WHITE, OP_RQ, OP_PERIODIC, OP_SUM, OP_PRODUCT = 1, 2, 3, 4, 5, 6, 7


def _p2s(v):
    # Kernel::P2S: operator<< with setprecision(digits10 + 1) == %.16g  (include/Kernel.h:127-132)
    return "%.16g" % float(v)


class Kernel:
    name = "Kernel"
    op = 0
    nparams = 0

    def GetParameters(self):
        return list(self._params)

    def SetParameters(self, params):
        params = [float(p) for p in params]
        if len(params) != self.GetNumberOfParameters():
            raise ValueError(f"{self.name}::SetParameters: wrong number of parameters.")
        self._params = params

    def GetNumberOfParameters(self):
        return self.nparams

    def ToString(self):
        return self.name + "(" + "".join(_p2s(p) + "," for p in self._params) + ")"

    def lower(self):
        """-> (ops, params) post-order program for gprb200_set_kernel."""
        return [self.op], list(self._params)

    def __eq__(self, other):
        # include/Kernel.h:93-114: same string form and parameters within 10 eps
        if not isinstance(other, Kernel) or self.ToString() != other.ToString():
            return False
        a, b = self.GetParameters(), other.GetParameters()
        return len(a) == len(b) and all(abs(x - y) <= 10 * 2.220446049250313e-16 for x, y in zip(a, b))

    def __ne__(self, other):
        return not self.__eq__(other)

    __hash__ = None


class GaussianKernel(Kernel):
    """k = scale^2 exp(-0.5 r^2 / sigma^2); parameters [sigma, scale] (include/Kernel.h:483-489)."""
    name, op, nparams = "GaussianKernel", OP_GAUSSIAN, 2

    def __init__(self, sigma, scale=1.0):
        sigma, scale = float(sigma), float(scale)
        if sigma == 0:
            raise ValueError("GaussianKernel: sigma has to be positive")  # include/Kernel.h:529
        if scale == 0:
            raise ValueError("GaussianKernel: scale has to be positive")  # include/Kernel.h:530
        self._params = [sigma, scale]


class GaussianExpKernel(Kernel):
    """Gaussian kernel with log-space parameters [sigma, scale] (include/Kernel.h:580-608)."""
    name, op, nparams = "GaussianExpKernel", OP_GAUSSIAN_EXP, 2

    def __init__(self, sigma, scale=1.0):
        self._params = [float(sigma), float(scale)]


class WhiteKernel(Kernel):
    """k = scale^2 [||x - y|| == 0]; parameters [scale] (include/Kernel.h:695-721)."""
    name, op, nparams = "WhiteKernel", OP_WHITE, 1

    def __init__(self, scale):
        self._params = [float(scale)]


class RationalQuadraticKernel(Kernel):
    """k = scale^2 (1 + r^2 / (2 alpha sigma^2))^-alpha; [scale, sigma, alpha] (include/Kernel.h:794-820)."""
    name, op, nparams = "RationalQuadraticKernel", OP_RQ, 3

    def __init__(self, scale, sigma, alpha):
        self._params = [float(scale), float(sigma), float(alpha)]


class PeriodicKernel(Kernel):
    """k = scale^2 exp(-0.5 sum_d sin^2(b (x_d - y_d)) / sigma^2); [scale, b, sigma] (include/Kernel.h:912-961)."""
    name, op, nparams = "PeriodicKernel", OP_PERIODIC, 3

    def __init__(self, scale, b, sigma):
        scale, b, sigma = float(scale), float(b), float(sigma)
        if scale == 0:
            raise ValueError("PeriodicKernel: scale parameter has to be positive.")
        if b == 0:
            raise ValueError("PeriodicKernel: period length parameter has to be positive.")
        if sigma == 0:
            raise ValueError("PeriodicKernel: sigma parameter has to be positive.")
        self._params = [scale, b, sigma]


class _Binary(Kernel):
    def __init__(self, k1, k2):
        self.k1, self.k2 = k1, k2

    def GetKernel1(self):
        return self.k1

    def GetKernel2(self):
        return self.k2

    def GetNumberOfParameters(self):
        return self.k1.GetNumberOfParameters() + self.k2.GetNumberOfParameters()

    def GetParameters(self):  # k1's block first (include/Kernel.h:187-191)
        return self.k1.GetParameters() + self.k2.GetParameters()

    def SetParameters(self, params):
        params = list(params)
        if len(params) != self.GetNumberOfParameters():
            raise ValueError(f"{self.name}::SetParameters: wrong number of parameters.")
        n1 = self.k1.GetNumberOfParameters()
        self.k1.SetParameters(params[:n1])
        self.k2.SetParameters(params[n1:])

    def ToString(self):
        return f"{self.name}({self.k1.ToString()},{self.k2.ToString()})"

    def lower(self):
        o1, p1 = self.k1.lower()
        o2, p2 = self.k2.lower()
        return o1 + o2 + [self.op], p1 + p2


class SumKernel(_Binary):
    name, op = "SumKernel", OP_SUM


class ProductKernel(_Binary):
    name, op = "ProductKernel", OP_PRODUCT


_LEAVES = {c.name: c for c in (GaussianKernel, GaussianExpKernel, WhiteKernel, RationalQuadraticKernel, PeriodicKernel)}
_BINARY = {c.name: c for c in (SumKernel, ProductKernel)}


class KernelFactory:
    """String -> kernel tree, accepting what Kernel::ToString emits and the JSON `kernel_string`
    of the apps, e.g. "GaussianKernel(35, 30,)" (include/KernelFactory.h:83-178)."""

    @staticmethod
    def GetKernel(kernel_string):
        k, pos = KernelFactory._parse(kernel_string.strip(), 0)
        return k

    @staticmethod
    def _parse(s, pos):
        lp = s.find("(", pos)
        if lp < 0:
            raise ValueError("KernelFactory::GetKernel: failed to tokanize kernel name string")
        name = s[pos:lp].strip()
        pos = lp + 1
        if name in _BINARY:
            k1, pos = KernelFactory._parse(s, pos)
            while pos < len(s) and s[pos] in ", ":
                pos += 1
            k2, pos = KernelFactory._parse(s, pos)
            while pos < len(s) and s[pos] in ", ":
                pos += 1
            if pos >= len(s) or s[pos] != ")":
                raise ValueError("KernelFactory::GetKernel: failed to tokanize  sum kernel name string")
            return _BINARY[name](k1, k2), pos + 1
        if name not in _LEAVES:
            raise ValueError("KernelFactory::GetKernel: failed to load kernel.")
        rp = s.find(")", pos)
        if rp < 0:
            raise ValueError("KernelFactory::GetKernel: failed to load kernel.")
        toks = [t for t in s[pos:rp].split(",") if t.strip() != ""]
        cls = _LEAVES[name]
        if len(toks) != cls.nparams:
            raise ValueError(f"{name}::Load: wrong number of kernel parameters.")
        return cls(*[float(t) for t in toks]), rp + 1


def GetGeneralKernel(params):
    """13-parameter composite ((G + G*P) + RQ) + (G + W)  (include/KernelUtils.h:45-89)."""
    p = list(params)
    if len(p) != 13:
        raise ValueError("Wrong number of arguments.")
    k1 = GaussianKernel(p[1], p[0])
    k2 = ProductKernel(GaussianKernel(p[3], p[2]), PeriodicKernel(p[4], p[5], p[6]))
    k3 = RationalQuadraticKernel(p[7], p[8], p[9])
    k4 = SumKernel(GaussianKernel(p[11], p[10]), WhiteKernel(p[12]))
    return SumKernel(SumKernel(SumKernel(k1, k2), k3), k4)
