"""PCA kernels in the order an ncu capture wants them (tools/ncu_r01c.sh): stored basis -> reconstruct x3, project x3,
then a small fit (center, split-K Gram, Jacobi steps, rotation GEMM). D = 3 * 64^3, nF = 32."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import gpr_b200  # noqa: E402
from gpr_b200.pca import PCA  # noqa: E402

D, n, nF = 3 * 64 ** 3, 64, 32
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev)
g.manual_seed(1)
X = torch.randn(D, n, generator=g, device=dev, dtype=torch.float64) * torch.logspace(0, -2, n, device=dev, dtype=torch.float64)
h = gpr_b200.Handle(0)
U = torch.linalg.qr(X).Q.cpu().numpy()
p = PCA.from_arrays(np.zeros(D), np.geomspace(1, 1e-2, n), U, nF, handle=h)
W = torch.randn(nF, 1, generator=g, device=dev, dtype=torch.float64)
out = torch.empty(D, device=dev, dtype=torch.float64)
F = torch.empty(nF, device=dev, dtype=torch.float64)
col = X[:, 0].contiguous()
for _ in range(3):
    p.reconstruct_dev(W.data_ptr(), 1, 0, 1, out.data_ptr(), 1)
for _ in range(3):
    p.project_dev(col.data_ptr(), 1, 1, nF, F.data_ptr(), 1)
torch.cuda.synchronize()
q = PCA.fit_dev(X.data_ptr(), n, D, n, nF, h)
print("fit_ms", q.stat("fit_ms"), "sweeps", q.stat("sweeps"), "sigma", q.GetEigenvalues()[:3])
q.close()
p.close()
h.close()
