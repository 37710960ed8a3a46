"""TEST INFRASTRUCTURE: a numpy stand-in for gpr_b200.dist.DeviceOps so that the multi-GPU host logic (ownership,
2-D schedule, look-ahead ordering, collectives) runs under gloo on CPU with world_size 2, 3 and 4. Block operations are
plain numpy/scipy on CPU tensors; the kernel matrix comes from the oracle. Never imported by the product."""
import contextlib

import numpy as np
import scipy.linalg as sla
import torch

from oracle import gpr_oracle as orc


class CpuOps:
    device = torch.device("cpu")

    def __init__(self, kernel_string, X, sigma):
        self.k = orc.parse_kernel(kernel_string)
        self.X = np.ascontiguousarray(X, dtype=np.float64)
        self.sigma = float(sigma)
        self.log = []  # (stream, op) trace, lets tests assert the schedule

    def set_samples(self, X, sigma):
        self.X = np.ascontiguousarray(X, dtype=np.float64)
        self.sigma = float(sigma)

    def empty(self, rows, cols):
        return torch.full((rows, cols), float("nan"), dtype=torch.float64)

    def zeros(self, rows, cols):
        return torch.zeros((rows, cols), dtype=torch.float64)

    def eye(self, n):
        return torch.eye(n, dtype=torch.float64)

    def from_host(self, a):
        if isinstance(a, torch.Tensor):
            return a
        return torch.as_tensor(np.ascontiguousarray(a, dtype=np.float64))

    def new_info(self):
        return torch.zeros(1, dtype=torch.int32)

    def read_info(self, info):
        return int(info.item())

    @contextlib.contextmanager
    def on(self, which):
        yield

    def record(self, name, which):
        self.log.append((which, "record", name))

    def wait(self, name, which):
        self.log.append((which, "wait", name))

    def forget(self, name):
        pass

    def gram_block(self, row0, col0, out):
        r, c = out.shape
        K = orc.kernel_matrix(self.k, self.X[row0:row0 + r], self.X[col0:col0 + c])
        for i in range(max(row0, col0), min(row0 + r, col0 + c)):
            K[i - row0, i - col0] += self.sigma ** 2
        out.copy_(torch.as_tensor(K))

    def potrf_block(self, P, invL, col0, info, which="panel"):
        self.log.append((which, "potrf", tuple(P.shape)))
        w = P.shape[1]
        A = P[:w].numpy().copy()
        S = np.tril(A) + np.tril(A, -1).T
        try:
            L = np.linalg.cholesky(S)
        except np.linalg.LinAlgError:
            if int(info.item()) == 0:           # LAPACK INFO: order of the first leading minor that is not SPD
                bad = next(j for j in range(1, w + 1) if np.linalg.eigvalsh(S[:j, :j]).min() <= 0)
                info.fill_(col0 + bad)
            return
        P[:w] = torch.as_tensor(np.tril(L) + np.triu(A, 1))
        if P.shape[0] > w:
            P[w:] = torch.as_tensor(sla.solve_triangular(L, P[w:].numpy().T, lower=True).T)

    def gemm(self, form, A, B, Cm, alpha, beta, which="main"):
        self.log.append((which, "gemm" + form, tuple(Cm.shape)))
        a, b = A.numpy(), B.numpy()
        prod = a @ b.T if form == "NT" else (a @ b if form == "NN" else a.T @ b)
        Cm.copy_(torch.as_tensor(alpha * prod + (beta * Cm.numpy() if beta != 0 else 0.0)))

    def trsm(self, L, invL, B, transposed, which="main"):
        Lm = np.tril(L.numpy())
        B.copy_(torch.as_tensor(sla.solve_triangular(Lm, B.numpy(), lower=True, trans="T" if transposed else "N")))

    def logdet(self, L):
        return float(2.0 * np.sum(np.log(np.diag(L.numpy()))))

    def inv_blocks(self, w):
        return torch.zeros(1, 1, dtype=torch.float64)

    def pad_cols(self, M):
        return (M + 1) // 2 * 2

    def synchronize(self):
        pass

    def close(self):
        pass


class CpuLikOps:
    """numpy stand-in for gpr_b200.dist.DeviceLikOps (ShardedLikelihood's local operations)."""
    device = torch.device("cpu")

    def __init__(self, kernel_string, X, Y, sigma):
        self.k0 = orc.parse_kernel(kernel_string)
        self.X = np.ascontiguousarray(X, dtype=np.float64)
        self.Y = np.ascontiguousarray(Y, dtype=np.float64)
        self.sigma = float(sigma)
        self.N, self.M = self.Y.shape
        self.P = orc.num_params(self.k0)

    def refactor(self, params, mode):
        self.k = orc.set_params(self.k0, list(params))
        K = orc.kernel_matrix(self.k, self.X, self.X) + self.sigma ** 2 * np.eye(self.N)
        self.L = np.linalg.cholesky(K)
        self.a = sla.cho_solve((self.L, True), self.Y)
        logdet = 2.0 * np.sum(np.log(np.diag(self.L)))
        q = np.sum(self.Y * self.a, axis=0)
        return -0.5 * q - 0.5 * logdet - 0.5 * self.N * np.log(2 * np.pi)

    def alpha(self):
        return torch.as_tensor(self.a)

    def zeros(self, rows, cols):
        return torch.zeros((rows, cols), dtype=torch.float64)

    def trtri_cols(self, first, stride, cb, ncl):
        Wf = sla.solve_triangular(self.L, np.eye(self.N), lower=True)
        W = np.zeros((self.N, ncl))
        q = 0
        for j in range(first, (self.N + cb - 1) // cb, stride):
            w = min(cb, self.N - j * cb)
            W[:, q * cb:q * cb + w] = Wf[:, j * cb:j * cb + w]
            q += 1
        return torch.as_tensor(W)

    def atb(self, A, B, out):
        out.copy_(torch.as_tensor(A.numpy().T @ B.numpy()))

    def outer_minus(self, o, m, S, c0, c1, scale):
        S[:, c0:c1] = scale * (torch.as_tensor(self.a[o:o + m] @ self.a[c0:c1].T) - S[:, c0:c1])

    def contract_rows(self, o, m, ncols, S):
        D = orc.kernel_derivative(self.k, self.X[o:o + m], self.X[:ncols])      # (P, m, ncols)
        return np.einsum("pij,ij->p", D, S.numpy()[:, :ncols])

    def close(self):
        pass
