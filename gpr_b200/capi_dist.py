"""ctypes binding of include/gprb200_dist.h (libgprb200_dist.so): GP Learn over the GPUs of one node with the whole
schedule - ownership, streams, NCCL - in C++ under the C-ABI. One DistHandle per rank (process or thread); the
128-byte NCCL id made by rank 0 travels to the peers by the caller's own means (`exchange_id` uses an initialised
torch.distributed group of any backend, e.g. gloo; a file or socket does as well).

Reproduces, for N beyond one GPU, the result of GaussianProcess::ComputeRegressionVectors (reference
lib/GaussianProcess.cpp:642-672). The torch.distributed form of the same schedule is gpr_b200.dist.BlockCyclicGP."""
import ctypes as C
import os

import numpy as np

from .capi import GprError, LIB_PATH, load_library

DIST_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libgprb200_dist.so")
ID_BYTES = 128

_PROTOS = {
    "gprb200_dist_get_unique_id": (C.c_int, [C.c_void_p]),
    "gprb200_dist_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int]),
    "gprb200_dist_destroy": (C.c_int, [C.c_void_p]),
    "gprb200_dist_last_error": (C.c_char_p, [C.c_void_p]),
    "gprb200_dist_set_kernel": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int]),
    "gprb200_dist_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "gprb200_dist_learn": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int64, C.c_double, C.c_int64,
                                     C.c_void_p, C.POINTER(C.c_double)]),
    "gprb200_dist_solve": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "gprb200_dist_plan": (C.c_int, [C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_int64]),
    "gprb200_dist_phase_ms": (C.c_double, [C.c_void_p, C.c_int]),
    "gprb200_dist_counter": (C.c_double, [C.c_void_p, C.c_int]),
}
_lib = None


def load_dist_library():
    """dlopen libgprb200_dist.so (after libgprb200.so, which it is layered on). No fallback: raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    load_library()
    # libgprb200_dist.so needs "libnccl.so.2". torch ships its own (newer) one under the same soname: whichever is mapped
    # first serves both, and torch does not import against the older system library - so let torch map its copy first.
    try:
        import torch  # noqa: F401
    except ImportError:
        pass
    if not os.path.exists(DIST_LIB_PATH):
        raise ImportError(f"{DIST_LIB_PATH} not found: build it with `make -C gpr_b200/csrc` (needs nccl.h / libnccl)")
    lib = C.CDLL(DIST_LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in _PROTOS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def get_unique_id():
    """Rank 0: the 128 bytes every rank passes to DistHandle."""
    lib = load_dist_library()
    buf = C.create_string_buffer(ID_BYTES)
    rc = lib.gprb200_dist_get_unique_id(buf)
    if rc != 0:
        raise GprError(rc, lib.gprb200_dist_last_error(None).decode())
    return buf.raw


def exchange_id(group=None):
    """The NCCL id of rank 0 on every rank of an initialised torch.distributed group (any backend)."""
    import torch.distributed as dist
    box = [get_unique_id() if dist.get_rank(group) == 0 else None]
    dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    return box[0]


def plan(N, nb, Pr, Pc, rank, mirror_rows=True, mirror_cols=False):
    """Host-only layout of one rank (gprb200_dist_plan): dict with H, Wtot, bytes, flops, row / column offsets, stair."""
    lib = load_dist_library()
    summary = np.zeros(6)
    nblk = (N + nb - 1) // nb
    rows, cols = np.full(nblk, -2, dtype=np.int32), np.full(nblk, -2, dtype=np.int32)
    rc = lib.gprb200_dist_plan(N, nb, Pr, Pc, rank, int(mirror_rows), int(mirror_cols), summary.ctypes.data, rows.ctypes.data,
                               cols.ctypes.data, None, 0)
    if rc != 0:
        raise GprError(rc, lib.gprb200_dist_last_error(None).decode())
    stair = np.zeros(int(summary[5]), dtype=np.int32)
    lib.gprb200_dist_plan(N, nb, Pr, Pc, rank, int(mirror_rows), int(mirror_cols), summary.ctypes.data, None, None,
                          stair.ctypes.data, stair.size)
    return {"H": int(summary[0]), "Wtot": int(summary[1]), "bytes": summary[2], "flops": summary[3], "row_offset": rows,
            "col_offset": cols, "stair": stair}


def _arr(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class DistHandle:
    """One rank of a Pr x Pc grid. grid=None: the library's default (8 -> 4 x 2)."""

    def __init__(self, device, rank=0, world=1, unique_id=None, grid=None):
        self.lib = load_dist_library()
        self.h = C.c_void_p()
        pr, pc = (0, 0) if grid is None else (int(grid[0]), int(grid[1]))
        if world > 1 and (unique_id is None or len(unique_id) != ID_BYTES):
            raise ValueError("DistHandle: a %d-byte unique id from get_unique_id() is needed when world > 1" % ID_BYTES)
        idbuf = C.create_string_buffer(unique_id, ID_BYTES) if unique_id is not None else None
        rc = self.lib.gprb200_dist_create(C.byref(self.h), int(device), int(rank), int(world), idbuf, pr, pc)
        if rc != 0:
            raise GprError(rc, self.lib.gprb200_dist_last_error(None).decode())
        self.rank, self.world = rank, world
        self.N = self.M = 0

    def _ck(self, rc):
        if rc != 0:
            raise GprError(rc, self.lib.gprb200_dist_last_error(self.h).decode())

    def set_kernel(self, ops, params):
        o = np.ascontiguousarray(ops, dtype=np.int32)
        p = _arr(params)
        self._ck(self.lib.gprb200_dist_set_kernel(self.h, o.ctypes.data, len(o), p.ctypes.data, len(p)))

    def set_option(self, name, value):
        self._ck(self.lib.gprb200_dist_set_option(self.h, name.encode(), int(value)))

    def learn(self, X, Y, sigma, nb=1024, want_alpha=True):
        """Collective. Returns (alpha N x M or None, logdet). Raises GprError(code > 0) for a non-SPD matrix (INFO)."""
        X = _arr(X)
        self.N, d = X.shape
        alpha = None
        yp, ap, M = None, None, 0
        if Y is not None:
            Y = _arr(Y).reshape(self.N, -1)
            M = Y.shape[1]
            yp = Y.ctypes.data
            if want_alpha:
                alpha = np.empty((self.N, M))
                ap = alpha.ctypes.data
        self.M = M
        ld = C.c_double()
        self._ck(self.lib.gprb200_dist_learn(self.h, X.ctypes.data, self.N, d, yp, M, float(sigma), int(nb), ap, C.byref(ld)))
        return alpha, ld.value

    def solve(self, Y):
        Y = _arr(Y).reshape(self.N, -1)
        alpha = np.empty_like(Y)
        self._ck(self.lib.gprb200_dist_solve(self.h, Y.ctypes.data, Y.shape[1], alpha.ctypes.data))
        return alpha

    def phase_ms(self):
        return {k: self.lib.gprb200_dist_phase_ms(self.h, i) for i, k in enumerate(("gram", "factor", "solve"))}

    def counters(self):
        return {k: self.lib.gprb200_dist_counter(self.h, i) for i, k in enumerate(("collectives", "bytes", "gemm_flops", "launches", "stall_ms", "busy_ms", "update_flops", "Pr", "Pc"))}

    def close(self):
        if self.h:
            self.lib.gprb200_dist_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
