"""Multi-GPU GP Learn/Predict: one process per GPU, torch.distributed (NCCL over
NVLink 5 / NVSwitch) for the plumbing, every numerical step through the C-ABI of include/gprb200.h.

The reference (agiger/GPR) is a single-process CPU library with no distributed code at all
(/root/reference: grep for mpi|nccl finds nothing), so everything here is new; what it must reproduce is the result of
GaussianProcess::ComputeRegressionVectors (reference lib/GaussianProcess.cpp:642-672) and Predict (:55-61).

Two partitionings:

* ``ShardedOutputsGP`` - the many-output model (BASELINE config C2, M = 786432 columns): the output columns of
  Y / alpha and the predicted field are independent, so each rank factors the (small) kernel matrix itself and owns
  M / g columns. No data-path collective; an optional all_gather assembles a full prediction.
* ``BlockCyclicGP`` - N beyond one GPU's sweet spot (config C5, N = 131072, K = 137 GB): block-cyclic Cholesky over
  a 1 x g process grid (block column j of width nb lives on rank j mod g, lower triangle only), right-looking with
  a one-panel look-ahead: the owner factors block column k, broadcasts it (NCCL), every rank updates the block
  columns it owns with one tensor-core GEMM each. On NVSwitch every peer is equidistant, so the 1 x g grid's
  larger broadcast volume (4 N^2 bytes per GPU in total, 68.7 GB at N = 131072, ~0.2 s at measured NCCL rates
  against ~3.4 s of FP64 work per GPU at g = 8) is cheaper than a Pr x Pc grid's second, transposed broadcast and
  distributed panel solve; the panel factorisation and its broadcast run on a side stream under the trailing update.

Host logic (ownership, schedule, collectives) is independent of the device: ``ops`` is the object that executes
local block operations. ``DeviceOps`` (default) calls the CUDA library; the CPU tests in tests/test_dist_gloo.py
inject a numpy stand-in to run the same schedule under gloo with world_size 2.
"""
import ctypes

import numpy as np
import torch
import torch.distributed as dist


def owned_blocks(nblk, rank, world):
    """Block columns owned by `rank` in the 1 x world block-cyclic layout."""
    return list(range(rank, nblk, world))


def owner_of(j, world):
    return j % world


def column_shard(M, rank, world):
    """[begin, end) of the output columns owned by `rank`: contiguous, sizes differ by at most one."""
    base, extra = divmod(M, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


class DeviceOps:
    """Local block operations on CUDA tensors through libgprb200.so. Two C-ABI handles: one bound to the main
    stream (model state: kernel + samples, trailing updates), one to the panel stream (panel factorisation)."""

    def __init__(self, device_index, ops, params, X, sigma):
        from .capi import Handle
        self.device = torch.device("cuda", device_index)
        torch.cuda.set_device(self.device)
        self.main_stream = torch.cuda.Stream(self.device)
        self.panel_stream = torch.cuda.Stream(self.device, priority=-1)
        self.h = Handle(device_index)
        self.h.set_stream(self.main_stream.cuda_stream)
        self.hp = Handle(device_index)
        self.hp.set_stream(self.panel_stream.cuda_stream)
        self.h.set_kernel(ops, params)
        Xd = torch.as_tensor(np.ascontiguousarray(X, dtype=np.float64)).to(self.device)
        self.h._ck(self.h.lib.gprb200_set_train_dev(self.h.h, Xd.data_ptr(), Xd.shape[0], Xd.shape[1], None, 0,
                                                     ctypes.c_double(sigma)))
        self.h.N, self.h.d, self.h.M = Xd.shape[0], Xd.shape[1], 0
        self.h.synchronize()
        self._events = {}

    # -- memory ---------------------------------------------------------------------------------------------------
    def empty(self, rows, cols):
        return torch.empty((rows, cols), dtype=torch.float64, device=self.device)

    def zeros(self, rows, cols):
        return torch.zeros((rows, cols), dtype=torch.float64, device=self.device)

    def from_host(self, a):
        return torch.as_tensor(np.ascontiguousarray(a, dtype=np.float64)).to(self.device)

    # -- streams / events ---------------------------------------------------------------------------------------------
    def on(self, which):
        return torch.cuda.stream(self.panel_stream if which == "panel" else self.main_stream)

    def record(self, name, which):
        ev = torch.cuda.Event()
        ev.record(self.panel_stream if which == "panel" else self.main_stream)
        self._events[name] = ev

    def wait(self, name, which):
        ev = self._events.get(name)
        if ev is not None:
            (self.panel_stream if which == "panel" else self.main_stream).wait_event(ev)

    def forget(self, name):
        self._events.pop(name, None)

    def _handle(self, which):
        return self.hp if which == "panel" else self.h

    # -- numerics ---------------------------------------------------------------------------------------------------
    def gram_block(self, row0, col0, out):
        h = self.h
        h._ck(h.lib.gprb200_gram_block_dev(h.h, row0, out.shape[0], col0, out.shape[1], out.data_ptr(), out.stride(0)))

    def potrf_panel(self, P, invL, which="panel"):
        h = self._handle(which)
        info = ctypes.c_int(0)
        h._ck(h.lib.gprb200_potrf_panel_dev(h.h, P.data_ptr(), P.stride(0), P.shape[0], P.shape[1], invL.data_ptr(),
                                            ctypes.byref(info)))
        return info.value

    def gemm(self, form, A, B, Cm, alpha, beta, which="main"):
        """form 'NT': C = alpha A B^T + beta C; 'NN': A B; 'TN': A^T B."""
        h = self._handle(which)
        code = {"NT": 0, "NN": 1, "TN": 2}[form]
        M, N = Cm.shape
        K = A.shape[1] if form != "TN" else A.shape[0]
        h._ck(h.lib.gprb200_gemm_dev(h.h, code, M, N, K, ctypes.c_double(alpha), A.data_ptr(), A.stride(0), B.data_ptr(),
                                     B.stride(0), ctypes.c_double(beta), Cm.data_ptr(), Cm.stride(0), 0))

    def trsm(self, L, invL, B, transposed, which="main"):
        h = self._handle(which)
        h._ck(h.lib.gprb200_trsm_dev(h.h, L.data_ptr(), L.stride(0), invL.data_ptr(), L.shape[1], B.data_ptr(),
                                     B.stride(0), B.shape[1], 1 if transposed else 0))

    def logdet(self, L):
        v = ctypes.c_double()
        self.h._ck(self.h.lib.gprb200_logdet_dev(self.h.h, L.data_ptr(), L.stride(0), L.shape[1], ctypes.byref(v)))
        return v.value

    def inv_blocks(self, w):
        return self.empty(((w + 127) // 128) * 128, 128)

    def pad_cols(self, M):
        """Right-hand sides are stored with an even, 16-byte aligned leading dimension."""
        return (M + 1) // 2 * 2

    def synchronize(self):
        torch.cuda.synchronize(self.device)

    def close(self):
        self.h.close()
        self.hp.close()


class BlockCyclicGP:
    """GP Learn for N beyond one GPU: block-cyclic Cholesky of K + sigma^2 I over the ranks of `group`, then
    alpha = (K + sigma^2 I)^-1 Y by distributed forward / backward substitution (Y has few columns here; the
    many-column case is ShardedOutputsGP). Every rank ends with the full alpha and log det.

    Replaces, for large N, the reference's single-process InvertKernelMatrix + C * Y
    (lib/GaussianProcess.cpp:531-560, :661)."""

    def __init__(self, ops, N, nb=1024, group=None):
        self.ops = ops
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.N = int(N)
        self.nb = int(nb)
        if self.nb % 128 != 0:
            raise ValueError("BlockCyclicGP: nb must be a multiple of 128 (the Cholesky diagonal block)")
        self.nblk = (self.N + self.nb - 1) // self.nb
        self.mine = owned_blocks(self.nblk, self.rank, self.world)
        self.cols = {}      # j -> (N - j nb) x w_j tensor: block column j of K, later of L, rows j nb .. N
        self.invl = {}      # j -> inverted 128 x 128 diagonal blocks of L_jj
        self.factored = False
        self.logdet = None
        self.stats = {"bcast_bytes": 0, "gemm_flops": 0.0}

    # -- helpers ---------------------------------------------------------------------------------------------------
    def _width(self, j):
        return min(self.nb, self.N - j * self.nb)

    def _bcast(self, t, src):
        if self.world > 1:
            dist.broadcast(t, src=dist.get_global_rank(self.group, src) if self.group is not None else src,
                           group=self.group)
        self.stats["bcast_bytes"] += t.numel() * 8

    # -- Gram: every rank builds exactly the block columns it owns (no exchange) ------------------------------------
    def build_gram(self):
        ops = self.ops
        with ops.on("main"):
            for j in self.mine:
                r0 = j * self.nb
                blk = ops.empty(self.N - r0, self._width(j))
                ops.gram_block(r0, r0, blk)
                self.cols[j] = blk
                self.invl[j] = ops.inv_blocks(self._width(j))
        self.factored = False

    # -- factorisation ------------------------------------------------------------------------------------------------
    def _update(self, j, k, Pk, which):
        """A[j nb:, block j] -= P_k[rows of block column j] P_k[rows of block j]^T  (one GEMM, K = w_k)."""
        off = (j - k) * self.nb
        rows = Pk[off:]
        top = Pk[off:off + self._width(j)]
        self.ops.gemm("NT", rows, top, self.cols[j], -1.0, 1.0, which)
        self.stats["gemm_flops"] += 2.0 * rows.shape[0] * top.shape[0] * Pk.shape[1]

    def factor(self):
        """Right-looking block-cyclic Cholesky with one-panel look-ahead. Returns LAPACK-style info (0 = ok)."""
        ops, nb, nblk, world, rank = self.ops, self.nb, self.nblk, self.world, self.rank
        recv = [None, None]
        if world > 1:
            with ops.on("main"):
                recv = [ops.empty(self.N, nb), ops.empty(self.N, nb)]
        info = 0
        ops.record("gram", "main")
        with ops.on("panel"):
            ops.wait("gram", "panel")
            if owner_of(0, world) == rank:
                info = ops.potrf_panel(self.cols[0], self.invl[0], "panel")
        for k in range(nblk):
            own_k = owner_of(k, world) == rank
            rows_k = self.N - k * nb
            with ops.on("panel"):
                if own_k:
                    Pk = self.cols[k]
                else:
                    ops.wait("main_step_%d" % (k - 2), "panel")  # receive buffer k % 2 was last read at step k - 2
                    Pk = recv[k % 2][:rows_k, :self._width(k)]
                self._bcast(Pk, owner_of(k, world))
                ops.record("panel_%d" % k, "panel")
                # look-ahead: block column k + 1 gets this panel first, off the main stream
                lookahead = k + 1 < nblk and owner_of(k + 1, world) == rank
                if lookahead:
                    ops.wait("main_step_%d" % (k - 1), "panel")  # its updates with panels < k ran on the main stream
                    self._update(k + 1, k, Pk, "panel")
            with ops.on("main"):
                ops.wait("panel_%d" % k, "main")
                for j in self.mine:
                    if j > k + 1:
                        self._update(j, k, Pk, "main")
                ops.record("main_step_%d" % k, "main")
            if lookahead:
                # factored at once (the call returns when the panel is done; the main stream already holds this
                # step's trailing update, so the GPU stays busy while the host waits here)
                with ops.on("panel"):
                    bad = ops.potrf_panel(self.cols[k + 1], self.invl[k + 1], "panel")
                if bad and not info:
                    info = bad + (k + 1) * nb
            ops.forget("main_step_%d" % (k - 2))
            ops.forget("panel_%d" % (k - 1))
        # A rank that meets a non-SPD panel keeps following the schedule (the failed block is left as it was, so the
        # remaining steps compute garbage but nobody hangs in a broadcast); all ranks agree on INFO here.
        ops.synchronize()
        if world > 1:
            flag = torch.tensor([info], dtype=torch.int64, device=self._flag_device())
            dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=self.group)
            info = int(flag.item())
        self.factored = info == 0
        if self.factored:
            part = 0.0
            for j in self.mine:
                part += ops.logdet(self.cols[j][:self._width(j)])
            t = torch.tensor([part], dtype=torch.float64, device=self._flag_device())
            if world > 1:
                dist.all_reduce(t, group=self.group)
            self.logdet = float(t.item())
        return info

    def _flag_device(self):
        return getattr(self.ops, "device", torch.device("cpu"))

    # -- alpha = (K + sigma^2 I)^-1 Y ------------------------------------------------------------------------------
    def solve(self, Y):
        """Y: N x M host array or tensor (few columns), the same on every rank. Returns alpha (N x M tensor on the ops'
        device, the same on every rank). Forward: the owner of block column k solves its diagonal block and forms
        the update of the rows below, broadcasts both ((N - k nb) x M values); backward: the owner reduces against
        the already known tail and broadcasts x_k (nb x M values)."""
        if not self.factored:
            raise RuntimeError("BlockCyclicGP.solve: factor() first")
        ops, nb, nblk, world, rank = self.ops, self.nb, self.nblk, self.world, self.rank
        Yh = Y.detach().cpu().numpy() if isinstance(Y, torch.Tensor) else np.asarray(Y, dtype=np.float64)
        M = Yh.shape[1]
        ld = ops.pad_cols(M)
        with ops.on("main"):
            b = ops.zeros(self.N, ld)
            b[:, :M] = ops.from_host(Yh)
            msg = ops.empty(self.N, ld)
            # forward: L z = y
            for k in range(nblk):
                r0, w = k * nb, self._width(k)
                rows = self.N - r0
                buf = msg[:rows]
                if owner_of(k, world) == rank:
                    Lk = self.cols[k]
                    buf[:w] = b[r0:r0 + w]
                    ops.trsm(Lk[:w], self.invl[k], buf[:w], False)
                    if rows > w:
                        ops.gemm("NN", Lk[w:], buf[:w], buf[w:], 1.0, 0.0)
                self._bcast(buf, owner_of(k, world))
                b[r0:r0 + w] = buf[:w]
                if rows > w:
                    b[r0 + w:] -= buf[w:]
            # backward: L^T x = z
            for k in range(nblk - 1, -1, -1):
                r0, w = k * nb, self._width(k)
                rows = self.N - r0
                buf = msg[:w]
                if owner_of(k, world) == rank:
                    Lk = self.cols[k]
                    buf.copy_(b[r0:r0 + w])
                    if rows > w:
                        ops.gemm("TN", Lk[w:], b[r0 + w:], buf, -1.0, 1.0)
                    ops.trsm(Lk[:w], self.invl[k], buf, True)
                self._bcast(buf, owner_of(k, world))
                b[r0:r0 + w] = buf
        return b[:, :M]

    def learn(self, Y):
        """Gram + factorisation + solve; returns (alpha, logdet, info)."""
        self.build_gram()
        info = self.factor()
        if info:
            return None, None, info
        return self.solve(Y), self.logdet, 0


class ShardedOutputsGP:
    """Many-output GP (BASELINE config C2): output columns are independent, so rank r owns columns
    column_shard(M, r, g) of Y, alpha and every prediction. K is small (N = 4096: 134 MB, 2.3e10 flop) and is
    factored on every rank, which is cheaper than broadcasting the factor. No collective on the data path;
    predict(..., gather=True) all_gathers the column slices when one rank needs the whole field."""

    def __init__(self, handle, group=None):
        self.h = handle
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.M = 0
        self.shard = (0, 0)

    def set_train(self, X, Y_full_or_shard, sigma, M_total=None):
        """Y may be the full N x M matrix (this rank keeps its column slice) or already this rank's slice (then
        pass M_total)."""
        Y = np.asarray(Y_full_or_shard)
        if M_total is None:
            self.M = Y.shape[1]
            b, e = column_shard(self.M, self.rank, self.world)
            Y = Y[:, b:e]
        else:
            self.M = int(M_total)
            b, e = column_shard(self.M, self.rank, self.world)
            if Y.shape[1] != e - b:
                raise ValueError("ShardedOutputsGP: slice has %d columns, rank %d owns %d" % (Y.shape[1], self.rank, e - b))
        self.shard = (b, e)
        self.h.set_train(X, np.ascontiguousarray(Y), sigma)

    def learn(self):
        self.h.learn(0)

    def predict(self, Xs, gather=False):
        """Mean of this rank's output columns for the test points Xs (T x d); gather=True returns T x M on every rank."""
        mean = self.h.predict(Xs)
        if not gather or self.world == 1:
            return mean
        t = torch.as_tensor(np.ascontiguousarray(mean))
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(self.group) == "nccl" else t.device
        widths = [column_shard(self.M, r, self.world) for r in range(self.world)]
        wmax = max(e - b for b, e in widths)
        pad = torch.zeros(t.shape[0], wmax, dtype=torch.float64, device=dev)
        pad[:, :t.shape[1]] = t.to(dev)
        parts = [torch.empty_like(pad) for _ in range(self.world)]
        dist.all_gather(parts, pad, group=self.group)
        return torch.cat([p[:, :e - b] for p, (b, e) in zip(parts, widths)], dim=1).cpu().numpy()
