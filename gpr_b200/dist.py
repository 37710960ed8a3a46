"""Multi-GPU GP Learn/Predict: one process per GPU, torch.distributed (NCCL over
NVLink 5 / NVSwitch) for the plumbing, every numerical step through the C-ABI of include/gprb200.h.

The reference (agiger/GPR) is a single-process CPU library with no distributed code at all
(/root/reference: grep for mpi|nccl finds nothing), so everything here is new; what it must reproduce is the result of
GaussianProcess::ComputeRegressionVectors (reference lib/GaussianProcess.cpp:642-672), Predict (:55-61),
GetCredibleInterval (:103-114) and GaussianLogLikelihood::GetValueAndParameterDerivatives (include/Likelihood.h:231-285).

Partitionings:

* ``ShardedOutputsGP`` - the many-output model (BASELINE config C2, M = 786432 columns): the output columns of
  Y / alpha and the predicted field are independent, so each rank factors the (small) kernel matrix itself and owns
  M / g columns; test points for the credible interval are split T / g with L replicated. No data-path collective;
  optional all_gathers assemble a full prediction.
* ``BlockCyclicGP`` - N beyond one GPU's sweet spot (config C5, N = 131072, K = 137 GB): 2-D block-cyclic Cholesky
  over a Pr x Pc process grid, lower triangle only: block (i, j), i >= j, of nb x nb lives on rank
  (i mod Pr) * Pc + (j mod Pc). Right-looking with a one-panel look-ahead; per step k
      diagonal owner : L_kk = chol(A_kk) on the device, W = inv(L_kk)                      (single GPU, critical path)
      grid column k  : W broadcast down the column (Pr ranks), every rank solves ITS rows: X_i = A_ik W^T
      grid rows      : every rank of the column broadcasts its X rows along its grid row (Pc ranks)
      grid columns   : the blocks that are a column operand L_jk travel down each grid column (Pr roots)
      everyone       : A_ij -= X_i L_jk^T for the owned blocks, one tensor-core GEMM per owned block column.
  1 x g is the degenerate case (no column communicator, the layout of round 1). INFO stays on the device for the whole
  factorisation and is reduced once at the end: the host never waits inside the loop.
* ``ShardedLikelihood`` - the likelihood gradient 1/2 tr((alpha alpha^T - C) dK_p) with row blocks of A = alpha alpha^T - C
  and the matching dK rows per rank, partial sums all-reduced (P doubles).

Host logic (ownership, schedule, collectives) is independent of the device: ``ops`` is the object that executes
local block operations. ``DeviceOps`` (default) calls the CUDA library; the CPU tests in tests/test_dist_gloo.py
inject a numpy stand-in to run the same schedule under gloo with world_size 2, 3 and 4 (2 x 2).
"""
import bisect
import ctypes
import time

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


def default_grid(world):
    """Pr x Pc with Pr <= Pc, Pr the largest power of two with Pr^2 <= world that divides it (8 -> 2 x 4, 4 -> 2 x 2)."""
    pr = 1
    while (2 * pr) * (2 * pr) <= world and world % (2 * pr) == 0:
        pr *= 2
    return pr, world // pr


def block_owner(i, j, Pr, Pc):
    """Rank that owns block (i, j) of the Pr x Pc block-cyclic layout."""
    return (i % Pr) * Pc + (j % Pc)


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
        self._events = {}
        self.set_samples(X, sigma)

    def set_samples(self, X, sigma):
        """(Re)install the training inputs: N x d host array (the H2D copy of the end-to-end path) or device tensor."""
        Xd = X if isinstance(X, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(X, dtype=np.float64))
        with torch.cuda.stream(self.main_stream):
            Xd = Xd.to(self.device, non_blocking=True).contiguous()
            self.h._ck(self.h.lib.gprb200_set_train_dev(self.h.h, Xd.data_ptr(), Xd.shape[0], Xd.shape[1], None, 0,
                                                         ctypes.c_double(sigma)))
        self.h.N, self.h.d, self.h.M = Xd.shape[0], Xd.shape[1], 0
        self._X = Xd

    # -- memory ---------------------------------------------------------------------------------------------------
    def empty(self, rows, cols):
        return torch.empty((rows, cols), dtype=torch.float64, device=self.device)

    def zeros(self, rows, cols):
        return torch.zeros((rows, cols), dtype=torch.float64, device=self.device)

    def eye(self, n):
        return torch.eye(n, dtype=torch.float64, device=self.device)

    def from_host(self, a):
        if isinstance(a, torch.Tensor):
            return a.to(self.device, non_blocking=True)
        return torch.as_tensor(np.ascontiguousarray(a, dtype=np.float64)).to(self.device, non_blocking=True)

    def new_info(self):
        return torch.zeros(1, dtype=torch.int32, device=self.device)

    def read_info(self, info):
        return int(info.item())

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

    def potrf_block(self, P, invL, col0, info, which="panel"):
        """Cholesky of the top w x w block of P (rows x w) and the rows below solved against it, no host sync: the
        first failing leading minor (col0 + index, 1-based) is left in the device int `info`."""
        h = self._handle(which)
        h._ck(h.lib.gprb200_potrf_panel_async_dev(h.h, P.data_ptr(), P.stride(0), P.shape[0], P.shape[1], invL.data_ptr(),
                                                  int(col0), info.data_ptr()))

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
    """GP Learn for N beyond one GPU: 2-D block-cyclic Cholesky of K + sigma^2 I over the ranks of `group` (Pr x Pc
    grid, default 1 x world), then alpha = (K + sigma^2 I)^-1 Y by distributed forward / backward substitution (Y has
    few columns here; the many-column case is ShardedOutputsGP). Every rank ends with the full alpha and log det.

    Replaces, for large N, the reference's single-process InvertKernelMatrix + C * Y
    (lib/GaussianProcess.cpp:531-560, :661)."""

    def __init__(self, ops, N, nb=1024, grid=None, group=None, mirror_rows=False, mirror_cols=False):
        self.ops = ops
        # block-to-grid maps: plain cyclic, or mirrored (0 1 .. P-1 P-1 .. 1 0), which evens out the lower triangle's bias
        # towards high grid rows; libgprb200_dist.so mirrors its row map by default
        self.mirror_rows, self.mirror_cols = bool(mirror_rows), bool(mirror_cols)
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.N = int(N)
        self.nb = int(nb)
        if self.nb % 128 != 0:
            raise ValueError("BlockCyclicGP: nb must be a multiple of 128 (the Cholesky diagonal block)")
        self.Pr, self.Pc = (1, self.world) if grid is None else (int(grid[0]), int(grid[1]))
        if self.Pr * self.Pc != self.world:
            raise ValueError("BlockCyclicGP: grid %d x %d does not match world size %d" % (self.Pr, self.Pc, self.world))
        self.pr, self.pc = divmod(self.rank, self.Pc)
        self.nblk = (self.N + self.nb - 1) // self.nb
        self.hgt = [min(self.nb, self.N - i * self.nb) for i in range(self.nblk)]   # block heights = widths
        self.myrows = [i for i in range(self.nblk) if self._rowof(i) == self.pr]       # block rows stacked in every local panel
        self.roff = {}
        o = 0
        for i in self.myrows:
            self.roff[i] = o
            o += self.hgt[i]
        self.H = o
        self.mine = [j for j in range(self.nblk) if self._colof(j) == self.pc]         # owned block columns
        self._ranks = list(range(self.world))
        if group is not None and self.world > 1:
            self._ranks = dist.get_process_group_ranks(group)
        self.row_group, self.col_group = self._make_groups()
        self.cols = {}      # j -> stacked blocks (i, j), i in myrows, i >= j: (H - off(j)) x even(w_j); K, later L
        self.invl = {}      # j -> inverted 128 x 128 diagonal blocks of L_jj (diagonal owner only)
        self.factored = False
        self.logdet = None
        self.info = None
        self._bufs = None
        self.stats = {"bcast_bytes": 0, "gemm_flops": 0.0, "collectives": 0}

    # -- grid helpers -----------------------------------------------------------------------------------------------
    @staticmethod
    def _cyc(i, P, mirror):
        if not mirror or P == 1:
            return i % P
        t = i % (2 * P)
        return t if t < P else 2 * P - 1 - t

    def _rowof(self, i):
        return self._cyc(i, self.Pr, self.mirror_rows)

    def _colof(self, j):
        return self._cyc(j, self.Pc, self.mirror_cols)

    def _grank(self, r, c):
        return self._ranks[r * self.Pc + c]

    def _make_groups(self):
        """Row communicators (Pc ranks, same grid row) and column communicators (Pr ranks, same grid column). Every
        rank creates every group, in the same order (torch.distributed.new_group is collective)."""
        if self.world == 1:
            return None, None
        row = col = None
        if self.Pc > 1:
            if self.Pr == 1:
                row = self.group if self.group is not None else dist.group.WORLD
            else:
                for r in range(self.Pr):
                    g = dist.new_group([self._grank(r, c) for c in range(self.Pc)])
                    if r == self.pr:
                        row = g
        if self.Pr > 1:
            if self.Pc == 1:
                col = self.group if self.group is not None else dist.group.WORLD
            else:
                for c in range(self.Pc):
                    g = dist.new_group([self._grank(r, c) for r in range(self.Pr)])
                    if c == self.pc:
                        col = g
        return row, col

    def _first_row_at_or_after(self, j):
        """Offset (in rows of a stacked panel) of the first owned block row i >= j, and that i (or None)."""
        s = bisect.bisect_left(self.myrows, j)
        if s == len(self.myrows):
            return self.H, None
        i = self.myrows[s]
        return self.roff[i], i

    def _bcast(self, t, src, group):
        dist.broadcast(t, src=src, group=group)
        self.stats["bcast_bytes"] += t.numel() * 8
        self.stats["collectives"] += 1

    # -- Gram: every rank builds exactly the blocks it owns (no exchange) -------------------------------------------
    def _alloc(self):
        ops = self.ops
        with ops.on("main"):
            for j in self.mine:
                off, first = self._first_row_at_or_after(j)
                if first is None:
                    continue
                w = self.hgt[j]
                self.cols[j] = ops.empty(self.H - off, (w + 1) // 2 * 2)    # even leading dimension (16-byte rows)
                if first == j:
                    self.invl[j] = ops.inv_blocks(w)
            nb = self.nb
            self._bufs = {"row": [ops.empty(max(self.H, 1), nb) for _ in range(2)], "W": [ops.empty(nb, nb) for _ in range(2)],
                          "eye": ops.eye(nb), "col": None}
            if self.Pr > 1:
                cnt = [sum(self.hgt[j] for j in self.mine if self._rowof(j) == r2) for r2 in range(self.Pr)]
                self._bufs["col"] = [[ops.empty(max(c, 1), nb) for c in cnt] for _ in range(2)]
            self.info = ops.new_info()

    def build_gram(self):
        ops = self.ops
        if self._bufs is None:
            self._alloc()
        with ops.on("main"):
            self.info.zero_()
            for j in self.mine:
                if j not in self.cols:
                    continue
                off, _ = self._first_row_at_or_after(j)
                w = self.hgt[j]
                if self.Pr == 1:    # the rows of a block column are contiguous in the matrix: one launch
                    ops.gram_block(j * self.nb, j * self.nb, self.cols[j][:, :w])
                    continue
                for i in self.myrows:
                    if i >= j:
                        o = self.roff[i] - off
                        ops.gram_block(i * self.nb, j * self.nb, self.cols[j][o:o + self.hgt[i], :w])
        self.factored = False

    # -- factorisation ------------------------------------------------------------------------------------------------
    def _panel_factor(self, k):
        """Ranks of grid column k mod Pc, panel stream: L_kk, W = inv(L_kk) down the column, X_i = A_ik W^T for the
        owned rows below; X is left in the row-panel buffer k % 2 AND in cols[k] (the factor, needed by solve)."""
        ops, nb = self.ops, self.nb
        if self._colof(k) != self.pc:
            return
        w = self.hgt[k]
        diag = (self._rowof(k) == self.pr)
        if diag:
            blk = self.cols[k][:w, :w]
            ops.potrf_block(blk, self.invl[k], k * nb, self.info, "panel")
        if k == self.nblk - 1:
            return                      # nothing below the last diagonal block
        W = self._bufs["W"][k % 2]
        if diag:
            W.copy_(self._bufs["eye"])
            ops.trsm(blk, self.invl[k], W, False, "panel")
        if self.Pr > 1:
            self._bcast(W, self._grank(self._rowof(k), self.pc), self.col_group)
        o1, first = self._first_row_at_or_after(k + 1)
        if first is None:
            return
        off_k, _ = self._first_row_at_or_after(k)
        below = self.cols[k][o1 - off_k:, :w]
        X = self._bufs["row"][k % 2][:self.H - o1]
        ops.gemm("NT", below, W, X, 1.0, 0.0, "panel")
        self.stats["gemm_flops"] += 2.0 * below.shape[0] * w * w
        below.copy_(X)

    def _distribute(self, k):
        """All ranks, panel stream: row broadcast of the solved panel rows, then the column operands down the columns."""
        if k == self.nblk - 1:
            return
        o1, first = self._first_row_at_or_after(k + 1)
        X = self._bufs["row"][k % 2][:self.H - o1]
        if self.Pc > 1 and first is not None:
            self._bcast(X, self._grank(self.pr, self._colof(k)), self.row_group)
        if self.Pr > 1:
            for r2 in range(self.Pr):
                blocks = [j for j in self.mine if j > k and self._rowof(j) == r2]
                if not blocks:
                    continue
                buf = self._bufs["col"][k % 2][r2][:sum(self.hgt[j] for j in blocks)]
                if r2 == self.pr:       # this rank holds those block rows in X: pack them
                    pos = 0
                    for j in blocks:
                        buf[pos:pos + self.hgt[j]].copy_(X[self.roff[j] - o1:self.roff[j] - o1 + self.hgt[j]])
                        pos += self.hgt[j]
                self._bcast(buf, self._grank(r2, self.pc), self.col_group)

    def _col_operand(self, j, k):
        """L_jk (h_j x nb) as received at step k."""
        if self.Pr == 1:
            o1, _ = self._first_row_at_or_after(k + 1)
            return self._bufs["row"][k % 2][self.roff[j] - o1:self.roff[j] - o1 + self.hgt[j]]
        r2 = self._rowof(j)
        pos = sum(self.hgt[q] for q in self.mine if k < q < j and self._rowof(q) == r2)
        return self._bufs["col"][k % 2][r2][pos:pos + self.hgt[j]]

    def _update(self, j, k, which):
        """Owned blocks of block column j: A_ij -= X_i L_jk^T, i >= j (one GEMM, K = nb)."""
        if j not in self.cols:
            return
        o1, _ = self._first_row_at_or_after(k + 1)
        oj, first = self._first_row_at_or_after(j)
        if first is None:
            return
        rows = self._bufs["row"][k % 2][oj - o1:self.H - o1]
        B = self._col_operand(j, k)
        self.ops.gemm("NT", rows, B, self.cols[j][:, :self.hgt[j]], -1.0, 1.0, which)
        self.stats["gemm_flops"] += 2.0 * rows.shape[0] * B.shape[0] * self.nb

    def factor(self):
        """Right-looking 2-D block-cyclic Cholesky with one-panel look-ahead. Returns LAPACK-style info (0 = ok).
        No host synchronisation inside the loop: INFO is read once at the end."""
        ops, nblk = self.ops, self.nblk
        ops.record("gram", "main")
        with ops.on("panel"):
            ops.wait("gram", "panel")
            self._panel_factor(0)
        for k in range(nblk):
            nxt_mine = k + 1 < nblk and self._colof(k + 1) == self.pc
            with ops.on("panel"):
                ops.wait("main_step_%d" % (k - 2), "panel")    # the buffers k % 2 were last read at step k - 2
                self._distribute(k)
                ops.record("panel_%d" % k, "panel")
                if nxt_mine:
                    # look-ahead: block column k + 1 gets this panel first, off the main stream, and is factored at
                    # once so that its broadcasts overlap everybody's trailing update
                    ops.wait("main_step_%d" % (k - 1), "panel")    # its updates with panels < k ran on the main stream
                    self._update(k + 1, k, "panel")
                    self._panel_factor(k + 1)
            with ops.on("main"):
                ops.wait("panel_%d" % k, "main")
                for j in self.mine:
                    if j > k + 1 or (j == k + 1 and not nxt_mine):
                        self._update(j, k, "main")
                ops.record("main_step_%d" % k, "main")
            ops.forget("main_step_%d" % (k - 2))
            ops.forget("panel_%d" % (k - 1))
        # A rank that meets a non-SPD block keeps following the schedule (the failed block is left as it was, so the
        # remaining steps compute garbage but nobody hangs in a broadcast); all ranks agree on INFO here.
        ops.synchronize()
        info = ops.read_info(self.info)
        if self.world > 1:
            big = 2 ** 62
            flag = torch.tensor([info if info > 0 else big], dtype=torch.int64, device=self._flag_device())
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
            info = int(flag.item())
            info = 0 if info == big else info
        self.factored = info == 0
        if self.factored:
            part = 0.0
            for j in self.mine:
                if j in self.invl:
                    part += ops.logdet(self.cols[j][:self.hgt[j], :self.hgt[j]])
            t = torch.tensor([part], dtype=torch.float64, device=self._flag_device())
            if self.world > 1:
                dist.all_reduce(t, group=self.group)
            self.logdet = float(t.item())
        return info

    def _flag_device(self):
        return getattr(self.ops, "device", torch.device("cpu"))

    # -- alpha = (K + sigma^2 I)^-1 Y ------------------------------------------------------------------------------
    def solve(self, Y):
        """Y: N x M host array or device tensor (few columns), the same on every rank. Returns alpha (N x M tensor on
        the ops' device, the same on every rank).

        Forward L z = y, block row k: the ranks of grid row k hold the partial sums  sum_{j<k} L_kj z_j  of the blocks
        they own -> ONE small reduce (nb x M doubles) to the diagonal owner, which solves with L_kk and sends z_k down
        grid column k, whose ranks add L_ik z_k to their partial sums with one skinny GEMM over the stacked column.
        Backward L^T x = z mirrors it: the column's ranks form  sum_{i>k} L_ik^T x_i  with one skinny GEMM, reduce down
        the column, the owner solves with L_kk^T and sends x_k along grid row k. Per block step: at most two
        collectives of nb x M doubles and a handful of small launches, no host synchronisation, nothing of size
        (N - k nb) x M is moved; the cost is latency (2 nblk steps), independent of the number of ranks."""
        if not self.factored:
            raise RuntimeError("BlockCyclicGP.solve: factor() first")
        ops, nb, nblk = self.ops, self.nb, self.nblk
        with ops.on("main"):
            Yd = ops.from_host(Y)
            M = Yd.shape[1]
            ld = ops.pad_cols(M)
            y = ops.zeros(self.N, ld)
            y[:, :M].copy_(Yd)
            acc = ops.zeros(max(self.H, 1), ld)     # stacked over myrows: sum_j L_ij z_j (forward), then x_i (backward)
            zown = {}                               # z_k on the diagonal owner
            msg = [ops.zeros(nb, ld) for _ in range(2)]
            # ---- forward
            for k in range(nblk):
                w = self.hgt[k]
                in_row, in_col = (self._rowof(k) == self.pr), (self._colof(k) == self.pc)
                diag = in_row and in_col
                buf = msg[k % 2][:w]
                if in_row:
                    part = acc[self.roff[k]:self.roff[k] + w]
                    if self.Pc > 1:
                        dist.reduce(part, dst=self._grank(self.pr, self._colof(k)), group=self.row_group)
                        self.stats["collectives"] += 1
                    if diag:
                        torch.sub(y[k * nb:k * nb + w], part, out=buf)
                        ops.trsm(self.cols[k][:w, :w], self.invl[k], buf, False)
                        zown[k] = buf.clone()
                if not in_col or k == nblk - 1:
                    continue
                if self.Pr > 1:
                    self._bcast(msg[k % 2], self._grank(self._rowof(k), self.pc), self.col_group)
                o1, first = self._first_row_at_or_after(k + 1)
                if first is not None:
                    off_k, _ = self._first_row_at_or_after(k)
                    ops.gemm("NN", self.cols[k][o1 - off_k:, :w], buf, acc[o1:self.H], 1.0, 1.0)
            # ---- backward: acc now collects x_i for the owned rows
            for k in range(nblk - 1, -1, -1):
                w = self.hgt[k]
                in_row, in_col = (self._rowof(k) == self.pr), (self._colof(k) == self.pc)
                diag = in_row and in_col
                buf = msg[k % 2][:w]
                if in_col:
                    o1, first = self._first_row_at_or_after(k + 1)
                    if first is not None:
                        off_k, _ = self._first_row_at_or_after(k)
                        ops.gemm("TN", self.cols[k][o1 - off_k:, :w], acc[o1:self.H], buf, 1.0, 0.0)
                    else:
                        buf.zero_()
                    if self.Pr > 1:
                        dist.reduce(msg[k % 2], dst=self._grank(self._rowof(k), self.pc), group=self.col_group)
                        self.stats["collectives"] += 1
                    if diag:
                        torch.sub(zown[k], buf, out=buf)
                        ops.trsm(self.cols[k][:w, :w], self.invl[k], buf, True)
                if in_row:
                    if self.Pc > 1:
                        self._bcast(msg[k % 2], self._grank(self.pr, self._colof(k)), self.row_group)
                    acc[self.roff[k]:self.roff[k] + w].copy_(buf)
            # ---- every rank gets the whole alpha: each block row is contributed by grid column 0 of its grid row
            out = ops.zeros(self.N, ld)
            if self.pc == 0 or self.Pr == 1:    # with Pr = 1 every rank already holds every block row
                for i in self.myrows:
                    out[i * nb:i * nb + self.hgt[i]].copy_(acc[self.roff[i]:self.roff[i] + self.hgt[i]])
            if self.Pr > 1:
                dist.all_reduce(out, group=self.group)
                self.stats["collectives"] += 1
        return out[:, :M]

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
        # a rank may own no output column (M < g): it still holds the samples and factors K for its test points
        self.h.set_train(X, np.ascontiguousarray(Y) if e > b else None, sigma)

    def learn(self):
        if self.shard[1] > self.shard[0]:
            self.h.learn(0)
        else:
            self.h.factorize()

    def credible_interval(self, Xs, gather=True):
        """GetCredibleInterval (lib/GaussianProcess.cpp:103-114; called once per test point by
        apps/GaussianProcessPredict.cpp:185-193) for T test points: the points are independent, so rank r evaluates
        the slice column_shard(T, r, g) against its own replica of L (K is small and factored on every rank). No
        data-path collective; gather=True all_gathers the T scalars so every rank returns all of them."""
        Xs = np.ascontiguousarray(np.atleast_2d(Xs), dtype=np.float64)
        T = Xs.shape[0]
        b, e = column_shard(T, self.rank, self.world)
        local = self.h.credible_interval(Xs[b:e]) if e > b else np.empty(0)
        if not gather or self.world == 1:
            return local
        cuts = [column_shard(T, r, self.world) for r in range(self.world)]
        wmax = max(hi - lo for lo, hi in cuts)
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(self.group) == "nccl" else torch.device("cpu")
        pad = torch.zeros(wmax, dtype=torch.float64, device=dev)
        pad[:e - b] = torch.as_tensor(local).to(dev)
        parts = [torch.empty_like(pad) for _ in range(self.world)]
        dist.all_gather(parts, pad, group=self.group)
        return torch.cat([p[:hi - lo] for p, (lo, hi) in zip(parts, cuts)]).cpu().numpy()

    def predict(self, Xs, gather=False):
        """Mean of this rank's output columns for the test points Xs (T x d); gather=True returns T x M on every rank."""
        if self.shard[1] > self.shard[0]:
            mean = self.h.predict(Xs)
        else:
            mean = np.empty((np.atleast_2d(Xs).shape[0], 0))
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


class _DeviceView:
    """zero-copy torch view of a device buffer owned by the library (CUDA array interface)"""

    def __init__(self, ptr, shape, strides=None):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f8", "data": (int(ptr), False), "version": 2,
                                         "strides": strides}


class DeviceLikOps:
    """Local operations of ShardedLikelihood on one GPU: a C-ABI handle that holds the replicated model (samples,
    labels, factor) and the device building blocks (triangular solve, FP64 tensor-core GEMM, fused dK contraction)."""

    def __init__(self, device_index, ops, params, X, Y, sigma):
        from .capi import Handle
        self.device = torch.device("cuda", device_index)
        torch.cuda.set_device(self.device)
        self.h = Handle(device_index)
        self.h.set_stream(torch.cuda.current_stream(self.device).cuda_stream)
        self.h.set_kernel(ops, params)
        self.h.set_train(X, Y, sigma)
        self.N, self.M, self.P = self.h.N, self.h.M, self.h.P
        self._L = self._invL = self._alpha = None

    def refactor(self, params, mode):
        """New kernel parameters -> Gram + Cholesky + alpha on THIS rank (replicated); returns the M likelihood values."""
        self.h.set_kernel_params(params)
        value = self.h.loglik(mode, True, False, False)[0]
        lp, ld, ip = self.h.factor_dev()
        self._L = torch.as_tensor(_DeviceView(lp, (self.N, ld)), device=self.device)
        self._invL = ip
        ap, lda = self.h.alpha_dev()
        self._alpha = torch.as_tensor(_DeviceView(ap, (self.N, lda)), device=self.device)
        return value

    def alpha(self):
        return self._alpha[:, :self.M]

    def zeros(self, rows, cols):
        return torch.zeros((rows, cols), dtype=torch.float64, device=self.device)

    def trtri_cols(self, first, stride, cb, ncl):
        """Columns of W = inv(L) for the cyclic column blocks first, first + stride, ... (width cb), compact: N x ncl."""
        h, L = self.h, self._L
        W = torch.zeros((self.N, ncl), dtype=torch.float64, device=self.device)
        h._ck(h.lib.gprb200_trtri_cols_dev(h.h, L.data_ptr(), L.stride(0), self._invL, self.N, cb, first, stride,
                                           W.data_ptr(), W.stride(0)))
        return W

    def atb(self, A, B, out):
        """out = A^T B on the FP64 tensor path (A: K x m, B: K x n)."""
        h = self.h
        h._ck(h.lib.gprb200_gemm_dev(h.h, 2, out.shape[0], out.shape[1], A.shape[0], ctypes.c_double(1.0), A.data_ptr(),
                                     A.stride(0), B.data_ptr(), B.stride(0), ctypes.c_double(0.0), out.data_ptr(),
                                     out.stride(0), 0))

    def outer_minus(self, o, m, S, c0, c1, scale):
        """S[:, c0:c1] <- scale (alpha[o:o+m] alpha[c0:c1]^T - S[:, c0:c1]), summed over the output columns (C is
        subtracted ONCE, include/Likelihood.h:280-282)."""
        h, a = self.h, self._alpha
        h._ck(h.lib.gprb200_gemm_dev(h.h, 0, m, c1 - c0, self.M, ctypes.c_double(scale), a.data_ptr() + 8 * o * a.stride(0),
                                     a.stride(0), a.data_ptr() + 8 * c0 * a.stride(0), a.stride(0), ctypes.c_double(-scale),
                                     S.data_ptr() + 8 * c0, S.stride(0), 0))

    def contract_rows(self, o, m, ncols, S):
        return self.h.dgram_contract_rows_dev(o, m, ncols, S.data_ptr(), S.stride(0))

    def close(self):
        self.h.close()


class ShardedLikelihood:
    """GaussianLogLikelihood value and parameter gradient (include/Likelihood.h:231-285) over g GPUs for a model whose
    factor fits one GPU (N <= 32768; BASELINE config C4 is N = 16384).

    The gradient 1/2 tr((alpha alpha^T - C) dK_p) reads ALL of C = (K + sigma^2 I)^-1; forming C from the factor is
    2/3 of the iteration on one GPU (TRTRI + LAUUM, 2 N^3 / 3 flop). Here the factor is replicated (every rank factors
    K itself: N^3 / 3, cheaper than shipping L) and the inverse and the contraction are sharded by ROW BLOCKS of
    A = alpha alpha^T - C, block i (width m) on rank i mod g:
        1. W = inv(L): each rank inverts its own column blocks j (gprb200_trtri_cols_dev, long-K GEMMs)     ~2 N^3 / (3 g)
        2. one all_gather of the compact column blocks (NVLink), one permuted copy into a row-major W     8 N^2 bytes
        3. lower part of C[R_i, :] = W[o_i:, R_i]^T W[o_i:, :o_i + m], ONE tensor-core GEMM per row block     N^3 / (3 g)
        4. A[R_i, :] = alpha[R_i] alpha^T - C[R_i, :] (strictly-lower blocks weighted twice: A is symmetric), fused
           contraction with dK_p(x_i, x_j) generated on the fly
        5. all_reduce of the P partial sums.
    No rank ever holds all of C. ops: DeviceLikOps, or the numpy stand-in of the CPU tests."""

    def __init__(self, ops, block=512, group=None):
        self.ops = ops
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        if block % 128 != 0:
            raise ValueError("ShardedLikelihood: block must be a multiple of 128")
        self.m = int(block)
        N = ops.N
        self.nblk = (N + self.m - 1) // self.m
        self.width = [min(self.m, N - j * self.m) for j in range(self.nblk)]
        self.mine = list(range(self.rank, self.nblk, self.world))
        self.stats = {"gathered_bytes": 0}
        self.profile = False        # True: synchronise after each phase and accumulate phase times in stats
        self._W = None

    def value_and_gradient(self, params, mode=1):
        """Returns (value[M], grad[P]), identical on every rank."""
        ops, m, N, g = self.ops, self.m, self.ops.N, self.world
        t0 = time.perf_counter()
        value = ops.refactor(params, mode)
        self._tick("factor", t0)
        # 1 + 2: W = inv(L): every rank inverts its own cyclic column blocks (compact N x Q m), one all_gather, one
        # permuted copy into the row-major N x (Q g m) matrix
        t0 = time.perf_counter()
        Q = (self.nblk + g - 1) // g
        ncl = Q * m
        Wloc = ops.trtri_cols(self.rank, g, m, ncl)
        if g > 1:
            Wg = torch.empty((g, N, ncl), dtype=torch.float64, device=Wloc.device)
            if Wloc.is_cuda:
                dist.all_gather_into_tensor(Wg, Wloc, group=self.group)    # one NCCL all-gather, no staging copies
            else:
                dist.all_gather(list(Wg.unbind(0)), Wloc, group=self.group)
            self.stats["gathered_bytes"] += Wg.numel() * 8
        else:
            Wg = Wloc.unsqueeze(0)
        if self._W is None:
            self._W = ops.zeros(N, Q * g * m)
        W = self._W
        W.view(N, Q, g, m).copy_(Wg.view(g, N, Q, m).permute(1, 2, 0, 3))
        del Wg, Wloc
        self._tick("trtri_gather", t0)
        # 3 + 4: lower part of the owned row blocks of C = W^T W (S is symmetric: strictly-lower blocks count twice)
        t0 = time.perf_counter()
        part = np.zeros(ops.P)
        for i in self.mine:
            oi, wi = i * m, self.width[i]
            nc = oi + wi
            S = ops.zeros(wi, (nc + 1) // 2 * 2)
            ops.atb(W[oi:, oi:oi + wi], W[oi:, :nc], S[:, :nc])          # C[R_i, :nc]: k >= o_i, one GEMM
            if oi > 0:
                ops.outer_minus(oi, wi, S, 0, oi, 2.0)
            ops.outer_minus(oi, wi, S, oi, nc, 1.0)
            part += ops.contract_rows(oi, wi, nc, S)
        self._tick("lauum_contract", t0)
        t = torch.as_tensor(part)
        if g > 1:
            dev = getattr(ops, "device", torch.device("cpu"))
            t = t.to(dev)
            dist.all_reduce(t, group=self.group)
            t = t.cpu()
        return value, 0.5 * t.numpy()

    def _tick(self, name, t0):
        if self.profile:
            if hasattr(self.ops, "device") and self.ops.device.type == "cuda":
                torch.cuda.synchronize(self.ops.device)
            self.stats[name + "_ms"] = self.stats.get(name + "_ms", 0.0) + (time.perf_counter() - t0) * 1e3
