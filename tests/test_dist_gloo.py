"""CPU tests of the multi-GPU host logic (gpr_b200/dist.py) under gloo, world_size 2 and 3: the block-cyclic
Cholesky + distributed solve schedule reproduces the single-process oracle, ownership/shard helpers are exact
partitions, and the column-sharded prediction gather reassembles the full field."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tests.synth import KERNELS, make_problem  # noqa: E402


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, N, nb, M, kname, out, grid=None, mirror=(False, False)):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from gpr_b200 import dist as gd
        from tests.dist_cpu_ops import CpuOps
        from oracle import gpr_oracle as orc
        X, Y = make_problem(N, 3, M, seed=11)
        ops = CpuOps(KERNELS[kname], X, 1.0)
        gp = gd.BlockCyclicGP(ops, N, nb=nb, grid=grid, mirror_rows=mirror[0], mirror_cols=mirror[1])
        if any(mirror):    # mirrored maps: every lower block still has exactly one owner (checked over all ranks below)
            owned = torch.zeros(gp.nblk, gp.nblk, dtype=torch.int64)
            for j in gp.mine:
                for i in gp.myrows:
                    owned[i, j] = 1
            dist.all_reduce(owned)
            assert bool((owned == 1).all())
            if mirror[0] and grid[0] > 1:   # 0 1 .. P-1 P-1 .. 1 0
                assert [gp._rowof(i) for i in range(2 * grid[0])] == list(range(grid[0])) + list(range(grid[0] - 1, -1, -1))
        elif grid is None:
            assert gp.mine == list(range(rank, gp.nblk, world))
        else:   # every block (i, j), i >= j, has exactly one owner and this rank holds exactly its own
            pr, pc = divmod(rank, grid[1])
            assert gp.mine == [j for j in range(gp.nblk) if j % grid[1] == pc]
            assert gp.myrows == [i for i in range(gp.nblk) if i % grid[0] == pr]
            assert all(gd.block_owner(i, j, *grid) == rank for j in gp.mine for i in gp.myrows if i >= j)
        alpha, logdet, info = gp.learn(Y)
        assert info == 0
        o = orc.OracleGP(KERNELS[kname], X, Y, 1.0)
        o.learn()
        err = float(np.max(np.abs(alpha.numpy() - o.alpha)) / np.max(np.abs(o.alpha)))
        Kfull = o.kernel_matrix() + np.eye(N)
        ld_ref = float(np.linalg.slogdet(Kfull)[1])
        # schedule properties: panels are factored on the panel stream only, trailing updates on main
        potrf_streams = {s for s, op, *_ in ops.log if op == "potrf"}
        main_ops = {op for s, op, *_ in ops.log if s == "main"}
        res = {"err": err, "logdet_err": abs(logdet - ld_ref) / abs(ld_ref), "potrf_streams": sorted(potrf_streams),
               "main_has_potrf": "potrf" in main_ops, "bcast_bytes": gp.stats["bcast_bytes"],
               "collectives": gp.stats["collectives"]}
        # a second Learn on the same object (buffers are reused) gives the same answer
        alpha2, _, info2 = gp.learn(Y)
        res["relearn_same"] = bool(info2 == 0 and torch.equal(alpha2, alpha))
        # every rank must hold the same alpha
        t = alpha.clone()
        dist.broadcast(t, src=0)
        res["same_alpha"] = bool(torch.equal(t, alpha))
        if rank == 0:
            torch.save(res, out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,N,nb,M,kname", [(2, 700, 128, 3, "sum3"), (3, 1000, 256, 1, "rq")])
def test_block_cyclic_learn_matches_oracle(tmp_path, world, N, nb, M, kname):
    out = str(tmp_path / "res.pt")
    mp.spawn(_worker, args=(world, _free_port(), N, nb, M, kname, out), nprocs=world, join=True)
    res = torch.load(out)
    assert res["err"] <= 1e-9, res          # alpha vs the oracle's LU-inverse path (tolerance of BASELINE.json)
    assert res["logdet_err"] <= 1e-10, res
    assert res["same_alpha"] and res["relearn_same"]
    assert res["potrf_streams"] == ["panel"] and not res["main_has_potrf"]
    nblk = (N + nb - 1) // nb
    expect = sum((N - (k + 1) * nb) * nb * 8 for k in range(nblk - 1))   # every panel below its diagonal block, once
    assert res["bcast_bytes"] >= expect  # factor broadcasts (+ the solve's small messages)


@pytest.mark.parametrize("world,grid,N,nb,M,kname", [
    (2, (2, 1), 700, 128, 2, "sum3"),          # column communicator only
    (4, (2, 2), 1001, 128, 3, "rq"),           # the 2 x 2 schedule, ragged AND odd last block (1001 = 7 * 128 + 105)
])
def test_two_dimensional_grid_matches_oracle(tmp_path, world, grid, N, nb, M, kname):
    out = str(tmp_path / "res.pt")
    mp.spawn(_worker, args=(world, _free_port(), N, nb, M, kname, out, grid), nprocs=world, join=True)
    res = torch.load(out)
    assert res["err"] <= 1e-9, res
    assert res["logdet_err"] <= 1e-10, res
    assert res["same_alpha"] and res["relearn_same"]
    assert res["potrf_streams"] == ["panel"] and not res["main_has_potrf"]


@pytest.mark.parametrize("world,grid,mirror,N,nb,M,kname", [
    (2, (2, 1), (True, False), 1100, 128, 2, "sum3"),      # the default map of libgprb200_dist.so on two GPUs
    (4, (2, 2), (True, True), 1001, 128, 3, "prod"),       # both maps mirrored (consecutive block rows / panels on the same
                                                           # grid row / column), ragged and odd last block
])
def test_mirrored_block_maps_match_oracle(tmp_path, world, grid, mirror, N, nb, M, kname):
    out = str(tmp_path / "res.pt")
    mp.spawn(_worker, args=(world, _free_port(), N, nb, M, kname, out, grid, mirror), nprocs=world, join=True)
    res = torch.load(out)
    assert res["err"] <= 1e-9, res
    assert res["logdet_err"] <= 1e-10, res
    assert res["same_alpha"] and res["relearn_same"]
    assert res["potrf_streams"] == ["panel"] and not res["main_has_potrf"]


def test_single_process_block_cyclic_equals_oracle():
    from gpr_b200 import dist as gd
    from tests.dist_cpu_ops import CpuOps
    from oracle import gpr_oracle as orc
    X, Y = make_problem(300, 2, 2, seed=5)
    gp = gd.BlockCyclicGP(CpuOps(KERNELS["prod"], X, 1.0), 300, nb=128)
    alpha, logdet, info = gp.learn(Y)
    o = orc.OracleGP(KERNELS["prod"], X, Y, 1.0)
    o.learn()
    assert info == 0 and np.max(np.abs(alpha.numpy() - o.alpha)) <= 1e-9 * np.max(np.abs(o.alpha))


def test_partitions_are_exact():
    from gpr_b200 import dist as gd
    for world in (1, 2, 3, 4, 8):
        for M in (1, 7, 64, 786432, 786433):
            cuts = [gd.column_shard(M, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == M
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in cuts]
            assert max(sizes) - min(sizes) <= 1
        for nblk in (1, 5, 128):
            owned = sorted(j for r in range(world) for j in gd.owned_blocks(nblk, r, world))
            assert owned == list(range(nblk))
            assert all(gd.owner_of(j, world) == r for r in range(world) for j in gd.owned_blocks(nblk, r, world))


def test_non_spd_matrix_reports_info_on_every_rank():
    from gpr_b200 import dist as gd
    from tests.dist_cpu_ops import CpuOps
    X, Y = make_problem(256, 2, 1, seed=2)
    ops = CpuOps(KERNELS["gaussian"], X, 1.0)
    gp = gd.BlockCyclicGP(ops, 256, nb=128)
    gp.build_gram()
    gp.cols[1][:128] -= 50.0 * torch.eye(128, dtype=torch.float64)  # make the second diagonal block indefinite
    assert gp.factor() > 128
    with pytest.raises(RuntimeError):
        gp.solve(Y)


def _bad_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from gpr_b200 import dist as gd
        from tests.dist_cpu_ops import CpuOps
        X, Y = make_problem(512, 2, 1, seed=2)
        gp = gd.BlockCyclicGP(CpuOps(KERNELS["gaussian"], X, 1.0), 512, nb=128, grid=(2, 2))
        gp.build_gram()
        if 2 in gp.invl:    # the owner of diagonal block 2 makes it indefinite; nobody else knows until the end
            gp.cols[2][:128, :128] -= 50.0 * torch.eye(128, dtype=torch.float64)
        info = gp.factor()  # must return on every rank (no hang in a broadcast) with the same INFO
        torch.save(info, out + ".%d" % rank)
    finally:
        dist.destroy_process_group()


def test_non_spd_block_on_a_2x2_grid_gives_every_rank_the_same_info(tmp_path):
    out = str(tmp_path / "info")
    mp.spawn(_bad_worker, args=(4, _free_port(), out), nprocs=4, join=True)
    infos = [torch.load(out + ".%d" % r) for r in range(4)]
    assert len(set(infos)) == 1 and 256 < infos[0] <= 384, infos
