"""CPU tests (gloo) of the two sharded section-8(e) rows in gpr_b200/dist.py:
  * ShardedLikelihood - row blocks of A = alpha alpha^T - C per rank, partial dK contractions all-reduced - against the
    oracle's GaussianLogLikelihood gradient (include/Likelihood.h:231-285: P full N x N products and their traces);
  * the helper logic of ShardedOutputsGP.credible_interval (test points split T / g)."""
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


def _oracle_value_grad(kname, X, Y, params=None):
    from oracle import gpr_oracle as orc
    k = orc.parse_kernel(KERNELS[kname])
    if params is not None:
        k = orc.set_params(k, list(params))
    o = orc.OracleGP(k, X, Y, 1.0)
    o.learn()
    return o.loglik_value(clamp=False), o.loglik_grad()


def test_single_process_sharded_likelihood_equals_oracle():
    from gpr_b200 import dist as gd
    from tests.dist_cpu_ops import CpuLikOps
    from oracle import gpr_oracle as orc
    X, Y = make_problem(300, 3, 2, seed=4)
    for kname, block in (("rq", 128), ("sum3", 256)):
        ops = CpuLikOps(KERNELS[kname], X, Y, 1.0)
        sl = gd.ShardedLikelihood(ops, block=block)
        p = orc.get_params(orc.parse_kernel(KERNELS[kname]))
        v, g = sl.value_and_gradient(p)
        v0, g0 = _oracle_value_grad(kname, X, Y)
        assert np.max(np.abs(g - g0)) <= 1e-8 * np.max(np.abs(g0))
        assert np.max(np.abs(v - v0)) <= 1e-8 * np.max(np.abs(v0))


def _worker(rank, world, port, N, block, kname, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from gpr_b200 import dist as gd
        from tests.dist_cpu_ops import CpuLikOps
        from oracle import gpr_oracle as orc
        X, Y = make_problem(N, 3, 2, seed=6)
        ops = CpuLikOps(KERNELS[kname], X, Y, 1.0)
        sl = gd.ShardedLikelihood(ops, block=block)
        assert sl.mine == list(range(rank, sl.nblk, world))
        p = np.array(orc.get_params(orc.parse_kernel(KERNELS[kname]))) * 1.05
        v, g = sl.value_and_gradient(p)
        v0, g0 = _oracle_value_grad(kname, X, Y, p)
        t = torch.as_tensor(g.copy())
        dist.broadcast(t, src=0)
        res = {"gerr": float(np.max(np.abs(g - g0)) / np.max(np.abs(g0))), "verr": float(np.max(np.abs(v - v0)) / np.max(np.abs(v0))),
               "same": bool(np.array_equal(t.numpy(), g)), "gathered": sl.stats["gathered_bytes"]}
        if rank == 0:
            torch.save(res, out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,N,block,kname", [(2, 700, 128, "rq"), (3, 530, 128, "general13")])
def test_sharded_likelihood_gradient_matches_oracle(tmp_path, world, N, block, kname):
    out = str(tmp_path / "res.pt")
    mp.spawn(_worker, args=(world, _free_port(), N, block, kname, out), nprocs=world, join=True)
    res = torch.load(out)
    assert res["gerr"] <= 1e-8 and res["verr"] <= 1e-8, res      # BASELINE.json tolerance on the likelihood gradient
    assert res["same"] and res["gathered"] > 0


def test_test_point_shards_cover_every_point_once():
    from gpr_b200 import dist as gd
    for world in (1, 2, 3, 8):
        for T in (1, 5, 256, 1000):
            cuts = [gd.column_shard(T, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == T and all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
