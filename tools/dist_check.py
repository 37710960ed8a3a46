"""torchrun --nproc-per-node G tools/dist_check.py [N] [nb] [PrxPc]: block-cyclic Learn on G GPUs vs the single-GPU library
(same inputs; results must agree), with device timings (max over ranks)."""
import json, os, sys, time
import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpr_b200
from gpr_b200 import dist as gd

KS = "SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))"


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    nb = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
    grid = tuple(int(v) for v in sys.argv[3].split("x")) if len(sys.argv) > 3 else None
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rng = np.random.default_rng(1)
    X = rng.uniform(-1, 1, (N, 8))
    Y = (np.sin(3 * X[:, :1]) + 0.1 * rng.standard_normal((N, 1)))
    ops_, params = gpr_b200.KernelFactory.GetKernel(KS).lower()
    ops = gd.DeviceOps(local, ops_, params, X, 1.0)
    gp = gd.BlockCyclicGP(ops, N, nb=nb, grid=grid)
    out = {}
    for it in range(2):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        gp.build_gram(); torch.cuda.synchronize(); t1 = time.perf_counter()
        info = gp.factor(); torch.cuda.synchronize(); t2 = time.perf_counter()
        alpha = gp.solve(Y); torch.cuda.synchronize(); t3 = time.perf_counter()
        tm = torch.tensor([t1 - t0, t2 - t1, t3 - t2], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        out = {"N": N, "nb": nb, "world": world, "grid": "%dx%d" % (gp.Pr, gp.Pc), "info": info,
               "collectives": gp.stats["collectives"], "bcast_GB": gp.stats["bcast_bytes"] * 1e-9, "gram_s": tm[0].item(), "factor_s": tm[1].item(),
               "solve_s": tm[2].item(), "factor_TFs_total": N ** 3 / 3.0 / tm[1].item() * 1e-12, "logdet": gp.logdet}
    if rank == 0 and N <= 32768:
        h = gpr_b200.Handle(local)
        h.set_kernel(ops_, params); h.set_train(X, Y, 1.0); h.learn(0)
        ref = h.alpha()
        out["alpha_vs_single_gpu"] = float(np.max(np.abs(alpha.cpu().numpy() - ref)) / np.max(np.abs(ref)))
        out["logdet_vs_single_gpu"] = abs(gp.logdet - h.logdet()) / abs(h.logdet())
        h.close()
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
