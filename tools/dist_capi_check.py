"""torchrun --nproc-per-node G tools/dist_capi_check.py [N] [nb] [PrxPc] [repeats]: the C++/NCCL block-cyclic Learn
(libgprb200_dist.so) on G GPUs - torch.distributed (gloo) only carries the 128-byte NCCL id and the max over ranks of the
device timings - against the single-GPU library on the same inputs when N fits one GPU, else against sampled residual
rows evaluated with the library's own kernel vector. Prints one JSON line on rank 0 and writes
gpurun_out/dist_capi_<G>gpu_<N>.json."""
import json, os, sys, time
import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpr_b200
from gpr_b200 import capi_dist

KS = "SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))"


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    nb = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
    grid = tuple(int(v) for v in sys.argv[3].split("x")) if len(sys.argv) > 3 and "x" in sys.argv[3] else None
    reps = int(sys.argv[4]) if len(sys.argv) > 4 else 2
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    uid = None
    if world > 1:
        dist.init_process_group("gloo")
        uid = capi_dist.exchange_id()
    rng = np.random.default_rng(1)
    X = rng.uniform(-1, 1, (N, 8))
    Y = (np.sin(3 * X[:, :1]) + 0.1 * rng.standard_normal((N, 1)))
    ops_, params = gpr_b200.KernelFactory.GetKernel(KS).lower()
    d = capi_dist.DistHandle(local, rank, world, uid, grid)
    d.set_kernel(ops_, params)
    if os.environ.get("DIST_PROFILE"):
        d.set_option("profile", 1)
    for kv in os.environ.get("DIST_OPTS", "").split(","):
        if "=" in kv:
            d.set_option(kv.split("=")[0], int(kv.split("=")[1]))
    if os.environ.get("DIST_BAD"):
        # a duplicated sample and no noise: every rank must come back with the SAME LAPACK-style INFO and nobody may hang
        Xb = X.copy()
        dup = N // 2 + 3
        Xb[dup] = Xb[7]
        opsb, paramsb = gpr_b200.KernelFactory.GetKernel("GaussianKernel(2,1,)").lower()
        d.set_kernel(opsb, paramsb)
        code = 0
        try:
            d.learn(Xb, Y, 0.0, nb=nb)
        except gpr_b200.GprError as e:
            code = e.code
        codes = [None] * world
        if world > 1:
            dist.all_gather_object(codes, code)
        else:
            codes = [code]
        if rank == 0:
            print(json.dumps({"non_spd_info_per_rank": codes, "expected": dup + 1, "ok": all(c == dup + 1 for c in codes)}), flush=True)
        d.set_kernel(ops_, params)
    out = {}
    for it in range(reps):
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        alpha, logdet = d.learn(X, Y, 1.0, nb=nb)
        wall = time.perf_counter() - t0
        ph = d.phase_ms()
        tm = torch.tensor([ph["gram"], ph["factor"], ph["solve"], wall * 1e3, d.counters()["stall_ms"]], dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        c = d.counters()
        stall = tm[4].item()
        out = {"N": N, "nb": nb, "world": world, "grid": "%dx%d" % (c["Pr"], c["Pc"]), "impl": "libgprb200_dist.so (NCCL from C++)",
               "gram_s": tm[0].item() * 1e-3, "factor_s": tm[1].item() * 1e-3, "solve_s": tm[2].item() * 1e-3,
               "learn_wall_s_host_to_host": tm[3].item() * 1e-3, "factor_TFs_total": N ** 3 / 3.0 / (tm[1].item() * 1e-3) * 1e-12,
               "stall_ms_max_over_ranks": stall, "busy_ms": c["busy_ms"], "gemm_TFs_inside_updates_this_rank": (c["update_flops"] / c["busy_ms"] * 1e-9 if c["busy_ms"] > 0 else None),
               "collectives": c["collectives"], "exchanged_GB": c["bytes"] * 1e-9, "launches": c["launches"], "logdet": logdet}
    if world > 1:
        chk = torch.from_numpy(alpha.copy())
        dist.broadcast(chk, src=0)
        out_same = float((chk - torch.from_numpy(alpha)).abs().max())
        t = torch.tensor([out_same], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out["alpha_max_difference_between_ranks"] = t.item()
    if rank == 0:
        h = gpr_b200.Handle(local)
        h.set_kernel(ops_, params)
        if N <= 32768:
            h.set_train(X, Y, 1.0); h.learn(0)
            ref = h.alpha()
            out["alpha_vs_single_gpu"] = float(np.max(np.abs(alpha - ref)) / np.max(np.abs(ref)))
            out["logdet_vs_single_gpu"] = abs(logdet - h.logdet()) / abs(h.logdet())
        else:
            rows = rng.choice(N, 256, replace=False)
            h.set_train(X, Y, 1.0)
            Kx = h.kernel_vector(X[rows])
            res = Kx @ alpha + 1.0 * alpha[rows] - Y[rows]
            out["rel_residual_rows"] = float(np.max(np.abs(res)) / np.max(np.abs(Y)))
        h.close()
        print(json.dumps(out), flush=True)
        os.makedirs("gpurun_out", exist_ok=True)
        json.dump(out, open("gpurun_out/dist_capi_%dgpu_%d.json" % (world, N), "w"), indent=1)
    d.close()
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
