"""[torchrun --nproc-per-node G] tools/sharded_lik_probe.py [N] [block]: ShardedLikelihood phase times (C4 shape) against
the single-GPU gprb200_loglik on rank 0."""
import json, os, sys, time
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpr_b200
from gpr_b200 import dist as gd

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
block = int(sys.argv[2]) if len(sys.argv) > 2 else 512
rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rng = np.random.default_rng(4)
X = rng.uniform(-1, 1, (N, 8))
Y = np.sin(3.0 * X[:, :1]) + 0.1 * rng.standard_normal((N, 1))
ops, params = gpr_b200.KernelFactory.GetKernel("RationalQuadraticKernel(1,1.5,2,)").lower()
lo = gd.DeviceLikOps(local, ops, params, X, Y, 1.0)
sl = gd.ShardedLikelihood(lo, block=block)
p = np.asarray(params)
sl.value_and_gradient(p)
sl.value_and_gradient(p * 1.01)
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
t0 = time.perf_counter()
for it in range(3):
    v, g = sl.value_and_gradient(p * (1.0 + 0.01 * it))
torch.cuda.synchronize()
ms = (time.perf_counter() - t0) / 3 * 1e3
sl.profile = True
sl.stats = {"gathered_bytes": 0}
v, g = sl.value_and_gradient(p * 1.02)
out = {"N": N, "block": block, "world": world, "iteration_ms": ms, "phases": {k: round(x, 2) for k, x in sl.stats.items() if k.endswith("_ms")}}
if rank == 0:
    h = lo.h
    h.set_kernel_params(p * 1.02)
    h.loglik(1, True, True, False)
    t0 = time.perf_counter()
    v1, g1, _ = h.loglik(1, True, True, False)
    out["single_gpu_ms"] = (time.perf_counter() - t0) * 1e3
    out["grad_vs_1gpu"] = float(np.max(np.abs(g - g1)) / np.max(np.abs(g1)))
    print(json.dumps(out), flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
