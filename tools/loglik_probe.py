"""C4: GaussianLogLikelihood value + gradient at N=16384, RationalQuadratic kernel, per-iteration time and phases."""
import json, time, numpy as np, torch, gpr_b200
N, d = 16384, 8
rng = np.random.default_rng(4)
X = rng.uniform(-1, 1, (N, d)); Y = (np.sin(3 * X[:, :1]) + 0.1 * rng.standard_normal((N, 1)))
h = gpr_b200.Handle(0)
ops, params = gpr_b200.KernelFactory.GetKernel("RationalQuadraticKernel(1,1.5,2,)").lower()
h.set_kernel(ops, params); h.set_train(X, Y, 1.0)
for it in range(4):
    p = np.array(params) * (1 + 0.01 * it)
    h.set_kernel_params(p)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    v, g, _ = h.loglik(1, True, True, False)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    print(json.dumps({"iter": it, "wall_ms": round((t1 - t0) * 1e3, 1), "value": float(v[0]), "grad": [float(x) for x in g],
                      "phases": {k: round(x, 1) for k, x in h.phase_ms().items() if x}}), flush=True)
