"""One C4 log-likelihood + gradient evaluation (N = 16384, RQ) and the gpPredict loop body (Predict + credible
interval for ONE test point, N = 16384) for an ncu launch list."""
import sys
import numpy as np
sys.path.insert(0, ".")
import gpr_b200
N, d = 16384, 8
rng = np.random.default_rng(1)
X = rng.uniform(-1, 1, (N, d))
Y = np.sin(3.0 * X[:, :1]) + 0.1 * rng.standard_normal((N, 1))
h = gpr_b200.Handle(0)
ops, params = gpr_b200.KernelFactory.GetKernel("RationalQuadraticKernel(1,1.5,2,)").lower()
h.set_kernel(ops, params)
h.set_train(X, Y, 1.0)
print(h.loglik(1, True, True, False))
h.learn(0)
xs = rng.uniform(-1, 1, (1, d))
print("MARK predict")
for _ in range(3):
    print(h.predict(xs, want_ci=True))
print(h.predict_derivative(xs[0])[0])
print(h.posterior_cov(xs, xs))
h.close()
