import time, numpy as np, torch, gpr_b200
from bench import c3_inputs, C3_KERNEL
N=32768
X,Y=c3_inputs(N,0)
h=gpr_b200.Handle(0)
ops,params=gpr_b200.KernelFactory.GetKernel(C3_KERNEL).lower()
h.set_kernel(ops,params)
out=np.empty((N,1))
for it in range(4):
    t0=time.perf_counter(); h.set_train(X,Y,1.0); t1=time.perf_counter(); h.learn(0); h.synchronize(); t2=time.perf_counter()
    h.lib.gprb200_get_alpha(h.h, out.ctypes.data); t3=time.perf_counter()
    print("set_train %.1f ms  learn %.1f ms  get_alpha %.1f ms  phases %s"%((t1-t0)*1e3,(t2-t1)*1e3,(t3-t2)*1e3,{k:round(v,1) for k,v in h.phase_ms().items() if v}))
