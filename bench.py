#!/usr/bin/env python
"""bench.py — the reference's headline benchmark on B200 (BASELINE.json):
    "GP Learn s & Cholesky FP64 TFLOP/s at N=32k; predictions/sec at 1/2/4/8 GPUs"

A step = one GP Learn pass (Gram build + fp64 Cholesky + solve for alpha) of BASELINE config C3
(N = 32768, d = 8, Sum(Sum(Gaussian, Periodic), White) kernel, sigma_n = 1, fp64) through the C-ABI of
include/gprb200.h. `value` is useful fp64 TFLOP/s (N^3/3 + 2 N^2 M per model) of the whole job with the
inputs already resident in HBM; `e2e` is the same metric through the host-pointer entry points
(pinned host buffers, H2D + D2H inside the timed region). With N GPUs every rank learns an independent
model (the likelihood grid search of the reference, tests/GaussianLikelihoodTest.cpp:51-145, evaluates
independent hyper-parameter points): weak scaling, no data-path collective. The secondary "predict"
object reports predictions/s of BASELINE config C2 (N = 4096, d = 30, M = 786432 output columns,
alpha column-sharded over the ranks, one test point per call as apps/GaussianProcessPredict.cpp:185-193).

    python bench.py --gpus N --steps K --warmup W          (this implementation)
    python bench.py --impl reference ...                   (the reference's own CPU code, oracle/_ref)
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

C3_KERNEL = "SumKernel(SumKernel(GaussianKernel(2,1,),PeriodicKernel(1,2,1,)),WhiteKernel(0.1,))"
C3_N, C3_D, C3_M, C3_SIGMA = 32768, 8, 1, 1.0
C3_WORKLOAD = "C3: GP Learn (Gram + Cholesky + solve), Sum(Sum(Gaussian,Periodic),White), d=8, M=1, fp64"
C2_KERNEL = "GaussianKernel(8.215838362577491,1,)"  # sigma = 1.5 sqrt(30)
C2_N, C2_D, C2_M, C2_SIGMA = 4096, 30, 786432, 1.0
SEED = 0x6770725F62323030 % (2 ** 63)
HBM_FALLBACK_GBS = 6650.0  # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent
FP64_PEAK_TFLOPS = 37.1  # measured DMMA.8x8x4 issue rate on this pool's B200 (profiles/r01_fp64_probe.txt)
FP64_CUBLAS_TFLOPS = 36.2  # measured cuBLAS DGEMM 16384^3 (profiles/r01_dgemm_ceiling.json)


def learn_flops(N, M):
    return N ** 3 / 3.0 + 2.0 * N * N * M


def c3_inputs(N, seed):
    """X ~ U[-1,1]^8, y = smooth periodic + trend + noise."""
    rng = np.random.default_rng(SEED + seed)
    X = rng.uniform(-1, 1, (N, C3_D))
    y = np.sin(3.0 * X[:, 0]) + 0.5 * np.cos(2.0 * X[:, 1] + X[:, 2]) + 0.3 * X[:, 3] + 0.1 * rng.standard_normal(N)
    return X, y.reshape(N, C3_M)


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(",") for r in open(self.f.name) if r.strip()]
        os.unlink(self.f.name)
        sm, reasons, smax = [], set(), None
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for r in rows:
            try:
                sm.append(float(r[1]))
                smax = float(r[2])
                for i, nm in enumerate(names):
                    if r[5 + i].strip().lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                continue
        if sm:
            # "under load" = samples above half of max clock (idle samples at the edges are dropped)
            load = [v for v in sm if smax and v > 0.5 * smax] or sm
            out.update(sm_mhz=float(np.median(load)), sm_max_mhz=smax, reasons=sorted(reasons), samples=len(sm))
        return out


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


# ---------------------------------------------------------------------------------------------------
# reference arm: the reference's own CPU implementation (oracle/_ref), all host threads
# ---------------------------------------------------------------------------------------------------
def time_reference_learn(N, threads, seed=0):
    from oracle import ref
    ref.set_threads(threads)
    X, Y = c3_inputs(N, seed)
    t0 = time.perf_counter()
    g = ref.RefGP(C3_KERNEL, C3_SIGMA).add_samples(X, Y)
    g.initialize()  # Gram (OpenMP) + dgetrf/dgetri + C*Y : lib/GaussianProcess.cpp:642-672
    a = g.alpha()
    dt = time.perf_counter() - t0
    return dt, float(np.abs(a).max())


def run_reference(args):
    """The reference's own CPU code on the product arm's configuration: C3 at N = 32768 (same kernel, d, M, sigma_n).
    One Learn pass is ~2-3 minutes on the GPU box's host cores (2 N^3 flop in dgetrf/dgetri), so ONE pass is timed -
    the work is deterministic and identical from pass to pass - and the line says so: `steps` is what was timed,
    `steps_requested` what the command line asked for. --ref-n overrides N (the CPU tests use a small one)."""
    rank, _, world = dist_env()
    if rank != 0:
        return 0
    from oracle import ref
    if not ref.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libgpr_ref.so was not built"}))
        return 0
    threads = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(threads)
    N = args.ref_n or C3_N
    passes = args.ref_passes
    time_reference_learn(min(N, 1024), threads)   # library start-up (thread pools), not the workload
    t = []
    for s in range(passes):
        dt, _ = time_reference_learn(N, threads, seed=s)
        t.append(dt)
    total = sum(t)
    value = learn_flops(N, C3_M) * passes / total * 1e-12
    sample = (f"{passes} full pass(es) of the C3 workload at N={N} (the product arm's N={C3_N}): Gram (OpenMP) + LAPACK "
              f"dgetrf/dgetri inverse + C*Y, the reference's own lib/GaussianProcess.cpp via oracle/_ref; useful flops "
              f"N^3/3 + 2N^2M counted, {total / passes:.1f} s per pass; requested --steps {args.steps} --warmup "
              f"{args.warmup} are not repeated (minutes per pass, identical deterministic work)")
    line = {
        "impl": "reference", "metric": "gp_learn_fp64_tflops", "value": value, "unit": "TFLOP/s",
        "n_gpus": args.gpus, "steps": passes, "warmup": 0, "steps_requested": args.steps, "warmup_requested": args.warmup,
        "ms_per_step": total / passes * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": C3_WORKLOAD, "N": N, "useful_flops_per_model": learn_flops(N, C3_M)},
        "cpu_baseline": {"value": value, "unit": "TFLOP/s", "cores": threads, "kind": "reference", "sample": sample,
                         "seconds_per_pass": total / passes},
        "e2e": {"value": value, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if world > 1 or args.gpus > 1:
        line["config"]["note"] = ("the product arm at N > 1 GPUs runs C5 (N=131072, K = 137 GB, 4.5e15 flop by LU): it "
                                  "does not fit the host; the reference is timed on C3, the largest config it can run")
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------------
# this implementation (CUDA through the C-ABI)
# ---------------------------------------------------------------------------------------------------
def run_gpu(args):
    import torch
    import gpr_b200
    rank, local_rank, world = dist_env()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist_mod.init_process_group("nccl", device_id=dev)
        dist = dist_mod

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    N, d, M = args.n, C3_D, C3_M
    kernel = gpr_b200.KernelFactory.GetKernel(C3_KERNEL)
    ops, params = kernel.lower()
    # every rank learns its own model: same shape, different data and hyper-parameter point
    params = list(params)
    params[0] *= 1.0 + 0.01 * rank
    Xh, Yh = c3_inputs(N, seed=rank)
    Xp = torch.from_numpy(Xh).pin_memory()
    Yp = torch.from_numpy(Yh).pin_memory()
    Xd, Yd = Xp.to(dev), Yp.to(dev)
    alpha_host = torch.empty(N, M, dtype=torch.float64).pin_memory()

    h = gpr_b200.Handle(local_rank)
    stream = torch.cuda.current_stream(dev)
    h.set_stream(stream.cuda_stream)
    h.set_kernel(ops, params)
    h.set_option("profile", 1)
    h.set_train_dev(Xd.data_ptr(), N, d, Yd.data_ptr(), M, C3_SIGMA)

    # ---- device-resident Learn: value ---------------------------------------------------------------
    for _ in range(args.warmup):
        h.learn(0)
    barrier()
    sampler = ClockSampler(local_rank)
    l0 = h.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    phases = {"gram": 0.0, "potrf": 0.0, "solve": 0.0}
    tb_ms = tb_flops = tb_launches = 0.0
    e0.record(stream)
    for _ in range(args.steps):
        h.learn(0)
        ms = h.phase_ms()
        for k in phases:
            phases[k] += ms[k]
        tb_ms += h.get_stat("tb_ms")
        tb_flops += h.get_stat("tb_flops")
        tb_launches += h.get_stat("tb_launches")
    e1.record(stream)
    barrier()
    launches = h.launch_count() - l0
    clocks = sampler.stop()
    elapsed_ms = max_over_ranks(e0.elapsed_time(e1))
    value = world * learn_flops(N, M) * args.steps / elapsed_ms * 1e-9
    alpha_dev_check = h.alpha()

    # ---- end to end through the host-pointer C-ABI: pinned host X, Y -> alpha on the host -------------
    Xn, Yn = Xp.numpy(), Yp.numpy()
    for _ in range(min(args.warmup, 2)):
        h.set_train(Xn, Yn, C3_SIGMA)
        h.learn(0)
        h.lib.gprb200_get_alpha(h.h, alpha_host.data_ptr())
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        h.set_train(Xn, Yn, C3_SIGMA)                         # H2D of this step's inputs
        h.learn(0)
        h.lib.gprb200_get_alpha(h.h, alpha_host.data_ptr())   # D2H of the result
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
    barrier()
    e2e_value = world * learn_flops(N, M) * args.steps / e2e_ms * 1e-9
    e2e_ok = bool(np.allclose(alpha_host.numpy(), alpha_dev_check, rtol=1e-9, atol=1e-12))
    total_launches = int(sum_over_ranks(launches))

    # residual-based parity at full size (size-independent property): ||(K + s^2 I) alpha - y|| / ||y||
    resid = None
    if rank == 0 and not args.skip_checks:
        resid = residual_check(torch, gpr_b200, h, Xd, Yd, N, dev)
    h.close()
    del Xd, Yd
    torch.cuda.empty_cache()

    predict = None if args.skip_predict else run_predict(args, torch, gpr_b200, rank, local_rank, world, dev, barrier,
                                                         max_over_ranks)

    loglik = None
    if rank == 0 and not args.skip_loglik:
        loglik = run_loglik(torch, gpr_b200, local_rank)

    pca = None
    if rank == 0 and not args.skip_pca:
        pca = run_pca(torch, gpr_b200, local_rank)

    host_io = None
    if rank == 0 and not args.skip_hostio:
        host_io = run_host_io(torch, gpr_b200, local_rank)
    sparse = None
    if rank == 0 and not args.skip_sparse:
        sparse = run_sparse(torch, gpr_b200, local_rank)

    large = None
    if world > 1 and not args.skip_large:
        large = run_block_cyclic(args, torch, gpr_b200, rank, local_rank, world, dev, dist)

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.skip_cpu:
        cpu_baseline = run_cpu_baseline()

    if rank == 0:
        peaks = measured_peaks()
        line = {
            "metric": "gp_learn_fp64_tflops", "value": value, "unit": "TFLOP/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": elapsed_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {
                "workload": "C3: GP Learn (Gram + Cholesky + solve), Sum(Sum(Gaussian,Periodic),White), d=8, M=1, fp64",
                "N": N, "models_per_step": world, "useful_flops_per_model": learn_flops(N, M),
                "l2": "inputs larger than L2 (K is %.1f GB)" % (8.0 * N * N * 1e-9),
                "parallelism": "one independent model per GPU, no collective" if world > 1 else "single GPU",
            },
            "learn_s": elapsed_ms / args.steps * 1e-3,
            "cholesky_tflops": N ** 3 / 3.0 / (phases["potrf"] / args.steps) * 1e-9,
            "phase_ms": {k: v / args.steps for k, v in phases.items()},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "TFLOP/s", "h2d_bytes_per_step": int(8 * N * (d + M)),
                    "d2h_bytes_per_step": int(8 * N * M), "ms_per_step": e2e_ms / args.steps,
                    "matches_device_run": e2e_ok},
            "gpu_launches": total_launches,
            "roofline": {
                "kernel": "dgemm_kernel<128,64,64,16,NT,3 stages> (SYRK trailing update of the blocked Cholesky, "
                          "DMMA.8x8x4, 16 warps/SM)",
                "bound": "tensor", "achieved": tb_flops / tb_ms * 1e-9 if tb_ms else None, "peak": FP64_PEAK_TFLOPS,
                "unit": "TFLOP/s", "frac": (tb_flops / tb_ms * 1e-9 / FP64_PEAK_TFLOPS) if tb_ms else None,
                "traffic": TRAFFIC_TB_BYTES_PER_LAUNCH,
                "traffic_note": "ncu capture of the largest launch T^b(0) (31744^2/2 by K=1024, 28.62 ms): 7.61 GB read "
                                "+ 3.76 GB written against 7.79 GB algorithmic (panel + C read + C write); "
                                "profiles/r01b_ncu_summary.md",
                "launches": tb_launches, "avg_launch_ms": tb_ms / tb_launches if tb_launches else None,
                "share_of_step": tb_ms / (elapsed_ms) if elapsed_ms else None,
                "peak_source": "FP64 is absent from MEASURED_PEAKS.json; 37.1 = measured DMMA issue rate on this pool "
                               "(profiles/r01_fp64_probe.txt); cuBLAS DGEMM measured 36.2",
                "cholesky_frac_of_peak": N ** 3 / 3.0 / (phases["potrf"] / args.steps) * 1e-9 / FP64_PEAK_TFLOPS,
            },
            "residual": resid,
        }
        gram_ms = phases["gram"] / args.steps
        hbm_peak = (peaks or {}).get("hbm_gbs", 6650.0)
        # C3 kernel per pair: 2d (difference, square) + 2d (PERIODIC trig-feature dot product) + 2 x ~20 (two exponentials) + ~5
        gram_fp64_per_pair = 2 * d + 2 * d + 2 * 20 + 5
        gram_floor_ms = (N * (N + 1) / 2.0) * gram_fp64_per_pair / (FP64_PEAK_TFLOPS / 2.0 * 1e12) * 1e3
        line["gram"] = {
            "kernel": "gram_rows_kernel<1> (term form, lower triangle only, 4 N^2 bytes written)",
            "ms": gram_ms, "bytes_written": 4.0 * N * N, "achieved_gbs": 4.0 * N * N / gram_ms * 1e-6,
            "frac_of_hbm_peak": 4.0 * N * N / gram_ms * 1e-6 / hbm_peak, "bound": "fp64 alu",
            "fp64_instr_per_pair": gram_fp64_per_pair, "fp64_floor_ms": gram_floor_ms,
            "frac_of_fp64_floor": gram_floor_ms / gram_ms if gram_ms else None,
            "note": "FP64-ALU-bound, not HBM-bound, for this kernel family at d = 8: the DFMA pipe (half of the 37.1 TFLOP/s "
                    "FMA rate counted as instructions) needs fp64_floor_ms; HBM would need %.2f ms" % (4.0 * N * N / hbm_peak * 1e-6),
        }
        if predict:
            line["predict"] = predict
        if large:
            line["large_n"] = large
        if loglik:
            line["loglik"] = loglik
        if pca:
            line["pca_frame"] = pca
        if sparse:
            line["sparse"] = sparse
        if host_io:
            line["host_io"] = host_io
        if cpu_baseline:
            line["cpu_baseline"] = cpu_baseline
        if peaks:
            line["measured_peaks"] = {"hbm_gbs": peaks.get("hbm_gbs")}
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


# dram__bytes_read.sum + dram__bytes_write.sum of the trailing-update launch captured with ncu --set full
# (profiles/r01b_ncu_summary.md: T^b(0) at N=32768, NBO=1024: 7.608593 GB + 3.756411 GB).
TRAFFIC_TB_BYTES_PER_LAUNCH = 7.608593e9 + 3.756411e9


def residual_check(torch, gpr_b200, h, Xd, Yd, N, dev):
    """||(K + sigma^2 I) alpha - y|| / ||y|| at full size, K regenerated by the Gram kernel in row blocks."""
    alpha = torch.from_numpy(h.alpha()).to(dev)
    ldk = (N + 15) // 16 * 16
    K = torch.empty(N, ldk, dtype=torch.float64, device=dev)
    h._ck(h.lib.gprb200_gram_dev(h.h, K.data_ptr(), ldk, 1))
    h.synchronize()
    r = K[:, :N] @ alpha - Yd
    out = float(torch.linalg.norm(r) / torch.linalg.norm(Yd))
    del K
    return {"rel_residual_alpha": out, "note": "||(K+s^2 I) alpha - y|| / ||y|| at N=%d (cond <= N+1)" % N}


def run_predict(args, torch, gpr_b200, rank, local_rank, world, dev, barrier, max_over_ranks):
    """C2: N=4096, d=30, M=786432 columns sharded over the ranks; T = 1 prediction per call."""
    N, d = C2_N, C2_D
    M = args.predict_m
    Ml = M // world
    g = torch.Generator(device=dev)
    g.manual_seed(1234)
    X = torch.randn(N, d, dtype=torch.float64, device=dev, generator=g)
    g.manual_seed(99 + rank)
    W = torch.randn(d, Ml, dtype=torch.float64, device=dev, generator=g) / (d ** 0.5)
    Y = torch.tanh(X @ W)
    Y += torch.randn(N, Ml, dtype=torch.float64, device=dev, generator=g)
    del W
    h = gpr_b200.Handle(local_rank)
    stream = torch.cuda.current_stream(dev)
    h.set_stream(stream.cuda_stream)
    ops, params = gpr_b200.KernelFactory.GetKernel(C2_KERNEL).lower()
    h.set_kernel(ops, params)
    h.set_train_dev(X.data_ptr(), N, d, Y.data_ptr(), Ml, C2_SIGMA)
    del Y
    torch.cuda.empty_cache()
    t0 = time.perf_counter()
    h.learn(0)
    h.synchronize()
    learn_s = time.perf_counter() - t0
    ms = h.phase_ms()
    xs = torch.randn(args.predict_steps + 8, d, dtype=torch.float64, device=dev)
    mean = torch.empty(1, Ml, dtype=torch.float64, device=dev)
    for i in range(3):
        h.predict_dev(xs[i:i + 1].data_ptr(), 1, mean.data_ptr(), Ml)
    h.set_option("profile", 1)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = h.launch_count()
    e0.record(stream)
    for i in range(args.predict_steps):
        h.predict_dev(xs[3 + i:4 + i].data_ptr(), 1, mean.data_ptr(), Ml)
    e1.record(stream)
    barrier()
    dev_ms = max_over_ranks(e0.elapsed_time(e1))
    gemv_ms, gemv_bytes, gemv_n = h.get_stat("gemv_ms"), h.get_stat("gemv_bytes"), h.get_stat("gemv_launches")
    launches = h.launch_count() - l0
    h.set_option("profile", 0)
    # end to end: pinned host test point in, host mean out (host-pointer C-ABI)
    xs_h = xs.cpu().pin_memory().numpy()
    mean_h = torch.empty(1, Ml, dtype=torch.float64).pin_memory()
    for i in range(2):
        h._ck(h.lib.gprb200_predict(h.h, xs_h[i:i + 1].ctypes.data, 1, mean_h.data_ptr(), None))
    barrier()
    t0 = time.perf_counter()
    for i in range(args.predict_steps):
        h._ck(h.lib.gprb200_predict(h.h, xs_h[3 + i:4 + i].ctypes.data, 1, mean_h.data_ptr(), None))
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
    barrier()
    peaks = measured_peaks() or {}
    hbm = peaks.get("hbm_gbs", 6650.0)
    ach = gemv_bytes / gemv_ms * 1e-6 if gemv_ms else None
    out = {
        "metric": "predictions_per_s", "value": args.predict_steps / dev_ms * 1e3, "unit": "predictions/s",
        "n_gpus": world, "scaling": "strong (alpha columns sharded, no collective)",
        "config": {"workload": "C2: Predict, one test point per call", "N": N, "d": d, "M": M, "M_per_gpu": Ml,
                   "alpha_bytes_per_gpu": 8 * N * Ml, "l2": "alpha shard larger than L2"},
        "steps": args.predict_steps, "ms_per_prediction": dev_ms / args.predict_steps,
        "e2e": {"value": args.predict_steps / e2e_ms * 1e3, "unit": "predictions/s", "h2d_bytes_per_step": 8 * d,
                "d2h_bytes_per_step": 8 * Ml},
        "gpu_launches": launches,
        "roofline": {"kernel": "gemv_t_kernel (Kx^T alpha)", "bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s",
                     "frac": ach / hbm if ach else None, "traffic": TRAFFIC_GEMV_BYTES_PER_LAUNCH,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy, read+write)" if "hbm_gbs" in peaks else "fallback"},
        "learn": {"wall_s": learn_s, "solve_ms": ms["solve"], "solve_tflops": 2.0 * N * N * Ml / ms["solve"] * 1e-9,
                  "potrf_ms": ms["potrf"], "gram_ms": ms["gram"]},
    }
    h.close()
    return out


# profiles/r01b_ncu_summary.md: gemv_t_kernel at C2, 25.77 GB read (= 8NM) + the M-vector written.
TRAFFIC_GEMV_BYTES_PER_LAUNCH = 25.77e9 + 8.0 * 786432


def run_loglik(torch, gpr_b200, local_rank, N=16384, iters=3):
    """C4: one iteration of the MaximumLikelihood loop (include/GaussianProcessInference.h:84-156): set parameters,
    GaussianLogLikelihood value + gradient (include/Likelihood.h:231-285) at N = 16384, RationalQuadratic kernel.
    Per iteration: Gram, Cholesky (N^3/3), explicit inverse from the factor (2N^3/3), fused trace contraction."""
    rng = np.random.default_rng(SEED + 4)
    X = rng.uniform(-1, 1, (N, C3_D))
    Y = np.sin(3.0 * X[:, :1]) + 0.1 * rng.standard_normal((N, 1))
    h = gpr_b200.Handle(local_rank)
    ops, params = gpr_b200.KernelFactory.GetKernel("RationalQuadraticKernel(1,1.5,2,)").lower()
    h.set_kernel(ops, params)
    h.set_train(X, Y, 1.0)
    h.loglik(1, True, True, False)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for it in range(iters):
        h.set_kernel_params(np.asarray(params) * (1.0 + 0.01 * (it + 1)))   # P scalars cross the boundary per step
        v, g, _ = h.loglik(1, True, True, False)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / iters * 1e3
    ph = h.phase_ms()
    h.close()
    flops = N ** 3  # N^3/3 factor + 2N^3/3 inverse
    return {"metric": "loglik_iteration_ms", "value": ms, "N": N, "kernel": "RationalQuadraticKernel", "P": 3,
            "phase_ms": {k: ph[k] for k in ("gram", "potrf", "inverse", "solve", "loglik")},
            "tflops_factor_plus_inverse": flops / ((ph["potrf"] + ph["inverse"]) * 1e-3) * 1e-12,
            "reference_equivalent_flops": "2N^3 (LU inverse) + 2N^3/3 (long double det) + 3 x 2N^3 (trace GEMMs) = 3.8e13",
            "last_value": float(v[0]), "last_grad": [float(x) for x in g]}


def run_pca(torch, gpr_b200, local_rank, D=3 * 64 ** 3, n=256, nF=32, N=256, d=30, frames=40):
    """the per-frame path of gpPredict (apps/GaussianProcessPredict.cpp:188-203 +
    include/DataParser.h:471-495): gp->Predict(v) -> principal weights -> PCA GetReconstruction -> the 3 x 64^3
    displacement field, plus the training-time PCA fit (include/PCA.h:33-49) of n fields.
    Device-resident frames are timed with CUDA events on the handle's stream; e2e frames go through the host-pointer
    C-ABI (test point H2D, field D2H). Roofline: the reconstruct kernel streams 8 D nF bytes of basis per frame."""
    from gpr_b200.pca import PCA
    dev = torch.device("cuda", local_rank)
    g = torch.Generator(device=dev)
    g.manual_seed(SEED + 77)
    rank_ = 48
    grid = torch.linspace(0, 1, D, device=dev, dtype=torch.float64)
    modes = torch.stack([torch.sin((k + 1) * np.pi * grid + 0.3 * k) for k in range(rank_)], dim=1)
    amp = torch.logspace(0, -3, rank_, device=dev, dtype=torch.float64)
    coef = torch.randn(rank_, n, generator=g, device=dev, dtype=torch.float64) * amp[:, None]
    Xf = modes @ coef + 1e-4 * torch.randn(D, n, generator=g, device=dev, dtype=torch.float64) + 2.0
    del modes
    h = gpr_b200.Handle(local_rank)
    stream = torch.cuda.current_stream(dev)
    h.set_stream(stream.cuda_stream)
    torch.cuda.synchronize()
    p = PCA.fit_dev(Xf.data_ptr(), n, D, n, nF, h)          # warm-up fit (allocator, kernel attributes)
    p.close()
    l0 = h.launch_count()
    p = PCA.fit_dev(Xf.data_ptr(), n, D, n, nF, h)
    fit_ms, sweeps, fit_launches = p.stat("fit_ms"), p.stat("sweeps"), h.launch_count() - l0
    sigma = p.GetEigenvalues()
    # the GP that predicts the principal weights: inputs = d surrogate features, outputs = the training fields' weights
    F = torch.empty(nF, n, device=dev, dtype=torch.float64)
    p.project_dev(Xf.data_ptr(), n, n, nF, F.data_ptr(), n)
    rng = np.random.default_rng(SEED + 78)
    Xin = rng.standard_normal((N, d))
    Y = F.T.contiguous().cpu().numpy()[:N]
    ops, params = gpr_b200.KernelFactory.GetKernel("GaussianKernel(8,1,)").lower()
    h.set_kernel(ops, params)
    h.set_train(Xin, Y, 0.1)
    h.learn(0)
    Xs = torch.from_numpy(rng.standard_normal((frames, d))).to(dev)
    out = torch.empty(D, device=dev, dtype=torch.float64)
    W = torch.empty(frames, 32, device=dev, dtype=torch.float64)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for t in range(3):
        p.predict_reconstruct_dev(h, Xs[t].data_ptr(), 1, out.data_ptr(), 1)
    l0 = h.launch_count()
    e0.record(stream)
    for t in range(frames):
        p.predict_reconstruct_dev(h, Xs[t].data_ptr(), 1, out.data_ptr(), 1)
    e1.record(stream)
    torch.cuda.synchronize()
    frame_ms = e0.elapsed_time(e1) / frames
    frame_launches = (h.launch_count() - l0) / frames
    # the reconstruct kernel alone (the HBM stream): weights already on the device
    h.predict_dev(Xs.data_ptr(), frames, W.data_ptr(), 32)
    for t in range(3):
        p.reconstruct_dev(W[t].data_ptr(), 32, 1, 1, out.data_ptr(), 1)
    e0.record(stream)
    for t in range(frames):
        p.reconstruct_dev(W[t].data_ptr(), 32, 1, 1, out.data_ptr(), 1)
    e1.record(stream)
    torch.cuda.synchronize()
    rec_ms = e0.elapsed_time(e1) / frames
    rec_bytes = 8.0 * D * nF + 8.0 * D + 8.0 * D
    # DimensionalityReductionFast of one field (the input side of the frame, include/DataParser.h:225-229)
    Fo = torch.empty(nF, device=dev, dtype=torch.float64)
    col = Xf[:, 0].contiguous()
    for t in range(3):
        p.project_dev(col.data_ptr(), 1, 1, nF, Fo.data_ptr(), 1)
    e0.record(stream)
    for t in range(frames):
        p.project_dev(col.data_ptr(), 1, 1, nF, Fo.data_ptr(), 1)
    e1.record(stream)
    torch.cuda.synchronize()
    proj_ms = e0.elapsed_time(e1) / frames
    proj_bytes = 8.0 * D * nF + 8.0 * D + 8.0 * D
    # end to end: host test point in, host field out
    Xs_h = Xs.cpu().pin_memory()
    fields = torch.empty(D, dtype=torch.float64).pin_memory()
    for t in range(2):
        h._ck(h.lib.gprb200_predict_reconstruct(h.h, p.p, Xs_h[t].data_ptr(), 1, fields.data_ptr(), None))
    t0 = time.perf_counter()
    for t in range(frames):
        h._ck(h.lib.gprb200_predict_reconstruct(h.h, p.p, Xs_h[t].data_ptr(), 1, fields.data_ptr(), None))
    e2e_ms = (time.perf_counter() - t0) / frames * 1e3
    # parity property at full size: reconstruct(project(x)) reproduces a training field up to the discarded modes
    back = torch.empty(D, 4, device=dev, dtype=torch.float64)
    p.reconstruct_dev(F.data_ptr(), n, 0, 4, back.data_ptr(), 4)
    tail = float(np.sqrt(np.sum(sigma[nF:] ** 2)) / np.sqrt(np.sum(sigma ** 2)))
    rt = float((back - Xf[:, :4]).norm() / (Xf[:, :4] - Xf.mean(dim=1, keepdim=True)).norm())
    peaks = measured_peaks()
    hbm = peaks.get("hbm_gbs", HBM_FALLBACK_GBS) if peaks else HBM_FALLBACK_GBS
    ach = rec_bytes / (rec_ms * 1e-3) * 1e-9
    p.close()
    h.close()
    del Xf
    torch.cuda.empty_cache()
    return {"metric": "frames_per_s", "value": 1e3 / frame_ms, "unit": "frames/s",
            "config": {"workload": "gpPredict frame: Predict (N=%d, d=%d, M=%d) -> PCA GetReconstruction (D=%d)" % (N, d, nF, D),
                       "D": D, "n_train_fields": n, "nFeatures": nF},
            "ms_per_frame": frame_ms, "launches_per_frame": frame_launches,
            "e2e": {"value": 1e3 / e2e_ms, "unit": "frames/s", "ms_per_frame": e2e_ms, "h2d_bytes_per_step": 8 * d,
                    "d2h_bytes_per_step": 8 * D},
            "roofline": {"kernel": "pca_reconstruct_kernel<1>", "bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s",
                         "frac": ach / hbm, "traffic": None, "ms": rec_ms, "algorithmic_bytes": rec_bytes},
            "project": {"kernel": "pca_project_kernel<1>", "ms": proj_ms, "achieved_gbs": proj_bytes / (proj_ms * 1e-3) * 1e-9,
                        "frac": proj_bytes / (proj_ms * 1e-3) * 1e-9 / hbm, "algorithmic_bytes": proj_bytes},
            "fit": {"ms": fit_ms, "jacobi_sweeps": sweeps, "launches": fit_launches,
                    "note": "mean + 2 x (split-K DMMA Gram, parallel Jacobi, Z <- Z V) on D x n = %d x %d" % (D, n)},
            "round_trip_rel_err": rt, "discarded_energy": tail}


def run_host_io(torch, gpr_b200, local_rank, N=4096, d=30, M=196608):
    """G1 / G5 / G15 at the motion-model shape (a quarter of C2's 786432 columns: Y = 6.4 GB): label vectors stored one
    allocation per sample (the reference's std::vector<VectorType>) -> device, and the regression vectors back.
    Streamed (pinned ring, csrc/hostio.cu) against what the same host program does without it: gather the N x M
    matrix (ComputeLabelMatrixInternal, lib/GaussianProcess.cpp:621-639), then one pageable cudaMemcpy each way."""
    rng = np.random.default_rng(SEED + 7)
    X = rng.standard_normal((N, d))
    rows = [rng.standard_normal(M) for _ in range(N)]
    h = gpr_b200.Handle(local_rank)
    ops, params = gpr_b200.KernelFactory.GetKernel(C2_KERNEL).lower()
    h.set_kernel(ops, params)
    gb = 8.0 * N * M * 1e-9
    out = {"shape": [N, M], "gigabytes": gb}
    xr = [x for x in X]
    h.set_train_rows(xr, rows, C2_SIGMA)      # warm-up: pinned ring and device buffers allocated
    t0 = time.perf_counter()
    h.set_train_rows(xr, rows, C2_SIGMA)
    out["rows_to_device_s"] = time.perf_counter() - t0
    h.learn(0)
    A = np.empty((N, M))
    h._ck(h.lib.gprb200_get_alpha(h.h, A.ctypes.data))      # warm-up (page faults of the destination)
    t0 = time.perf_counter()
    h._ck(h.lib.gprb200_get_alpha(h.h, A.ctypes.data))
    out["alpha_to_host_s"] = time.perf_counter() - t0
    # Learn end to end (host vectors -> host alpha): three steps one after the other, then pipelined over column blocks
    t0 = time.perf_counter()
    h.set_train_rows(xr, rows, C2_SIGMA)
    h.learn(0)
    h._ck(h.lib.gprb200_get_alpha(h.h, A.ctypes.data))
    out["learn_e2e_sequential_s"] = time.perf_counter() - t0
    h.learn_rows(xr, rows, C2_SIGMA, A)       # warm-up: copy streams, second pinned ring
    t0 = time.perf_counter()
    h.learn_rows(xr, rows, C2_SIGMA, A)
    out["learn_e2e_pipelined_s"] = time.perf_counter() - t0
    out["learn_e2e_note"] = ("host label vectors -> Gram + Cholesky + solve of all %d columns -> host regression vectors; "
                             "pipelined = gprb200_learn_rows (upload, solve, download overlapped on three streams)" % M)
    h.set_option("stream_io", 0)
    t0 = time.perf_counter()
    Y = np.empty((N, M))
    for i in range(N):
        Y[i] = rows[i]
    t1 = time.perf_counter()
    h.set_train(X, Y, C2_SIGMA)
    out["gather_s"], out["gather_plus_pageable_to_device_s"] = t1 - t0, time.perf_counter() - t0
    h.learn(0)
    t0 = time.perf_counter()
    h._ck(h.lib.gprb200_get_alpha(h.h, A.ctypes.data))
    out["alpha_to_host_pageable_s"] = time.perf_counter() - t0
    h.close()
    out["rows_to_device_gbs"] = gb / out["rows_to_device_s"]
    out["alpha_to_host_gbs"] = gb / out["alpha_to_host_s"]
    out["speedup_to_device"] = out["gather_plus_pageable_to_device_s"] / out["rows_to_device_s"]
    out["speedup_to_host"] = out["alpha_to_host_pageable_s"] / out["alpha_to_host_s"]
    return out


def run_sparse(torch, gpr_b200, local_rank, n=262144, m=4096, d=8):
    """SparseGaussianProcess::PreComputeRegression (include/SparseGaussianProcess.h:274-312) at a
    size whose dense kernel matrix (n^2 * 8 B = 550 GB) does not exist: cross-Gram Knm (n x m), S = K + sigma^-2 Kmn Knm
    (n m^2 flop on the FP64 tensor path, lower triangle), two m x m Cholesky factorisations, alpha. Device-timed."""
    rng = np.random.default_rng(SEED + 91)
    X = rng.uniform(-1, 1, (n, d))
    Y = np.sin(3.0 * X[:, :1]) + 0.5 * np.cos(2.0 * X[:, 1:2] + X[:, 2:3]) + 0.1 * rng.standard_normal((n, 1))
    Xm = rng.uniform(-1, 1, (m, d))
    dev = torch.device("cuda", local_rank)
    h = gpr_b200.Handle(local_rank)
    stream = torch.cuda.current_stream(dev)
    h.set_stream(stream.cuda_stream)
    ops, params = gpr_b200.KernelFactory.GetKernel(C3_KERNEL).lower()
    h.set_kernel(ops, params)
    h.set_train(X, Y, C3_SIGMA)
    h._ck(h.lib.gprb200_sparse_set_inducing(h.h, Xm.ctypes.data, m, d, 1e-2))
    h._ck(h.lib.gprb200_sparse_learn(h.h))          # warm-up
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = h.launch_count()
    e0.record(stream)
    for _ in range(3):
        h._ck(h.lib.gprb200_sparse_learn(h.h))
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    launches = (h.launch_count() - l0) / 3
    flops = 1.0 * n * m * m + 2.0 * m ** 3 / 3.0 + 2.0 * n * m   # S is symmetric: only its lower triangle is formed
    Xs = rng.uniform(-1, 1, (256, d))
    mean, var = np.empty((256, 1)), np.empty(256)
    t0 = time.perf_counter()
    h._ck(h.lib.gprb200_sparse_predict(h.h, Xs.ctypes.data, 256, mean.ctypes.data, var.ctypes.data))
    pred_ms = (time.perf_counter() - t0) * 1e3
    truth = np.sin(3.0 * Xs[:, 0]) + 0.5 * np.cos(2.0 * Xs[:, 1] + Xs[:, 2])
    # SparseGaussianLogLikelihood::GetValueAndParameterDerivatives (include/SparseLikelihood.h:287-409): one optimiser
    # iteration = relearn + beta + two m x m inverses + Knm S^-1 (2 n m^2 flop) + the two fused dK contractions
    P = len(params)
    val, grad = np.empty(1), np.empty(P)
    h._ck(h.lib.gprb200_sparse_loglik(h.h, 1, val.ctypes.data, grad.ctypes.data, None))   # warm-up
    torch.cuda.synchronize()
    l0 = h.launch_count()
    e0.record(stream)
    for _ in range(3):
        h._ck(h.lib.gprb200_sparse_loglik(h.h, 1, val.ctypes.data, grad.ctypes.data, None))
    e1.record(stream)
    torch.cuda.synchronize()
    lik_ms = e0.elapsed_time(e1) / 3
    lik_launches = (h.launch_count() - l0) / 3
    lik_flops = flops + 2.0 * n * m * m + 2.0 * (2.0 * m ** 3 / 3.0)
    h.close()
    loglik = {"ms_per_iteration": lik_ms, "launches": lik_launches, "P": P, "value": float(val[0]),
              "grad": [float(g) for g in grad], "tflops": lik_flops / lik_ms * 1e-9,
              "useful_flops": "learn + 2 n m^2 (Knm S^-1) + 2 x 2 m^3/3 (K^-1, S^-1); the dK contractions are ALU work on top",
              "reference_equivalent": "P + 1 dense n x n matrices (%.0f GB each) and O(P n^3) = %.1e flop"
                                      % (8.0 * n * n * 1e-9, 2.0 * P * 3 * float(n) ** 3)}
    return {"metric": "sparse_learn_ms", "value": ms, "n": n, "m": m, "d": d, "kernel": C3_KERNEL, "loglik": loglik,
            "tflops": flops / ms * 1e-9, "frac_of_fp64_peak": flops / ms * 1e-9 / FP64_PEAK_TFLOPS, "launches": launches,
            "useful_flops": "n m^2 (lower triangle of S) + 2 m^3/3 (two Cholesky) + 2 n m (Kmn Y)",
            "dense_equivalent": "K would be %.0f GB" % (8.0 * n * n * 1e-9),
            "predict_256_with_variance_ms": pred_ms, "mean_abs_err_vs_truth": float(np.abs(mean[:, 0] - truth).mean()),
            "variance_min": float(var.min())}


def run_block_cyclic(args, torch, gpr_b200, rank, local_rank, world, dev, dist):
    """C5: ONE model of N = 131072 samples factored across all ranks (block-cyclic Cholesky, NCCL panel broadcasts,
    gpr_b200/dist.py) + distributed solve. Strong scaling of a single large problem; device-timed, max over ranks."""
    from gpr_b200 import dist as gd
    N, nb = args.large_n, args.large_nb
    rng = np.random.default_rng(SEED + 131)
    X = rng.uniform(-1, 1, (N, C3_D))
    Y = (np.sin(3.0 * X[:, :1]) + 0.5 * np.cos(2.0 * X[:, 1:2] + X[:, 2:3]) + 0.1 * rng.standard_normal((N, 1)))
    ops_, params = gpr_b200.KernelFactory.GetKernel(C3_KERNEL).lower()
    ops = gd.DeviceOps(local_rank, ops_, params, X, C3_SIGMA)
    # warm-up on a small problem with the same code path (NCCL communicator, allocator, kernel attributes)
    warm = gd.BlockCyclicGP(ops, 8192, nb=nb)
    warm.learn(Y[:8192])
    del warm
    torch.cuda.empty_cache()
    gp = gd.BlockCyclicGP(ops, N, nb=nb)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    dist.barrier()
    torch.cuda.synchronize()
    with ops.on("main"):
        ev[0].record()
    gp.build_gram()
    with ops.on("main"):
        ev[1].record()
    info = gp.factor()
    with ops.on("main"):
        ev[2].record()
    alpha = gp.solve(Y) if info == 0 else None
    with ops.on("main"):
        ev[3].record()
    torch.cuda.synchronize()
    t = torch.tensor([ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3]),
                      ev[0].elapsed_time(ev[3])], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    gram_ms, factor_ms, solve_ms, total_ms = (float(v) for v in t.tolist())
    # residual on a sample of rows: ||(K + s^2 I) alpha - y|| over 512 rows, K rows regenerated by the Gram kernel
    resid = None
    if info == 0 and rank == 0:
        rows = 512
        Kr = ops.empty(rows, N)
        with ops.on("main"):
            ops.gram_block(N - rows, 0, Kr)
            r = Kr @ alpha - ops.from_host(Y[N - rows:])
            resid = float(torch.linalg.norm(r) / np.linalg.norm(Y[N - rows:]))
    out = {"metric": "gp_learn_large_n_s", "N": N, "nb": nb, "n_gpus": world, "info": info,
           "learn_s": total_ms * 1e-3, "gram_s": gram_ms * 1e-3, "cholesky_s": factor_ms * 1e-3, "solve_s": solve_ms * 1e-3,
           "cholesky_tflops_total": N ** 3 / 3.0 / factor_ms * 1e-9,
           "cholesky_frac_of_peak": N ** 3 / 3.0 / factor_ms * 1e-9 / (FP64_PEAK_TFLOPS * world),
           "scaling": "strong (one N x N problem over all ranks)", "logdet": gp.logdet,
           "bcast_GB_per_rank": gp.stats["bcast_bytes"] * 1e-9, "rel_residual_rows": resid,
           "layout": "block-cyclic over a 1 x %d grid, lower triangle only, %.1f GB per GPU" % (world, 4.0 * N * N / world * 1e-9)}
    ops.close()
    return out


def run_cpu_baseline():
    """The reference's own CPU code (oracle/_ref) on a bounded sample of the same workload, all host threads."""
    try:
        from oracle import ref
        if not ref.available():
            return {"value": None, "unit": "TFLOP/s", "cores": 0, "kind": "reference",
                    "sample": "unavailable: oracle/_ref/libgpr_ref.so missing"}
        threads = os.cpu_count() or 1
        os.environ["OMP_NUM_THREADS"] = str(threads)
        N = 10240  # ~10 s on 16 cores; throughput of dgetrf/dgetri is flat in N from here on
        time_reference_learn(1024, threads)
        dt, _ = time_reference_learn(N, threads)
        return {"value": learn_flops(N, C3_M) / dt * 1e-12, "unit": "TFLOP/s", "cores": threads, "kind": "reference",
                "seconds": dt,
                "sample": f"one C3 Learn pass at N={N} (full workload N={C3_N}); reference lib/GaussianProcess.cpp via "
                          f"oracle/_ref: OpenMP Gram + dgetrf/dgetri + C*Y; useful flops N^3/3 + 2N^2M counted"}
    except Exception as e:  # the baseline is reported, never required
        return {"value": None, "unit": "TFLOP/s", "cores": 0, "kind": "reference", "sample": f"failed: {e}"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gpu", choices=["gpu", "reference"])
    ap.add_argument("--n", type=int, default=C3_N)
    ap.add_argument("--ref-n", type=int, default=0, help="reference arm: N (default: the product arm's 32768)")
    ap.add_argument("--ref-passes", type=int, default=1, help="reference arm: full passes timed")
    ap.add_argument("--predict-m", type=int, default=C2_M)
    ap.add_argument("--predict-steps", type=int, default=20)
    ap.add_argument("--skip-predict", action="store_true")
    ap.add_argument("--skip-large", action="store_true")
    ap.add_argument("--skip-loglik", action="store_true")
    ap.add_argument("--skip-pca", action="store_true")
    ap.add_argument("--skip-sparse", action="store_true")
    ap.add_argument("--skip-hostio", action="store_true")
    ap.add_argument("--large-n", type=int, default=131072)
    ap.add_argument("--large-nb", type=int, default=1024, help="block-column width of the block-cyclic Cholesky")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-checks", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
