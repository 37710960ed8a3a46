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

    env = {"torch": torch, "gpr_b200": gpr_b200, "rank": rank, "local_rank": local_rank, "world": world, "dev": dev,
           "dist": dist, "barrier": barrier, "max_over_ranks": max_over_ranks, "sum_over_ranks": sum_over_ranks}
    detail = {}
    if world == 1:
        line = headline_c3(args, env, detail)
    else:
        line = headline_c5(args, env, detail)
        if not args.skip_replicas:
            rep = headline_c3(args, env, {}, steps=min(args.steps, 3), warmup=3, checks=False)
            if rank == 0:
                detail["replicas"] = rep
                line["config"]["replicas"] = {"what": "one independent C3 model per GPU, no collective (the round-1 headline)",
                                              "tflops_total": rep["value"], "ms_per_step": rep["ms_per_step"]}
    torch.cuda.empty_cache()
    if rank == 0:
        sys.stderr.write("bench headline: " + json.dumps(line) + "\n")
        sys.stderr.flush()

    predict = None if args.skip_predict else run_predict(args, torch, gpr_b200, rank, local_rank, world, dev, barrier,
                                                         max_over_ranks)
    if predict and rank == 0:
        detail["predict"] = predict
        line["config"]["predict"] = {"workload": predict["config"]["workload"], "M": predict["config"]["M"],
                                     "predictions_per_s": predict["value"], "e2e_predictions_per_s": predict["e2e"]["value"],
                                     "hbm_gbs": predict["roofline"]["achieved"], "hbm_frac": predict["roofline"]["frac"],
                                     "sharding": "alpha columns M / %d per GPU, no collective" % world}
    if not args.skip_predict_ci:
        pci = run_predict_ci(args, env)
        if pci and rank == 0:
            detail["predict_ci"] = pci
            line["config"]["predict_ci"] = {k: pci[k] for k in ("points_per_s", "ms_per_point_t1", "T", "N", "sharding")}
    if world > 1 and not args.skip_loglik:
        sl = run_sharded_loglik(args, env)
        if sl and rank == 0:
            detail["loglik_sharded"] = sl
            line["config"]["loglik_sharded"] = {k: sl[k] for k in ("iteration_ms", "N", "collective", "grad_vs_1gpu")}
    if rank == 0 and world == 1:
        for name, flag, fn in (("loglik", args.skip_loglik, run_loglik), ("pca_frame", args.skip_pca, run_pca),
                               ("host_io", args.skip_hostio, run_host_io), ("sparse", args.skip_sparse, run_sparse)):
            if not flag:
                detail[name] = fn(torch, gpr_b200, local_rank)
        if not args.skip_predict_loop:
            detail["gp_predict_loop"] = run_predict_loop()
            if "ms_per_point" in detail["gp_predict_loop"]:
                line["config"]["gp_predict_loop"] = {k: detail["gp_predict_loop"][k] for k in
                                                     ("N", "M", "ms_per_point", "predict_ms_per_point", "interval_ms_per_point")}
        if "loglik" in detail:
            line["config"]["loglik_iteration_ms"] = detail["loglik"]["value"]
            line["config"]["loglik_phase_ms"] = detail["loglik"]["phase_ms"]
        if not args.skip_cpu:
            line["cpu_baseline"] = run_cpu_baseline()
    if rank == 0:
        peaks = measured_peaks()
        if peaks:
            line["measured_peaks"] = {"hbm_gbs": peaks.get("hbm_gbs")}
        # the contract line stays compact (the driver keeps `config` and `roofline`); the full sub-objects go to stderr
        # and, when the directory exists, to gpurun_out/bench_detail_<n>gpu.json
        sys.stderr.write("bench detail: " + json.dumps(detail) + "\n")
        try:
            if os.path.isdir(os.path.join(ROOT, "gpurun_out")):
                json.dump({"line": line, "detail": detail},
                          open(os.path.join(ROOT, "gpurun_out", "bench_detail_%dgpu.json" % world), "w"), indent=1)
        except Exception:
            pass
        print(json.dumps(line))
        sys.stdout.flush()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def oracle_rows(kernel_string, X, rows, chunk=16):
    """k(x_r, x_j) for the sampled rows on the CPU (numpy restatement of include/Kernel.h, pinned to oracle/_ref): the
    checker's side of the residual checks, independent of the device kernel functor."""
    from oracle import gpr_oracle as orc
    k = orc.parse_kernel(kernel_string)
    return np.concatenate([orc.kernel_matrix(k, X[rows[i:i + chunk]], X) for i in range(0, len(rows), chunk)], axis=0)


def residual_rows(kernel_string, X, Y, alpha, sigma, nrows=192, seed=5):
    """max |K_r alpha + sigma^2 alpha_r - y_r| / max |y| over sampled rows r, kernel rows from the oracle."""
    rng = np.random.default_rng(seed)
    rows = np.sort(rng.choice(X.shape[0], nrows, replace=False))
    Kr = oracle_rows(kernel_string, X, rows)
    r = Kr @ alpha + sigma * sigma * alpha[rows] - Y[rows]
    return float(np.abs(r).max() / np.abs(Y).max())


def headline_c3(args, env, detail, steps=None, warmup=None, checks=True):
    """C3 Learn on every rank (at 1 GPU: the headline; at N > 1: the independent-replica sub-object)."""
    torch, gpr_b200 = env["torch"], env["gpr_b200"]
    rank, local_rank, world, dev = env["rank"], env["local_rank"], env["world"], env["dev"]
    steps = steps or args.steps
    warmup = warmup or args.warmup
    N, d, M = args.n, C3_D, C3_M
    kernel = gpr_b200.KernelFactory.GetKernel(C3_KERNEL)
    ops, params = kernel.lower()
    params = list(params)
    params[0] *= 1.0 + 0.01 * rank          # every rank learns its own model: different data and hyper-parameter point
    Xh, Yh = c3_inputs(N, seed=rank)
    Xd, Yd = torch.from_numpy(Xh).to(dev), torch.from_numpy(Yh).to(dev)
    alpha_host = np.empty((N, M))            # pageable, as the reference API's own Eigen storage

    h = gpr_b200.Handle(local_rank)
    stream = torch.cuda.current_stream(dev)
    h.set_stream(stream.cuda_stream)
    h.set_kernel(ops, params)
    h.set_option("profile", 1)
    h.set_train_dev(Xd.data_ptr(), N, d, Yd.data_ptr(), M, C3_SIGMA)
    for _ in range(warmup):
        h.learn(0)
    env["barrier"]()
    sampler = ClockSampler(local_rank)
    l0 = h.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    phases = {"gram": 0.0, "potrf": 0.0, "solve": 0.0}
    tb_ms = tb_flops = tb_launches = 0.0
    e0.record(stream)
    for _ in range(steps):
        h.learn(0)
        ms = h.phase_ms()
        for k in phases:
            phases[k] += ms[k]
        tb_ms += h.get_stat("tb_ms")
        tb_flops += h.get_stat("tb_flops")
        tb_launches += h.get_stat("tb_launches")
    e1.record(stream)
    env["barrier"]()
    launches = h.launch_count() - l0
    clocks = sampler.stop()
    elapsed_ms = env["max_over_ranks"](e0.elapsed_time(e1))
    value = world * learn_flops(N, M) * steps / elapsed_ms * 1e-9
    alpha_dev_check = h.alpha()

    # end to end through the host-pointer C-ABI: PAGEABLE host X, Y (what the reference API owns) -> alpha on the host
    e2e_steps = steps
    for _ in range(2):
        h.set_train(Xh, Yh, C3_SIGMA)
        h.learn(0)
        h.lib.gprb200_get_alpha(h.h, alpha_host.ctypes.data)
    env["barrier"]()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        h.set_train(Xh, Yh, C3_SIGMA)                              # H2D of this step's inputs
        h.learn(0)
        h.lib.gprb200_get_alpha(h.h, alpha_host.ctypes.data)       # D2H of the result
    torch.cuda.synchronize()
    e2e_ms = env["max_over_ranks"]((time.perf_counter() - t0) * 1e3)
    env["barrier"]()
    e2e_value = world * learn_flops(N, M) * e2e_steps / e2e_ms * 1e-9
    e2e_ok = bool(np.allclose(alpha_host, alpha_dev_check, rtol=1e-9, atol=1e-12))
    total_launches = int(env["sum_over_ranks"](launches))
    resid = None
    if rank == 0 and checks and not args.skip_checks:
        resid = {"rel_residual_rows": residual_rows(kernel_string_with(params), Xh, Yh, alpha_host, C3_SIGMA),
                 "note": "max |K_r alpha + s^2 alpha_r - y_r| / max |y| on 192 sampled rows of the N=%d problem, kernel rows "
                         "evaluated by the CPU oracle (not by the device functor)" % N}
    gram = None
    if rank == 0 and checks:
        gram = gram_evidence(torch, h, N, d, phases["gram"] / steps, dev)
    h.close()
    del Xd, Yd
    torch.cuda.empty_cache()
    chol_tf = N ** 3 / 3.0 / (phases["potrf"] / steps) * 1e-9
    line = {
        "metric": "gp_learn_fp64_tflops", "value": value, "unit": "TFLOP/s", "n_gpus": world,
        "steps": steps, "warmup": warmup, "ms_per_step": elapsed_ms / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {
            "workload": C3_WORKLOAD, "N": N, "d": d, "M": M, "models_per_step": world,
            "useful_flops_per_model": learn_flops(N, M),
            "l2": "inputs larger than L2 (K is %.1f GB)" % (8.0 * N * N * 1e-9),
            "parallelism": "one independent model per GPU, no collective" if world > 1 else "single GPU",
            "learn_s": elapsed_ms / steps * 1e-3, "cholesky_tflops": chol_tf,
            "phase_ms": {k: v / steps for k, v in phases.items()},
        },
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "TFLOP/s", "h2d_bytes_per_step": int(8 * N * (d + M)),
                "d2h_bytes_per_step": int(8 * N * M), "ms_per_step": e2e_ms / e2e_steps, "matches_device_run": e2e_ok,
                "host_buffers": "pageable numpy arrays"},
        "gpu_launches": total_launches,
        "roofline": {
            "kernel": "dgemm_kernel<128,64,64,16,NT,3 stages> (SYRK trailing update of the blocked Cholesky, DMMA.8x8x4)",
            "bound": "tensor", "achieved": tb_flops / tb_ms * 1e-9 if tb_ms else None, "peak": FP64_PEAK_TFLOPS,
            "unit": "TFLOP/s", "frac": (tb_flops / tb_ms * 1e-9 / FP64_PEAK_TFLOPS) if tb_ms else None,
            "traffic": TRAFFIC_TB_BYTES_PER_LAUNCH,
            "traffic_note": "ncu --set full of the largest launch T^b(0): 7.61 GB read + 3.76 GB written vs 7.79 GB "
                            "algorithmic (profiles/r01b_ncu_summary.md)",
            "launches": tb_launches, "avg_launch_ms": tb_ms / tb_launches if tb_launches else None,
            "share_of_step": tb_ms / elapsed_ms if elapsed_ms else None,
            "peak_source": "FP64 is absent from MEASURED_PEAKS.json; 37.1 = measured DMMA issue rate on this pool "
                           "(profiles/r01_fp64_probe.txt); cuBLAS DGEMM measured 36.2",
            "cholesky_frac_of_peak": chol_tf / FP64_PEAK_TFLOPS,
        },
        "residual": resid,
    }
    if gram:
        line["config"]["gram"] = gram
    return line


def kernel_string_with(params):
    """C3 kernel string with the given parameter vector (order: Gaussian sigma, scale; Periodic scale, b, sigma; White)."""
    p = [repr(float(v)) for v in params]
    return "SumKernel(SumKernel(GaussianKernel(%s,%s,),PeriodicKernel(%s,%s,%s,)),WhiteKernel(%s,))" % tuple(p)


def gram_evidence(torch, h, N, d, learn_gram_ms, dev):
    """Achieved HBM write GB/s of the Gram paths the API runs (north_star subsystem 1): the lower-triangle build inside
    Learn and the SYMMETRIC full-K build of ComputeKernelMatrix (gprb200_gram_dev: N^2/2 pairs evaluated, mirrored)."""
    peaks = measured_peaks() or {}
    hbm = peaks.get("hbm_gbs", HBM_FALLBACK_GBS)
    ldk = (N + 15) // 16 * 16
    K = torch.empty(N, ldk, dtype=torch.float64, device=dev)
    out = {"learn_lower_triangle": {"ms": learn_gram_ms, "bytes": 4.0 * N * N, "gbs": 4.0 * N * N / learn_gram_ms * 1e-6,
                                    "frac_of_hbm": 4.0 * N * N / learn_gram_ms * 1e-6 / hbm}}
    for _ in range(2):
        h._ck(h.lib.gprb200_gram_dev(h.h, K.data_ptr(), ldk, 1))
    t = []
    for _ in range(3):
        h._ck(h.lib.gprb200_gram_dev(h.h, K.data_ptr(), ldk, 1))
        t.append(h.phase_ms()["gram"])
    ms = min(t)
    out["symmetric_full_K"] = {"kernel": "C3 Sum(Sum(Gaussian,Periodic),White)", "ms": ms, "bytes": 8.0 * N * N,
                               "gbs": 8.0 * N * N / ms * 1e-6, "frac_of_hbm": 8.0 * N * N / ms * 1e-6 / hbm}
    del K
    return out


def headline_c5(args, env, detail):
    """N > 1 GPUs: ONE model of N = 131072 samples (BASELINE config C5; K = 137 GB) over all ranks - Gram in the
    factorisation's layout, 2-D block-cyclic Cholesky with NCCL panel broadcasts, distributed solve (gpr_b200/dist.py).
    Strong scaling: the problem is fixed, value = useful flops / time (max over ranks, CUDA events)."""
    torch, gpr_b200, dist = env["torch"], env["gpr_b200"], env["dist"]
    rank, local_rank, world, dev = env["rank"], env["local_rank"], env["world"], env["dev"]
    from gpr_b200 import dist as gd
    N, nb, M = args.large_n, args.large_nb, C3_M
    grid = tuple(int(v) for v in args.grid.split("x")) if args.grid else gd.default_grid(world)
    rng = np.random.default_rng(SEED + 131)
    X = rng.uniform(-1, 1, (N, C3_D))
    Y = (np.sin(3.0 * X[:, :1]) + 0.5 * np.cos(2.0 * X[:, 1:2] + X[:, 2:3]) + 0.1 * rng.standard_normal((N, 1)))
    ops_, params = gpr_b200.KernelFactory.GetKernel(C3_KERNEL).lower()
    impl = args.c5_impl
    if impl == "capi":
        # the whole schedule under the C-ABI (libgprb200_dist.so, NCCL from C++): every step takes HOST X, Y and returns a
        # HOST alpha, so one loop gives both the device-timed value (the library's CUDA events) and the end-to-end time
        from gpr_b200 import capi_dist
        dh = capi_dist.DistHandle(local_rank, rank, world, capi_dist.exchange_id(), grid if args.grid else None)
        dh.set_kernel(ops_, params)
        grid = (int(dh.counters()["Pr"]), int(dh.counters()["Pc"]))
        dh.learn(X[:8192], Y[:8192], C3_SIGMA, nb=nb)          # communicators, kernel attributes
        for _ in range(1 if args.large_quick_warmup else args.warmup):
            dh.learn(X, Y, C3_SIGMA, nb=nb)
        env["barrier"]()
        sampler = ClockSampler(local_rank)
        tsum = np.zeros(4)
        wall, launches, info, logdet, alpha_host = 0.0, 0, 0, None, None
        stats = {"bcast_bytes": 0.0, "collectives": 0.0}
        for _ in range(args.steps):
            t0 = time.perf_counter()
            alpha_host, logdet = dh.learn(X, Y, C3_SIGMA, nb=nb)
            wall += time.perf_counter() - t0
            ph, c = dh.phase_ms(), dh.counters()
            tsum += np.array([ph["gram"], ph["factor"], ph["solve"], ph["gram"] + ph["factor"] + ph["solve"]])
            launches += int(c["launches"])
            stats["bcast_bytes"] += c["bytes"]
            stats["collectives"] += c["collectives"]
        env["barrier"]()
        clocks = sampler.stop()
        t = torch.tensor(tsum, dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        gram_ms, factor_ms, solve_ms, total_ms = (float(v) / args.steps for v in t.tolist())
        e2e_steps = args.steps
        e2e_ms = env["max_over_ranks"](wall * 1e3) / e2e_steps
        per_learn = args.steps
        alpha = None
        dh.close()
    else:
        ops = gd.DeviceOps(local_rank, ops_, params, X, C3_SIGMA)
        warm = gd.BlockCyclicGP(ops, 8192, nb=nb, grid=grid)     # communicators, allocator, kernel attributes
        warm.learn(Y[:8192])
        del warm
        torch.cuda.empty_cache()
        gp = gd.BlockCyclicGP(ops, N, nb=nb, grid=grid)
        Yd = torch.from_numpy(Y).to(dev)
        for _ in range(1 if args.large_quick_warmup else args.warmup):
            gp.learn(Yd)
        env["barrier"]()
        sampler = ClockSampler(local_rank)
        l0 = ops.h.launch_count() + ops.hp.launch_count()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        tsum = np.zeros(4)
        info, alpha = 0, None
        for _ in range(args.steps):
            with ops.on("main"):
                ev[0].record()
            gp.build_gram()
            with ops.on("main"):
                ev[1].record()
            info = gp.factor()
            with ops.on("main"):
                ev[2].record()
            alpha = gp.solve(Yd) if info == 0 else None
            with ops.on("main"):
                ev[3].record()
            torch.cuda.synchronize()
            tsum += np.array([ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3]),
                              ev[0].elapsed_time(ev[3])])
        env["barrier"]()
        clocks = sampler.stop()
        launches = ops.h.launch_count() + ops.hp.launch_count() - l0
        t = torch.tensor(tsum, dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        gram_ms, factor_ms, solve_ms, total_ms = (float(v) / args.steps for v in t.tolist())
        stats = dict(gp.stats)
        # end to end: host X, Y in, host alpha out (pageable), every step
        e2e_steps = min(args.steps, 2)
        alpha_host = None
        env["barrier"]()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            ops.set_samples(X, C3_SIGMA)
            gp.build_gram()
            if gp.factor() == 0:
                alpha_host = gp.solve(Y).cpu().numpy()
        torch.cuda.synchronize()
        e2e_ms = env["max_over_ranks"]((time.perf_counter() - t0) * 1e3) / e2e_steps
        per_learn = max(1, args.steps + e2e_steps + 1)
        logdet = gp.logdet
        ops.close()
        del gp, Yd
    resid = None
    if info == 0 and rank == 0 and not args.skip_checks:
        resid = residual_rows(C3_KERNEL, X, Y, alpha_host if alpha_host is not None else alpha.cpu().numpy(), C3_SIGMA, nrows=96)
    total_launches = int(env["sum_over_ranks"](launches))
    flops = learn_flops(N, M)
    chol_tf = N ** 3 / 3.0 / factor_ms * 1e-9
    line = {
        "metric": "gp_learn_fp64_tflops", "value": flops / total_ms * 1e-9, "unit": "TFLOP/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {
            "workload": "C5: GP Learn (Gram + Cholesky + solve) of ONE N=131072 model over all GPUs, "
                        "Sum(Sum(Gaussian,Periodic),White), d=8, M=1, fp64",
            "N": N, "d": C3_D, "M": M, "nb": nb, "grid": "%dx%d" % grid,
            "layout": "2-D block-cyclic %d x %d, lower triangle only, %.1f GB per GPU" % (grid[0], grid[1], 4.0 * N * N / world * 1e-9),
            "parallelism": "Pr x Pc block-cyclic Cholesky (row map mirrored), NCCL broadcasts on row / column communicators, "
                           "one-panel look-ahead, one staircase GEMM per trailing update",
            "useful_flops": flops, "l2": "inputs larger than L2 (%.0f GB per GPU)" % (4.0 * N * N / world * 1e-9),
            "learn_s": total_ms * 1e-3, "gram_s": gram_ms * 1e-3, "cholesky_s": factor_ms * 1e-3, "solve_s": solve_ms * 1e-3,
            "cholesky_tflops_total": chol_tf, "cholesky_frac_of_peak": chol_tf / (FP64_PEAK_TFLOPS * world),
            "bcast_GB_per_rank": stats["bcast_bytes"] * 1e-9 / per_learn,
            "collectives_per_learn": stats["collectives"] / per_learn,
            "impl": "libgprb200_dist.so: schedule and NCCL calls in C++ under the C-ABI (include/gprb200_dist.h)" if impl == "capi"
                    else "gpr_b200.dist.BlockCyclicGP: torch.distributed schedule over the single-GPU C-ABI",
            "info": info, "logdet": logdet, "rel_residual_rows": resid,
            "residual_note": "96 sampled rows, kernel rows evaluated by the CPU oracle",
            "warmup_note": "DEVELOPMENT RUN: 1 full-size warm-up pass only" if args.large_quick_warmup
                           else "%d full-size warm-up passes" % args.warmup,
        },
        "clocks": clocks,
        "e2e": {"value": flops / e2e_ms * 1e-9, "unit": "TFLOP/s", "h2d_bytes_per_step": int(8 * N * (C3_D + M)) * world,
                "d2h_bytes_per_step": int(8 * N * M) * world, "ms_per_step": e2e_ms, "steps": e2e_steps,
                "host_buffers": "pageable numpy arrays, every rank uploads X and Y and downloads alpha"},
        "gpu_launches": total_launches,
        "roofline": {
            "kernel": "dgemm_kernel (block-column trailing updates of the block-cyclic Cholesky, DMMA.8x8x4) per GPU",
            "bound": "tensor", "achieved": chol_tf / world, "peak": FP64_PEAK_TFLOPS, "unit": "TFLOP/s",
            "frac": chol_tf / world / FP64_PEAK_TFLOPS, "traffic": None,
            "note": "achieved = N^3/3 / cholesky_s / n_gpus: the whole factorisation per GPU, panel chain, broadcasts "
                    "and load imbalance included (not a single-kernel figure)",
            "peak_source": "37.1 = measured DMMA issue rate (profiles/r01_fp64_probe.txt)",
        },
    }
    torch.cuda.empty_cache()
    return line


# dram__bytes_read.sum + dram__bytes_write.sum of the trailing-update launch captured with ncu --set full
# (profiles/r01b_ncu_summary.md: T^b(0) at N=32768, NBO=1024: 7.608593 GB + 3.756411 GB).
TRAFFIC_TB_BYTES_PER_LAUNCH = 7.608593e9 + 3.756411e9


def run_predict_ci(args, env, N=16384, d=8, T=4096):
    """Section 8(e) row 3 and the loop body of apps/GaussianProcessPredict.cpp:185-193: Predict + GetCredibleInterval.
    Test points are independent: T / g per GPU, L replicated, no data-path collective. Reports batched points/s over
    all ranks and the latency of ONE point per call (mean + interval), both through the host-pointer C-ABI."""
    torch, gpr_b200 = env["torch"], env["gpr_b200"]
    rank, local_rank, world = env["rank"], env["local_rank"], env["world"]
    from gpr_b200 import dist as gd
    rng = np.random.default_rng(SEED + 16)
    X = rng.uniform(-1, 1, (N, d))
    Y = np.sin(3.0 * X[:, :1]) + 0.1 * rng.standard_normal((N, 1))
    Xs = rng.uniform(-1, 1, (T, d))
    h = gpr_b200.Handle(local_rank)
    ops, params = gpr_b200.KernelFactory.GetKernel(C3_KERNEL).lower()
    h.set_kernel(ops, params)
    g = gd.ShardedOutputsGP(h)
    g.set_train(X, Y, C3_SIGMA)
    g.learn()
    b, e = gd.column_shard(T, rank, world)
    g.credible_interval(Xs[:64], gather=False)
    batch_ms = None
    for _ in range(3):      # a 10-30 ms host-timed region, max over ranks: the best of three passes
        env["barrier"]()
        t0 = time.perf_counter()
        mean = g.predict(Xs[b:e])                      # this rank's output columns (may be none when M < g)
        ci = g.credible_interval(Xs, gather=False)     # this rank's test points
        torch.cuda.synchronize()
        ms = env["max_over_ranks"]((time.perf_counter() - t0) * 1e3)
        batch_ms = ms if batch_ms is None else min(batch_ms, ms)
    one_ms = None
    if g.shard[1] > g.shard[0]:
        for i in range(3):
            h.predict(Xs[i:i + 1], want_ci=True)
        t0 = time.perf_counter()
        reps = 20
        for i in range(reps):
            h.predict(Xs[i:i + 1], want_ci=True)
        one_ms = (time.perf_counter() - t0) * 1e3 / reps
    h.close()
    return {"N": N, "d": d, "T": T, "points_per_s": T / batch_ms * 1e3, "batch_ms": batch_ms, "ms_per_point_t1": one_ms,
            "sharding": "test points T / %d per GPU, L replicated, no collective" % world,
            "finite": bool(np.isfinite(ci).all() and np.isfinite(mean).all())}


def run_sharded_loglik(args, env, N=16384, iters=2):
    """Section 8(e) row 4 at BASELINE config C4 (N = 16384, RationalQuadratic, P = 3): one log-likelihood + gradient
    iteration with the inverse and the dK contraction sharded by row blocks over the ranks (gpr_b200.dist.ShardedLikelihood;
    collective: all_gather of inv(L) column blocks + all_reduce of P doubles), against the single-GPU gprb200_loglik."""
    torch, gpr_b200 = env["torch"], env["gpr_b200"]
    rank, local_rank, world = env["rank"], env["local_rank"], env["world"]
    from gpr_b200 import dist as gd
    rng = np.random.default_rng(SEED + 4)
    X = rng.uniform(-1, 1, (N, C3_D))
    Y = np.sin(3.0 * X[:, :1]) + 0.1 * rng.standard_normal((N, 1))
    ops, params = gpr_b200.KernelFactory.GetKernel("RationalQuadraticKernel(1,1.5,2,)").lower()
    lo = gd.DeviceLikOps(local_rank, ops, params, X, Y, 1.0)
    sl = gd.ShardedLikelihood(lo, block=512)
    p = np.asarray(params)
    sl.value_and_gradient(p, mode=1)
    env["barrier"]()
    t0 = time.perf_counter()
    for it in range(iters):
        v, g = sl.value_and_gradient(p * (1.0 + 0.01 * (it + 1)), mode=1)
    torch.cuda.synchronize()
    ms = env["max_over_ranks"]((time.perf_counter() - t0) * 1e3) / iters
    one = None
    if rank == 0:
        h = lo.h
        h.set_kernel_params(p * (1.0 + 0.01 * iters))
        h.loglik(1, True, True, False)
        t0 = time.perf_counter()
        v1, g1, _ = h.loglik(1, True, True, False)
        torch.cuda.synchronize()
        one = {"ms": (time.perf_counter() - t0) * 1e3, "grad": [float(x) for x in g1]}
    lo.close()
    if rank != 0:
        return None
    return {"N": N, "P": 3, "n_gpus": world, "iteration_ms": ms, "single_gpu_iteration_ms": one["ms"],
            "collective": "all_gather of inv(L) column blocks (%.2f GB per iteration) + all_reduce of P doubles"
                          % (sl.stats["gathered_bytes"] / (iters + 1) * 1e-9),
            "grad": [float(x) for x in g], "grad_vs_1gpu": float(np.max(np.abs(g - np.array(one["grad"]))) / np.max(np.abs(g)))}


def run_predict_loop(N=16384, d=8, M=4, T=200):
    """The loop body of the reference's gpPredict (apps/GaussianProcessPredict.cpp:185-193) through the host C++ API
    (libgplib.so): one gp->Predict(x) and one gp->GetCredibleInterval(x) per test point (tests/cxx/predict_loop_bench.cpp)."""
    exe = os.path.join(ROOT, "tests", "cxx", "bin", "predict_loop_bench")
    if not os.path.exists(exe):
        return {"unavailable": "tests/cxx/bin/predict_loop_bench not built"}
    try:
        p = subprocess.run([exe, str(N), str(d), str(M), str(T)], capture_output=True, text=True, timeout=300)
        if p.returncode != 0:
            return {"unavailable": "predict_loop_bench failed: " + p.stderr[-300:]}
        out = json.loads(p.stdout.strip().splitlines()[-1])
        out["what"] = "gp->Predict(x) + gp->GetCredibleInterval(x), one test point per call, through libgplib.so"
        return out
    except Exception as e:
        return {"unavailable": "predict_loop_bench: %s" % e}


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
    ap.add_argument("--grid", default="", help="PrxPc process grid of the block-cyclic Cholesky (default: dist.default_grid)")
    ap.add_argument("--c5-impl", choices=["capi", "torch"], default="capi",
                    help="N>1 headline: capi = libgprb200_dist.so (schedule + NCCL in C++), torch = gpr_b200.dist.BlockCyclicGP")
    ap.add_argument("--large-quick-warmup", action="store_true",
                    help="development only: ONE full-size warm-up pass at N > 1 GPUs instead of --warmup of them")
    ap.add_argument("--skip-replicas", action="store_true")
    ap.add_argument("--skip-predict-ci", action="store_true")
    ap.add_argument("--skip-predict-loop", action="store_true")
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
