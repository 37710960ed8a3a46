// extern "C" boundary (include/gprb200.h): handle state machine + drivers that sequence the kernels.
// Mirrors the control flow of the reference's GaussianProcess<T> (lib/GaussianProcess.cpp) —
// Initialize -> ComputeRegressionVectors (:642-672), Predict (:55-61), operator() (:85-99),
// GetCredibleInterval (:103-114), PredictDerivative (:65-81) — and of GaussianLogLikelihood
// (include/Likelihood.h:166-344), with device-resident state instead of Eigen matrices.
#include "common.cuh"
#include "kfunc.cuh"
#include <math.h>
#include <float.h>
#include <string.h>
#include <vector>
#include <new>
#include <atomic>
#include <thread>

using namespace gprb200;

namespace gprb200 {
// gram.cu extras
int kpair_values(Ctx* c, const double* Xa, const double* Xb, int64_t T, double* d_out);
int dgram_build_one(Ctx* c, double* out, int64_t ldo, int p_select);
__global__ void ci_finish_kernel(const double* kself, const double* q, int64_t T, double* out, int mode) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    const double v = kself[t] - q[t];
    out[t] = mode ? 2.0 * sqrt(fmax(0.0, v)) : v;  // mode 1: GetCredibleInterval, lib/GaussianProcess.cpp:112
}
// A(dd, n) = -(x[dd] - X[n][dd]) * kx[n] for dd < d ; A(d, n) = kx[n]   (lib/GaussianProcess.cpp:77-79, :697-706)
__global__ void deriv_lhs_kernel(const double* x, const double* X, const double* kx, int64_t N, int d, double* A,
                                 int64_t lda) {
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const double k = kx[n];
    for (int dd = 0; dd < d; ++dd) A[(int64_t)dd * lda + n] = -(x[dd] - X[n * d + dd]) * k;
    A[(int64_t)d * lda + n] = k;
}
}  // namespace gprb200

// the opaque C handle is the context itself

static thread_local std::string g_create_error;

namespace {

struct PhaseTimer {
    Ctx* c;
    int phase;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    PhaseTimer(Ctx* c_, int p) : c(c_), phase(p) {
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        cudaEventRecord(e0, c->stream);
    }
    // stop() synchronises on e1 only when asked for the number; learn() itself stays asynchronous
    void stop() { cudaEventRecord(e1, c->stream); }
    void resolve() {
        cudaEventSynchronize(e1);
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, e0, e1) == cudaSuccess) c->phase_ms[phase] = ms;
        release();
    }
    void release() {
        if (e0) cudaEventDestroy(e0);
        if (e1) cudaEventDestroy(e1);
        e0 = e1 = nullptr;
    }
    ~PhaseTimer() { release(); }   // early returns between construction and resolve() must not leak the events
    PhaseTimer(const PhaseTimer&) = delete;
    PhaseTimer& operator=(const PhaseTimer&) = delete;
};

template <typename T>
int dev_alloc(Ctx* c, T** p, size_t count) {
    if (*p) {
        cudaFree(*p);
        *p = nullptr;
    }
    if (count == 0) return 0;
    GP_CUDA(c, cudaMalloc((void**)p, sizeof(T) * count));
    return 0;
}

void free_model(Ctx* c) {
    cudaFree(c->dX); c->dX = nullptr;
    cudaFree(c->dY); c->dY = nullptr;
    cudaFree(c->dA); c->dA = nullptr;
    cudaFree(c->dL); c->dL = nullptr;
    cudaFree(c->dInvL); c->dInvL = nullptr;
    cudaFree(c->dC); c->dC = nullptr;
    cudaFree(c->dW); c->dW = nullptr;
    c->have_winv = false;
    c->have_train = c->have_factor = c->have_alpha = c->have_core = c->core_is_loaded = false;
}

int use_device(Ctx* c) {
    GP_CUDA(c, cudaSetDevice(c->device));
    return 0;
}

int h2d_2d(Ctx* c, double* dst, int64_t ldd, const double* src, int64_t rows, int64_t cols) {
    if (rows <= 0 || cols <= 0) return 0;
    GP_CUDA(c, cudaMemcpy2DAsync(dst, sizeof(double) * ldd, src, sizeof(double) * cols, sizeof(double) * cols, rows,
                                 cudaMemcpyHostToDevice, c->stream));
    return 0;
}
int d2h_2d(Ctx* c, double* dst, const double* src, int64_t lds, int64_t rows, int64_t cols) {
    if (rows <= 0 || cols <= 0) return 0;
    GP_CUDA(c, cudaMemcpy2DAsync(dst, sizeof(double) * cols, src, sizeof(double) * lds, sizeof(double) * cols, rows,
                                 cudaMemcpyDeviceToHost, c->stream));
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

struct TempBuf {  // stream-ordered scratch
    Ctx* c;
    double* p = nullptr;
    explicit TempBuf(Ctx* c_) : c(c_) {}
    int alloc(size_t count) {
        GP_CUDA(c, cudaMallocAsync((void**)&p, sizeof(double) * (count ? count : 1), c->stream));
        return 0;
    }
    ~TempBuf() {
        if (p) cudaFreeAsync(p, c->stream);
    }
};

int require(Ctx* c, bool cond, const char* msg) { return cond ? 0 : fail(c, -2, msg); }

constexpr int64_t STREAM_IO_MIN_BYTES = 64ll << 20;

// Yrows != nullptr: one host pointer per label vector (the reference's std::vector<VectorType>), Y ignored
// defer_labels: allocate the label matrix but leave the upload to the caller (gprb200_learn_rows streams it in column blocks)
int install_train(Ctx* c, const double* X, int64_t N, int d, const double* Y, int64_t M, double sigma, bool on_device,
                  const double* const* Yrows = nullptr, bool defer_labels = false) {
    if (Yrows) Y = Yrows[0];   // "labels present"
    GP_TRY(require(c, N > 0 && d > 0 && X != nullptr, "set_train: need N > 0 samples of dimension d > 0"));
    GP_TRY(require(c, N < (1ll << 31) - 256 && M < (1ll << 31) - 256, "set_train: N and M must fit in 31 bits"));
    // Same shapes as the resident model (the hyper-parameter loop, repeated Learn calls): keep every device buffer.
    // cudaFree / cudaMalloc of the 8.6 GB factor cost up to ~0.45 s per call at N = 32768 (tools/e2e_probe.py).
    const bool reuse = c->have_train && c->N == N && c->d == d && c->M == (Y ? M : 0) && c->dX &&
                       ((Y && M > 0) == (c->dY != nullptr));
    if (!reuse) free_model(c);
    c->have_factor = c->have_alpha = c->have_core = c->core_is_loaded = false;
    sparse_invalidate(c);
    c->N = N; c->d = d; c->M = (Y ? M : 0); c->sigma = sigma;
    c->ldy = round_up(c->M > 0 ? c->M : 1, 16);
    c->ldk = round_up(N, 16);
    if (!reuse) GP_TRY(dev_alloc(c, &c->dX, (size_t)N * d));
    const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    GP_CUDA(c, cudaMemcpyAsync(c->dX, X, sizeof(double) * N * d, kind, c->stream));
    if (Y && M > 0) {
        if (!reuse) GP_TRY(dev_alloc(c, &c->dY, (size_t)N * c->ldy));
        if (c->ldy != M) GP_CUDA(c, cudaMemsetAsync(c->dY, 0, sizeof(double) * N * c->ldy, c->stream));
        const bool big = !on_device && c->opt_stream_io && (int64_t)sizeof(double) * N * M >= STREAM_IO_MIN_BYTES;
        if (defer_labels) {
        } else if (Yrows || big) {
            GP_TRY(stream_h2d(c, c->dY, c->ldy, Yrows, Yrows ? nullptr : Y, M, N, M));
        } else {
            GP_CUDA(c, cudaMemcpy2DAsync(c->dY, sizeof(double) * c->ldy, Y, sizeof(double) * M, sizeof(double) * M, N, kind,
                                         c->stream));
        }
    }
    c->have_train = true;
    return 0;
}

int ensure_factor(Ctx* c) {
    if (c->have_factor) return 0;
    GP_TRY(require(c, c->have_train, "factorize: no training samples (set_train first)"));
    GP_TRY(require(c, c->have_kernel, "factorize: no kernel (set_kernel first)"));
    const int64_t N = c->N;
    if (!c->dL) GP_TRY(dev_alloc(c, &c->dL, (size_t)N * c->ldk));
    const int64_t nblk = (N + NB - 1) / NB;
    if (!c->dInvL) GP_TRY(dev_alloc(c, &c->dInvL, (size_t)nblk * NB * NB));
    {
        PhaseTimer t(c, GPRB200_T_GRAM);
        GP_CUDA(c, cudaMemsetAsync(c->dInfo + 1, 0, sizeof(int), c->stream));
        GP_TRY(gram_build(c, c->dX, N, c->dX, N, c->dL, c->ldk, true, c->opt_gram_lower_only != 0,
                          c->sigma * c->sigma, c->dInfo + 1));
        t.stop();
        int bad = 0;  // the finite check (lib/GaussianProcess.cpp:399) is fused into the Gram kernel
        GP_CUDA(c, cudaMemcpyAsync(&bad, c->dInfo + 1, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
        GP_CUDA(c, cudaStreamSynchronize(c->stream));
        t.resolve();
        if (bad)
            return fail(c, -3,
                        "GaussianProcess::ComputeKernelMatrixInternal: kernel matrix contains entries which are not "
                        "finite.");
    }
    int info = 0;
    {
        PhaseTimer t(c, GPRB200_T_POTRF);
        GP_TRY(potrf_blocked(c, c->dL, c->ldk, N, c->dInvL, &info));
        t.stop();
        t.resolve();
    }
    if (info != 0)
        return fail(c, info,
                    "potrf: K + sigma^2 I is not positive definite (leading minor of order " + std::to_string(info) +
                        ")");
    GP_TRY(logdet_from_factor(c, c->dL, c->ldk, N, &c->logdet));
    c->have_factor = true;
    c->have_winv = false;
    c->ci_small_calls = 0;
    if (!c->core_is_loaded) c->have_core = false;
    return 0;
}

int ensure_alpha(Ctx* c) {
    if (c->have_alpha) return 0;
    GP_TRY(ensure_factor(c));
    GP_TRY(require(c, c->dY != nullptr && c->M > 0, "solve_alpha: no label vectors"));
    if (!c->dA) GP_TRY(dev_alloc(c, &c->dA, (size_t)c->N * c->ldy));
    PhaseTimer t(c, GPRB200_T_SOLVE);
    GP_CUDA(c, cudaMemcpyAsync(c->dA, c->dY, sizeof(double) * c->N * c->ldy, cudaMemcpyDeviceToDevice, c->stream));
    GP_TRY(trsm_lower_fwd(c, c->dL, c->ldk, c->dInvL, c->N, c->dA, c->ldy, c->M));
    GP_TRY(trsm_lower_bwd(c, c->dL, c->ldk, c->dInvL, c->N, c->dA, c->ldy, c->M));
    t.stop();
    t.resolve();
    c->have_alpha = true;
    return 0;
}

int ensure_core(Ctx* c) {
    if (c->have_core) return 0;
    GP_TRY(ensure_factor(c));
    if (!c->dC) GP_TRY(dev_alloc(c, &c->dC, (size_t)c->N * c->ldk));
    PhaseTimer t(c, GPRB200_T_INVERSE);
    GP_TRY(inverse_from_factor(c, c->dL, c->ldk, c->dInvL, c->N, c->dC, c->ldk));
    t.stop();
    t.resolve();
    c->have_core = true;
    c->core_is_loaded = false;
    return 0;
}

void invalidate(Ctx* c) {
    c->have_factor = false;
    c->have_alpha = false;
    c->have_core = false;
    c->core_is_loaded = false;
    sparse_invalidate(c);
}

// q[t] = Kx_t^T (K + s^2 I)^-1 Ky_t for T pairs; dKa/dKb: T x N cross-Gram (ld ldkx), may alias.
int quad_forms(Ctx* c, const double* dXa, const double* dXb, int64_t T, double* d_q) {
    const int64_t N = c->N;
    const bool same = (dXa == dXb);
    if (c->have_core && !c->have_factor) {
        // loaded model (Load keeps the stored core matrix, lib/GaussianProcess.cpp:208): q = rowdot(Ka C, Kb)
        const int64_t ldkx = round_up(N, 16);
        TempBuf Ka(c), Kb(c), W(c);
        GP_TRY(Ka.alloc((size_t)T * ldkx));
        GP_TRY(W.alloc((size_t)T * ldkx));
        GP_TRY(gram_build(c, dXa, T, c->dX, N, Ka.p, ldkx, false, false, 0.0));
        const double* kb = Ka.p;
        if (!same) {
            GP_TRY(Kb.alloc((size_t)T * ldkx));
            GP_TRY(gram_build(c, dXb, T, c->dX, N, Kb.p, ldkx, false, false, 0.0));
            kb = Kb.p;
        }
        GemmArgs g{};
        g.A = Ka.p; g.lda = ldkx; g.B = c->dC; g.ldb = c->ldk; g.C = W.p; g.ldc = ldkx;
        g.M = (int)T; g.N = (int)N; g.K = (int)N; g.alpha = 1.0; g.beta = 0.0;
        GP_TRY(gemm(c, GEMM_NN, g));
        GP_TRY(rowdot(c, W.p, ldkx, kb, ldkx, T, N, d_q));
        return 0;
    }
    GP_TRY(ensure_factor(c));
    if (T <= 2 && c->opt_ci_winv && N > 2 * NB) {
        // One or two test points per call (the loop of apps/GaussianProcessPredict.cpp:185-193): the forward
        // substitution below is a latency-bound chain of N / 128 dependent steps per call. From the third such call
        // on, inv(L) is formed once (N^3 / 3 flop) and every interval becomes one row-parallel, HBM-bound triangular
        // matrix-vector product (trmv_lower_kernel).
        ++c->ci_small_calls;
        if (!c->have_winv && (c->opt_ci_winv >= 2 || c->ci_small_calls >= 3)) {
            if (!c->dW) GP_TRY(dev_alloc(c, &c->dW, (size_t)N * c->ldk));
            GP_TRY(trtri_cols(c, c->dL, c->ldk, c->dInvL, N, 512, 0, 1, c->dW, c->ldk));
            c->have_winv = true;
        }
        if (c->have_winv) {
            const int64_t ldkx = round_up(N, 16);
            TempBuf Kx(c), V(c);
            GP_TRY(Kx.alloc((size_t)2 * T * ldkx));
            GP_TRY(V.alloc((size_t)N * 4));
            GP_TRY(gram_build(c, dXa, T, c->dX, N, Kx.p, ldkx, false, false, 0.0));
            GP_TRY(trmv_lower(c, c->dW, c->ldk, N, Kx.p, ldkx, (int)T, V.p, 2));
            const double* vb = V.p;
            if (!same) {
                GP_TRY(gram_build(c, dXb, T, c->dX, N, Kx.p + T * ldkx, ldkx, false, false, 0.0));
                GP_TRY(trmv_lower(c, c->dW, c->ldk, N, Kx.p + T * ldkx, ldkx, (int)T, V.p + 2 * N, 2));
                vb = V.p + 2 * N;
            }
            GP_TRY(coldot(c, V.p, 2, vb, 2, N, T, d_q));
            return 0;
        }
    }
    // V = L^-1 Kx^T with Kx^T generated directly as N x T; q = column dots
    const int64_t ldv = round_up(T, 16);
    TempBuf Va(c), Vb(c);
    GP_TRY(Va.alloc((size_t)N * ldv));
    GP_TRY(gram_build(c, c->dX, N, dXa, T, Va.p, ldv, false, false, 0.0));
    GP_TRY(trsm_lower_fwd(c, c->dL, c->ldk, c->dInvL, N, Va.p, ldv, T));
    const double* vb = Va.p;
    if (!same) {
        GP_TRY(Vb.alloc((size_t)N * ldv));
        GP_TRY(gram_build(c, c->dX, N, dXb, T, Vb.p, ldv, false, false, 0.0));
        GP_TRY(trsm_lower_fwd(c, c->dL, c->ldk, c->dInvL, N, Vb.p, ldv, T));
        vb = Vb.p;
    }
    GP_TRY(coldot(c, Va.p, ldv, vb, ldv, N, T, d_q));
    return 0;
}

int predict_impl(Ctx* c, const double* dXs, int64_t T, double* d_mean, int64_t ld_mean, double* d_ci) {
    GP_TRY(require(c, c->have_kernel, "predict: no kernel"));
    GP_TRY(require(c, c->have_train, "predict: gaussian process is not initialized"));
    if (d_mean && !c->have_alpha) GP_TRY(ensure_alpha(c));   // the credible interval alone needs only the factor
    const int64_t N = c->N, M = c->M;
    // Predict stays asynchronous (a per-frame call must not cost a host synchronisation): the phase time is taken
    // with two persistent events and resolved only when gprb200_phase_ms asks for it.
    if (!c->ev0) {
        GP_CUDA(c, cudaEventCreate(&c->ev0));
        GP_CUDA(c, cudaEventCreate(&c->ev1));
    }
    GP_CUDA(c, cudaEventRecord(c->ev0, c->stream));
    if (d_mean) {
        const int64_t ldkx = round_up(N, 16);
        TempBuf Kx(c);
        GP_TRY(Kx.alloc((size_t)T * ldkx));
        GP_TRY(gram_build(c, dXs, T, c->dX, N, Kx.p, ldkx, false, false, 0.0));
        if (T <= 2) {
            for (int64_t i = 0; i < T; ++i)
                GP_TRY(gemv_t(c, Kx.p + i * ldkx, c->dA, c->ldy, N, M, d_mean + i * ld_mean));
        } else {
            const bool direct = ((ld_mean & 1) == 0) && ((reinterpret_cast<uintptr_t>(d_mean) & 15) == 0);
            TempBuf tmp(c);
            double* out = d_mean;
            int64_t ldo = ld_mean;
            if (!direct) {
                ldo = round_up(M, 16);
                GP_TRY(tmp.alloc((size_t)T * ldo));
                out = tmp.p;
            }
            GemmArgs g{};
            g.A = Kx.p; g.lda = ldkx; g.B = c->dA; g.ldb = c->ldy; g.C = out; g.ldc = ldo;
            g.M = (int)T; g.N = (int)M; g.K = (int)N; g.alpha = 1.0; g.beta = 0.0;
            GP_TRY(gemm(c, GEMM_NN, g));
            if (!direct) GP_TRY(copy2d(c, d_mean, ld_mean, out, ldo, T, M));
        }
    }
    if (d_ci) {
        TempBuf q(c), ks(c);
        GP_TRY(q.alloc((size_t)T));
        GP_TRY(ks.alloc((size_t)T));
        GP_TRY(quad_forms(c, dXs, dXs, T, q.p));
        GP_TRY(kpair_values(c, dXs, dXs, T, ks.p));
        ci_finish_kernel<<<(unsigned)((T + 255) / 256), 256, 0, c->stream>>>(ks.p, q.p, T, d_ci, 1);
        ++c->launches;
        GP_CUDA(c, cudaGetLastError());
    }
    GP_CUDA(c, cudaEventRecord(c->ev1, c->stream));
    c->predict_time_pending = true;
    return 0;
}

}  // namespace

// =============================================================================================
extern "C" {

const char* gprb200_version(void) { return "gprb200 0.1 (sm_100a)"; }

int gprb200_create(gprb200_handle* h, int device) {
    if (!h) return -2;
    *h = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        g_create_error = std::string("gprb200_create: no CUDA device (") + cudaGetErrorString(e) +
                         "); this library has no CPU fallback";
        return -1;
    }
    if (device < 0 || device >= ndev) {
        g_create_error = "gprb200_create: invalid device index";
        return -2;
    }
    gprb200_ctx* c = new (std::nothrow) gprb200_ctx();
    if (!c) return -1;
    c->device = device;
    cudaSetDevice(device);
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    if (prop.major != 10) {
        g_create_error = std::string("gprb200_create: device '") + prop.name +
                         "' is not sm_100; this library is built for B200 (sm_100a) only";
        delete c;
        return -1;
    }
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
        g_create_error = "gprb200_create: cudaStreamCreate failed";
        delete c;
        return -1;
    }
    c->own_stream = true;
    {
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        if (cudaStreamCreateWithPriority(&c->panel_stream, cudaStreamNonBlocking, hi) != cudaSuccess)
            c->panel_stream = nullptr;  // look-ahead is then disabled
        if (cudaStreamCreateWithPriority(&c->aux_stream, cudaStreamNonBlocking, hi) != cudaSuccess) c->aux_stream = nullptr;
        if (cudaEventCreateWithFlags(&c->ev_pcrit, cudaEventDisableTiming) != cudaSuccess) c->ev_pcrit = nullptr;
        if (cudaEventCreateWithFlags(&c->ev_paux, cudaEventDisableTiming) != cudaSuccess) c->ev_paux = nullptr;
    }
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
        uint64_t thr = UINT64_MAX;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    }
    cudaMalloc((void**)&c->dInfo, sizeof(int) * 4);
    cudaMalloc((void**)&c->dScratch, 4096);
    c->scratch_bytes = 4096;
    *h = c;
    return 0;
}

int gprb200_destroy(gprb200_handle h) {
    if (!h) return 0;
    gprb200_ctx* c = h;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    sparse_destroy(c);
    hostio_destroy(c);
    free_model(c);
    cudaFree(c->dInfo);
    cudaFree(c->dScratch);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->own_stream) cudaStreamDestroy(c->stream);
    if (c->panel_stream) cudaStreamDestroy(c->panel_stream);
    if (c->aux_stream) cudaStreamDestroy(c->aux_stream);
    if (c->ev_pcrit) cudaEventDestroy(c->ev_pcrit);
    if (c->ev_paux) cudaEventDestroy(c->ev_paux);
    delete c;
    return 0;
}

const char* gprb200_last_error(gprb200_handle h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int gprb200_set_stream(gprb200_handle h, void* cuda_stream) {
    if (!h) return -2;
    GP_TRY(use_device(h));
    cudaStreamSynchronize(h->stream);
    if (h->own_stream) cudaStreamDestroy(h->stream);
    h->stream = (cudaStream_t)cuda_stream;
    h->own_stream = false;
    return 0;
}

int gprb200_synchronize(gprb200_handle h) {
    if (!h) return -2;
    GP_TRY(use_device(h));
    GP_CUDA(h, cudaStreamSynchronize(h->stream));
    return 0;
}

int64_t gprb200_launch_count(gprb200_handle h) { return h ? h->launches : -1; }

int gprb200_phase_ms(gprb200_handle h, double* ms) {
    if (!h || !ms) return -2;
    if (h->predict_time_pending && h->ev1) {
        float t = 0.f;
        if (cudaEventSynchronize(h->ev1) == cudaSuccess && cudaEventElapsedTime(&t, h->ev0, h->ev1) == cudaSuccess)
            h->phase_ms[GPRB200_T_PREDICT] = t;
        h->predict_time_pending = false;
    }
    for (int i = 0; i < GPRB200_T_COUNT; ++i) ms[i] = h->phase_ms[i];
    return 0;
}

int gprb200_get_stat(gprb200_handle h, const char* name, double* out) {
    if (!h || !name || !out) return -2;
    struct { const char* n; double v; } st[] = {
        {"tb_ms", h->prof_tb_ms}, {"tb_flops", h->prof_tb_flops}, {"tb_launches", h->prof_tb_launches},
        {"gemv_ms", h->prof_gemv_ms}, {"gemv_bytes", h->prof_gemv_bytes}, {"gemv_launches", h->prof_gemv_launches},
        {"logdet", h->logdet}};
    for (auto& e : st)
        if (!strcmp(name, e.n)) {
            *out = e.v;
            return 0;
        }
    return fail(h, -2, std::string("get_stat: unknown statistic '") + name + "'");
}

int gprb200_set_option(gprb200_handle h, const char* name, int64_t value) {
    if (!h || !name) return -2;
    struct { const char* n; int64_t* p; } opts[] = {
        {"gram_lower_only", &h->opt_gram_lower_only}, {"gram_mma", &h->opt_gram_mma}, {"gram_mma_min_d", &h->opt_gram_mma_min_d}, {"gram_terms", &h->opt_gram_terms}, {"stream_io", &h->opt_stream_io}, {"nbo", &h->opt_nbo}, {"lookahead", &h->opt_lookahead}, {"chol_left", &h->opt_chol_left},
        {"gemm_stages", &h->opt_gemm_stages}, {"gemm_tma", &h->opt_gemm_tma}, {"gemm_tma_min_k", &h->opt_gemm_tma_min_k}, {"gemm_tma_min_tiles", &h->opt_gemm_tma_min_tiles}, {"ci_winv", &h->opt_ci_winv}, {"panel_split", &h->opt_panel_split}, {"gemm_small", &h->opt_gemm_small}, {"gemm_variant", &h->opt_gemm_variant}, {"skinny", &h->opt_skinny}, {"potrf_tile", &h->opt_potrf_tile}, {"solve_left_min_m", &h->opt_solve_left_min_m}, {"inverse_long_k", &h->opt_inverse_long_k}, {"trsv_chain", &h->opt_trsv_chain},
        {"profile", &h->opt_profile}};
    if (!strcmp(name, "profile")) h->prof_gemv_ms = h->prof_gemv_bytes = h->prof_gemv_launches = 0.0;
    for (auto& o : opts)
        if (!strcmp(name, o.n)) {
            *o.p = value;
            return 0;
        }
    return fail(h, -2, std::string("set_option: unknown option '") + name + "'");
}

int gprb200_set_kernel(gprb200_handle h, const int32_t* ops, int nops, const double* params, int nparams) {
    if (!h) return -2;
    Ctx* c = h;
    GP_TRY(require(c, ops && nops > 0 && nops <= GPRB200_MAX_OPS, "set_kernel: 1..GPRB200_MAX_OPS ops required"));
    GP_TRY(require(c, nparams >= 0 && nparams <= GPRB200_MAX_PARAMS && (params || nparams == 0),
                   "set_kernel: too many kernel parameters"));
    KProg p{};
    p.nops = nops;
    p.nparams = nparams;
    for (int i = 0; i < nops; ++i) p.op[i] = (int8_t)ops[i];
    for (int i = 0; i < nparams; ++i) p.par[i] = params[i];
    DevProg tmp;
    GP_TRY(lower_program(p, &tmp, &c->err));
    c->prog = p;
    c->have_kernel = true;
    invalidate(c);
    return 0;
}

int gprb200_set_kernel_params(gprb200_handle h, const double* params, int nparams) {
    if (!h) return -2;
    Ctx* c = h;
    GP_TRY(require(c, c->have_kernel, "set_kernel_params: no kernel program installed"));
    GP_TRY(require(c, nparams == c->prog.nparams && params, "set_kernel_params: wrong number of parameters"));
    for (int i = 0; i < nparams; ++i) c->prog.par[i] = params[i];
    invalidate(c);
    return 0;
}

int gprb200_set_train(gprb200_handle h, const double* X, int64_t N, int d, const double* Y, int64_t M, double sigma) {
    if (!h) return -2;
    GP_TRY(use_device(h));
    GP_TRY(install_train(h, X, N, d, Y, M, sigma, false));
    GP_CUDA(h, cudaStreamSynchronize(h->stream));  // host buffers may be reused by the caller
    return 0;
}

int gprb200_set_train_rows(gprb200_handle h, const double* const* Xrows, int64_t N, int d, const double* const* Yrows,
                           int64_t M, double sigma) {
    if (!h) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, Xrows != nullptr && N > 0 && d > 0, "set_train_rows: need N > 0 samples of dimension d > 0"));
    std::vector<double> X((size_t)N * d);   // the samples are small (N x d); only the labels are streamed
    for (int64_t i = 0; i < N; ++i) {
        GP_TRY(require(c, Xrows[i] != nullptr && (!Yrows || M <= 0 || Yrows[i] != nullptr), "set_train_rows: null row pointer"));
        memcpy(X.data() + i * d, Xrows[i], sizeof(double) * d);
    }
    GP_TRY(install_train(c, X.data(), N, d, nullptr, M, sigma, false, (Yrows && M > 0) ? Yrows : nullptr));
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

// Learn straight from the reference's storage to the host result, pipelined over column blocks of the label matrix:
//   uploader thread : label rows -> pinned ring 0 -> dY block            (copy stream 0)
//   this thread     : K = LL^T once, then per block  alpha_b = L^-T L^-1 Y_b    (compute stream)
//   downloader thread: alpha block -> pinned ring 1 -> alpha_out           (copy stream 1)
// The output columns are independent (lib/GaussianProcess.cpp:661 is one GEMM over all of them), so block b+1 crosses
// PCIe one way and block b-1 the other way while block b is solved; Y and alpha stay resident for Predict.
int gprb200_learn_rows(gprb200_handle h, const double* const* Xrows, int64_t N, int d, const double* const* Yrows, int64_t M,
                       double sigma, double* alpha_out) {
    if (!h) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, Xrows && Yrows && N > 0 && d > 0 && M > 0, "learn_rows: need samples and label vectors"));
    std::vector<double> X((size_t)N * d);
    for (int64_t i = 0; i < N; ++i) {
        GP_TRY(require(c, Xrows[i] != nullptr && Yrows[i] != nullptr, "learn_rows: null row pointer"));
        memcpy(X.data() + i * d, Xrows[i], sizeof(double) * d);
    }
    GP_TRY(install_train(c, X.data(), N, d, nullptr, M, sigma, false, Yrows, true));
    for (int r = 0; r < 2; ++r)
        if (!c->io_stream[r]) GP_CUDA(c, cudaStreamCreateWithFlags(&c->io_stream[r], cudaStreamNonBlocking));
    if (!c->dA) GP_TRY(dev_alloc(c, &c->dA, (size_t)N * c->ldy));
    // column blocks of ~512 MB
    int64_t wc = ((512ll << 20) / (8 * N)) / 16 * 16;
    if (wc < 16) wc = 16;
    if (wc > M) wc = M;
    const int64_t B = (M + wc - 1) / wc;
    std::vector<cudaEvent_t> evU((size_t)B), evS((size_t)B);
    for (int64_t b = 0; b < B; ++b) {
        GP_CUDA(c, cudaEventCreateWithFlags(&evU[(size_t)b], cudaEventDisableTiming));
        GP_CUDA(c, cudaEventCreateWithFlags(&evS[(size_t)b], cudaEventDisableTiming));
    }
    std::atomic<int64_t> uploaded(0), solved(0);
    std::atomic<int> rc_up(0), rc_down(0);
    GP_CUDA(c, cudaStreamSynchronize(c->stream));   // dY allocated and padded before the copy stream writes into it
    std::thread up([&] {
        cudaSetDevice(c->device);
        for (int64_t b = 0; b < B; ++b) {
            const int64_t c0 = b * wc, w = (M - c0 < wc) ? M - c0 : wc;
            int r = stream_h2d(c, c->dY + c0, c->ldy, Yrows, nullptr, 0, N, w, c->io_stream[0], c0);
            if (r == 0 && cudaEventRecord(evU[(size_t)b], c->io_stream[0]) != cudaSuccess) r = -1;
            if (r != 0) {
                rc_up.store(r);
                uploaded.store(B);   // unblock the solver; it checks rc_up
                return;
            }
            uploaded.store(b + 1);
        }
    });
    std::thread down([&] {
        cudaSetDevice(c->device);
        if (!alpha_out) return;
        for (int64_t b = 0; b < B; ++b) {
            while (solved.load() <= b) std::this_thread::yield();
            if (solved.load() > B) return;   // the solver failed
            const int64_t c0 = b * wc, w = (M - c0 < wc) ? M - c0 : wc;
            int r = (cudaStreamWaitEvent(c->io_stream[1], evS[(size_t)b], 0) == cudaSuccess) ? 0 : -1;
            if (r == 0) r = stream_d2h(c, alpha_out + c0, M, c->dA + c0, c->ldy, N, w, c->io_stream[1]);
            if (r != 0) {
                rc_down.store(r);
                return;
            }
        }
    });
    int rc = ensure_factor(c);   // Gram + Cholesky while the first block is on its way
    for (int64_t b = 0; b < B && rc == 0; ++b) {
        while (uploaded.load() <= b) std::this_thread::yield();
        if (rc_up.load() != 0) {
            rc = rc_up.load();
            break;
        }
        const int64_t c0 = b * wc, w = (M - c0 < wc) ? M - c0 : wc;
        if (cudaStreamWaitEvent(c->stream, evU[(size_t)b], 0) != cudaSuccess ||
            cudaMemcpy2DAsync(c->dA + c0, sizeof(double) * c->ldy, c->dY + c0, sizeof(double) * c->ldy, sizeof(double) * w,
                              (size_t)N, cudaMemcpyDeviceToDevice, c->stream) != cudaSuccess) {
            rc = fail(c, -1, "learn_rows: CUDA error while staging a label block");
            break;
        }
        rc = trsm_lower_fwd(c, c->dL, c->ldk, c->dInvL, N, c->dA + c0, c->ldy, w);
        if (rc == 0) rc = trsm_lower_bwd(c, c->dL, c->ldk, c->dInvL, N, c->dA + c0, c->ldy, w);
        if (rc == 0 && cudaEventRecord(evS[(size_t)b], c->stream) != cudaSuccess) rc = fail(c, -1, "learn_rows: event record failed");
        if (rc == 0) solved.store(b + 1);
    }
    if (rc != 0) solved.store(B + 1);   // release the downloader
    up.join();
    down.join();
    cudaStreamSynchronize(c->io_stream[0]);
    cudaStreamSynchronize(c->io_stream[1]);
    cudaStreamSynchronize(c->stream);
    for (int64_t b = 0; b < B; ++b) {
        cudaEventDestroy(evU[(size_t)b]);
        cudaEventDestroy(evS[(size_t)b]);
    }
    if (rc == 0 && rc_up.load() != 0) rc = rc_up.load();
    if (rc == 0 && rc_down.load() != 0) rc = rc_down.load();
    if (rc != 0) return rc;
    c->have_alpha = true;
    return 0;
}

int gprb200_set_train_dev(gprb200_handle h, const double* dX, int64_t N, int d, const double* dY, int64_t M,
                          double sigma) {
    if (!h) return -2;
    GP_TRY(use_device(h));
    return install_train(h, dX, N, d, dY, M, sigma, true);
}

int gprb200_set_sigma(gprb200_handle h, double sigma) {
    if (!h) return -2;
    h->sigma = sigma;
    invalidate(h);
    return 0;
}

int gprb200_set_state(gprb200_handle h, const double* X, int64_t N, int d, const double* Y, const double* alpha,
                      int64_t M, const double* core, double sigma) {
    if (!h) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, alpha != nullptr && M > 0, "set_state: regression vectors required"));
    GP_TRY(install_train(c, X, N, d, Y, M, sigma, false));
    c->M = M;
    c->ldy = round_up(M, 16);
    GP_TRY(dev_alloc(c, &c->dA, (size_t)N * c->ldy));
    if (c->ldy != M) GP_CUDA(c, cudaMemsetAsync(c->dA, 0, sizeof(double) * N * c->ldy, c->stream));
    GP_TRY(h2d_2d(c, c->dA, c->ldy, alpha, N, M));
    c->have_alpha = true;
    if (core) {
        GP_TRY(dev_alloc(c, &c->dC, (size_t)N * c->ldk));
        GP_TRY(h2d_2d(c, c->dC, c->ldk, core, N, N));
        c->have_core = true;
        c->core_is_loaded = true;
    }
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

int gprb200_gram_dev(gprb200_handle h, double* dK, int64_t ldk, int add_noise) {
    if (!h) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, c->have_train && c->have_kernel, "gram: need samples and a kernel"));
    PhaseTimer t(c, GPRB200_T_GRAM);
    GP_TRY(gram_build(c, c->dX, c->N, c->dX, c->N, dK, ldk, true, false, add_noise ? c->sigma * c->sigma : 0.0));
    t.stop();
    t.resolve();
    return 0;
}

int gprb200_gram(gprb200_handle h, double* K_out, int add_noise) {
    if (!h) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, c->have_train && c->have_kernel && K_out, "gram: need samples and a kernel"));
    TempBuf K(c);
    GP_TRY(K.alloc((size_t)c->N * c->ldk));
    GP_TRY(gprb200_gram_dev(h, K.p, c->ldk, add_noise));
    int bad = 0;
    GP_TRY(check_finite(c, K.p, c->N, c->N, c->ldk, false, &bad));
    GP_TRY(d2h_2d(c, K_out, K.p, c->ldk, c->N, c->N));
    if (bad)
        return fail(c, -3,
                    "GaussianProcess::ComputeKernelMatrixInternal: kernel matrix contains entries which are not finite.");
    return 0;
}

int gprb200_factorize(gprb200_handle h) {
    if (!h) return -2;
    GP_TRY(use_device(h));
    h->have_factor = false;
    h->have_alpha = false;
    if (!h->core_is_loaded) h->have_core = false;
    return ensure_factor(h);
}

int gprb200_solve_alpha(gprb200_handle h) {
    if (!h) return -2;
    GP_TRY(use_device(h));
    h->have_alpha = false;
    return ensure_alpha(h);
}

int gprb200_learn(gprb200_handle h, unsigned flags) {
    if (!h) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    invalidate(c);
    GP_TRY(ensure_factor(c));
    GP_TRY(ensure_alpha(c));
    if (flags & GPRB200_WANT_CORE) GP_TRY(ensure_core(c));
    return 0;
}

int gprb200_get_alpha(gprb200_handle h, double* alpha) {
    if (!h || !alpha) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(ensure_alpha(c));
    if (c->opt_stream_io && (int64_t)sizeof(double) * c->N * c->M >= STREAM_IO_MIN_BYTES)
        return stream_d2h(c, alpha, c->M, c->dA, c->ldy, c->N, c->M);
    return d2h_2d(c, alpha, c->dA, c->ldy, c->N, c->M);
}

int gprb200_get_alpha_dev(gprb200_handle h, const double** d_alpha, int64_t* ld) {
    if (!h || !d_alpha || !ld) return -2;
    GP_TRY(use_device(h));
    GP_TRY(ensure_alpha(h));
    *d_alpha = h->dA;
    *ld = h->ldy;
    return 0;
}

int gprb200_get_factor_dev(gprb200_handle h, const double** d_L, int64_t* ld, const double** d_invL) {
    if (!h || !d_L || !ld) return -2;
    GP_TRY(use_device(h));
    GP_TRY(ensure_factor(h));
    *d_L = h->dL;
    *ld = h->ldk;
    if (d_invL) *d_invL = h->dInvL;
    return 0;
}

int gprb200_dgram_contract_rows_dev(gprb200_handle h, int64_t row0, int64_t nrows, int64_t ncols, const double* dS,
                                    int64_t lds, double* out) {
    if (!h || !dS || !out) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, c->have_train && c->have_kernel, "dgram_contract_rows: need samples and a kernel"));
    GP_TRY(require(c, row0 >= 0 && nrows > 0 && row0 + nrows <= c->N && ncols > 0 && ncols <= c->N && lds >= ncols,
                   "dgram_contract_rows: rows / columns outside the training set or lds < ncols"));
    return dgram_contract_xy(c, c->dX + row0 * c->d, nrows, c->dX, ncols, dS, lds, out);
}

int gprb200_get_core(gprb200_handle h, double* C) {
    if (!h || !C) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(ensure_core(c));
    return d2h_2d(c, C, c->dC, c->ldk, c->N, c->N);
}

int gprb200_get_factor(gprb200_handle h, double* L) {
    if (!h || !L) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(ensure_factor(c));
    TempBuf t(c);
    GP_TRY(t.alloc((size_t)c->N * c->ldk));
    GP_TRY(copy2d(c, t.p, c->ldk, c->dL, c->ldk, c->N, c->N));
    GP_TRY(zero_upper(c, t.p, c->ldk, c->N));
    return d2h_2d(c, L, t.p, c->ldk, c->N, c->N);
}

int gprb200_get_logdet(gprb200_handle h, double* logdet) {
    if (!h || !logdet) return -2;
    GP_TRY(use_device(h));
    GP_TRY(ensure_factor(h));
    *logdet = h->logdet;
    return 0;
}

int gprb200_spd_inverse(gprb200_handle h, const double* A, int64_t n, double* Ainv, double* logdet) {
    if (!h || !A || !Ainv || n <= 0) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    const int64_t ld = round_up(n, 16), nblk = (n + NB - 1) / NB;
    TempBuf L(c), invL(c), C(c);
    GP_TRY(L.alloc((size_t)n * ld));
    GP_TRY(invL.alloc((size_t)nblk * NB * NB));
    GP_TRY(C.alloc((size_t)n * ld));
    GP_TRY(h2d_2d(c, L.p, ld, A, n, n));
    int info = 0;
    GP_TRY(potrf_blocked(c, L.p, ld, n, invL.p, &info));
    if (info != 0)
        return fail(c, info, "InvertKernelMatrix: matrix is not positive definite (leading minor of order " +
                                 std::to_string(info) + ")");
    if (logdet) GP_TRY(logdet_from_factor(c, L.p, ld, n, logdet));
    GP_TRY(inverse_from_factor(c, L.p, ld, invL.p, n, C.p, ld));
    return d2h_2d(c, Ainv, C.p, ld, n, n);
}

int gprb200_gemm_dev(gprb200_handle h, int form, int64_t M, int64_t N, int64_t K, double alpha, const double* dA,
                     int64_t lda, const double* dB, int64_t ldb, double beta, double* dC, int64_t ldc, int lower) {
    if (!h) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, form >= 0 && form <= 3 && M >= 0 && N >= 0 && K >= 0 && M < (1ll << 31) && N < (1ll << 31) &&
                          K < (1ll << 31), "gemm_dev: bad form or extents"));
    GemmArgs g{};
    g.A = dA; g.lda = lda; g.B = dB; g.ldb = ldb; g.C = dC; g.ldc = ldc;
    g.M = (int)M; g.N = (int)N; g.K = (int)K; g.alpha = alpha; g.beta = beta; g.lower = lower ? 1 : 0;
    if (form == 3) {    // NT with a lower-triangular B: the k loop of a tile stops at its last column
        GP_TRY(require(c, !lower, "gemm_dev: form 3 has no lower-only variant"));
        g.ktri = 4;
        return gemm(c, GEMM_NT, g);
    }
    return gemm(c, (GemmForm)form, g);
}

int gprb200_gemm_stair_dev(gprb200_handle h, int64_t M, int64_t N, int64_t K, double alpha, const double* dA, int64_t lda,
                           const double* dB, int64_t ldb, double beta, double* dC, int64_t ldc, const int* dStair, int64_t row0) {
    if (!h) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, M >= 0 && N >= 0 && K >= 0 && M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31) && row0 >= 0 &&
                          row0 + M < (1ll << 31), "gemm_stair_dev: bad extents"));
    GP_TRY(require(c, dStair != nullptr, "gemm_stair_dev: null stair table (use gprb200_gemm_dev for a full C)"));
    GemmArgs g{};
    g.A = dA; g.lda = lda; g.B = dB; g.ldb = ldb; g.C = dC; g.ldc = ldc;
    g.M = (int)M; g.N = (int)N; g.K = (int)K; g.alpha = alpha; g.beta = beta;
    g.set_stair(dStair, (int)row0);
    return gemm(c, GEMM_NT, g);
}

int gprb200_gram_block_dev(gprb200_handle h, int64_t row0, int64_t nrows, int64_t col0, int64_t ncols, double* dOut,
                           int64_t ldo) {
    if (!h || !dOut) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, c->have_train && c->have_kernel, "gram_block: need samples and a kernel"));
    GP_TRY(require(c, row0 >= 0 && col0 >= 0 && nrows > 0 && ncols > 0 && row0 + nrows <= c->N && col0 + ncols <= c->N,
                   "gram_block: block outside the N x N kernel matrix"));
    GP_TRY(gram_build(c, c->dX + row0 * c->d, nrows, c->dX + col0 * c->d, ncols, dOut, ldo, false, false, 0.0));
    // sigma^2 on the global diagonal (AddNoiseToKernelMatrix, lib/GaussianProcess.cpp:375-381)
    const int64_t lo = row0 > col0 ? row0 : col0;
    const int64_t hi = (row0 + nrows < col0 + ncols) ? row0 + nrows : col0 + ncols;
    if (hi > lo && c->sigma != 0.0)
        GP_TRY(add_diag(c, dOut + (lo - row0) * ldo + (lo - col0), ldo, hi - lo, c->sigma * c->sigma));
    return 0;
}

int gprb200_gram_pair_dev(gprb200_handle h, const double* dXa, int64_t na, const double* dXb, int64_t nb, double* dOut,
                          int64_t ldo) {
    if (!h || !dOut || !dXa || !dXb) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, c->have_train && c->have_kernel, "gram_pair: need a kernel and samples (for the input dimension)"));
    GP_TRY(require(c, na > 0 && nb > 0 && ldo >= nb, "gram_pair: bad extents"));
    return gram_build(c, dXa, na, dXb, nb, dOut, ldo, false, false, 0.0);
}

int gprb200_potrf_panel_dev(gprb200_handle h, double* dP, int64_t ld, int64_t rows, int64_t w, double* dInvL, int* info) {
    if (!h || !dP || !dInvL) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, rows >= w && w > 0 && (ld & 1) == 0, "potrf_panel: need rows >= w > 0 and an even leading dimension"));
    int inf = 0;
    GP_TRY(potrf_panel(c, dP, ld, rows, w, dInvL, 0, &inf));
    if (info) *info = inf;
    if (inf != 0)
        return fail(c, inf, "potrf_panel: block column is not positive definite (local leading minor of order " +
                                std::to_string(inf) + ")");
    return 0;
}

int gprb200_potrf_panel_async_dev(gprb200_handle h, double* dP, int64_t ld, int64_t rows, int64_t w, double* dInvL,
                                  int64_t col0, int* dInfo) {
    if (!h || !dP || !dInvL || !dInfo) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, rows >= w && w > 0 && (ld & 1) == 0 && col0 >= 0 && col0 < (1ll << 31) - w,
                   "potrf_panel_async: need rows >= w > 0, an even leading dimension and 0 <= col0 < 2^31"));
    return potrf_panel_async(c, dP, ld, rows, w, dInvL, (int)col0, dInfo);
}

int gprb200_trsm_dev(gprb200_handle h, const double* dL, int64_t ldl, const double* dInvL, int64_t n, double* dB,
                     int64_t ldb, int64_t M, int transposed) {
    if (!h || !dL || !dInvL || !dB) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    return transposed ? trsm_lower_bwd(c, dL, ldl, dInvL, n, dB, ldb, M) : trsm_lower_fwd(c, dL, ldl, dInvL, n, dB, ldb, M);
}

int gprb200_trtri_cols_dev(gprb200_handle h, const double* dL, int64_t ldl, const double* dInvL, int64_t n, int64_t col_block,
                           int64_t first, int64_t stride, double* dW, int64_t ldw) {
    if (!h || !dL || !dInvL || !dW) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, n > 0 && col_block > 0 && col_block % NB == 0 && first >= 0 && stride > 0 && (ldw & 1) == 0 &&
                          (ldl & 1) == 0, "trtri_cols: col_block must be a multiple of 128, leading dimensions even"));
    return trtri_cols(c, dL, ldl, dInvL, n, col_block, first, stride, dW, ldw);
}

int gprb200_logdet_dev(gprb200_handle h, const double* dL, int64_t ld, int64_t n, double* logdet) {
    if (!h || !dL || !logdet) return -2;
    GP_TRY(use_device(h));
    return logdet_from_factor(h, dL, ld, n, logdet);
}

int gprb200_predict_dev(gprb200_handle h, const double* dXs, int64_t T, double* d_mean, int64_t ld_mean,
                        double* d_ci) {
    if (!h) return -2;
    GP_TRY(use_device(h));
    GP_TRY(require(h, T > 0 && dXs, "predict: need T > 0 test points"));
    return predict_impl(h, dXs, T, d_mean, ld_mean, d_ci);
}

int gprb200_predict(gprb200_handle h, const double* Xs, int64_t T, double* mean, double* ci) {
    if (!h) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, T > 0 && Xs, "predict: need T > 0 test points"));
    GP_TRY(require(c, c->have_train, "GaussianProcess::ComputeKernelVectorInternal: gaussian process is not initialized."));
    TempBuf xs(c), mu(c), civ(c);
    GP_TRY(xs.alloc((size_t)T * c->d));
    GP_CUDA(c, cudaMemcpyAsync(xs.p, Xs, sizeof(double) * T * c->d, cudaMemcpyHostToDevice, c->stream));
    const int64_t ldm = round_up(c->M > 0 ? c->M : 1, 16);
    if (mean) GP_TRY(mu.alloc((size_t)T * ldm));
    if (ci) GP_TRY(civ.alloc((size_t)T));
    GP_TRY(predict_impl(c, xs.p, T, mean ? mu.p : nullptr, ldm, ci ? civ.p : nullptr));
    if (mean) GP_TRY(d2h_2d(c, mean, mu.p, ldm, T, c->M));
    if (ci) GP_TRY(d2h_2d(c, ci, civ.p, T, 1, T));
    return 0;
}

int gprb200_kernel_vector(gprb200_handle h, const double* Xs, int64_t T, double* Kx) {
    if (!h || !Xs || !Kx || T <= 0) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, c->have_train && c->have_kernel, "kernel_vector: need samples and a kernel"));
    const int64_t ldkx = round_up(c->N, 16);
    TempBuf xs(c), k(c);
    GP_TRY(xs.alloc((size_t)T * c->d));
    GP_TRY(k.alloc((size_t)T * ldkx));
    GP_CUDA(c, cudaMemcpyAsync(xs.p, Xs, sizeof(double) * T * c->d, cudaMemcpyHostToDevice, c->stream));
    GP_TRY(gram_build(c, xs.p, T, c->dX, c->N, k.p, ldkx, false, false, 0.0));
    return d2h_2d(c, Kx, k.p, ldkx, T, c->N);
}

int gprb200_posterior_cov(gprb200_handle h, const double* Xa, const double* Xb, int64_t T, double* cov) {
    if (!h || !Xa || !Xb || !cov || T <= 0) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, c->have_train && c->have_kernel, "operator(): gaussian process is not initialized"));
    TempBuf xa(c), xb(c), q(c), ks(c), out(c);
    GP_TRY(xa.alloc((size_t)T * c->d));
    GP_TRY(xb.alloc((size_t)T * c->d));
    GP_TRY(q.alloc((size_t)T));
    GP_TRY(ks.alloc((size_t)T));
    GP_TRY(out.alloc((size_t)T));
    // gp(x, x) (GetCredibleInterval, lib/GaussianProcess.cpp:108): one kernel vector and one solve instead of two
    const bool same = (Xa == Xb) || memcmp(Xa, Xb, sizeof(double) * T * c->d) == 0;
    GP_CUDA(c, cudaMemcpyAsync(xa.p, Xa, sizeof(double) * T * c->d, cudaMemcpyHostToDevice, c->stream));
    const double* xbp = xa.p;
    if (!same) {
        GP_CUDA(c, cudaMemcpyAsync(xb.p, Xb, sizeof(double) * T * c->d, cudaMemcpyHostToDevice, c->stream));
        xbp = xb.p;
    }
    GP_TRY(quad_forms(c, xa.p, xbp, T, q.p));
    GP_TRY(kpair_values(c, xa.p, xbp, T, ks.p));
    ci_finish_kernel<<<(unsigned)((T + 255) / 256), 256, 0, c->stream>>>(ks.p, q.p, T, out.p, 0);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return d2h_2d(c, cov, out.p, T, 1, T);
}

int gprb200_predict_derivative(gprb200_handle h, const double* x, double* mean, double* D) {
    if (!h || !x) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, c->have_train && c->have_kernel, "PredictDerivative: gaussian process is not initialized"));
    GP_TRY(ensure_alpha(c));
    const int64_t N = c->N, M = c->M;
    const int d = c->d;
    const int64_t ldn = round_up(N, 16), ldm = round_up(M, 16);
    TempBuf xs(c), kx(c), A(c), out(c);
    GP_TRY(xs.alloc((size_t)d));
    GP_TRY(kx.alloc((size_t)ldn));
    GP_TRY(A.alloc((size_t)(d + 1) * ldn));
    GP_TRY(out.alloc((size_t)(d + 1) * ldm));
    GP_CUDA(c, cudaMemcpyAsync(xs.p, x, sizeof(double) * d, cudaMemcpyHostToDevice, c->stream));
    GP_TRY(gram_build(c, xs.p, 1, c->dX, N, kx.p, ldn, false, false, 0.0));
    deriv_lhs_kernel<<<(unsigned)((N + 255) / 256), 256, 0, c->stream>>>(xs.p, c->dX, kx.p, N, d, A.p, ldn);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    GemmArgs g{};
    g.A = A.p; g.lda = ldn; g.B = c->dA; g.ldb = c->ldy; g.C = out.p; g.ldc = ldm;
    g.M = d + 1; g.N = (int)M; g.K = (int)N; g.alpha = 1.0; g.beta = 0.0;
    GP_TRY(gemm(c, GEMM_NN, g));
    if (D) GP_TRY(d2h_2d(c, D, out.p, ldm, d, M));
    if (mean) GP_TRY(d2h_2d(c, mean, out.p + (int64_t)d * ldm, ldm, 1, M));
    return 0;
}

int gprb200_dgram(gprb200_handle h, double* D_out) {
    if (!h || !D_out) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, c->have_train && c->have_kernel, "dgram: need samples and a kernel"));
    const int P = c->prog.nparams;
    PhaseTimer t(c, GPRB200_T_DGRAM);
    TempBuf D(c);
    GP_TRY(D.alloc((size_t)P * c->N * c->ldk));
    GP_TRY(dgram_build(c, D.p, c->ldk));
    t.stop();
    t.resolve();
    return d2h_2d(c, D_out, D.p, c->ldk, (int64_t)P * c->N, c->N);
}

int gprb200_quadform(gprb200_handle h, double* q) {
    if (!h || !q) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(ensure_alpha(c));
    GP_TRY(require(c, c->dY != nullptr, "quadform: label vectors are not resident"));
    TempBuf t(c);
    GP_TRY(t.alloc((size_t)c->M));
    GP_TRY(coldot(c, c->dY, c->ldy, c->dA, c->ldy, c->N, c->M, t.p));
    return d2h_2d(c, q, t.p, c->M, 1, c->M);
}

int gprb200_loglik(gprb200_handle h, int mode, double* value, double* grad, double* jac) {
    if (!h) return -2;
    Ctx* c = h;
    GP_TRY(use_device(c));
    GP_TRY(require(c, c->have_train && c->have_kernel && c->dY, "loglik: need samples, labels and a kernel"));
    // the reference rebuilds and re-inverts K on every call (include/Likelihood.h:237 -> lib/GaussianProcess.cpp:513-528)
    invalidate(c);
    GP_TRY(ensure_factor(c));
    GP_TRY(ensure_alpha(c));
    const int64_t N = c->N, M = c->M;
    const int P = c->prog.nparams;
    PhaseTimer t(c, GPRB200_T_LOGLIK);
    if (value) {
        std::vector<double> q((size_t)M);
        GP_TRY(gprb200_quadform(h, q.data()));
        long double cp;
        if (mode == GPRB200_LOGLIK_TRUE) {
            cp = -0.5L * (long double)c->logdet;
        } else {
            // include/Likelihood.h:77-79 returns the long-double determinant through a TScalarType (double):
            // it overflows to +inf / underflows to 0 there, and :180-188 clamps those cases.
            const long double det_ld = expl((long double)c->logdet);
            const double det_d = (double)det_ld;
            const long double det = (long double)det_d;
            if (det <= LDBL_MIN) cp = -0.5L * logl(LDBL_MIN);
            else if (det > LDBL_MAX) cp = -0.5L * logl(LDBL_MAX);
            else cp = -0.5L * logl(det);
        }
        const double ct = -(double)N / 2.0 * log(2.0 * M_PI);
        double sum = 0.0;
        for (int64_t k = 0; k < M; ++k) {
            value[k] = (double)((long double)(-0.5 * q[k]) + (cp + (long double)ct));
            sum += value[k];
        }
        if (isinf(sum))
            return fail(c, -4, "GaussianLogLikelihood::GetValueAndParameterDerivatives: likelihood is infinite.");
    }
    if (grad || jac) {
        GP_TRY(ensure_core(c));
        if (grad) {
            // S = alpha alpha^T - C, then grad[p] = 0.5 sum_ij S_ij dK_p(i,j)  (include/Likelihood.h:280-282)
            TempBuf S(c);
            GP_TRY(S.alloc((size_t)N * c->ldk));
            GP_TRY(copy2d(c, S.p, c->ldk, c->dC, c->ldk, N, N));
            GemmArgs g{};
            g.A = c->dA; g.lda = c->ldy; g.B = c->dA; g.ldb = c->ldy; g.C = S.p; g.ldc = c->ldk;
            g.M = (int)N; g.N = (int)N; g.K = (int)M; g.alpha = 1.0; g.beta = -1.0;
            GP_TRY(gemm(c, GEMM_NT, g));
            GP_TRY(dgram_contract(c, S.p, c->ldk, grad));
            for (int p = 0; p < P; ++p) grad[p] *= 0.5;
        }
        if (jac) {
            // jac[i,p] = 0.5 (alpha_i^T dK_p alpha_i - tr(C dK_p))          (include/Likelihood.h:335-341)
            std::vector<double> trc((size_t)P), qi((size_t)M);
            GP_TRY(dgram_contract(c, c->dC, c->ldk, trc.data()));
            TempBuf Dp(c), G(c), qd(c);
            GP_TRY(Dp.alloc((size_t)N * c->ldk));
            GP_TRY(G.alloc((size_t)N * c->ldy));
            GP_TRY(qd.alloc((size_t)M));
            for (int p = 0; p < P; ++p) {
                GP_TRY(dgram_build_one(c, Dp.p, c->ldk, p));
                GemmArgs g{};
                g.A = Dp.p; g.lda = c->ldk; g.B = c->dA; g.ldb = c->ldy; g.C = G.p; g.ldc = c->ldy;
                g.M = (int)N; g.N = (int)M; g.K = (int)N; g.alpha = 1.0; g.beta = 0.0;
                GP_TRY(gemm(c, GEMM_NN, g));
                GP_TRY(coldot(c, c->dA, c->ldy, G.p, c->ldy, N, M, qd.p));
                GP_TRY(d2h_2d(c, qi.data(), qd.p, M, 1, M));
                for (int64_t i = 0; i < M; ++i) jac[i * P + p] = 0.5 * (qi[(size_t)i] - trc[(size_t)p]);
            }
        }
    }
    t.stop();
    t.resolve();
    return 0;
}

}  // extern "C"
