// K6: prediction kernels.
//   gemv_t : out[m] = sum_n kx[n] * A[n][m]  -- Predict for ONE test point, the call gpPredict makes per
//            frame (reference lib/GaussianProcess.cpp:55-61, apps/GaussianProcessPredict.cpp:185-193).
//            Reads all of alpha once: 8 N M bytes, purely HBM-bound (C2: 25.77 GB per prediction).
// Layout: alpha is N x M row-major, so consecutive threads own consecutive output columns and every
// warp-level load is a contiguous 512-byte run; each thread keeps UNROLL independent 16-byte loads
// in flight. Rows are split across blockIdx.y so the grid is several waves of 148 SMs even when M
// is small; partials are summed in a fixed order (deterministic).
#include "common.cuh"
#include <mutex>
#include <string>

namespace gprb200 {

constexpr int GV_THREADS = 256;
constexpr int GV_UNROLL = 8;

__device__ __forceinline__ double2 ld_stream(const double2* p) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}

// partial[split][m] = sum_{n in split} kx[n] A[n][m]; thread owns columns (2 c, 2 c + 1)
__global__ void __launch_bounds__(GV_THREADS) gemv_t_kernel(const double* __restrict__ kx, const double* __restrict__ A,
                                                             int64_t lda, int64_t N, int64_t M, int64_t rows_per_split,
                                                             double* __restrict__ partial, int64_t ldp) {
    extern __shared__ double skx[];
    const int64_t n_begin = (int64_t)blockIdx.y * rows_per_split;
    const int64_t n_end = min(N, n_begin + rows_per_split);
    for (int64_t i = threadIdx.x; i < n_end - n_begin; i += GV_THREADS) skx[i] = kx[n_begin + i];
    __syncthreads();
    const int64_t col = ((int64_t)blockIdx.x * GV_THREADS + threadIdx.x) * 2;
    if (col >= M) return;
    double s0 = 0.0, s1 = 0.0;
    if (col + 1 < M) {
        const double* base = A + col;
        int64_t n = n_begin;
        for (; n + GV_UNROLL <= n_end; n += GV_UNROLL) {
            double2 v[GV_UNROLL];
#pragma unroll
            for (int u = 0; u < GV_UNROLL; ++u) v[u] = ld_stream(reinterpret_cast<const double2*>(base + (n + u) * lda));
#pragma unroll
            for (int u = 0; u < GV_UNROLL; ++u) {
                const double k = skx[n - n_begin + u];
                s0 = fma(k, v[u].x, s0);
                s1 = fma(k, v[u].y, s1);
            }
        }
        for (; n < n_end; ++n) {
            const double2 v = ld_stream(reinterpret_cast<const double2*>(base + n * lda));
            const double k = skx[n - n_begin];
            s0 = fma(k, v.x, s0);
            s1 = fma(k, v.y, s1);
        }
        partial[(int64_t)blockIdx.y * ldp + col] = s0;
        partial[(int64_t)blockIdx.y * ldp + col + 1] = s1;
    } else {
        for (int64_t n = n_begin; n < n_end; ++n) s0 = fma(skx[n - n_begin], A[n * lda + col], s0);
        partial[(int64_t)blockIdx.y * ldp + col] = s0;
    }
}

__global__ void gemv_reduce_kernel(const double* partial, int64_t ldp, int splits, int64_t M, double* out) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    double s = 0.0;
    for (int k = 0; k < splits; ++k) s += partial[(int64_t)k * ldp + m];
    out[m] = s;
}

int gemv_t(Ctx* c, const double* kx, const double* A, int64_t lda, int64_t N, int64_t M, double* out) {
    if (M <= 0) return 0;
    const int64_t col_blocks = (M + 2 * GV_THREADS - 1) / (2 * GV_THREADS);
    // aim for >= 8 waves of 148 SMs x 8 resident CTAs, but keep >= 256 rows per split
    int64_t splits = (148 * 8 * 8 + col_blocks - 1) / col_blocks;
    const int64_t max_splits = (N + 255) / 256;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    if (splits > 65535) splits = 65535;
    int64_t rows_per_split = (N + splits - 1) / splits;
    const int64_t cap_rows = 4096;  // shared-memory slice of kx: 32 KB
    if (rows_per_split > cap_rows) rows_per_split = cap_rows;
    splits = (N + rows_per_split - 1) / rows_per_split;
    const int64_t ldp = round_up(M, 2);
    const size_t need = sizeof(double) * splits * ldp;
    double* partial = nullptr;
    if (splits == 1) {
        partial = out;
    } else {
        GP_CUDA(c, cudaMallocAsync(&partial, need, c->stream));
    }
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (c->opt_profile) {
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        cudaEventRecord(e0, c->stream);
    }
    gemv_t_kernel<<<dim3((unsigned)col_blocks, (unsigned)splits), GV_THREADS, sizeof(double) * rows_per_split,
                    c->stream>>>(kx, A, lda, N, M, rows_per_split, partial, splits == 1 ? 0 : ldp);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    if (c->opt_profile) {
        cudaEventRecord(e1, c->stream);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, e0, e1) == cudaSuccess) {
            c->prof_gemv_ms += ms;
            c->prof_gemv_bytes += 8.0 * (double)N * (double)M;  // algorithmic: alpha is read once
            c->prof_gemv_launches += 1.0;
        }
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
    }
    if (splits > 1) {
        gemv_reduce_kernel<<<(unsigned)((M + 255) / 256), 256, 0, c->stream>>>(partial, ldp, (int)splits, M, out);
        ++c->launches;
        GP_CUDA(c, cudaGetLastError());
        GP_CUDA(c, cudaFreeAsync(partial, c->stream));
    }
    return 0;
}

// ---- trmv_lower: V (N x T, T <= 2) = W Kx^T for a lower-triangular W (the explicit inverse factor) ---------------------
// Single-point credible interval (GetCredibleInterval, lib/GaussianProcess.cpp:103-114, called once per test point by
// apps/GaussianProcessPredict.cpp:188): v = inv(L) kx used to be a forward substitution, a chain of N / 128 dependent
// block steps (0.76 ms at N = 16384, latency-bound). With W = inv(L) formed once it is a matrix-vector product with no
// dependence between rows: one warp per row, 16-byte loads, only the non-zero prefix k <= i of each row is read
// (4 N^2 bytes in total: HBM-bound, 1.07 GB = ~0.15 ms at N = 16384). kx is T x ldk (the cross-Gram layout).
template <int T>
__global__ void __launch_bounds__(256) trmv_lower_kernel(const double* __restrict__ W, int64_t ldw, int64_t N,
                                                          const double* __restrict__ kx, int64_t ldk, double* __restrict__ V,
                                                          int64_t ldv) {
    const int lane = threadIdx.x & 31;
    // long rows first: the last CTAs to start get the short rows
    const int64_t i = N - 1 - ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5));
    if (i < 0) return;
    const double* w = W + i * ldw;
    double s[T];
#pragma unroll
    for (int t = 0; t < T; ++t) s[t] = 0.0;
    const int64_t n2 = (i + 1) & ~(int64_t)1;   // even prefix, 16-byte loads (W rows are 16-byte aligned, ldw even)
    for (int64_t k = lane * 2; k < n2; k += 64) {
        const double2 a = ld_stream(reinterpret_cast<const double2*>(w + k));
#pragma unroll
        for (int t = 0; t < T; ++t) {
            const double2 b = *reinterpret_cast<const double2*>(kx + t * ldk + k);
            s[t] = fma(a.x, b.x, s[t]);
            s[t] = fma(a.y, b.y, s[t]);
        }
    }
    if (lane == 0 && n2 <= i) {
#pragma unroll
        for (int t = 0; t < T; ++t) s[t] = fma(w[i], kx[t * ldk + i], s[t]);
    }
#pragma unroll
    for (int t = 0; t < T; ++t) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s[t] += __shfl_xor_sync(0xffffffffu, s[t], o);
        if (lane == 0) V[i * ldv + t] = s[t];
    }
}

int trmv_lower(Ctx* c, const double* W, int64_t ldw, int64_t N, const double* kx, int64_t ldk, int T, double* V, int64_t ldv) {
    if (T < 1 || T > 2) return fail(c, -2, "trmv_lower: 1 or 2 vectors");
    if ((ldw & 1) || (ldk & 1) || (reinterpret_cast<uintptr_t>(W) & 15) || (reinterpret_cast<uintptr_t>(kx) & 15))
        return fail(c, -2, "trmv_lower: operands must be 16-byte aligned with even leading dimensions");
    const unsigned grid = (unsigned)((N + 7) / 8);
    if (T == 1) trmv_lower_kernel<1><<<grid, 256, 0, c->stream>>>(W, ldw, N, kx, ldk, V, ldv);
    else trmv_lower_kernel<2><<<grid, 256, 0, c->stream>>>(W, ldw, N, kx, ldk, V, ldv);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}

// ---- allocation-free variant of gemv_t (gprb200_gemv_t_noalloc_dev, include/gprb200.h) -------------------------------
// The row-split partial sums live in a file-scope device array instead of a cudaMallocAsync buffer, for callers that
// cannot allocate (CUDA graph capture, tight serving loops). GV_SCRATCH_DOUBLES doubles = 32 MiB are held by every GPU
// that loads the library until the process exits. In HOST code the array's name is only a stand-in: the device
// address is taken with cudaGetSymbolAddress once per device (with that device current) and passed to the kernels.
constexpr int64_t GV_SCRATCH_DOUBLES = 4ll << 20;
static __device__ __align__(256) double g_gemv_scratch[GV_SCRATCH_DOUBLES];
constexpr int GV_MAX_DEVICES = 64;
static double* g_gemv_scratch_ptr[GV_MAX_DEVICES] = {nullptr};
static std::mutex g_gemv_scratch_mutex;

// same split geometry as gemv_t above (so both entry points sum in the same order)
static void gemv_geometry(int64_t N, int64_t M, int64_t& col_blocks, int64_t& splits, int64_t& rows_per_split) {
    col_blocks = (M + 2 * GV_THREADS - 1) / (2 * GV_THREADS);
    splits = (148 * 8 * 8 + col_blocks - 1) / col_blocks;
    const int64_t max_splits = (N + 255) / 256;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    if (splits > 65535) splits = 65535;
    rows_per_split = (N + splits - 1) / splits;
    if (rows_per_split > 4096) rows_per_split = 4096;
    splits = (N + rows_per_split - 1) / rows_per_split;
}

static int gemv_t_noalloc(Ctx* c, const double* kx, const double* A, int64_t lda, int64_t N, int64_t M, double* out) {
    if (c->device < 0 || c->device >= GV_MAX_DEVICES) return fail(c, -2, "gemv_t_noalloc: device index out of range");
    GP_CUDA(c, cudaSetDevice(c->device));
    int64_t col_blocks, splits, rows_per_split;
    gemv_geometry(N, M, col_blocks, splits, rows_per_split);
    const int64_t ldp = round_up(M, 2);
    if (splits > 65535 || col_blocks > 0x7fffffff)
        return fail(c, -2, "gemv_t_noalloc: N / M exceed the launch grid");
    if (splits > 1 && splits * ldp > GV_SCRATCH_DOUBLES)
        return fail(c, -2, "gemv_t_noalloc: " + std::to_string(splits) + " row splits x " + std::to_string(ldp) +
                               " columns = " + std::to_string(splits * ldp) + " partial sums exceed the static scratch of " +
                               std::to_string(GV_SCRATCH_DOUBLES) + " doubles (32 MiB); use gprb200_predict_dev");
    double* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_gemv_scratch_mutex);
        if (!g_gemv_scratch_ptr[c->device]) {   // set-up, once per device, this device current
            double* p = nullptr;
            GP_CUDA(c, cudaGetSymbolAddress((void**)&p, g_gemv_scratch));
            g_gemv_scratch_ptr[c->device] = p;
        }
        scratch = g_gemv_scratch_ptr[c->device];
    }
    // nothing in the scratch has to start at zero: every partial the reduction reads is written by this call's
    // gemv_t_kernel first (row splits x columns < M); the contents of the previous call are simply overwritten
    double* partial = (splits == 1) ? out : scratch;
    gemv_t_kernel<<<dim3((unsigned)col_blocks, (unsigned)splits), GV_THREADS, sizeof(double) * rows_per_split,
                    c->stream>>>(kx, A, lda, N, M, rows_per_split, partial, splits == 1 ? 0 : ldp);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    if (splits > 1) {
        gemv_reduce_kernel<<<(unsigned)((M + 255) / 256), 256, 0, c->stream>>>(partial, ldp, (int)splits, M, out);
        ++c->launches;
        GP_CUDA(c, cudaGetLastError());
    }
    return 0;
}

}  // namespace gprb200

extern "C" int gprb200_gemv_t_noalloc_dev(gprb200_handle h, const double* dKx, const double* dA, int64_t lda, int64_t N,
                                          int64_t M, double* dOut) {
    if (!h) return -2;
    gprb200::Ctx* c = h;
    if (!dKx || !dA || !dOut || N <= 0 || M <= 0 || lda < M)
        return gprb200::fail(c, -2, "gemv_t_noalloc: need device pointers, N > 0, M > 0 and lda >= M");
    if ((lda & 1) || (reinterpret_cast<uintptr_t>(dA) & 15))
        return gprb200::fail(c, -2, "gemv_t_noalloc: A must be 16-byte aligned with an even leading dimension");
    return gprb200::gemv_t_noalloc(c, dKx, dA, lda, N, M, dOut);
}
