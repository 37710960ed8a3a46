// Skinny products C (M x n, n <= 8) = alpha op(A) B + beta C for the few-right-hand-side case (M = 1 output column
// is the common GP regression; BASELINE configs C3-C5). With n <= 8 a tensor-core tile would waste 8-64x of its
// columns and the product is bound by reading A once from HBM (8 M K bytes), so these are streaming kernels:
//   NN: A is M x K row-major  -> one warp per row, lanes stride k (each row is one contiguous run), shuffle-reduce
//   TN: A is K x M row-major  -> one thread per output row m (coalesced across m), k split over blockIdx.y,
//                                partials combined in a fixed order (deterministic)
// They carry the triangular solves alpha = L^-T L^-1 y of the blocked Cholesky (reference: alpha = C * Y,
// lib/GaussianProcess.cpp:661) when Y has few columns, and the distributed solve of gpr_b200/dist.py.
#include "common.cuh"

namespace gprb200 {

constexpr int SK_THREADS = 256;

template <int NC>
__global__ void __launch_bounds__(SK_THREADS) skinny_nn_kernel(GemmArgs g) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (SK_THREADS / 32) + warp;
    if (row >= g.M) return;
    const double* a = g.A + row * g.lda;
    double acc[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) acc[c] = 0.0;
    for (int k = lane; k < g.K; k += 32) {
        const double av = a[k];
        const double* b = g.B + (int64_t)k * g.ldb;
#pragma unroll
        for (int c = 0; c < NC; ++c)
            if (c < g.N) acc[c] = fma(av, b[c], acc[c]);
    }
#pragma unroll
    for (int c = 0; c < NC; ++c)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], o);
    if (lane == 0) {
        double* crow = g.C + row * g.ldc;
#pragma unroll
        for (int c = 0; c < NC; ++c)
            if (c < g.N) {
                double v = g.alpha * acc[c];
                if (g.beta != 0.0) v = fma(g.beta, crow[c], v);
                crow[c] = v;
            }
    }
}

// partial[split][m][c] (or C directly when there is one split)
template <int NC>
__global__ void __launch_bounds__(SK_THREADS) skinny_tn_kernel(GemmArgs g, int k_per_split, double* partial) {
    const int64_t m = (int64_t)blockIdx.x * SK_THREADS + threadIdx.x;
    const int k0 = blockIdx.y * k_per_split;
    const int k1 = min(g.K, k0 + k_per_split);
    __shared__ double sb[64][NC];
    double acc[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) acc[c] = 0.0;
    for (int kb = k0; kb < k1; kb += 64) {
        const int kn = min(64, k1 - kb);
        __syncthreads();
        for (int e = threadIdx.x; e < kn * NC; e += SK_THREADS) {
            const int kk = e / NC, c = e - kk * NC;
            sb[kk][c] = (c < g.N) ? g.B[(int64_t)(kb + kk) * g.ldb + c] : 0.0;
        }
        __syncthreads();
        if (m < g.M) {
            const double* a = g.A + (int64_t)kb * g.lda + m;
#pragma unroll 8
            for (int kk = 0; kk < kn; ++kk) {
                const double av = a[(int64_t)kk * g.lda];
#pragma unroll
                for (int c = 0; c < NC; ++c) acc[c] = fma(av, sb[kk][c], acc[c]);
            }
        }
    }
    if (m >= g.M) return;
    if (partial) {
        double* p = partial + ((int64_t)blockIdx.y * g.M + m) * NC;
#pragma unroll
        for (int c = 0; c < NC; ++c) p[c] = acc[c];
    } else {
        double* crow = g.C + m * g.ldc;
#pragma unroll
        for (int c = 0; c < NC; ++c)
            if (c < g.N) {
                double v = g.alpha * acc[c];
                if (g.beta != 0.0) v = fma(g.beta, crow[c], v);
                crow[c] = v;
            }
    }
}

template <int NC>
__global__ void skinny_tn_reduce_kernel(GemmArgs g, int splits, const double* partial) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= g.M) return;
    double* crow = g.C + m * g.ldc;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        if (c >= g.N) break;
        double s = 0.0;
        for (int sp = 0; sp < splits; ++sp) s += partial[((int64_t)sp * g.M + m) * NC + c];
        double v = g.alpha * s;
        if (g.beta != 0.0) v = fma(g.beta, crow[c], v);
        crow[c] = v;
    }
}

template <int NC>
static int launch_skinny(Ctx* c, GemmForm form, const GemmArgs& g, cudaStream_t st) {
    if (form == GEMM_NN) {
        const int64_t blocks = (g.M + SK_THREADS / 32 - 1) / (SK_THREADS / 32);
        skinny_nn_kernel<NC><<<(unsigned)blocks, SK_THREADS, 0, st>>>(g);
        ++c->launches;
        GP_CUDA(c, cudaGetLastError());
        return 0;
    }
    // TN: enough k-splits to fill the machine (~4 CTAs per SM), each at least 256 deep
    const int64_t mblocks = (g.M + SK_THREADS - 1) / SK_THREADS;
    int64_t splits = (148 * 4 + mblocks - 1) / mblocks;
    const int64_t max_splits = (g.K + 255) / 256;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    if (splits > 65535) splits = 65535;
    int kps = (int)((g.K + splits - 1) / splits);
    kps = (kps + 63) / 64 * 64;
    splits = (g.K + kps - 1) / kps;
    double* partial = nullptr;
    if (splits > 1) GP_CUDA(c, cudaMallocAsync((void**)&partial, sizeof(double) * splits * g.M * NC, st));
    skinny_tn_kernel<NC><<<dim3((unsigned)mblocks, (unsigned)splits), SK_THREADS, 0, st>>>(g, kps, partial);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    if (splits > 1) {
        skinny_tn_reduce_kernel<NC><<<(unsigned)((g.M + 255) / 256), 256, 0, st>>>(g, (int)splits, partial);
        ++c->launches;
        GP_CUDA(c, cudaGetLastError());
        GP_CUDA(c, cudaFreeAsync(partial, st));
    }
    return 0;
}

// C must not alias A or B (true for every caller: the solves write row blocks disjoint from the ones they read).
int gemm_skinny(Ctx* c, GemmForm form, const GemmArgs& g, cudaStream_t st) {
    if (g.N <= 1) return launch_skinny<1>(c, form, g, st);
    if (g.N <= 2) return launch_skinny<2>(c, form, g, st);
    if (g.N <= 4) return launch_skinny<4>(c, form, g, st);
    return launch_skinny<8>(c, form, g, st);
}

}  // namespace gprb200
