// PCA on the device: the reference's PCA<T> (include/PCA.h) is what turns the GP's predicted
// principal weights into the 3 x 64^3 = 786 432-element displacement field, once per frame
// (include/DataParser.h:211-286, :471-495; apps/GaussianProcessPredict.cpp:188-203).
//
//   fit          X (D x n, one sample per column) -> mean, sigma = s / sqrt(n), U (thin, D x n)       PCA.h:33-49
//                "snapshot" form: G = Xc^T Xc (n x n, FP64 DMMA, split over D), parallel two-sided Jacobi on G,
//                Z <- Xc V; a second Gram/Jacobi pass on the rotated (nearly orthogonal) columns restores the
//                accuracy the first Gram squared away; U = Z diag(1/s).
//   project      F = basis^T (X - mean), basis = U diag(1/sigma)                                      PCA.h:92-108
//   reconstruct  X = (U diag(sigma))[:, :nFeatures] W + mean                                          PCA.h:110-115
//
// project / reconstruct are HBM-bound streams over a D x nFeatures matrix (8 D nFeatures bytes per call).
// Singular vectors are defined up to sign; nothing here fixes a sign convention (neither does Eigen).
#include <algorithm>
#include <cmath>
#include <numeric>
#include <vector>

#include "../../include/gprb200.h"
#include "common.cuh"

namespace gprb200 {

struct PcaModel {
    Ctx* c = nullptr;
    int64_t D = 0, n = 0, nF = 0;
    int64_t ldu = 0, ldb = 0;
    double* dMean = nullptr;   // D
    double* dSigma = nullptr;  // n
    double* dU = nullptr;      // D x ldu
    double* dBinv = nullptr;   // D x ldb : (U diag(sigma))[:, :nF]   (m_basisInverse, PCA.h:47-48)
    double* dBasis = nullptr;  // D x ldb : (U diag(1/sigma))[:, :nF] (m_basisTranspose^T, PCA.h:67-69)
    int sweeps = 0, passes = 2;
    double fit_ms = 0.0;
};

namespace {

constexpr int PT = 256;

template <class T>
int pca_alloc(Ctx* c, T** p, size_t count) {
    GP_CUDA(c, cudaMalloc((void**)p, sizeof(T) * (count ? count : 1)));
    return 0;
}
void pca_free_all(PcaModel* m) {
    cudaFree(m->dMean); cudaFree(m->dSigma); cudaFree(m->dU); cudaFree(m->dBinv); cudaFree(m->dBasis);
    m->dMean = m->dSigma = m->dU = m->dBinv = m->dBasis = nullptr;
    m->D = m->n = m->nF = 0;
}

// ---- fit: centring -------------------------------------------------------------------------------------------
// one warp per row: mean_i = sum_j X[i][j] / n, Z[i][j] = X[i][j] - mean_i
__global__ void __launch_bounds__(PT) pca_center_kernel(const double* __restrict__ X, int64_t ldx, int64_t D, int n,
                                                        double* __restrict__ mean, double* __restrict__ Z, int64_t ldz) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (PT / 32) + (threadIdx.x >> 5);
    if (row >= D) return;
    const double* x = X + row * ldx;
    double s = 0.0;
    for (int j = lane; j < n; j += 32) s += x[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const double mu = s / (double)n;
    if (lane == 0) mean[row] = mu;
    double* z = Z + row * ldz;
    for (int j = lane; j < n; j += 32) z[j] = x[j] - mu;
}

// G = sum_z P_z, symmetrised; written into the n2 x n2 Jacobi work matrix (zero padded). One warp per element:
// lanes stride the batch, fixed-order shuffle tree (deterministic; the batch can be ~300 deep for small n).
__global__ void __launch_bounds__(256) pca_gram_reduce_kernel(const double* __restrict__ P, int batch, int n, int64_t ldp,
                                                              int64_t strideP, double* __restrict__ G, int n2) {
    const int lane = threadIdx.x & 31;
    const int64_t e = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (e >= (int64_t)n2 * n2) return;
    const int i = (int)(e / n2), j = (int)(e % n2);
    double v = 0.0;
    if (i < n && j < n) {
        for (int z = lane; z < batch; z += 32)
            v += P[z * strideP + (int64_t)i * ldp + j] + P[z * strideP + (int64_t)j * ldp + i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        v *= 0.5;
    }
    if (lane == 0) G[e] = v;
}

// ---- parallel two-sided Jacobi on the n2 x n2 symmetric matrix A (n2 even), eigenvectors accumulated in V -----
// Round-robin (chess tournament) ordering: step s pairs up all n2 indices into n2/2 disjoint (p, q); the n2/2
// rotations commute, so A <- J^T A J is applied to every 2 x 2 block independently.
__device__ __forceinline__ void jacobi_pair(int s, int t, int n2, int& p, int& q) {
    const int m1 = n2 - 1;
    int a, b;
    if (t == 0) {
        a = m1;
        b = s;
    } else {
        a = (s + t) % m1;
        b = (s - t + m1) % m1;
    }
    p = min(a, b);
    q = max(a, b);
}

// rotation that annihilates A[p][q]: A <- J^T A J with J = [[c, s], [-s, c]] in the (p, q) plane
__device__ __forceinline__ double2 jacobi_rotation(const double* __restrict__ A, int n2, int n, int p, int q) {
    double c = 1.0, s = 0.0;
    if (q < n) {
        const double apq = A[(int64_t)p * n2 + q], app = A[(int64_t)p * n2 + p], aqq = A[(int64_t)q * n2 + q];
        if (apq != 0.0 && fabs(apq) > 1e-18 * sqrt(fabs(app * aqq))) {
            const double theta = (aqq - app) / (2.0 * apq);
            const double tt = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
            c = 1.0 / sqrt(tt * tt + 1.0);
            s = tt * c;
        }
    }
    return make_double2(c, s);
}

// One round-robin step in ONE launch: every thread derives the two rotations it needs from the previous matrix Ain
// (read-only during the launch) and writes its 2 x 2 block of Aout; V is rotated in place (its threads read Ain only).
__global__ void __launch_bounds__(PT) pca_jacobi_step_kernel(const double* __restrict__ Ain, double* __restrict__ Aout,
                                                             double* __restrict__ V, int n2, int n, int step) {
    const int m = n2 / 2;
    const int64_t e = (int64_t)blockIdx.x * PT + threadIdx.x;
    const int64_t nA = (int64_t)m * m;
    if (e < nA) {
        const int ta = (int)(e / m), tb = (int)(e % m);
        int pa, qa, pb, qb;
        jacobi_pair(step, ta, n2, pa, qa);
        jacobi_pair(step, tb, n2, pb, qb);
        const double2 ra = jacobi_rotation(Ain, n2, n, pa, qa);
        const double2 rb = (ta == tb) ? ra : jacobi_rotation(Ain, n2, n, pb, qb);
        const double* a0 = Ain + (int64_t)pa * n2;
        const double* a1 = Ain + (int64_t)qa * n2;
        const double b00 = a0[pb], b01 = a0[qb], b10 = a1[pb], b11 = a1[qb];
        const double n00 = rb.x * b00 - rb.y * b01, n01 = rb.y * b00 + rb.x * b01;
        const double n10 = rb.x * b10 - rb.y * b11, n11 = rb.y * b10 + rb.x * b11;
        double r00 = ra.x * n00 - ra.y * n10, r10 = ra.y * n00 + ra.x * n10;
        double r01 = ra.x * n01 - ra.y * n11, r11 = ra.y * n01 + ra.x * n11;
        if (ta == tb && ra.y != 0.0) r01 = r10 = 0.0;  // the annihilated pair
        double* o0 = Aout + (int64_t)pa * n2;
        double* o1 = Aout + (int64_t)qa * n2;
        o0[pb] = r00; o0[qb] = r01; o1[pb] = r10; o1[qb] = r11;
        return;
    }
    const int64_t f = e - nA;
    if (f >= (int64_t)n2 * m) return;
    const int i = (int)(f / m), tb = (int)(f % m);
    int pb, qb;
    jacobi_pair(step, tb, n2, pb, qb);
    const double2 rb = jacobi_rotation(Ain, n2, n, pb, qb);
    if (rb.y == 0.0) return;
    double* v = V + (int64_t)i * n2;
    const double vp = v[pb], vq = v[qb];
    v[pb] = rb.x * vp - rb.y * vq;
    v[qb] = rb.y * vp + rb.x * vq;
}

// out[0] = sum of squares off the diagonal, out[1] = on it (one block: fixed reduction order)
__global__ void __launch_bounds__(1024) pca_offnorm_kernel(const double* __restrict__ A, int n2, double* __restrict__ out) {
    __shared__ double so[1024], sd[1024];
    double off = 0.0, dg = 0.0;
    const int64_t tot = (int64_t)n2 * n2;
    for (int64_t e = threadIdx.x; e < tot; e += 1024) {
        const double v = A[e];
        if (e / n2 == e % n2) dg += v * v;
        else off += v * v;
    }
    so[threadIdx.x] = off;
    sd[threadIdx.x] = dg;
    __syncthreads();
    for (int w = 512; w > 0; w >>= 1) {
        if (threadIdx.x < w) {
            so[threadIdx.x] += so[threadIdx.x + w];
            sd[threadIdx.x] += sd[threadIdx.x + w];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        out[0] = so[0];
        out[1] = sd[0];
    }
}

__global__ void pca_diag_kernel(const double* __restrict__ A, int n2, int n, double* __restrict__ lam) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) lam[i] = A[(int64_t)i * n2 + i];
}

// W[i][k] = V[i][perm[k]] * scale[k]  (n x n into ldw)
__global__ void pca_build_w_kernel(const double* __restrict__ V, int n2, int n, const int* __restrict__ perm,
                                   const double* __restrict__ scale, double* __restrict__ W, int64_t ldw) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (int64_t)n * n) return;
    const int i = (int)(e / n), k = (int)(e % n);
    W[(int64_t)i * ldw + k] = V[(int64_t)i * n2 + perm[k]] * scale[k];
}

// Binv = (U diag(sigma))[:, :nF], Basis = (U diag(1/sigma))[:, :nF]; numerically null modes give zero columns
__global__ void pca_pack_kernel(const double* __restrict__ U, int64_t ldu, const double* __restrict__ sigma, int64_t D,
                                int nF, double null_below, double* __restrict__ Binv, double* __restrict__ Basis, int64_t ldb) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= D * nF) return;
    const int64_t i = e / nF;
    const int k = (int)(e % nF);
    const double u = U[i * ldu + k], s = sigma[k];
    Binv[i * ldb + k] = u * s;
    Basis[i * ldb + k] = (s > null_below) ? u / s : 0.0;
}

// ---- project: F[k][t] = scale_k * sum_i B[i][k] (X[i][t] - mean[i]) ------------------------------------------
// grid (ceil(K / 32), splits): HBM stream over B (8 D K bytes). A group of 16 lanes reads one row's 32 modes with
// one 16-byte load per lane, PU rows per group in flight; every lane keeps two mode sums per column.
// partial[split][k][t], combined in a fixed order by pca_project_reduce_kernel.
template <int NC>
__global__ void __launch_bounds__(PT) pca_project_kernel(const double* __restrict__ B, int64_t ldb, int K,
                                                         const double* __restrict__ X, int64_t ldx, int T,
                                                         const double* __restrict__ mean, int64_t D, int64_t rows_per_split,
                                                         double* __restrict__ partial) {
    constexpr int PU = (NC == 1) ? 8 : 4;  // rows in flight per group
    const int tid = threadIdx.x, sub = tid & 15, grp = tid >> 4;
    const int k = blockIdx.x * 32 + 2 * sub;
    const int64_t r0 = (int64_t)blockIdx.y * rows_per_split;
    const int64_t r1 = min(D, r0 + rows_per_split);
    double acc[2][NC];
#pragma unroll
    for (int t = 0; t < NC; ++t) acc[0][t] = acc[1][t] = 0.0;
    const bool live0 = k < K, live1 = k + 1 < K;
    for (int64_t i0 = r0 + grp * PU; i0 < r1; i0 += (PT / 16) * PU) {
        double2 b[PU];
        double xc[PU][NC];
        if (i0 + PU <= r1 && live1 && T == NC) {
            // whole group of rows, both modes, all columns: every load is issued before the first use (with the per-row
            // bounds checks below, the subtraction after each row's loads kept ONE row in flight per thread: 40 % of HBM)
            double mu[PU];
#pragma unroll
            for (int u = 0; u < PU; ++u) b[u] = __ldcs(reinterpret_cast<const double2*>(B + (i0 + u) * ldb + k));
#pragma unroll
            for (int u = 0; u < PU; ++u) mu[u] = mean[i0 + u];
#pragma unroll
            for (int u = 0; u < PU; ++u)
#pragma unroll
                for (int t = 0; t < NC; ++t) xc[u][t] = X[(i0 + u) * ldx + t];
#pragma unroll
            for (int u = 0; u < PU; ++u)
#pragma unroll
                for (int t = 0; t < NC; ++t) xc[u][t] -= mu[u];
        } else {
#pragma unroll
            for (int u = 0; u < PU; ++u) {
                const int64_t i = i0 + u;
                b[u] = make_double2(0.0, 0.0);
                if (i < r1) {
                    const double* src = B + i * ldb + k;  // ldb even, k even: 16-byte aligned
                    if (live1) b[u] = __ldcs(reinterpret_cast<const double2*>(src));
                    else if (live0) b[u].x = __ldcs(src);
                    const double mu = mean[i];
#pragma unroll
                    for (int t = 0; t < NC; ++t) xc[u][t] = (t < T) ? X[i * ldx + t] - mu : 0.0;
                } else {
#pragma unroll
                    for (int t = 0; t < NC; ++t) xc[u][t] = 0.0;
                }
            }
        }
#pragma unroll
        for (int u = 0; u < PU; ++u)
#pragma unroll
            for (int t = 0; t < NC; ++t) {
                acc[0][t] = fma(b[u].x, xc[u][t], acc[0][t]);
                acc[1][t] = fma(b[u].y, xc[u][t], acc[1][t]);
            }
    }
    __shared__ double red[PT / 16][32][NC];
#pragma unroll
    for (int t = 0; t < NC; ++t) {
        red[grp][2 * sub][t] = acc[0][t];
        red[grp][2 * sub + 1][t] = acc[1][t];
    }
    __syncthreads();
    if (tid < 32 && blockIdx.x * 32 + tid < K) {
#pragma unroll
        for (int t = 0; t < NC; ++t) {
            double s = 0.0;
#pragma unroll
            for (int g = 0; g < PT / 16; ++g) s += red[g][tid][t];
            partial[((int64_t)blockIdx.y * K + blockIdx.x * 32 + tid) * NC + t] = s;
        }
    }
}

// one warp per mode k: lanes stride the splits, fixed-order shuffle tree (deterministic)
template <int NC>
__global__ void __launch_bounds__(256) pca_project_reduce_kernel(const double* __restrict__ partial, int splits, int K, int T,
                                                                 const double* __restrict__ sigma, double null_below,
                                                                 double* __restrict__ F, int64_t ldf) {
    const int lane = threadIdx.x & 31;
    const int k = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (k >= K) return;
    double s[NC];
#pragma unroll
    for (int t = 0; t < NC; ++t) s[t] = 0.0;
    for (int z = lane; z < splits; z += 32) {
        const double* p = partial + ((int64_t)z * K + k) * NC;
#pragma unroll
        for (int t = 0; t < NC; ++t) s[t] += p[t];
    }
#pragma unroll
    for (int t = 0; t < NC; ++t) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s[t] += __shfl_xor_sync(0xffffffffu, s[t], o);
        if (lane == 0 && t < T) {
            double v = s[t];
            if (sigma) v = (sigma[k] > null_below) ? v / sigma[k] : 0.0;
            F[(int64_t)k * ldf + t] = v;
        }
    }
}

// ---- reconstruct: out[i][t] = mean[i] + sum_k Binv[i][k] w(k, t) ------------------------------------------------
// HBM stream over Binv (8 D nF bytes). A group of 16 lanes owns a row: one 16-byte load per lane covers 32 modes,
// RU rows per group are in flight at once, the dot products are folded with 4 shuffle steps. The weights sit in
// shared memory. w(k, t) = W[k*ldw + t], or W[t*ldw + k] when they come straight out of Predict (T x M row-major).
constexpr int RT = 256;   // threads per CTA = 16 row groups
constexpr int RU = 4;     // rows in flight per group
constexpr int RROWS = (RT / 16) * RU;
constexpr int RMAXF = 512;  // modes held in shared memory per pass
template <int NC>
__global__ void __launch_bounds__(RT) pca_reconstruct_kernel(const double* __restrict__ Binv, int64_t ldb, int nF,
                                                             const double* __restrict__ W, int64_t ldw, int w_transposed,
                                                             int T, const double* __restrict__ mean, int64_t D,
                                                             double* __restrict__ out, int64_t ldo) {
    __shared__ double sw[NC][RMAXF];
    const int tid = threadIdx.x, sub = tid & 15, grp = tid >> 4;
    const int64_t row0 = (int64_t)blockIdx.x * RROWS + grp * RU;
    double acc[RU][NC];
#pragma unroll
    for (int u = 0; u < RU; ++u)
#pragma unroll
        for (int t = 0; t < NC; ++t) acc[u][t] = 0.0;
    for (int f0 = 0; f0 < nF; f0 += RMAXF) {
        const int fn = min(RMAXF, nF - f0);
        __syncthreads();
        for (int e = tid; e < NC * RMAXF; e += RT) {
            const int t = e / RMAXF, k = e % RMAXF;
            double v = 0.0;
            if (k < fn && t < T) v = w_transposed ? W[(int64_t)t * ldw + f0 + k] : W[(int64_t)(f0 + k) * ldw + t];
            sw[t][k] = v;
        }
        __syncthreads();
        for (int k0 = 0; k0 < fn; k0 += 32) {
            const int k = k0 + 2 * sub;  // ldb is even and Binv 16-byte aligned: (row, k) starts a double2
            double2 b[RU];
#pragma unroll
            for (int u = 0; u < RU; ++u) {
                b[u] = make_double2(0.0, 0.0);
                if (row0 + u < D && k < fn) {
                    const double* src = Binv + (row0 + u) * ldb + f0 + k;
                    if (k + 1 < fn) b[u] = __ldcs(reinterpret_cast<const double2*>(src));
                    else b[u].x = __ldcs(src);
                }
            }
            const int kc = min(k, RMAXF - 2);
#pragma unroll
            for (int t = 0; t < NC; ++t) {
                const double w0 = sw[t][kc], w1 = sw[t][kc + 1];
#pragma unroll
                for (int u = 0; u < RU; ++u) acc[u][t] = fma(b[u].y, w1, fma(b[u].x, w0, acc[u][t]));
            }
        }
    }
#pragma unroll
    for (int u = 0; u < RU; ++u)
#pragma unroll
        for (int t = 0; t < NC; ++t) {
            double v = acc[u][t];
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            acc[u][t] = v;
        }
    if (sub < RU) {  // lane u of the group writes row u (consecutive rows -> adjacent addresses)
        const int64_t row = row0 + sub;
        if (row < D) {
            const double mu = mean[row];
#pragma unroll
            for (int t = 0; t < NC; ++t) {
                double v = 0.0;
#pragma unroll
                for (int u = 0; u < RU; ++u)
                    if (u == sub) v = acc[u][t];
                if (t < T) out[row * ldo + t] = mu + v;
            }
        }
    }
}

template <int NC>
int project_chunk(PcaModel* m, const double* B, int64_t ldb, int K, const double* sigma, double null_below,
                  const double* dX, int64_t ldx, int T, double* dF, int64_t ldf) {
    Ctx* c = m->c;
    const int kblocks = (K + 31) / 32;
    int64_t splits = std::max<int64_t>(1, (148 * 3 + kblocks - 1) / kblocks);  // one resident wave (3 CTAs/SM at 82 regs)
    splits = std::min<int64_t>(splits, std::max<int64_t>(1, m->D / 256));
    const int64_t rps = (m->D + splits - 1) / splits;
    splits = (m->D + rps - 1) / rps;
    double* partial = nullptr;
    GP_CUDA(c, cudaMallocAsync((void**)&partial, sizeof(double) * splits * K * NC, c->stream));
    pca_project_kernel<NC><<<dim3((unsigned)kblocks, (unsigned)splits), PT, 0, c->stream>>>(B, ldb, K, dX, ldx, T, m->dMean,
                                                                                           m->D, rps, partial);
    pca_project_reduce_kernel<NC><<<(K + 7) / 8, 256, 0, c->stream>>>(partial, (int)splits, K, T, sigma, null_below,
                                                                              dF, ldf);
    c->launches += 2;
    GP_CUDA(c, cudaGetLastError());
    GP_CUDA(c, cudaFreeAsync(partial, c->stream));
    return 0;
}

// features (rows_out x T, ldf) from device X (D x T, ldx). rows_out <= nF uses the compact basis, otherwise U / sigma.
int project_dev(PcaModel* m, const double* dX, int64_t ldx, int64_t T, int64_t rows_out, double* dF, int64_t ldf) {
    Ctx* c = m->c;
    if (!m->dU) return fail(c, -2, "PCA: no basis (fit or set first)");
    if (rows_out <= 0 || rows_out > m->n) return fail(c, -2, "PCA::DimensionalityReduction: number of features out of range");
    const bool compact = rows_out <= m->nF;
    const double* B = compact ? m->dBasis : m->dU;
    const int64_t ldb = compact ? m->ldb : m->ldu;
    const double* sg = compact ? nullptr : m->dSigma;
    for (int64_t t0 = 0; t0 < T; t0 += 8) {
        const int tn = (int)std::min<int64_t>(8, T - t0);
        const double* x = dX + t0;
        double* f = dF + t0;
        if (tn <= 1) GP_TRY(project_chunk<1>(m, B, ldb, (int)rows_out, sg, 0.0, x, ldx, tn, f, ldf));
        else if (tn <= 4) GP_TRY(project_chunk<4>(m, B, ldb, (int)rows_out, sg, 0.0, x, ldx, tn, f, ldf));
        else GP_TRY(project_chunk<8>(m, B, ldb, (int)rows_out, sg, 0.0, x, ldx, tn, f, ldf));
    }
    return 0;
}

int reconstruct_dev(PcaModel* m, const double* dW, int64_t ldw, int transposed, int64_t nrows_w, int64_t T, double* dOut,
                    int64_t ldo) {
    Ctx* c = m->c;
    if (!m->dBinv) return fail(c, -2, "PCA: no basis (fit or set first)");
    if (nrows_w != m->nF) return fail(c, -2, "PCA::GetReconstruction: weights must have nFeatures rows");
    const unsigned blocks = (unsigned)((m->D + RROWS - 1) / RROWS);
    for (int64_t t0 = 0; t0 < T; t0 += 8) {
        const int tn = (int)std::min<int64_t>(8, T - t0);
        const double* w = transposed ? dW + t0 * ldw : dW + t0;
        double* o = dOut + t0;
        if (tn <= 1)
            pca_reconstruct_kernel<1><<<blocks, RT, 0, c->stream>>>(m->dBinv, m->ldb, (int)m->nF, w, ldw, transposed, tn,
                                                                   m->dMean, m->D, o, ldo);
        else if (tn <= 4)
            pca_reconstruct_kernel<4><<<blocks, RT, 0, c->stream>>>(m->dBinv, m->ldb, (int)m->nF, w, ldw, transposed, tn,
                                                                   m->dMean, m->D, o, ldo);
        else
            pca_reconstruct_kernel<8><<<blocks, RT, 0, c->stream>>>(m->dBinv, m->ldb, (int)m->nF, w, ldw, transposed, tn,
                                                                   m->dMean, m->D, o, ldo);
        ++c->launches;
        GP_CUDA(c, cudaGetLastError());
    }
    return 0;
}

int pack_basis(PcaModel* m) {
    Ctx* c = m->c;
    m->ldb = round_up(m->nF > 0 ? m->nF : 1, 2);
    cudaFree(m->dBinv); cudaFree(m->dBasis);
    m->dBinv = m->dBasis = nullptr;
    if (m->nF <= 0) return 0;
    GP_TRY(pca_alloc(c, &m->dBinv, (size_t)m->D * m->ldb));
    GP_TRY(pca_alloc(c, &m->dBasis, (size_t)m->D * m->ldb));
    const int64_t tot = m->D * m->nF;
    pca_pack_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, c->stream>>>(m->dU, m->ldu, m->dSigma, m->D, (int)m->nF, 0.0,
                                                                         m->dBinv, m->dBasis, m->ldb);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}

// ---- fit ---------------------------------------------------------------------------------------------------------
// Jacobi eigen-decomposition of the n2 x n2 matrix in dA (destroyed), dB = same-size work matrix (the steps
// ping-pong between the two); on return *dRes points at the one whose diagonal holds the eigenvalues, dV = vectors.
int jacobi_eig(PcaModel* m, double* dA, double* dB, double* dV, int n2, int n, double* dNorm, int* sweeps_out,
               double** dRes) {
    Ctx* c = m->c;
    GP_TRY(set_identity(c, dV, n2, n2));
    const int half = n2 / 2;
    const int64_t work = (int64_t)half * half + (int64_t)n2 * half;
    const unsigned ablocks = (unsigned)((work + PT - 1) / PT);
    const double tol = (double)n2 * 1e-16;
    bool converged = false;
    int sweep = 0;
    double *cur = dA, *nxt = dB;
    for (; sweep < 40; ++sweep) {
        for (int step = 0; step < n2 - 1; ++step) {
            pca_jacobi_step_kernel<<<ablocks, PT, 0, c->stream>>>(cur, nxt, dV, n2, n, step);
            std::swap(cur, nxt);
        }
        c->launches += (int64_t)(n2 - 1);
        pca_offnorm_kernel<<<1, 1024, 0, c->stream>>>(cur, n2, dNorm);
        ++c->launches;
        GP_CUDA(c, cudaGetLastError());
        double nrm[2];
        GP_CUDA(c, cudaMemcpyAsync(nrm, dNorm, sizeof(nrm), cudaMemcpyDeviceToHost, c->stream));
        GP_CUDA(c, cudaStreamSynchronize(c->stream));
        if (!(nrm[0] == nrm[0]) || !(nrm[1] == nrm[1])) return fail(c, -3, "PCA: non-finite values in the data matrix");
        if (converged) { ++sweep; break; }                    // the extra sweep after reaching the tolerance
        if (nrm[0] <= tol * tol * nrm[1]) converged = true;   // off(A) <= n eps ||A||_F
        if (nrm[0] == 0.0) { ++sweep; break; }
    }
    *sweeps_out = sweep;
    *dRes = cur;
    return 0;
}

int fit_dev(PcaModel* m, const double* dX, int64_t ldx, int64_t D, int64_t n, int64_t nF) {
    Ctx* c = m->c;
    if (D <= 0 || n <= 0 || !dX) return fail(c, -2, "PCA: need a D x n data matrix");
    if (n > 8192) return fail(c, -2, "PCA: more than 8192 samples is not supported (n x n Jacobi on one GPU)");
    if (n > D) return fail(c, -2, "PCA: expects D >= n (features x samples)");
    if (nF < 0 || nF > n) return fail(c, -2, "PCA: nFeatures must be in [0, n]");
    pca_free_all(m);
    cudaEvent_t e0, e1;
    GP_CUDA(c, cudaEventCreate(&e0));
    GP_CUDA(c, cudaEventCreate(&e1));
    GP_CUDA(c, cudaEventRecord(e0, c->stream));
    m->D = D; m->n = n; m->nF = nF;
    m->ldu = round_up(n, 16);
    const int n2 = (int)(n + (n & 1));
    // split D so that tiles x batch fills the machine; every slab has the same depth (zero rows pad the last one)
    const int64_t tiles = ((n + 127) / 128) * ((n + 63) / 64);
    int64_t batch = std::max<int64_t>(1, (148 * 4 + tiles - 1) / tiles);
    batch = std::min<int64_t>(batch, std::max<int64_t>(1, D / 2048));
    const int64_t slab = round_up((D + batch - 1) / batch, 16);
    batch = (D + slab - 1) / slab;
    const int64_t Dpad = batch * slab;
    const int64_t ldg = round_up(n, 16);

    GP_TRY(pca_alloc(c, &m->dMean, (size_t)D));
    GP_TRY(pca_alloc(c, &m->dSigma, (size_t)n));
    GP_TRY(pca_alloc(c, &m->dU, (size_t)Dpad * m->ldu));
    double *dZt = nullptr, *dPart = nullptr, *dA = nullptr, *dB = nullptr, *dV = nullptr, *dW = nullptr, *dNorm = nullptr, *dScale = nullptr;
    int* dPerm = nullptr;
    int rc = 0;
    auto body = [&]() -> int {
        GP_TRY(pca_alloc(c, &dZt, (size_t)Dpad * m->ldu));
        GP_TRY(pca_alloc(c, &dPart, (size_t)batch * n * ldg));
        GP_TRY(pca_alloc(c, &dA, (size_t)n2 * n2));
        GP_TRY(pca_alloc(c, &dB, (size_t)n2 * n2));
        GP_TRY(pca_alloc(c, &dV, (size_t)n2 * n2));
        GP_TRY(pca_alloc(c, &dW, (size_t)n * ldg));
        GP_TRY(pca_alloc(c, &dNorm, 2));
        GP_TRY(pca_alloc(c, &dScale, (size_t)n));
        GP_TRY(pca_alloc(c, &dPerm, (size_t)n));
        const int passes = std::max(1, m->passes);
        double* zin = (passes % 2 == 0) ? m->dU : dZt;  // ping-pong so that the last product lands in dU
        double* zout = (passes % 2 == 0) ? dZt : m->dU;
        GP_CUDA(c, cudaMemsetAsync(m->dU, 0, sizeof(double) * Dpad * m->ldu, c->stream));
        GP_CUDA(c, cudaMemsetAsync(dZt, 0, sizeof(double) * Dpad * m->ldu, c->stream));
        pca_center_kernel<<<(unsigned)((D + PT / 32 - 1) / (PT / 32)), PT, 0, c->stream>>>(dX, ldx, D, (int)n, m->dMean, zin,
                                                                                         m->ldu);
        ++c->launches;
        GP_CUDA(c, cudaGetLastError());
        std::vector<double> lam((size_t)n), scale((size_t)n);
        std::vector<int> perm((size_t)n);
        m->sweeps = 0;
        for (int pass = 0; pass < passes; ++pass) {
            const bool last = (pass == passes - 1);
            GemmArgs g{};  // P_z = Z_z^T Z_z
            g.A = zin; g.lda = m->ldu; g.B = zin; g.ldb = m->ldu; g.C = dPart; g.ldc = ldg;
            g.M = (int)n; g.N = (int)n; g.K = (int)slab; g.alpha = 1.0; g.beta = 0.0;
            g.batch = (int)batch; g.sA = slab * m->ldu; g.sB = slab * m->ldu; g.sC = n * ldg;
            GP_TRY(gemm(c, GEMM_TN, g));
            pca_gram_reduce_kernel<<<(unsigned)(((int64_t)n2 * n2 + 7) / 8), 256, 0, c->stream>>>(dPart, (int)batch, (int)n,
                                                                                                    ldg, n * ldg, dA, n2);
            ++c->launches;
            GP_CUDA(c, cudaGetLastError());
            int sw = 0;
            double* dEig = nullptr;
            GP_TRY(jacobi_eig(m, dA, dB, dV, n2, (int)n, dNorm, &sw, &dEig));
            m->sweeps += sw;
            pca_diag_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(dEig, n2, (int)n, dScale);
            ++c->launches;
            GP_CUDA(c, cudaMemcpyAsync(lam.data(), dScale, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
            GP_CUDA(c, cudaStreamSynchronize(c->stream));
            std::iota(perm.begin(), perm.end(), 0);
            std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) { return lam[(size_t)a] > lam[(size_t)b]; });
            const double lam0 = std::max(lam[(size_t)perm[0]], 0.0);
            const double null_below = (double)n * 2.3e-16 * lam0;  // eigenvalues of the Gram below its rounding floor
            std::vector<double> sig((size_t)n);
            for (int64_t k = 0; k < n; ++k) {
                const double l = lam[(size_t)perm[(size_t)k]];
                sig[(size_t)k] = std::sqrt(std::max(l, 0.0) / (double)n);
                scale[(size_t)k] = !last ? 1.0 : (l > null_below && l > 0.0 ? 1.0 / std::sqrt(l) : 0.0);
            }
            GP_CUDA(c, cudaMemcpyAsync(dPerm, perm.data(), sizeof(int) * n, cudaMemcpyHostToDevice, c->stream));
            GP_CUDA(c, cudaMemcpyAsync(dScale, scale.data(), sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
            if (last) GP_CUDA(c, cudaMemcpyAsync(m->dSigma, sig.data(), sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
            pca_build_w_kernel<<<(unsigned)((n * n + 255) / 256), 256, 0, c->stream>>>(dV, n2, (int)n, dPerm, dScale, dW, ldg);
            ++c->launches;
            GP_CUDA(c, cudaGetLastError());
            GemmArgs u{};  // Z_out = Z_in W
            u.A = zin; u.lda = m->ldu; u.B = dW; u.ldb = ldg; u.C = zout; u.ldc = m->ldu;
            u.M = (int)Dpad; u.N = (int)n; u.K = (int)n; u.alpha = 1.0; u.beta = 0.0;
            GP_TRY(gemm(c, GEMM_NN, u));
            GP_CUDA(c, cudaStreamSynchronize(c->stream));  // perm / scale / sig are host vectors reused by the next pass
            std::swap(zin, zout);
        }
        return pack_basis(m);
    };
    rc = body();
    cudaStreamSynchronize(c->stream);
    cudaFree(dZt); cudaFree(dPart); cudaFree(dA); cudaFree(dB); cudaFree(dV); cudaFree(dW); cudaFree(dNorm); cudaFree(dScale);
    cudaFree(dPerm);
    if (rc == 0) {
        cudaEventRecord(e1, c->stream);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        m->fit_ms = ms;
    } else {
        pca_free_all(m);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return rc;
}

int pca_use_device(PcaModel* m) {
    GP_CUDA(m->c, cudaSetDevice(m->c->device));
    return 0;
}

struct Scratch {  // stream-ordered device scratch
    Ctx* c;
    double* p = nullptr;
    explicit Scratch(Ctx* c_) : c(c_) {}
    int alloc(size_t count) {
        GP_CUDA(c, cudaMallocAsync((void**)&p, sizeof(double) * (count ? count : 1), c->stream));
        return 0;
    }
    ~Scratch() {
        if (p) cudaFreeAsync(p, c->stream);
    }
};

}  // namespace
}  // namespace gprb200

using namespace gprb200;

struct gprb200_pca : gprb200::PcaModel {};

extern "C" {

int gprb200_pca_create(gprb200_handle h, gprb200_pca_handle* out) {
    if (!h || !out) return -2;
    gprb200_pca* p = new (std::nothrow) gprb200_pca();
    if (!p) return fail(h, -1, "pca_create: out of host memory");
    p->c = h;
    *out = p;
    return 0;
}

int gprb200_pca_destroy(gprb200_pca_handle p) {
    if (!p) return 0;
    cudaSetDevice(p->c->device);
    cudaStreamSynchronize(p->c->stream);
    pca_free_all(p);
    delete p;
    return 0;
}

int gprb200_pca_set_option(gprb200_pca_handle p, const char* name, int64_t value) {
    if (!p || !name) return -2;
    if (std::string(name) == "passes") {
        if (value < 1 || value > 4) return fail(p->c, -2, "pca option passes must be in [1, 4]");
        p->passes = (int)value;
        return 0;
    }
    return fail(p->c, -2, std::string("unknown pca option: ") + name);
}

int gprb200_pca_fit_dev(gprb200_pca_handle p, const double* dX, int64_t ldx, int64_t D, int64_t n, int64_t n_features) {
    if (!p) return -2;
    GP_TRY(pca_use_device(p));
    return fit_dev(p, dX, ldx, D, n, n_features);
}

int gprb200_pca_fit(gprb200_pca_handle p, const double* X, int64_t D, int64_t n, int64_t n_features) {
    if (!p) return -2;
    Ctx* c = p->c;
    GP_TRY(pca_use_device(p));
    if (!X || D <= 0 || n <= 0) return fail(c, -2, "PCA: need a D x n data matrix");
    double* dX = nullptr;
    GP_TRY(pca_alloc(c, &dX, (size_t)D * n));
    cudaError_t e = cudaMemcpyAsync(dX, X, sizeof(double) * D * n, cudaMemcpyHostToDevice, c->stream);
    int rc = (e == cudaSuccess) ? fit_dev(p, dX, n, D, n, n_features) : fail(c, -100 - (int)e, cudaGetErrorString(e));
    cudaStreamSynchronize(c->stream);
    cudaFree(dX);
    return rc;
}

int gprb200_pca_set(gprb200_pca_handle p, const double* mean, const double* sigma, const double* U, int64_t D, int64_t n,
                    int64_t n_features) {
    if (!p) return -2;
    Ctx* c = p->c;
    GP_TRY(pca_use_device(p));
    if (!mean || !sigma || !U || D <= 0 || n <= 0) return fail(c, -2, "PCA: need mean (D), sigma (n) and U (D x n)");
    if (n_features < 0 || n_features > n) return fail(c, -2, "PCA: nFeatures must be in [0, n]");
    pca_free_all(p);
    p->D = D; p->n = n; p->nF = n_features;
    p->ldu = round_up(n, 16);
    GP_TRY(pca_alloc(c, &p->dMean, (size_t)D));
    GP_TRY(pca_alloc(c, &p->dSigma, (size_t)n));
    GP_TRY(pca_alloc(c, &p->dU, (size_t)D * p->ldu));
    GP_CUDA(c, cudaMemcpyAsync(p->dMean, mean, sizeof(double) * D, cudaMemcpyHostToDevice, c->stream));
    GP_CUDA(c, cudaMemcpyAsync(p->dSigma, sigma, sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
    GP_CUDA(c, cudaMemcpy2DAsync(p->dU, sizeof(double) * p->ldu, U, sizeof(double) * n, sizeof(double) * n, (size_t)D,
                                 cudaMemcpyHostToDevice, c->stream));
    GP_TRY(pack_basis(p));
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

int gprb200_pca_dims(gprb200_pca_handle p, int64_t* D, int64_t* n, int64_t* n_features) {
    if (!p) return -2;
    if (D) *D = p->D;
    if (n) *n = p->n;
    if (n_features) *n_features = p->nF;
    return 0;
}

int gprb200_pca_get(gprb200_pca_handle p, double* mean, double* sigma, double* U) {
    if (!p) return -2;
    Ctx* c = p->c;
    GP_TRY(pca_use_device(p));
    if (!p->dU) return fail(c, -2, "PCA: no basis (fit or set first)");
    if (mean) GP_CUDA(c, cudaMemcpyAsync(mean, p->dMean, sizeof(double) * p->D, cudaMemcpyDeviceToHost, c->stream));
    if (sigma) GP_CUDA(c, cudaMemcpyAsync(sigma, p->dSigma, sizeof(double) * p->n, cudaMemcpyDeviceToHost, c->stream));
    if (U)
        GP_CUDA(c, cudaMemcpy2DAsync(U, sizeof(double) * p->n, p->dU, sizeof(double) * p->ldu, sizeof(double) * p->n,
                                     (size_t)p->D, cudaMemcpyDeviceToHost, c->stream));
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

int gprb200_pca_get_stat(gprb200_pca_handle p, const char* name, double* out) {
    if (!p || !name || !out) return -2;
    const std::string s(name);
    if (s == "fit_ms") *out = p->fit_ms;
    else if (s == "sweeps") *out = p->sweeps;
    else return fail(p->c, -2, "unknown pca stat: " + s);
    return 0;
}

int gprb200_pca_project_dev(gprb200_pca_handle p, const double* dX, int64_t ldx, int64_t T, int64_t rows_out, double* dF,
                            int64_t ldf) {
    if (!p) return -2;
    GP_TRY(pca_use_device(p));
    if (!dX || !dF || T <= 0) return fail(p->c, -2, "PCA::DimensionalityReduction: need T > 0 columns");
    return project_dev(p, dX, ldx, T, rows_out, dF, ldf);
}

int gprb200_pca_project(gprb200_pca_handle p, const double* X, int64_t T, int64_t rows_out, double* F) {
    if (!p) return -2;
    Ctx* c = p->c;
    GP_TRY(pca_use_device(p));
    if (!X || !F || T <= 0) return fail(c, -2, "PCA::DimensionalityReduction: need T > 0 columns");
    if (!p->dU) return fail(c, -2, "PCA: no basis (fit or set first)");
    Scratch x(c), f(c);
    GP_TRY(x.alloc((size_t)p->D * T));
    GP_TRY(f.alloc((size_t)rows_out * T));
    GP_CUDA(c, cudaMemcpyAsync(x.p, X, sizeof(double) * p->D * T, cudaMemcpyHostToDevice, c->stream));
    GP_TRY(project_dev(p, x.p, T, T, rows_out, f.p, T));
    GP_CUDA(c, cudaMemcpyAsync(F, f.p, sizeof(double) * rows_out * T, cudaMemcpyDeviceToHost, c->stream));
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

int gprb200_pca_reconstruct_dev(gprb200_pca_handle p, const double* dW, int64_t ldw, int transposed, int64_t n_weights,
                                int64_t T, double* dOut, int64_t ldo) {
    if (!p) return -2;
    GP_TRY(pca_use_device(p));
    if (!dW || !dOut || T <= 0) return fail(p->c, -2, "PCA::GetReconstruction: need T > 0 columns");
    return reconstruct_dev(p, dW, ldw, transposed, n_weights, T, dOut, ldo);
}

int gprb200_pca_reconstruct(gprb200_pca_handle p, const double* W, int64_t n_weights, int64_t T, double* Xout) {
    if (!p) return -2;
    Ctx* c = p->c;
    GP_TRY(pca_use_device(p));
    if (!W || !Xout || T <= 0) return fail(c, -2, "PCA::GetReconstruction: need T > 0 columns");
    Scratch w(c), o(c);
    GP_TRY(w.alloc((size_t)n_weights * T));
    GP_TRY(o.alloc((size_t)p->D * T));
    GP_CUDA(c, cudaMemcpyAsync(w.p, W, sizeof(double) * n_weights * T, cudaMemcpyHostToDevice, c->stream));
    GP_TRY(reconstruct_dev(p, w.p, T, 0, n_weights, T, o.p, T));
    GP_CUDA(c, cudaMemcpyAsync(Xout, o.p, sizeof(double) * p->D * T, cudaMemcpyDeviceToHost, c->stream));
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

// Predict -> GetReconstruction without leaving the device: test points in, displacement fields out.
// apps/GaussianProcessPredict.cpp:188-203 (gp->Predict per frame) + include/DataParser.h:471-495 (inversePca).
int gprb200_predict_reconstruct_dev(gprb200_handle h, gprb200_pca_handle p, const double* dXs, int64_t T, double* dFields,
                                    int64_t ldo, double* dWeights, int64_t ldw) {
    if (!h || !p) return -2;
    if (p->c != h) return fail(h, -2, "predict_reconstruct: the PCA object belongs to another handle");
    Ctx* c = h;
    if (c->M != p->nF) return fail(c, -2, "predict_reconstruct: the GP must have nFeatures output columns");
    Scratch mu(c);
    double* w = dWeights;
    int64_t ld = ldw;
    if (!w) {
        ld = round_up(c->M, 16);
        GP_TRY(mu.alloc((size_t)T * ld));
        w = mu.p;
    }
    GP_TRY(gprb200_predict_dev(h, dXs, T, w, ld, nullptr));
    return reconstruct_dev(p, w, ld, 1, c->M, T, dFields, ldo);
}

int gprb200_predict_reconstruct(gprb200_handle h, gprb200_pca_handle p, const double* Xs, int64_t T, double* fields,
                                double* weights) {
    if (!h || !p || !Xs || T <= 0) return -2;
    Ctx* c = h;
    GP_TRY(pca_use_device(p));
    if (!c->have_train) return fail(c, -2, "GaussianProcess::ComputeKernelVectorInternal: gaussian process is not initialized.");
    Scratch xs(c), out(c), mu(c);
    const int64_t ld = round_up(c->M > 0 ? c->M : 1, 16);
    GP_TRY(xs.alloc((size_t)T * c->d));
    GP_TRY(out.alloc((size_t)p->D * T));
    GP_TRY(mu.alloc((size_t)T * ld));
    GP_CUDA(c, cudaMemcpyAsync(xs.p, Xs, sizeof(double) * T * c->d, cudaMemcpyHostToDevice, c->stream));
    GP_TRY(gprb200_predict_reconstruct_dev(h, p, xs.p, T, out.p, T, mu.p, ld));
    if (fields) GP_CUDA(c, cudaMemcpyAsync(fields, out.p, sizeof(double) * p->D * T, cudaMemcpyDeviceToHost, c->stream));
    if (weights)
        GP_CUDA(c, cudaMemcpy2DAsync(weights, sizeof(double) * c->M, mu.p, sizeof(double) * ld, sizeof(double) * c->M,
                                     (size_t)T, cudaMemcpyDeviceToHost, c->stream));
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

}  // extern "C"
