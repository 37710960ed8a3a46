// K4 / K5: triangular solves with the Cholesky factor for many right-hand sides, as GEMMs.
//   forward   L Z = B        (B: N x M row-major, overwritten)
//   backward  L^T X = Z
// Together they replace alpha = C * Y with the explicit inverse C (reference
// lib/GaussianProcess.cpp:661) and, with B = I, rebuild C itself when the API has to expose it
// (-CoreMatrix.txt, lib/GaussianProcess.cpp:152-153; trace term of include/Likelihood.h:280-282).
// Diagonal blocks were inverted by potrf_tile_kernel, so each block step is
//   Z_k = inv(L_kk) B_k ;  B_i -= L_ik Z_k (i > k)        -- two GEMMs, all rows/columns in parallel.
#include "common.cuh"

namespace gprb200 {

// Right-looking form: every step updates all remaining rows (row-parallel; right for few RHS).
static int fwd_right(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* B, int64_t ldb,
                     int64_t M) {
    for (int64_t k0 = 0; k0 < N; k0 += NB) {
        const int n = (int)((N - k0 < NB) ? (N - k0) : NB);
        const double* invk = invL + (k0 / NB) * (int64_t)NB * NB;
        double* Bk = B + k0 * ldb;
        GemmArgs d{};  // Z_k = inv(L_kk) * B_k, in place: one row tile, every CTA owns its columns
        d.A = invk; d.lda = NB; d.B = Bk; d.ldb = ldb; d.C = Bk; d.ldc = ldb;
        d.M = n; d.N = (int)M; d.K = n; d.alpha = 1.0; d.beta = 0.0;
        GP_TRY(gemm(c, GEMM_NN, d));
        const int64_t m = N - k0 - n;
        if (m <= 0) break;
        GemmArgs u{};  // B_i -= L_ik Z_k
        u.A = L + (k0 + n) * ldl + k0; u.lda = ldl; u.B = Bk; u.ldb = ldb; u.C = B + (k0 + n) * ldb; u.ldc = ldb;
        u.M = (int)m; u.N = (int)M; u.K = n; u.alpha = -1.0; u.beta = 1.0;
        GP_TRY(gemm(c, GEMM_NN, u));
    }
    return 0;
}

// Left-looking form: block row k gathers all earlier blocks in ONE long-K GEMM
// (B_k -= L[k, 0:k0] Z[0:k0]); C is read/written once per step instead of once per earlier block.
// Parallel over the M right-hand-side columns, so it is chosen when M is large (C2: M = 786 432).
static int fwd_left(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* B, int64_t ldb,
                    int64_t M) {
    for (int64_t k0 = 0; k0 < N; k0 += NB) {
        const int n = (int)((N - k0 < NB) ? (N - k0) : NB);
        const double* invk = invL + (k0 / NB) * (int64_t)NB * NB;
        double* Bk = B + k0 * ldb;
        if (k0 > 0) {
            GemmArgs u{};
            u.A = L + k0 * ldl; u.lda = ldl; u.B = B; u.ldb = ldb; u.C = Bk; u.ldc = ldb;
            u.M = n; u.N = (int)M; u.K = (int)k0; u.alpha = -1.0; u.beta = 1.0;
            GP_TRY(gemm(c, GEMM_NN, u));
        }
        GemmArgs d{};
        d.A = invk; d.lda = NB; d.B = Bk; d.ldb = ldb; d.C = Bk; d.ldc = ldb;
        d.M = n; d.N = (int)M; d.K = n; d.alpha = 1.0; d.beta = 0.0;
        GP_TRY(gemm(c, GEMM_NN, d));
    }
    return 0;
}


// ---- few right-hand sides (M <= 8): one fused streaming kernel per 128-row block step -----------------------------
// The GEMM formulation above needs two dependent launches per block step and fills 1/64 of a tensor-core tile; with
// M <= 8 the solve is bound by reading L once (4 N^2 bytes per sweep), so each step is ONE kernel:
//   forward  step k: rows below block k get  b_r -= L[r, block k] . z_k  (one warp per row, a row segment of L is
//            one contiguous 1 KB run); the CTA that owns the next 128 rows then applies inv(L_{k+1,k+1}) to them,
//            so z_{k+1} is ready for the next launch.
//   backward step k: columns left of block k get z_j -= sum_r L[block k row r, j] x_k[r] (one thread per column,
//            coalesced across columns); the CTA holding block k-1's columns then applies inv(L_{k-1,k-1})^T.
constexpr int TS_THREADS = 512;
constexpr int TS_ROWS = NB;  // rows per CTA in the forward step
constexpr size_t TS_SMEM = sizeof(double) * NB * NB;  // the owner CTA stages one inverted diagonal block

__device__ __forceinline__ void ts_cp_async16(void* smem, const void* gmem) {
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sa), "l"(gmem));
}
// whole 128 x 128 inverted block -> shared memory, asynchronously (issued before the CTA's streaming work so the
// 128 KB fetch overlaps it); ts_wait_inv() before use.
__device__ __forceinline__ void ts_prefetch_inv(double* sinv, const double* __restrict__ invk) {
    for (int ch = threadIdx.x; ch < NB * NB / 2; ch += TS_THREADS) ts_cp_async16(sinv + ch * 2, invk + ch * 2);
    asm volatile("cp.async.commit_group;\n" ::);
}
__device__ __forceinline__ void ts_wait_inv() {
    asm volatile("cp.async.wait_group 0;\n" ::);
    __syncthreads();
}

// z (n x NC) = inv(L_kk) b  or inv(L_kk)^T b, in place in B (row-major, ldb); sinv = the block in shared memory.
template <int NC>
__device__ __forceinline__ void apply_inv_block(const double* sinv, int n, double* B, int64_t ldb, int M, bool transposed,
                                                double* sb /* [NB][NC] */, double* so /* [NB][NC] */) {
    const int tid = threadIdx.x;
    for (int e = tid; e < n * NC; e += TS_THREADS) {
        const int r = e / NC, c = e - r * NC;
        sb[e] = (c < M) ? B[(int64_t)r * ldb + c] : 0.0;
    }
    __syncthreads();
    if (!transposed) {
        // out[i] = sum_{j <= i} invL[i][j] b[j]: warp per row, lanes stride j (conflict-free), shuffle-reduce
        const int warp = tid >> 5, lane = tid & 31;
        for (int i = warp; i < n; i += TS_THREADS / 32) {
            double acc[NC];
#pragma unroll
            for (int c = 0; c < NC; ++c) acc[c] = 0.0;
            for (int j = lane; j <= i; j += 32) {
                const double v = sinv[i * NB + j];
#pragma unroll
                for (int c = 0; c < NC; ++c) acc[c] = fma(v, sb[j * NC + c], acc[c]);
            }
#pragma unroll
            for (int c = 0; c < NC; ++c)
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], o);
            if (lane == 0) {
#pragma unroll
                for (int c = 0; c < NC; ++c) so[i * NC + c] = acc[c];
            }
        }
    } else {
        // out[i] = sum_{j >= i} invL[j][i] b[j]: thread per column i (conflict-free across i)
        if (tid < n) {
            double acc[NC];
#pragma unroll
            for (int c = 0; c < NC; ++c) acc[c] = 0.0;
            for (int j = tid; j < n; ++j) {
                const double v = sinv[j * NB + tid];
#pragma unroll
                for (int c = 0; c < NC; ++c) acc[c] = fma(v, sb[j * NC + c], acc[c]);
            }
#pragma unroll
            for (int c = 0; c < NC; ++c) so[tid * NC + c] = acc[c];
        }
    }
    __syncthreads();
    for (int e = tid; e < n * NC; e += TS_THREADS) {
        const int r = e / NC, c = e - r * NC;
        if (c < M) B[(int64_t)r * ldb + c] = so[e];
    }
}

// k0 < 0: prologue (only z_0 = inv(L_00) b_0).
template <int NC>
__global__ void __launch_bounds__(TS_THREADS) trsv_fwd_step_kernel(const double* __restrict__ L, int64_t ldl,
                                                                    const double* __restrict__ invL, int64_t N,
                                                                    double* B, int64_t ldb, int M, int64_t k0) {
    extern __shared__ __align__(16) double sinv[];
    __shared__ double sz[NB * NC];
    __shared__ double so[NB * NC];
    const int tid = threadIdx.x;
    if (k0 < 0) {
        ts_prefetch_inv(sinv, invL);
        ts_wait_inv();
        apply_inv_block<NC>(sinv, (int)min((int64_t)NB, N), B, ldb, M, false, sz, so);
        return;
    }
    const int n = NB;  // block k is full whenever rows remain below it
    const int64_t k1 = k0 + n;
    if (blockIdx.x == 0) ts_prefetch_inv(sinv, invL + (k1 / NB) * (int64_t)NB * NB);
    for (int e = tid; e < n * NC; e += TS_THREADS) {
        const int r = e / NC, c = e - r * NC;
        sz[e] = (c < M) ? B[(k0 + r) * ldb + c] : 0.0;
    }
    __syncthreads();
    const int64_t r_begin = k1 + (int64_t)blockIdx.x * TS_ROWS;
    const int64_t r_end = min(N, r_begin + TS_ROWS);
    const int warp = tid >> 5, lane = tid & 31;
    constexpr int WARPS = TS_THREADS / 32;
    // two rows per warp iteration: 4 independent 16-byte loads in flight per lane
    for (int64_t r = r_begin + warp; r < r_end; r += 2 * WARPS) {
        const int64_t r2 = r + WARPS;
        const bool two = r2 < r_end;
        const double* la = L + r * ldl + k0;
        const double* lb = L + (two ? r2 : r) * ldl + k0;
        double2 va[NB / 64], vb[NB / 64];
#pragma unroll
        for (int q = 0; q < NB / 64; ++q) {
            va[q] = *reinterpret_cast<const double2*>(la + q * 64 + lane * 2);
            vb[q] = *reinterpret_cast<const double2*>(lb + q * 64 + lane * 2);
        }
        double acca[NC], accb[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) acca[c] = accb[c] = 0.0;
#pragma unroll
        for (int q = 0; q < NB / 64; ++q) {
            const int j = q * 64 + lane * 2;
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                acca[c] = fma(va[q].y, sz[(j + 1) * NC + c], fma(va[q].x, sz[j * NC + c], acca[c]));
                accb[c] = fma(vb[q].y, sz[(j + 1) * NC + c], fma(vb[q].x, sz[j * NC + c], accb[c]));
            }
        }
#pragma unroll
        for (int c = 0; c < NC; ++c)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                acca[c] += __shfl_xor_sync(0xffffffffu, acca[c], o);
                accb[c] += __shfl_xor_sync(0xffffffffu, accb[c], o);
            }
        if (lane == 0) {
#pragma unroll
            for (int c = 0; c < NC; ++c)
                if (c < M) {
                    B[r * ldb + c] -= acca[c];
                    if (two) B[r2 * ldb + c] -= accb[c];
                }
        }
    }
    if (blockIdx.x == 0) {  // this CTA owns the next diagonal block's rows: finish z_{k+1}
        ts_wait_inv();
        const int n1 = (int)min((int64_t)NB, N - k1);
        apply_inv_block<NC>(sinv, n1, B + k1 * ldb, ldb, M, false, sz, so);
    }
}

// k0 = first row of block k, n = its height. prologue: only x_last = inv(L)^T z_last.
template <int NC>
__global__ void __launch_bounds__(TS_THREADS) trsv_bwd_step_kernel(const double* __restrict__ L, int64_t ldl,
                                                                    const double* __restrict__ invL, double* B,
                                                                    int64_t ldb, int M, int64_t k0, int n, int prologue) {
    extern __shared__ __align__(16) double sinv[];
    __shared__ double sx[NB * NC];
    __shared__ double so[NB * NC];
    const int tid = threadIdx.x;
    if (prologue) {
        ts_prefetch_inv(sinv, invL + (k0 / NB) * (int64_t)NB * NB);
        ts_wait_inv();
        apply_inv_block<NC>(sinv, n, B + k0 * ldb, ldb, M, true, sx, so);
        return;
    }
    // the CTA whose columns include block k-1 (columns [k0 - NB, k0), inside one TS_THREADS-aligned chunk) also
    // finishes x_{k-1}
    const int64_t kprev = k0 - NB;
    const bool owner = (int64_t)blockIdx.x == kprev / TS_THREADS;
    if (owner) ts_prefetch_inv(sinv, invL + (kprev / NB) * (int64_t)NB * NB);
    for (int e = tid; e < n * NC; e += TS_THREADS) {
        const int r = e / NC, c = e - r * NC;
        sx[e] = (c < M) ? B[(k0 + r) * ldb + c] : 0.0;
    }
    __syncthreads();
    const int64_t j = (int64_t)blockIdx.x * TS_THREADS + tid;
    if (j < k0) {
        const double* lcol = L + k0 * ldl + j;
        double acc[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) acc[c] = 0.0;
#pragma unroll 16
        for (int r = 0; r < n; ++r) {
            const double v = lcol[(int64_t)r * ldl];
#pragma unroll
            for (int c = 0; c < NC; ++c) acc[c] = fma(v, sx[r * NC + c], acc[c]);
        }
#pragma unroll
        for (int c = 0; c < NC; ++c)
            if (c < M) B[j * ldb + c] -= acc[c];
    }
    if (owner) {
        ts_wait_inv();
        apply_inv_block<NC>(sinv, NB, B + kprev * ldb, ldb, M, true, sx, so);
    }
}

template <int NC>
static int trsv_fwd(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* B, int64_t ldb, int M) {
    GP_CUDA(c, cudaFuncSetAttribute(trsv_fwd_step_kernel<NC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TS_SMEM));
    trsv_fwd_step_kernel<NC><<<1, TS_THREADS, TS_SMEM, c->stream>>>(L, ldl, invL, N, B, ldb, M, -1);
    ++c->launches;
    for (int64_t k0 = 0; k0 + NB < N; k0 += NB) {
        const int64_t rows = N - k0 - NB;
        trsv_fwd_step_kernel<NC><<<(unsigned)((rows + TS_ROWS - 1) / TS_ROWS), TS_THREADS, TS_SMEM, c->stream>>>(
            L, ldl, invL, N, B, ldb, M, k0);
        ++c->launches;
    }
    GP_CUDA(c, cudaGetLastError());
    return 0;
}

template <int NC>
static int trsv_bwd(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* B, int64_t ldb, int M) {
    const int64_t nblk = (N + NB - 1) / NB;
    const int64_t klast = (nblk - 1) * NB;
    GP_CUDA(c, cudaFuncSetAttribute(trsv_bwd_step_kernel<NC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TS_SMEM));
    trsv_bwd_step_kernel<NC><<<1, TS_THREADS, TS_SMEM, c->stream>>>(L, ldl, invL, B, ldb, M, klast, (int)(N - klast), 1);
    ++c->launches;
    for (int64_t k0 = klast; k0 > 0; k0 -= NB) {
        const int n = (int)((N - k0 < NB) ? (N - k0) : NB);
        trsv_bwd_step_kernel<NC><<<(unsigned)((k0 + TS_THREADS - 1) / TS_THREADS), TS_THREADS, TS_SMEM, c->stream>>>(
            L, ldl, invL, B, ldb, M, k0, n, 0);
        ++c->launches;
    }
    GP_CUDA(c, cudaGetLastError());
    return 0;
}

// ---- chained TRSV: the whole sweep in ONE launch (M <= 2) -------------------------------------------------------------
// One CTA per 128-row block. A CTA takes its block index from an atomic ticket, so a block only ever waits for
// blocks whose CTAs started earlier (no deadlock whatever the residency), streams its off-diagonal blocks of L while
// the blocks before it are still being solved, and publishes x_j with a flag. The dependent chain per block is one
// flag poll, one 128-value load, one 128 x 128 block product and the pre-inverted diagonal block from shared memory:
// ~3 us instead of one ~25 us launch per block (the step kernels above). Deterministic: fixed summation order.
//   forward  x_j = inv(L_jj)   (b_j - sum_{k<j} L_jk x_k)     lane = 4 columns of L_jk, warp = 8 rows, shuffle fold
//   backward x_j = inv(L_jj)^T (z_j - sum_{k>j} L_kj^T x_k)   lane = 4 columns (outputs), warp = 8 rows, smem fold
constexpr int TC_THREADS = 512;
constexpr int TC_LD = NB + 1;

__device__ __forceinline__ void tc_wait(volatile int* flag) {
    if ((threadIdx.x & 31) == 0)
        while (*flag == 0) {}
    __syncwarp();
    __threadfence();
}

template <int NC, bool BWD>
__global__ void __launch_bounds__(TC_THREADS, 1)
trsv_chain_kernel(const double* __restrict__ L, int64_t ldl, const double* __restrict__ invL, int64_t N, double* B,
                  int64_t ldb, int nblk, int* ticket, volatile int* flags) {
    extern __shared__ double tcs[];
    double* S = tcs;                       // [NB][TC_LD]: S[i][t] multiplies y[i] into output t
    double* ys = S + NB * TC_LD;           // [NB][NC]
    double* red = ys + NB * NC;            // [16][NB][NC] (backward) / [4][NB][NC] (final product partials)
    __shared__ int s_ticket;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_ticket = atomicAdd(ticket, 1);
    __syncthreads();
    const int j = BWD ? (nblk - 1 - s_ticket) : s_ticket;
    const int64_t j0 = (int64_t)j * NB;
    const int n = (int)min((int64_t)NB, N - j0);
    {   // diagonal inverse into shared memory: forward needs inv[t][i] at S[i][t], backward inv[i][t]
        const double* inv = invL + (int64_t)j * NB * NB;
        for (int e = tid; e < NB * NB; e += TC_THREADS) {
            const int r = e / NB, cidx = e - r * NB;
            const double v = inv[e];
            if (BWD) S[r * TC_LD + cidx] = v;
            else S[cidx * TC_LD + r] = v;
        }
    }
    const int r8 = warp * 8, c4 = lane * 4;
    if (!BWD) {
        double acc[8][NC];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int c = 0; c < NC; ++c) acc[i][c] = 0.0;
        for (int k = 0; k < j; ++k) {
            double2 lo[8], hi[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                lo[i] = hi[i] = make_double2(0.0, 0.0);
                if (r8 + i < n) {
                    const double2* src = reinterpret_cast<const double2*>(L + (j0 + r8 + i) * ldl + (int64_t)k * NB + c4);
                    lo[i] = __ldg(src);
                    hi[i] = __ldg(src + 1);
                }
            }
            tc_wait(flags + k);
            double x[4][NC];
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int c = 0; c < NC; ++c) x[q][c] = __ldcg(B + ((int64_t)k * NB + c4 + q) * ldb + c);
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int c = 0; c < NC; ++c)
                    acc[i][c] = fma(hi[i].y, x[3][c], fma(hi[i].x, x[2][c], fma(lo[i].y, x[1][c], fma(lo[i].x, x[0][c], acc[i][c]))));
        }
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                double v = acc[i][c];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0) ys[(r8 + i) * NC + c] = v;
            }
        __syncthreads();
    } else {
        double acc[4][NC];
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int c = 0; c < NC; ++c) acc[q][c] = 0.0;
        for (int k = nblk - 1; k > j; --k) {
            const int64_t k0 = (int64_t)k * NB;
            double2 lo[8], hi[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                lo[i] = hi[i] = make_double2(0.0, 0.0);
                if (k0 + r8 + i < N) {
                    const double2* src = reinterpret_cast<const double2*>(L + (k0 + r8 + i) * ldl + j0 + c4);
                    lo[i] = __ldg(src);
                    hi[i] = __ldg(src + 1);
                }
            }
            tc_wait(flags + k);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (k0 + r8 + i < N) {
#pragma unroll
                    for (int c = 0; c < NC; ++c) {
                        const double xv = __ldcg(B + (k0 + r8 + i) * ldb + c);
                        acc[0][c] = fma(lo[i].x, xv, acc[0][c]);
                        acc[1][c] = fma(lo[i].y, xv, acc[1][c]);
                        acc[2][c] = fma(hi[i].x, xv, acc[2][c]);
                        acc[3][c] = fma(hi[i].y, xv, acc[3][c]);
                    }
                }
            }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int c = 0; c < NC; ++c) red[(warp * NB + c4 + q) * NC + c] = acc[q][c];
        __syncthreads();
        if (tid < NB) {
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                double v = 0.0;
#pragma unroll
                for (int w = 0; w < TC_THREADS / 32; ++w) v += red[(w * NB + tid) * NC + c];
                ys[tid * NC + c] = v;
            }
        }
        __syncthreads();
    }
    // y = b_j - sum (rows >= n of a ragged last block stay zero)
    if (tid < NB) {
#pragma unroll
        for (int c = 0; c < NC; ++c) ys[tid * NC + c] = (tid < n) ? B[(j0 + tid) * ldb + c] - ys[tid * NC + c] : 0.0;
    }
    __syncthreads();
    // x_j[t] = sum_i S[i][t] y[i]: four threads per output (32 terms each), folded through shared memory
    {
        const int t = tid & (NB - 1), part = tid >> 7;
        double v[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) v[c] = 0.0;
#pragma unroll 8
        for (int i = part * 32; i < part * 32 + 32; ++i) {
            const double sv = S[i * TC_LD + t];
#pragma unroll
            for (int c = 0; c < NC; ++c) v[c] = fma(sv, ys[i * NC + c], v[c]);
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) red[(part * NB + t) * NC + c] = v[c];
    }
    __syncthreads();
    if (tid < n) {
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const double v = (red[(0 * NB + tid) * NC + c] + red[(1 * NB + tid) * NC + c]) +
                             (red[(2 * NB + tid) * NC + c] + red[(3 * NB + tid) * NC + c]);
            B[(j0 + tid) * ldb + c] = v;
        }
        __threadfence();
    }
    __syncthreads();
    if (tid == 0) {
        __threadfence();  // cumulative: the CTA's stores of x_j (ordered by the barrier) become visible before the flag
        flags[j] = 1;
    }
}

template <int NC>
static int trsv_chain(Ctx* c, bool bwd, const double* L, int64_t ldl, const double* invL, int64_t N, double* B, int64_t ldb) {
    const int nblk = (int)((N + NB - 1) / NB);
    int* sync = nullptr;  // [0] ticket, [1 ..] flags
    GP_CUDA(c, cudaMallocAsync((void**)&sync, sizeof(int) * (nblk + 1), c->stream));
    GP_CUDA(c, cudaMemsetAsync(sync, 0, sizeof(int) * (nblk + 1), c->stream));
    const size_t smem = sizeof(double) * ((size_t)NB * TC_LD + (size_t)NB * NC + (size_t)(TC_THREADS / 32) * NB * NC);
    if (bwd) {
        auto kern = trsv_chain_kernel<NC, true>;
        GP_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<nblk, TC_THREADS, smem, c->stream>>>(L, ldl, invL, N, B, ldb, nblk, sync, sync + 1);
    } else {
        auto kern = trsv_chain_kernel<NC, false>;
        GP_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<nblk, TC_THREADS, smem, c->stream>>>(L, ldl, invL, N, B, ldb, nblk, sync, sync + 1);
    }
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    GP_CUDA(c, cudaFreeAsync(sync, c->stream));
    return 0;
}

static int trsv_dispatch(Ctx* c, bool bwd, const double* L, int64_t ldl, const double* invL, int64_t N, double* B,
                         int64_t ldb, int64_t M) {
#define GP_TRSV(NC) (bwd ? trsv_bwd<NC>(c, L, ldl, invL, N, B, ldb, (int)M) : trsv_fwd<NC>(c, L, ldl, invL, N, B, ldb, (int)M))
    if (c->opt_trsv_chain && (ldl & 3) == 0 && N > NB) {  // 32-byte row segments need ldl % 4 == 0
        if (M == 1) return trsv_chain<1>(c, bwd, L, ldl, invL, N, B, ldb);
        if (M == 2) return trsv_chain<2>(c, bwd, L, ldl, invL, N, B, ldb);
    }
    if (M <= 1) return GP_TRSV(1);
    if (M <= 2) return GP_TRSV(2);
    if (M <= 4) return GP_TRSV(4);
    return GP_TRSV(8);
#undef GP_TRSV
}

int trsm_lower_fwd(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* B, int64_t ldb,
                   int64_t M) {
    if (M <= 8 && c->opt_skinny && (ldl & 1) == 0) return trsv_dispatch(c, false, L, ldl, invL, N, B, ldb, M);
    return (M >= c->opt_solve_left_min_m) ? fwd_left(c, L, ldl, invL, N, B, ldb, M)
                                          : fwd_right(c, L, ldl, invL, N, B, ldb, M);
}

static int bwd_right(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* B, int64_t ldb,
                     int64_t M) {
    const int64_t nblk = (N + NB - 1) / NB;
    for (int64_t kb = nblk - 1; kb >= 0; --kb) {
        const int64_t k0 = kb * NB;
        const int n = (int)((N - k0 < NB) ? (N - k0) : NB);
        const double* invk = invL + kb * (int64_t)NB * NB;
        double* Bk = B + k0 * ldb;
        GemmArgs d{};  // X_k = inv(L_kk)^T * Z_k : A(m,kk) = invL[kk][m]
        d.A = invk; d.lda = NB; d.B = Bk; d.ldb = ldb; d.C = Bk; d.ldc = ldb;
        d.M = n; d.N = (int)M; d.K = n; d.alpha = 1.0; d.beta = 0.0;
        GP_TRY(gemm(c, GEMM_TN, d));
        if (k0 == 0) break;
        GemmArgs u{};  // Z_i -= L_ki^T X_k for i < k : A(m,kk) = L[k0 + kk][m]
        u.A = L + k0 * ldl; u.lda = ldl; u.B = Bk; u.ldb = ldb; u.C = B; u.ldc = ldb;
        u.M = (int)k0; u.N = (int)M; u.K = n; u.alpha = -1.0; u.beta = 1.0;
        GP_TRY(gemm(c, GEMM_TN, u));
    }
    return 0;
}

static int bwd_left(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* B, int64_t ldb,
                    int64_t M) {
    const int64_t nblk = (N + NB - 1) / NB;
    for (int64_t kb = nblk - 1; kb >= 0; --kb) {
        const int64_t k0 = kb * NB;
        const int n = (int)((N - k0 < NB) ? (N - k0) : NB);
        const double* invk = invL + kb * (int64_t)NB * NB;
        double* Bk = B + k0 * ldb;
        const int64_t below = N - k0 - n;
        if (below > 0) {
            GemmArgs u{};  // Z_k -= L[k0+n:, k-block]^T X[k0+n:] : A(m,kk) = L[k0 + n + kk][k0 + m]
            u.A = L + (k0 + n) * ldl + k0; u.lda = ldl; u.B = B + (k0 + n) * ldb; u.ldb = ldb; u.C = Bk; u.ldc = ldb;
            u.M = n; u.N = (int)M; u.K = (int)below; u.alpha = -1.0; u.beta = 1.0;
            GP_TRY(gemm(c, GEMM_TN, u));
        }
        GemmArgs d{};
        d.A = invk; d.lda = NB; d.B = Bk; d.ldb = ldb; d.C = Bk; d.ldc = ldb;
        d.M = n; d.N = (int)M; d.K = n; d.alpha = 1.0; d.beta = 0.0;
        GP_TRY(gemm(c, GEMM_TN, d));
    }
    return 0;
}

int trsm_lower_bwd(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* B, int64_t ldb,
                   int64_t M) {
    if (M <= 8 && c->opt_skinny && (ldl & 1) == 0) return trsv_dispatch(c, true, L, ldl, invL, N, B, ldb, M);
    return (M >= c->opt_solve_left_min_m) ? bwd_left(c, L, ldl, invL, N, B, ldb, M)
                                          : bwd_right(c, L, ldl, invL, N, B, ldb, M);
}

// K5, long-K form (default): C = W^T W with W = L^-1, both halves as GEMMs with a long contraction.
//   TRTRI, left-looking over super-rows S of 512 rows: W[S, :s0] <- L[S, :s0] W[:s0, :s0] is ONE product of depth
//          s0 whose k loop skips the zero part of the triangular W (GemmArgs::ktri = 1); inside the super-row
//          the four 128-row blocks are finished with short updates and the pre-inverted diagonal blocks.
//   LAUUM, one launch: C[m][n] = sum_{k >= max(m, n)} W[k][m] W[k][n], lower tiles only (ktri = 2).
// N^3/3 flop each, like the structured right-looking sweeps below, but those update all of C once per 128 columns
// (K = 128 per pass, ~N^3/48 bytes of read-modify-write traffic); here every output tile is written once.
static int inverse_long_k(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* C, int64_t ldc) {
    double* W = nullptr;
    GP_CUDA(c, cudaMallocAsync((void**)&W, sizeof(double) * N * ldc, c->stream));
    auto body = [&]() -> int {
        GP_TRY(fill2d(c, W, ldc, N, N, 0.0));
        constexpr int64_t WS = 512;
        for (int64_t s0 = 0; s0 < N; s0 += WS) {
            const int64_t ws = (N - s0 < WS) ? (N - s0) : WS;
            if (s0 > 0) {
                GemmArgs t{};
                t.A = L + s0 * ldl; t.lda = ldl; t.B = W; t.ldb = ldc; t.C = W + s0 * ldc; t.ldc = ldc;
                t.M = (int)ws; t.N = (int)s0; t.K = (int)s0; t.alpha = 1.0; t.beta = 0.0; t.ktri = 1;
                GP_TRY(gemm(c, GEMM_NN, t));
            }
            for (int64_t r0 = s0; r0 < s0 + ws; r0 += NB) {
                const int n = (int)((N - r0 < NB) ? (N - r0) : NB);
                const double* invk = invL + (r0 / NB) * (int64_t)NB * NB;
                double* Wr = W + r0 * ldc;
                if (r0 > s0) {  // X_r += L[r, s0:r0] W[s0:r0, :r0]
                    GemmArgs u{};
                    u.A = L + r0 * ldl + s0; u.lda = ldl; u.B = W + s0 * ldc; u.ldb = ldc; u.C = Wr; u.ldc = ldc;
                    u.M = n; u.N = (int)r0; u.K = (int)(r0 - s0); u.alpha = 1.0; u.beta = 1.0;
                    GP_TRY(gemm(c, GEMM_NN, u));
                }
                if (r0 > 0) {  // W[r, :r0] = -inv(L_rr) X_r, in place (one row tile: every CTA owns its columns)
                    GemmArgs d{};
                    d.A = invk; d.lda = NB; d.B = Wr; d.ldb = ldc; d.C = Wr; d.ldc = ldc;
                    d.M = n; d.N = (int)r0; d.K = n; d.alpha = -1.0; d.beta = 0.0;
                    GP_TRY(gemm(c, GEMM_NN, d));
                }
                GP_TRY(copy2d(c, Wr + r0, ldc, invk, NB, n, n));
            }
        }
        GemmArgs l{};
        l.A = W; l.lda = ldc; l.B = W; l.ldb = ldc; l.C = C; l.ldc = ldc;
        l.M = (int)N; l.N = (int)N; l.K = (int)N; l.alpha = 1.0; l.beta = 0.0; l.lower = 1; l.ktri = 2;
        GP_TRY(gemm(c, GEMM_TN, l));
        return symmetrize_from_lower(c, C, ldc, N);
    };
    const int rc = body();
    cudaFreeAsync(W, c->stream);
    return rc;
}

// W = inv(L) restricted to the cyclic column blocks j = first, first + stride, ... (width cb, a multiple of NB) and
// stored COMPACTLY: local column q * cb + t holds global column (first + q * stride) * cb + t. The columns of inv(L)
// are independent, so g ranks that take first = rank, stride = g share the N^3 / 3 flop of the TRTRI with no exchange
// (multi-GPU likelihood gradient, gpr_b200/dist.py: ShardedLikelihood). Same left-looking long-K sweep as the TRTRI half
// of inverse_long_k; the local columns with a global index below a row r form a prefix of the compact matrix, so every
// product is one GEMM on a contiguous column range; the k loop of a tile skips the zero part above its column block
// (GemmArgs::ktri = 3 maps the compact column of a tile to its global column).
int trtri_cols(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, int64_t cb, int64_t first,
               int64_t stride, double* W, int64_t ldw) {
    const int64_t nblk = (N + cb - 1) / cb;
    auto nloc = [&](int64_t x) {   // local columns whose global index is < x
        int64_t cnt = 0;
        for (int64_t j = first; j < nblk; j += stride) {
            const int64_t o = j * cb, w = (N - o < cb) ? (N - o) : cb;
            if (x <= o) break;
            cnt += (x - o < w) ? (x - o) : w;
        }
        return cnt;
    };
    const int64_t ncl = nloc(N);
    if (ncl <= 0) return 0;
    GP_TRY(fill2d(c, W, ldw, N, ncl, 0.0));
    constexpr int64_t WS = 512;
    for (int64_t s0 = 0; s0 < N; s0 += WS) {
        const int64_t ws = (N - s0 < WS) ? (N - s0) : WS;
        const int64_t n0 = nloc(s0);
        if (s0 > 0 && n0 > 0) {
            GemmArgs t{};
            t.A = L + s0 * ldl; t.lda = ldl; t.B = W; t.ldb = ldw; t.C = W + s0 * ldw; t.ldc = ldw;
            t.M = (int)ws; t.N = (int)n0; t.K = (int)s0; t.alpha = 1.0; t.beta = 0.0;
            t.ktri = 3; t.tri_cb = (int)cb; t.tri_first = (int)first; t.tri_stride = (int)stride;   // skip k < global column
            GP_TRY(gemm(c, GEMM_NN, t));
        }
        for (int64_t r0 = s0; r0 < s0 + ws; r0 += NB) {
            const int n = (int)((N - r0 < NB) ? (N - r0) : NB);
            const double* invk = invL + (r0 / NB) * (int64_t)NB * NB;
            double* Wr = W + r0 * ldw;
            const int64_t nr = nloc(r0);
            if (r0 > s0 && nr > 0) {  // X_r += L[r, s0:r0] W[s0:r0, :nr]
                GemmArgs u{};
                u.A = L + r0 * ldl + s0; u.lda = ldl; u.B = W + s0 * ldw; u.ldb = ldw; u.C = Wr; u.ldc = ldw;
                u.M = n; u.N = (int)nr; u.K = (int)(r0 - s0); u.alpha = 1.0; u.beta = 1.0;   // columns new in this super-row hold zeros
                GP_TRY(gemm(c, GEMM_NN, u));
            }
            if (r0 > 0 && nr > 0) {  // W[r, :nr] = -inv(L_rr) X_r, in place (one row tile: every CTA owns its columns)
                GemmArgs d{};
                d.A = invk; d.lda = NB; d.B = Wr; d.ldb = ldw; d.C = Wr; d.ldc = ldw;
                d.M = n; d.N = (int)nr; d.K = n; d.alpha = -1.0; d.beta = 0.0;
                GP_TRY(gemm(c, GEMM_NN, d));
            }
            const int64_t j = r0 / cb;   // the diagonal block inv(L_rr) itself, if this column block is a local one
            if (j >= first && (j - first) % stride == 0)
                GP_TRY(copy2d(c, Wr + nr, ldw, invk, NB, n, n));
        }
    }
    return 0;
}

// K5: C = (L L^T)^-1 from the factor, 2 N^3 / 3 flop (LAPACK dpotri's count; the reference calls dgetri / dpotri,
// include/LAPACKUtils.h:50, :65). Two structured sweeps over C = I:
//   forward  W = L^-1 is lower triangular: at block row k only columns [0, k0 + n) can be non-zero, so both GEMMs
//            of the step are restricted to them (N^3 / 3 flop instead of N^3);
//   backward C = L^-T W is symmetric: only its lower triangle is formed (row block i needs columns <= i), the
//            update of the rows above block k is a lower-tiles-only GEMM (N^3 / 3 flop), then mirrored.
int inverse_from_factor(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* C, int64_t ldc) {
    if (c->opt_inverse_long_k && N > 2 * NB) return inverse_long_k(c, L, ldl, invL, N, C, ldc);
    GP_TRY(set_identity(c, C, ldc, N));
    for (int64_t k0 = 0; k0 < N; k0 += NB) {
        const int n = (int)((N - k0 < NB) ? (N - k0) : NB);
        const int64_t cols = k0 + n;
        const double* invk = invL + (k0 / NB) * (int64_t)NB * NB;
        double* Bk = C + k0 * ldc;
        GemmArgs d{};
        d.A = invk; d.lda = NB; d.B = Bk; d.ldb = ldc; d.C = Bk; d.ldc = ldc;
        d.M = n; d.N = (int)cols; d.K = n; d.alpha = 1.0; d.beta = 0.0;
        GP_TRY(gemm(c, GEMM_NN, d));
        const int64_t m = N - k0 - n;
        if (m <= 0) break;
        GemmArgs u{};
        u.A = L + (k0 + n) * ldl + k0; u.lda = ldl; u.B = Bk; u.ldb = ldc; u.C = C + (k0 + n) * ldc; u.ldc = ldc;
        u.M = (int)m; u.N = (int)cols; u.K = n; u.alpha = -1.0; u.beta = 1.0;
        GP_TRY(gemm(c, GEMM_NN, u));
    }
    const int64_t nblk = (N + NB - 1) / NB;
    for (int64_t kb = nblk - 1; kb >= 0; --kb) {
        const int64_t k0 = kb * NB;
        const int n = (int)((N - k0 < NB) ? (N - k0) : NB);
        const int64_t cols = k0 + n;
        const double* invk = invL + kb * (int64_t)NB * NB;
        double* Bk = C + k0 * ldc;
        GemmArgs d{};
        d.A = invk; d.lda = NB; d.B = Bk; d.ldb = ldc; d.C = Bk; d.ldc = ldc;
        d.M = n; d.N = (int)cols; d.K = n; d.alpha = 1.0; d.beta = 0.0;
        GP_TRY(gemm(c, GEMM_TN, d));
        if (k0 == 0) break;
        GemmArgs u{};  // rows [0, k0), columns [0, k0), lower tiles only
        u.A = L + k0 * ldl; u.lda = ldl; u.B = Bk; u.ldb = ldc; u.C = C; u.ldc = ldc;
        u.M = (int)k0; u.N = (int)k0; u.K = n; u.alpha = -1.0; u.beta = 1.0; u.lower = 1;
        GP_TRY(gemm(c, GEMM_TN, u));
    }
    // the lower triangle is the result; mirror it so C == C^T holds exactly
    GP_TRY(symmetrize_from_lower(c, C, ldc, N));
    return 0;
}

}  // namespace gprb200
