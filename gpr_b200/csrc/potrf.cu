// K3: right-looking blocked fp64 Cholesky, A = L L^T, lower, row-major, in place.
// Replaces the reference's explicit LU inverse (include/LAPACKUtils.h:38-56, dgetrf+dgetri, 2 N^3
// flop) and its Cholesky variant (:59-73, dpotrf+dpotri) as the core of InvertKernelMatrix
// (lib/GaussianProcess.cpp:531-618): N^3/3 flop, and the factor is kept instead of the inverse.
//
// Per block step k (block size NB = 128):
//   1. potrf_tile_kernel : one CTA factors the NB x NB diagonal block in shared memory and also
//      forms inv(L_kk) (TRTRI) so that every triangular solve below becomes a GEMM;
//   2. panel solve       : A[i,k] <- A[i,k] inv(L_kk)^T          (GEMM NT, in place, 1 % of flops)
//   3. trailing update   : A[i,j] -= A[i,k] A[j,k]^T, i >= j > k (GEMM NT on lower tiles, 97 % of flops)
// INFO follows LAPACK dpotrf: k > 0 = leading minor of order k is not positive definite.
#include "common.cuh"
#include <math.h>
#include <vector>
#include <utility>

namespace gprb200 {

constexpr int PT_THREADS = 512;
constexpr int PT_LD = NB + 1;  // padded: column walks are conflict-free

// A: n x n diagonal block (lda), n <= NB. invL: NB x NB row-major (ld NB) receives inv(L) (lower,
// zero above the diagonal, zero-padded beyond n). info: set to col0 + j + 1 on the first failure.
__global__ void __launch_bounds__(PT_THREADS, 1)
potrf_tile_kernel(double* __restrict__ A, int64_t lda, int n, double* __restrict__ invL, int* info, int col0) {
    extern __shared__ double sm[];
    double* S = sm;                    // [NB][PT_LD]: strictly lower = L; upper incl. diagonal = inv(L)^T
    double* diag = sm + NB * PT_LD;    // L_jj
    double* colj = diag + NB;          // scaled column j
    __shared__ int bad;
    const int tid = threadIdx.x;
    if (tid == 0) bad = 0;

    for (int e = tid; e < n * n; e += PT_THREADS) {
        const int r = e / n, c = e - r * n;
        if (c <= r) S[r * PT_LD + c] = A[(int64_t)r * lda + c];
    }
    __syncthreads();

    // ---- unblocked right-looking Cholesky on the lower triangle
    const int tx = tid & 31, ty = tid >> 5;  // 32 x 16
    for (int j = 0; j < n; ++j) {
        const double ajj = S[j * PT_LD + j];
        if (!(ajj > 0.0) || !isfinite(ajj)) {  // uniform across the CTA: everybody reads the same value
            if (tid == 0) {
                bad = 1;
                atomicCAS(info, 0, col0 + j + 1);
            }
            break;
        }
        const double ljj = sqrt(ajj);
        for (int i = j + 1 + tid; i < n; i += PT_THREADS) {
            const double v = S[i * PT_LD + j] / ljj;
            colj[i] = v;
            S[i * PT_LD + j] = v;
        }
        if (tid == 0) diag[j] = ljj;
        __syncthreads();
        // rank-1 update of the trailing lower triangle
        for (int i = j + 1 + ty; i < n; i += PT_THREADS / 32) {
            const double li = colj[i];
            for (int k = j + 1 + tx; k <= i; k += 32) S[i * PT_LD + k] = fma(-li, colj[k], S[i * PT_LD + k]);
        }
        __syncthreads();
    }
    __syncthreads();
    if (bad) return;  // leave A untouched beyond this point; INFO tells the caller

    // write L back (diagonal from diag[])
    for (int e = tid; e < n * n; e += PT_THREADS) {
        const int r = e / n, c = e - r * n;
        if (c < r) A[(int64_t)r * lda + c] = S[r * PT_LD + c];
        else if (c == r) A[(int64_t)r * lda + c] = diag[r];
    }

    // ---- X = inv(L), stored transposed in the upper triangle: X(i,c) lives at S[c][i], c <= i.
    // Running right-hand side T starts as I; step j finalises row j of X and eliminates column j.
    for (int e = tid; e < n * n; e += PT_THREADS) {
        const int r = e / n, c = e - r * n;
        if (c >= r) S[r * PT_LD + c] = (c == r) ? 1.0 : 0.0;
    }
    __syncthreads();
    for (int j = 0; j < n; ++j) {
        const double ljj = diag[j];
        for (int cidx = tid; cidx <= j; cidx += PT_THREADS) S[cidx * PT_LD + j] /= ljj;  // X(j, c)
        __syncthreads();
        // T(i, c) -= L(i, j) X(j, c) for i > j, c <= j
        for (int i = j + 1 + ty; i < n; i += PT_THREADS / 32) {
            const double lij = S[i * PT_LD + j];
            for (int cidx = tx; cidx <= j; cidx += 32)
                S[cidx * PT_LD + i] = fma(-lij, S[cidx * PT_LD + j], S[cidx * PT_LD + i]);
        }
        __syncthreads();
    }
    for (int e = tid; e < NB * NB; e += PT_THREADS) {
        const int r = e / NB, c = e - r * NB;
        invL[e] = (r < n && c <= r) ? S[c * PT_LD + r] : 0.0;
    }
}

// ---- potrf_tile2_kernel: 128 x 128 diagonal block, blocked (32-wide) inside one CTA ---------------------------------
// Round-1 measurement: the unblocked kernel above needs ~512 CTA-wide barriers (~150 us per tile, 256 tiles per
// N = 32768 factorisation = the whole serial panel time). Here each 32-column step is
//   (1) diagonal 32 x 32 block factored AND inverted by one warp entirely in registers (shuffles, no barriers),
//   (2) rows below: X = A inv(L_dd)^T as a small product over all warps,
//   (3) trailing lower triangle: A -= X X^T,
// three barriers per step; inv(L) of the whole tile is then assembled block-wise from the inverted diagonal blocks
// (W_ij = -W_ii sum_k L_ik W_kj), two barriers per block distance.
constexpr int P2_THREADS = 256;
constexpr int P2_IB = 32;
constexpr int P2_LD = NB + 1;
constexpr size_t P2_SMEM = sizeof(double) * ((size_t)NB * P2_LD + 4 * P2_IB * (P2_IB + 1) + 3 * P2_IB * (P2_IB + 1) + NB);

__global__ void __launch_bounds__(P2_THREADS, 1)
potrf_tile2_kernel(double* __restrict__ A, int64_t lda, int n, double* __restrict__ invL, int* info, int col0) {
    extern __shared__ double sm2[];
    double* S = sm2;                                   // [NB][P2_LD]: lower = A then L; strict upper = inv(L)^T blocks
    double* Wd = S + NB * P2_LD;                       // [4][32][33]: inverses of the 32 x 32 diagonal blocks
    double* Tm = Wd + 4 * P2_IB * (P2_IB + 1);         // [3][32][33]: products in flight
    double* dW = Tm + 3 * P2_IB * (P2_IB + 1);         // [NB]: diagonal of inv(L)
    __shared__ int bad;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int WARPS = P2_THREADS / 32;
    constexpr unsigned FULL = 0xffffffffu;
    if (tid == 0) bad = 0;
    // rows / columns >= n are padded with the identity so that every block is a full 32 x 32 problem
    for (int e = tid; e < NB * NB; e += P2_THREADS) {
        const int r = e / NB, c = e - r * NB;
        double v = 0.0;
        if (r < n && c <= r) v = A[(int64_t)r * lda + c];
        else if (r >= n && c == r) v = 1.0;
        S[r * P2_LD + c] = v;
    }
    __syncthreads();

    for (int b = 0; b < NB / P2_IB; ++b) {
        const int c0 = b * P2_IB;
        if (c0 >= n) break;
        // ---- (1) warp 0: Cholesky of the diagonal block, lane i owns row i; then its inverse, lane j owns column j
        if (warp == 0) {
            // Shared-memory broadcast instead of one shuffle per (column, row) pair: the ncu source view of the first
            // version showed the other seven warps waiting at the barrier below for ~70 % of the kernel while this
            // warp ran SHFL -> DFMA pairs back to back with the full shuffle latency exposed (~60 cycles per pair).
            // Here a column is published once, read back with independent broadcast loads, then the FMAs run.
            double* colb = Tm;                 // [2][32] column broadcast buffers (Tm is free during this phase)
            double* rdb = Tm + 2 * P2_IB;      // [32] reciprocal diagonal
            double a[P2_IB];
#pragma unroll
            for (int k = 0; k < P2_IB; ++k) a[k] = (k <= lane) ? S[(c0 + lane) * P2_LD + c0 + k] : 0.0;
            int fail_at = -1;
            double rdiag = 1.0;  // lane j keeps 1 / l_jj: the dependent chain has one rsqrt and no division per column
#pragma unroll
            for (int j = 0; j < P2_IB; ++j) {
                const double djj = __shfl_sync(FULL, a[j], j);
                if (fail_at < 0 && (!(djj > 0.0) || !isfinite(djj))) fail_at = j;  // uniform across the warp
                const double rj = rsqrt(djj);
                if (lane == j) {
                    a[j] = djj * rj;  // l_jj = sqrt(d) = d * rsqrt(d)
                    rdiag = rj;
                } else if (lane > j) {
                    a[j] = a[j] * rj;
                }
                double* cb = colb + (j & 1) * P2_IB;
                cb[lane] = a[j];
                __syncwarp();
#pragma unroll
                for (int k = j + 1; k < P2_IB; ++k)
                    if (lane >= k) a[k] = fma(-a[j], cb[k], a[k]);  // broadcast loads, independent of the FMAs
            }
            if (fail_at >= 0) {
                if (lane == 0) {
                    bad = 1;
                    atomicCAS(info, 0, col0 + c0 + fail_at + 1);
                }
            } else {
                // L_dd back to shared memory first: the inverse reads row i of L as broadcast loads
#pragma unroll
                for (int k = 0; k < P2_IB; ++k)
                    if (k <= lane) S[(c0 + lane) * P2_LD + c0 + k] = a[k];
                rdb[lane] = rdiag;
                __syncwarp();
                // inverse: column j of W = inv(L_dd): w[j] = 1 / l_jj, w[i] = -(sum_{k=j}^{i-1} l_ik w[k]) / l_ii
                double w[P2_IB];
#pragma unroll
                for (int i = 0; i < P2_IB; ++i) {
                    const double* Li = S + (c0 + i) * P2_LD + c0;
                    const double ri = rdb[i];
                    // four chains: a dependent DFMA costs ~40 cycles on sm_100 (tools/fp64_probe.cu)
                    double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
#pragma unroll
                    for (int k = 0; k < i; ++k) {
                        if (k >= lane) {
                            if ((k & 3) == 0) acc0 = fma(Li[k], w[k], acc0);
                            else if ((k & 3) == 1) acc1 = fma(Li[k], w[k], acc1);
                            else if ((k & 3) == 2) acc2 = fma(Li[k], w[k], acc2);
                            else acc3 = fma(Li[k], w[k], acc3);
                        }
                    }
                    w[i] = (i == lane) ? ri : ((i > lane) ? -((acc0 + acc1) + (acc2 + acc3)) * ri : 0.0);
                }
#pragma unroll
                for (int i = 0; i < P2_IB; ++i) Wd[(b * P2_IB + i) * (P2_IB + 1) + lane] = w[i];
            }
        }
        __syncthreads();
        if (bad) return;  // A is left untouched; INFO tells the caller
        const int r0 = c0 + P2_IB;  // first row / column of the trailing part
        if (r0 >= NB) break;
        // ---- (2) rows below: X[r][j] = sum_{k <= j} A[r][c0 + k] W[j][k]; a warp owns a row, lane = j
        {
            const double* Wb = Wd + b * P2_IB * (P2_IB + 1);
            double wrow[P2_IB];
#pragma unroll
            for (int k = 0; k < P2_IB; ++k) wrow[k] = Wb[lane * (P2_IB + 1) + k];  // W[j = lane][k], zero for k > j
            for (int r = r0 + warp; r < NB; r += WARPS) {
                const double v = S[r * P2_LD + c0 + lane];
                double x0 = 0.0, x1 = 0.0, x2 = 0.0, x3 = 0.0;   // four independent chains (DFMA latency)
#pragma unroll
                for (int k = 0; k < P2_IB; k += 4) {
                    x0 = fma(__shfl_sync(FULL, v, k), wrow[k], x0);
                    x1 = fma(__shfl_sync(FULL, v, k + 1), wrow[k + 1], x1);
                    x2 = fma(__shfl_sync(FULL, v, k + 2), wrow[k + 2], x2);
                    x3 = fma(__shfl_sync(FULL, v, k + 3), wrow[k + 3], x3);
                }
                S[r * P2_LD + c0 + lane] = (x0 + x1) + (x2 + x3);
            }
        }
        __syncthreads();
        // ---- (3) trailing lower triangle: A[r][c] -= sum_k X[r][k] X[c][k]; lane owns a column (kept in registers)
        for (int cc = r0; cc < NB; cc += 32) {
            const int c = cc + lane;
            double colv[P2_IB];
#pragma unroll
            for (int k = 0; k < P2_IB; ++k) colv[k] = S[c * P2_LD + c0 + k];
            for (int r = cc + warp; r < NB; r += WARPS) {  // rows >= cc; entries with c > r are skipped
                const double* xr = S + r * P2_LD + c0;
                double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
                for (int k = 0; k < P2_IB; k += 4) {
                    a0 = fma(xr[k], colv[k], a0);
                    a1 = fma(xr[k + 1], colv[k + 1], a1);
                    a2 = fma(xr[k + 2], colv[k + 2], a2);
                    a3 = fma(xr[k + 3], colv[k + 3], a3);
                }
                if (c <= r) S[r * P2_LD + c] -= (a0 + a1) + (a2 + a3);
            }
        }
        __syncthreads();
    }

    // ---- L back to global (lower triangle incl. diagonal)
    for (int e = tid; e < n * n; e += P2_THREADS) {
        const int r = e / n, c = e - r * n;
        if (c <= r) A[(int64_t)r * lda + c] = S[r * P2_LD + c];
    }
    // ---- inv(L): diagonal blocks are Wd; block (i, j), i > j: W_ij = -W_ii (sum_{k=j}^{i-1} L_ik W_kj),
    // stored TRANSPOSED in the strict upper part of S: W_ij[t][c] lives at S[32 j + c][32 i + t]
    constexpr int NBLK = NB / P2_IB;
    for (int dist = 1; dist < NBLK; ++dist) {
        const int nprod = NBLK - dist;  // blocks (i, i - dist), i = dist .. NBLK-1
        // T = sum_k L_ik W_kj
        for (int o = tid; o < nprod * P2_IB * P2_IB; o += P2_THREADS) {
            const int p = o / (P2_IB * P2_IB), rc = o - p * P2_IB * P2_IB;
            const int r = rc / P2_IB, c = rc - r * P2_IB;
            const int i = dist + p, j = p;
            double b0 = 0.0, b1 = 0.0, b2 = 0.0, b3 = 0.0;   // four chains
            for (int k = j; k < i; ++k) {
                const double* Lik = S + (P2_IB * i + r) * P2_LD + P2_IB * k;
                if (k == j) {
                    const double* Wjj = Wd + j * P2_IB * (P2_IB + 1);
#pragma unroll
                    for (int t = 0; t < P2_IB; t += 4) {
                        b0 = fma(Lik[t], Wjj[t * (P2_IB + 1) + c], b0);
                        b1 = fma(Lik[t + 1], Wjj[(t + 1) * (P2_IB + 1) + c], b1);
                        b2 = fma(Lik[t + 2], Wjj[(t + 2) * (P2_IB + 1) + c], b2);
                        b3 = fma(Lik[t + 3], Wjj[(t + 3) * (P2_IB + 1) + c], b3);
                    }
                } else {
                    const double* Wkj = S + (P2_IB * j + c) * P2_LD + P2_IB * k;  // transposed storage
#pragma unroll
                    for (int t = 0; t < P2_IB; t += 4) {
                        b0 = fma(Lik[t], Wkj[t], b0);
                        b1 = fma(Lik[t + 1], Wkj[t + 1], b1);
                        b2 = fma(Lik[t + 2], Wkj[t + 2], b2);
                        b3 = fma(Lik[t + 3], Wkj[t + 3], b3);
                    }
                }
            }
            Tm[(p * P2_IB + r) * (P2_IB + 1) + c] = (b0 + b1) + (b2 + b3);
        }
        __syncthreads();
        // W_ij = -W_ii T
        for (int o = tid; o < nprod * P2_IB * P2_IB; o += P2_THREADS) {
            const int p = o / (P2_IB * P2_IB), rc = o - p * P2_IB * P2_IB;
            const int r = rc / P2_IB, c = rc - r * P2_IB;
            const int i = dist + p, j = p;
            const double* Wii = Wd + i * P2_IB * (P2_IB + 1);
            const double* T = Tm + p * P2_IB * (P2_IB + 1);
            double d0 = 0.0, d1 = 0.0, d2 = 0.0, d3 = 0.0;
#pragma unroll
            for (int t = 0; t < P2_IB; t += 4) {   // W_ii[r][t] = 0 for t > r
                d0 = fma(Wii[r * (P2_IB + 1) + t], T[t * (P2_IB + 1) + c], d0);
                d1 = fma(Wii[r * (P2_IB + 1) + t + 1], T[(t + 1) * (P2_IB + 1) + c], d1);
                d2 = fma(Wii[r * (P2_IB + 1) + t + 2], T[(t + 2) * (P2_IB + 1) + c], d2);
                d3 = fma(Wii[r * (P2_IB + 1) + t + 3], T[(t + 3) * (P2_IB + 1) + c], d3);
            }
            S[(P2_IB * j + c) * P2_LD + P2_IB * i + r] = -((d0 + d1) + (d2 + d3));
        }
        __syncthreads();
    }
    for (int e = tid; e < NB * NB; e += P2_THREADS) {
        const int r = e / NB, c = e - r * NB;
        double v = 0.0;
        if (r < n && c <= r) {
            const int bi = r / P2_IB, bj = c / P2_IB;
            v = (bi == bj) ? Wd[(bi * P2_IB + (r - bi * P2_IB)) * (P2_IB + 1) + (c - bj * P2_IB)] : S[c * P2_LD + r];
        }
        invL[e] = v;
    }
}

// ---- potrf_tile3_kernel: the diagonal tile with every block product on the FP64 tensor pipe -----------------------
// Round-1 launch lists: potrf_tile2_kernel costs 0.139 ms per 128 columns, N / 128 of them in a row on ONE SM, and is
// the exposed tail of every factorisation below N ~ 16384 (profiles/r01b_ncu_summary.md). Its time is dependent-DFMA
// latency (~40 cycles on sm_100): scalar FMA chains in the sub-panel, trailing and inv(L) products, and a 32-column
// serial diagonal phase with CUDA's generic rsqrt and a shared-memory round trip on the critical path. Here:
//   (1) warp 0 factors the 32 x 32 diagonal block in registers (lane = row) AND forms its inverse in the same loop
//       (lane = column of inv(L_dd), right-looking substitution: one FMA per remaining row, all independent), so the
//       inverse's FMAs fill the bubbles of the pivot chain. The chain per column is shuffle -> rsqrt (MUFU seed + one
//       cubic step, 4 dependent DFMA) -> scale -> next pivot (computed locally by the lane that owns it);
//   (2) X = A inv(L_dd)^T, (3) trailing A -= X X^T and the block-wise assembly of inv(L) (T = sum L_ik W_kj,
//       W_ij = -W_ii T) are DMMA.8x8x4 products out of shared memory. Leading dimensions are = 4 (mod 16) doubles, so
//       both fragment patterns (row * LD + k and k * LD + row over 4 x 4 lanes of a half-warp) hit 16 distinct 8-byte
//       bank slots.
// Measured (B200, tools/chol_r02.py): 0.139 -> 0.064 ms per tile; Cholesky N = 4096 5.46 -> 3.67 ms, 8192 13.4 -> 10.6,
// 16384 56.4 -> 52.9, 32768 373.4 -> 369.3.
constexpr int P3_THREADS = 256;
constexpr int P3_WARPS = P3_THREADS / 32;
constexpr int P3_IB = 32;
constexpr int P3_LD = NB + 4;      // 132
constexpr int P3_WLD = P3_IB + 4;  // 36
constexpr size_t P3_SMEM =
    sizeof(double) * ((size_t)NB * P3_LD + 4 * P3_IB * P3_WLD + 3 * P3_IB * P3_WLD + 2 * P3_IB);

// 1 / sqrt(d) for a normal positive d: MUFU.RSQ64H seed (relative error ~2^-22) and one third-order step
// y <- y + y e (1/2 + 3/8 e), e = 1 - d y^2 (error ~ e^3): four dependent DFMA instead of the generic routine.
__device__ __forceinline__ double rsqrt_pos(double d) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
    const double t = d * y;
    const double e = fma(-t, y, 1.0);
    const double ye = y * e;
    const double p = fma(0.375, e, 0.5);
    return fma(ye, p, y);
}

// One column of the fused diagonal phase. J is a template parameter so that every index into a[] / s[] is a
// compile-time constant: with a plain doubly unrolled loop nvcc left the long inner loops rolled and put both arrays in
// local memory (708 LDL / 985 STL in the first build).
template <int J>
__device__ __forceinline__ void diag_step(double (&a)[P3_IB], double (&s)[P3_IB], double& dnext, int& fail_at, double* cb,
                                          int lane) {
    constexpr unsigned FULL = 0xffffffffu;
    const double djj = dnext;
    // not positive (or not a normal number): leading minor c0 + J + 1 is not positive definite
    if (fail_at < 0 && !(djj >= 2.2250738585072014e-308 && djj <= 1.7976931348623157e308)) fail_at = J;
    const double y = rsqrt_pos(djj);
    const double aj = (lane == J) ? djj * y : a[J] * y;   // zero above the diagonal stays zero
    a[J] = aj;
    if (J + 1 < P3_IB) {
        // the next pivot needs only lane J+1's own entries: no shared-memory round trip on the chain
        const double pv = fma(-aj, aj, a[(J + 1) % P3_IB]);
        dnext = __shfl_sync(FULL, pv, (J + 1) % P3_IB);
    }
    double* cj = cb + (J & 1) * P3_IB;
    cj[lane] = aj;
    __syncwarp();
    const double wj = s[J] * y;   // W(J, lane); zero for lane > J
    s[J] = wj;
#pragma unroll
    for (int k = J + 1; k < P3_IB; ++k) {
        const double ck = cj[k];                       // L(k, J), broadcast load
        if (lane >= k) a[k] = fma(-aj, ck, a[k]);
        s[k] = fma(-ck, wj, s[k]);
    }
}
template <int... Js>
__device__ __forceinline__ void diag_all(std::integer_sequence<int, Js...>, double (&a)[P3_IB], double (&s)[P3_IB],
                                         double& dnext, int& fail_at, double* cb, int lane) {
    (diag_step<Js>(a, s, dnext, fail_at, cb, lane), ...);
}

__global__ void __launch_bounds__(P3_THREADS, 1)
potrf_tile3_kernel(double* __restrict__ A, int64_t lda, int n, double* __restrict__ invL, int* info, int col0) {
    extern __shared__ __align__(16) double sm3[];
    double* S = sm3;                                  // [NB][P3_LD]: lower = A then L; strict upper = inv(L)^T blocks
    double* Wd = S + NB * P3_LD;                      // [4][32][P3_WLD]: inverses of the diagonal blocks, row-major
    double* Tm = Wd + 4 * P3_IB * P3_WLD;             // [3][32][P3_WLD]: products in flight
    double* cb = Tm + 3 * P3_IB * P3_WLD;             // [2][32]: column broadcast of the diagonal phase
    __shared__ int bad;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int grp = lane >> 2, t4 = lane & 3;
    constexpr unsigned FULL = 0xffffffffu;
    if (tid == 0) bad = 0;
    // rows / columns >= n are padded with the identity so that every block is a full 32 x 32 problem
    if (n == NB && (lda & 1) == 0 && (reinterpret_cast<uintptr_t>(A) & 15) == 0) {
        // full tile: 32 independent 16-byte loads per thread, all issued before the first use (the scalar loop below
        // cost ~12 us of exposed global latency per tile in the ncu source view, 64 dependent round trips per thread)
        double2 v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            const int e = (i * P3_THREADS + tid) * 2;
            const int r = e >> 7, c = e & (NB - 1);
            v[i] = (c <= r) ? *reinterpret_cast<const double2*>(A + (int64_t)r * lda + c) : make_double2(0.0, 0.0);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            const int e = (i * P3_THREADS + tid) * 2;
            const int r = e >> 7, c = e & (NB - 1);
            if (c == r) v[i].y = 0.0;   // (r, r + 1) lies above the diagonal
            *reinterpret_cast<double2*>(S + r * P3_LD + c) = v[i];
        }
    } else {
        for (int e = tid; e < NB * NB; e += P3_THREADS) {
            const int r = e >> 7, c = e & (NB - 1);
            double v = 0.0;
            if (r < n && c <= r) v = A[(int64_t)r * lda + c];
            else if (r >= n && c == r) v = 1.0;
            S[r * P3_LD + c] = v;
        }
    }
    __syncthreads();

    for (int b = 0; b < NB / P3_IB; ++b) {
        const int c0 = b * P3_IB;
        if (c0 >= n) break;
        // ---- (1) warp 0: L_dd and inv(L_dd) in one pass
        if (warp == 0) {
            double a[P3_IB];   // row `lane` of the block: a[k] = A(lane, k), zero for k > lane
            double s[P3_IB];   // column `lane` of the running right-hand side of L W = I; s[i] becomes W(i, lane)
#pragma unroll
            for (int k = 0; k < P3_IB; ++k) {
                a[k] = (k <= lane) ? S[(c0 + lane) * P3_LD + c0 + k] : 0.0;
                s[k] = (k == lane) ? 1.0 : 0.0;
            }
            int fail_at = -1;
            double dnext = __shfl_sync(FULL, a[0], 0);
            diag_all(std::make_integer_sequence<int, P3_IB>{}, a, s, dnext, fail_at, cb, lane);
            if (fail_at >= 0) {
                if (lane == 0) {
                    bad = 1;
                    atomicCAS(info, 0, col0 + c0 + fail_at + 1);
                }
            } else {
#pragma unroll
                for (int k = 0; k < P3_IB; ++k)
                    if (k <= lane) S[(c0 + lane) * P3_LD + c0 + k] = a[k];
#pragma unroll
                for (int i = 0; i < P3_IB; ++i) Wd[(b * P3_IB + i) * P3_WLD + lane] = s[i];
            }
        }
        __syncthreads();
        if (bad) return;  // A is left untouched; INFO tells the caller
        const int r0 = c0 + P3_IB;  // first row / column of the trailing part
        if (r0 >= NB) break;
        // ---- (2) rows below: X = A[r0.., c0..c0+32) W_dd^T. A warp owns strips of 8 rows, so the in-place store is safe.
        {
            const double* Wb = Wd + b * P3_IB * P3_WLD;
            const int nstrips = (NB - r0) / 8;
            for (int st = warp; st < nstrips; st += P3_WARPS) {
                double* row = S + (r0 + st * 8 + grp) * P3_LD + c0;
                double af[8];
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) af[kk] = row[kk * 4 + t4];
                double acc[4][2];
#pragma unroll
                for (int nn = 0; nn < 4; ++nn) {
                    acc[nn][0] = acc[nn][1] = 0.0;
#pragma unroll
                    for (int kk = 0; kk < 2 * nn + 2; ++kk)   // W_dd(j, k) = 0 for k > j
                        dmma884(acc[nn][0], acc[nn][1], af[kk], Wb[(nn * 8 + grp) * P3_WLD + kk * 4 + t4]);
                }
                __syncwarp();
#pragma unroll
                for (int nn = 0; nn < 4; ++nn)
                    *reinterpret_cast<double2*>(row + nn * 8 + 2 * t4) = make_double2(acc[nn][0], acc[nn][1]);
            }
        }
        __syncthreads();
        // ---- (3) trailing lower triangle: A[r][c] -= sum_k X[r][k] X[c][k] in 16 x 16 warp tiles
        {
            const int nt = (NB - r0) / 16;
            const int ntiles = nt * (nt + 1) / 2;
            for (int t = warp; t < ntiles; t += P3_WARPS) {
                int ti = 0;
                while ((ti + 1) * (ti + 2) / 2 <= t) ++ti;
                const int tj = t - ti * (ti + 1) / 2;
                const double* xr = S + (r0 + ti * 16 + grp) * P3_LD + c0 + t4;
                const double* xc = S + (r0 + tj * 16 + grp) * P3_LD + c0 + t4;
                double* cr = S + (r0 + ti * 16 + grp) * P3_LD + r0 + tj * 16 + 2 * t4;
                double acc[2][2][2];
#pragma unroll
                for (int i = 0; i < 2; ++i)
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const double2 v = *reinterpret_cast<const double2*>(cr + i * 8 * P3_LD + j * 8);
                        acc[i][j][0] = v.x;
                        acc[i][j][1] = v.y;
                    }
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    const double a0 = -xr[kk * 4], a1 = -xr[8 * P3_LD + kk * 4];
                    const double b0 = xc[kk * 4], b1 = xc[8 * P3_LD + kk * 4];
                    dmma884(acc[0][0][0], acc[0][0][1], a0, b0);
                    dmma884(acc[0][1][0], acc[0][1][1], a0, b1);
                    dmma884(acc[1][0][0], acc[1][0][1], a1, b0);
                    dmma884(acc[1][1][0], acc[1][1][1], a1, b1);
                }
#pragma unroll
                for (int i = 0; i < 2; ++i)
#pragma unroll
                    for (int j = 0; j < 2; ++j)
                        if (ti != tj || j <= i)   // the upper 8 x 8 atom of a diagonal tile is not part of L
                            *reinterpret_cast<double2*>(cr + i * 8 * P3_LD + j * 8) = make_double2(acc[i][j][0], acc[i][j][1]);
            }
        }
        __syncthreads();
    }

    // ---- L back to global (lower triangle incl. diagonal)
    for (int e = tid; e < NB * NB; e += P3_THREADS) {
        const int r = e >> 7, c = e & (NB - 1);
        if (r < n && c <= r) A[(int64_t)r * lda + c] = S[r * P3_LD + c];
    }
    // ---- inv(L): diagonal blocks are Wd; block (i, j), i > j: W_ij = -W_ii (sum_{k=j}^{i-1} L_ik W_kj), stored
    // TRANSPOSED in the strict upper part of S: W_ij(r, c) lives at S[32 j + c][32 i + r]. Both products run as
    // 16 x 16 DMMA warp tiles (4 per 32 x 32 block).
    constexpr int NBLK = NB / P3_IB;
    for (int dist = 1; dist < NBLK; ++dist) {
        const int nprod = NBLK - dist;  // blocks (i, i - dist), i = dist .. NBLK-1
        for (int t = warp; t < nprod * 4; t += P3_WARPS) {
            const int p = t >> 2, m0 = ((t >> 1) & 1) * 16, n0 = (t & 1) * 16;
            const int i = dist + p, j = p;
            double acc[2][2][2] = {};
            for (int k = j; k < i; ++k) {
                const double* Lik = S + (P3_IB * i + m0 + grp) * P3_LD + P3_IB * k + t4;
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    double b0, b1;
                    if (k == j) {   // W_jj row-major in Wd: B[kr][c]
                        const double* Wjj = Wd + (j * P3_IB + kk * 4 + t4) * P3_WLD + n0 + grp;
                        b0 = Wjj[0];
                        b1 = Wjj[8];
                    } else {        // W_kj transposed in S: B[kr][c] at S[32 j + c][32 k + kr]
                        const double* Wkj = S + (P3_IB * j + n0 + grp) * P3_LD + P3_IB * k + kk * 4 + t4;
                        b0 = Wkj[0];
                        b1 = Wkj[8 * P3_LD];
                    }
                    const double a0 = Lik[kk * 4], a1 = Lik[8 * P3_LD + kk * 4];
                    dmma884(acc[0][0][0], acc[0][0][1], a0, b0);
                    dmma884(acc[0][1][0], acc[0][1][1], a0, b1);
                    dmma884(acc[1][0][0], acc[1][0][1], a1, b0);
                    dmma884(acc[1][1][0], acc[1][1][1], a1, b1);
                }
            }
            double* T = Tm + (p * P3_IB + m0 + grp) * P3_WLD + n0 + 2 * t4;
#pragma unroll
            for (int ii = 0; ii < 2; ++ii)
#pragma unroll
                for (int jj = 0; jj < 2; ++jj)
                    *reinterpret_cast<double2*>(T + ii * 8 * P3_WLD + jj * 8) = make_double2(acc[ii][jj][0], acc[ii][jj][1]);
        }
        __syncthreads();
        for (int t = warp; t < nprod * 4; t += P3_WARPS) {
            const int p = t >> 2, m0 = ((t >> 1) & 1) * 16, n0 = (t & 1) * 16;
            const int i = dist + p, j = p;
            const double* Wii = Wd + (i * P3_IB + m0 + grp) * P3_WLD + t4;
            const double* T = Tm + (p * P3_IB + t4) * P3_WLD + n0 + grp;
            double acc[2][2][2] = {};
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
                if (kk * 4 > m0 + 15) break;   // W_ii(r, t) = 0 for t > r
                const double a0 = -Wii[kk * 4], a1 = -Wii[8 * P3_WLD + kk * 4];
                const double b0 = T[kk * 4 * P3_WLD], b1 = T[kk * 4 * P3_WLD + 8];
                dmma884(acc[0][0][0], acc[0][0][1], a0, b0);
                dmma884(acc[0][1][0], acc[0][1][1], a0, b1);
                dmma884(acc[1][0][0], acc[1][0][1], a1, b0);
                dmma884(acc[1][1][0], acc[1][1][1], a1, b1);
            }
#pragma unroll
            for (int ii = 0; ii < 2; ++ii)
#pragma unroll
                for (int jj = 0; jj < 2; ++jj) {
                    double* dst = S + (P3_IB * j + n0 + jj * 8 + 2 * t4) * P3_LD + P3_IB * i + m0 + ii * 8 + grp;
                    dst[0] = acc[ii][jj][0];
                    dst[P3_LD] = acc[ii][jj][1];
                }
        }
        __syncthreads();
    }
    for (int e = tid; e < NB * NB; e += P3_THREADS) {
        const int r = e >> 7, c = e & (NB - 1);
        double v = 0.0;
        if (r < n && c <= r) {
            const int bi = r / P3_IB, bj = c / P3_IB;
            v = (bi == bj) ? Wd[(bi * P3_IB + (r - bi * P3_IB)) * P3_WLD + (c - bj * P3_IB)] : S[c * P3_LD + r];
        }
        invL[e] = v;
    }
}

__global__ void logdet_kernel(const double* L, int64_t ld, int64_t n, double* out) {
    // single CTA, fixed order: deterministic
    __shared__ double red[32];
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) s += log(L[i * ld + i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        out[0] = 2.0 * t;
    }
}

// One panel: columns [K0, K0 + W) of rows [K0, N), right-looking with NB-wide steps, on `st`.
// Per 128-column step the chain  tile -> panel solve -> update  used to run back to back on `st` (round-2 launch list at
// N = 4096: 64 + 24 + 40 us). Only the NEXT block column has to be up to date before the next tile can start, so the
// update is split (option "panel_split"):
//   st  : tile(k) | solve(k) | update of block column k+1 with column k | tile(k+1) | ...
//   aux :                       wait | update of block columns k+2.. with column k      | ...
// The two updates of one block column (by column k on aux, by column k+1 on st) read-modify-write the same memory, so
// st waits for aux's step k before it touches block column k+2; aux is one in-order stream, which orders the rest.
static int factor_panel(Ctx* c, double* A, int64_t ld, int64_t N, int64_t K0, int64_t W, double* invL, size_t smem,
                        cudaStream_t st, int* dinfo = nullptr, int info_base = 0) {
    if (!dinfo) dinfo = c->dInfo;
    cudaStream_t aux = (c->opt_panel_split && c->aux_stream && c->ev_pcrit && c->ev_paux) ? c->aux_stream : nullptr;
    bool aux_busy = false;
    for (int64_t k0 = K0; k0 < K0 + W; k0 += NB) {
        const int n = (int)((K0 + W - k0 < NB) ? (K0 + W - k0) : NB);
        double* Akk = A + k0 * ld + k0;
        double* invk = invL + (k0 / NB) * (int64_t)NB * NB;
        if (c->opt_potrf_tile == 3)
            potrf_tile3_kernel<<<1, P3_THREADS, P3_SMEM, st>>>(Akk, ld, n, invk, dinfo, info_base + (int)k0);
        else if (c->opt_potrf_tile == 2)
            potrf_tile2_kernel<<<1, P2_THREADS, P2_SMEM, st>>>(Akk, ld, n, invk, dinfo, info_base + (int)k0);
        else
            potrf_tile_kernel<<<1, PT_THREADS, smem, st>>>(Akk, ld, n, invk, dinfo, info_base + (int)k0);
        ++c->launches;
        GP_CUDA(c, cudaGetLastError());
        const int64_t m = N - k0 - n;
        if (m <= 0) break;
        double* below = A + (k0 + n) * ld + k0;  // rows below the diagonal block, this block column
        GemmArgs ps{};  // A[i,k] <- A[i,k] inv(L_kk)^T, in place
        ps.A = below; ps.lda = ld; ps.B = invk; ps.ldb = NB; ps.C = below; ps.ldc = ld;
        ps.M = (int)m; ps.N = n; ps.K = n; ps.alpha = 1.0; ps.beta = 0.0;
        GP_TRY(gemm(c, GEMM_NT, ps, st));
        const int64_t wrem = K0 + W - (k0 + n);  // panel columns still to be updated
        if (wrem <= 0) continue;
        GemmArgs pu{};
        pu.A = below; pu.lda = ld; pu.B = below; pu.ldb = ld; pu.C = A + (k0 + n) * ld + (k0 + n); pu.ldc = ld;
        pu.M = (int)m; pu.N = (int)wrem; pu.K = n; pu.alpha = -1.0; pu.beta = 1.0;
        const int64_t nxt = (wrem < NB) ? wrem : NB;
        if (!aux || wrem <= nxt || m <= nxt) {
            if (aux_busy) GP_CUDA(c, cudaStreamWaitEvent(st, c->ev_paux, 0));
            GP_TRY(gemm(c, GEMM_NT, pu, st));
            continue;
        }
        // critical: the next block column (all rows below); it last received aux's update of the previous step
        if (aux_busy) GP_CUDA(c, cudaStreamWaitEvent(st, c->ev_paux, 0));
        pu.N = (int)nxt;
        GP_TRY(gemm(c, GEMM_NT, pu, st));
        GP_CUDA(c, cudaEventRecord(c->ev_pcrit, st));
        GP_CUDA(c, cudaStreamWaitEvent(aux, c->ev_pcrit, 0));
        // the rest of the panel, rows from the block after next (what lies above is not part of L)
        GemmArgs pr{};
        pr.A = below + nxt * ld; pr.lda = ld; pr.B = below + nxt * ld; pr.ldb = ld;
        pr.C = A + (k0 + n + nxt) * ld + (k0 + n + nxt); pr.ldc = ld;
        pr.M = (int)(m - nxt); pr.N = (int)(wrem - nxt); pr.K = n; pr.alpha = -1.0; pr.beta = 1.0;
        GP_TRY(gemm(c, GEMM_NT, pr, aux));
        GP_CUDA(c, cudaEventRecord(c->ev_paux, aux));
        aux_busy = true;
    }
    if (aux_busy) GP_CUDA(c, cudaStreamWaitEvent(st, c->ev_paux, 0));  // the panel is complete on `st`
    return 0;
}

// Left-looking outer loop (option "chol_left"): block column J (width NBO) is brought up to date with ALL earlier
// panels in long-K GEMMs and is written once, instead of every trailing tile being read-modified-written once per
// panel as in the right-looking form below (the K = 1024 epilogue cost ~8 % of the tensor pipe in the round-1 ncu
// capture). The update of column J is split by what it depends on:
//     G_a(J): K range = panels 0 .. J-2   (independent of the panel being factored right now)
//     G_b(J): K range = panel J-1         (the only part on the critical path)
//   main stream : ... G_a(J) | wait panel(J-1) | G_b(J) | G_a(J+1) | wait panel(J) | G_b(J+1) ...
//   panel stream:                               wait G_b(J) | panel(J)
// so panel(J) (latency-bound) always runs under G_a(J+1) (the bulk of the flops, K up to N).
// MEASURED (B200, tools/left_ab.py): 404 ms vs 384 ms right-looking at N = 32768, 1266 vs 1223 ms at N = 49152: the
// column GEMMs have only (N - J0) / 128 x 16 tiles (3-4 waves of 296 resident CTAs, wave-quantisation loss) and the
// G_b -> panel chain is fully serial. Kept as an option (default off); the right-looking form is the product path.
static int update_column(Ctx* c, double* A, int64_t ld, int64_t N, int64_t J0, int64_t W, int64_t k_begin, int64_t k_end,
                         cudaStream_t st) {
    const int64_t K = k_end - k_begin;
    if (K <= 0) return 0;
    // top W x W block: symmetric, lower tiles only
    GemmArgs t{};
    t.A = A + J0 * ld + k_begin; t.lda = ld; t.B = t.A; t.ldb = ld; t.C = A + J0 * ld + J0; t.ldc = ld;
    t.M = (int)W; t.N = (int)W; t.K = (int)K; t.alpha = -1.0; t.beta = 1.0; t.lower = 1;
    GP_TRY(gemm(c, GEMM_NT, t, st));
    const int64_t m = N - J0 - W;
    if (m > 0) {
        GemmArgs b{};
        b.A = A + (J0 + W) * ld + k_begin; b.lda = ld; b.B = A + J0 * ld + k_begin; b.ldb = ld;
        b.C = A + (J0 + W) * ld + J0; b.ldc = ld;
        b.M = (int)m; b.N = (int)W; b.K = (int)K; b.alpha = -1.0; b.beta = 1.0;
        GP_TRY(gemm(c, GEMM_NT, b, st));
    }
    return 0;
}

static int potrf_left(Ctx* c, double* A, int64_t ld, int64_t N, double* invL, int64_t NBO, size_t smem) {
    cudaStream_t sm = c->stream, sp = c->panel_stream ? c->panel_stream : c->stream;
    const int64_t nJ = (N + NBO - 1) / NBO;
    std::vector<cudaEvent_t> ev_panel((size_t)nJ, nullptr), ev_gb((size_t)nJ, nullptr);
    int rc = 0;
    auto mk = [&](cudaEvent_t& e) { return cudaEventCreateWithFlags(&e, cudaEventDisableTiming); };
    cudaEvent_t ev_start = nullptr;
    mk(ev_start);
    cudaEventRecord(ev_start, sm);  // panel 0 waits for everything queued so far (the Gram build)
    cudaStreamWaitEvent(sp, ev_start, 0);
    rc = factor_panel(c, A, ld, N, 0, (N < NBO) ? N : NBO, invL, smem, sp);
    mk(ev_panel[0]);
    cudaEventRecord(ev_panel[0], sp);
    std::vector<cudaEvent_t> pev;
    std::vector<double> pflops;
    for (int64_t J = 1; J < nJ && rc == 0; ++J) {
        const int64_t J0 = J * NBO;
        const int64_t W = (N - J0 < NBO) ? (N - J0) : NBO;
        if (J >= 2) {
            cudaStreamWaitEvent(sm, ev_panel[(size_t)J - 2], 0);
            cudaEvent_t e0 = nullptr, e1 = nullptr;
            if (c->opt_profile) {
                cudaEventCreate(&e0);
                cudaEventCreate(&e1);
                cudaEventRecord(e0, sm);
            }
            rc = update_column(c, A, ld, N, J0, W, 0, J0 - NBO, sm);
            if (c->opt_profile) {
                cudaEventRecord(e1, sm);
                pev.push_back(e0);
                pev.push_back(e1);
                pflops.push_back((double)(J0 - NBO) * ((double)W * (W + 1) + 2.0 * (double)(N - J0 - W) * W));
            }
            if (rc) break;
        }
        cudaStreamWaitEvent(sm, ev_panel[(size_t)J - 1], 0);
        rc = update_column(c, A, ld, N, J0, W, J0 - NBO, J0, sm);
        if (rc) break;
        mk(ev_gb[(size_t)J]);
        cudaEventRecord(ev_gb[(size_t)J], sm);
        cudaStreamWaitEvent(sp, ev_gb[(size_t)J], 0);
        rc = factor_panel(c, A, ld, N, J0, W, invL, smem, sp);
        mk(ev_panel[(size_t)J]);
        cudaEventRecord(ev_panel[(size_t)J], sp);
    }
    if (ev_panel[(size_t)nJ - 1]) cudaStreamWaitEvent(sm, ev_panel[(size_t)nJ - 1], 0);  // rejoin the main stream
    if (c->opt_profile) {
        c->prof_tb_ms = c->prof_tb_flops = c->prof_tb_launches = 0.0;
        for (size_t i = 0; i + 1 < pev.size(); i += 2) {
            float ms = 0.f;
            cudaEventSynchronize(pev[i + 1]);
            if (cudaEventElapsedTime(&ms, pev[i], pev[i + 1]) == cudaSuccess) {
                c->prof_tb_ms += ms;
                c->prof_tb_flops += pflops[i / 2];
                c->prof_tb_launches += 1.0;
            }
            cudaEventDestroy(pev[i]);
            cudaEventDestroy(pev[i + 1]);
        }
    }
    for (auto e : ev_panel) if (e) cudaEventDestroy(e);
    for (auto e : ev_gb) if (e) cudaEventDestroy(e);
    cudaEventDestroy(ev_start);
    return rc;
}

// Two-level right-looking Cholesky with one-panel look-ahead:
//   main stream : ... T^b(K-1) | wait panel(K) | T^a(K) | T^b(K) ...
//   panel stream:      wait T^a(K-1) | panel(K)
// T^a(K) updates only the next panel's block column with panel K (so panel K+1 can start), T^b(K)
// the rest of the trailing matrix (K = NBO-deep SYRK on lower tiles, where ~97 % of the flops are).
int potrf_blocked(Ctx* c, double* A, int64_t ld, int64_t N, double* invL, int* info_host) {
    const size_t smem = sizeof(double) * (NB * PT_LD + 2 * NB);
    GP_CUDA(c, cudaFuncSetAttribute(potrf_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    GP_CUDA(c, cudaFuncSetAttribute(potrf_tile2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P2_SMEM));
    GP_CUDA(c, cudaFuncSetAttribute(potrf_tile3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P3_SMEM));
    GP_CUDA(c, cudaMemsetAsync(c->dInfo, 0, sizeof(int), c->stream));
    int64_t NBO = c->opt_nbo;
    // measured on B200 with the round-2 panel (tools/nbo_sweep.py; DMMA tile, 32 x 128 panel GEMMs, split panel step):
    // 384 wins at N = 8192 (9.19 ms), 768 at 12288-16384 (23.7 / 50.7 ms), 1024 from 24576 (158.7 / 364.3 ms)
    if (NBO <= 0) NBO = (N <= 10240) ? 384 : ((N <= 20480) ? 768 : 1024);
    if (NBO < NB) NBO = NB;
    NBO = NBO / NB * NB;
    if (c->opt_chol_left && c->panel_stream && N > 2 * NBO) {
        GP_TRY(potrf_left(c, A, ld, N, invL, NBO, smem));
        if (info_host) {
            GP_CUDA(c, cudaMemcpyAsync(info_host, c->dInfo, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
            GP_CUDA(c, cudaStreamSynchronize(c->stream));
        }
        return 0;
    }
    const bool la = c->opt_lookahead && c->panel_stream && N > NBO;
    cudaStream_t sm = c->stream, sp = la ? c->panel_stream : c->stream;
    cudaEvent_t ev_panel = nullptr, ev_ta = nullptr;
    if (la) {
        GP_CUDA(c, cudaEventCreateWithFlags(&ev_panel, cudaEventDisableTiming));
        GP_CUDA(c, cudaEventCreateWithFlags(&ev_ta, cudaEventDisableTiming));
        GP_CUDA(c, cudaEventRecord(ev_ta, sm));  // panel 0 waits for everything queued so far (the Gram build)
    }
    int rc = 0;
    std::vector<cudaEvent_t> pev;   // opt_profile: (start, stop) per trailing-update launch
    std::vector<double> pflops;
    for (int64_t K0 = 0; K0 < N && rc == 0; K0 += NBO) {
        const int64_t W = (N - K0 < NBO) ? (N - K0) : NBO;
        if (la) GP_CUDA(c, cudaStreamWaitEvent(sp, ev_ta, 0));
        rc = factor_panel(c, A, ld, N, K0, W, invL, smem, sp);
        if (rc) break;
        if (la) {
            GP_CUDA(c, cudaEventRecord(ev_panel, sp));
            GP_CUDA(c, cudaStreamWaitEvent(sm, ev_panel, 0));
        }
        const int64_t m = N - K0 - W;  // trailing size
        if (m <= 0) break;
        const double* P = A + (K0 + W) * ld + K0;  // m x W panel below the diagonal blocks
        const int64_t W2 = la ? ((m < NBO) ? m : NBO) : 0;
        if (W2 > 0) {
            GemmArgs ta{};  // next panel's block column (its top W2 x W2 block only needs the lower half; the waste is small)
            ta.A = P; ta.lda = ld; ta.B = P; ta.ldb = ld; ta.C = A + (K0 + W) * ld + (K0 + W); ta.ldc = ld;
            ta.M = (int)m; ta.N = (int)W2; ta.K = (int)W; ta.alpha = -1.0; ta.beta = 1.0;
            rc = gemm(c, GEMM_NT, ta, sm);
            if (rc) break;
            GP_CUDA(c, cudaEventRecord(ev_ta, sm));
        }
        const int64_t mb = m - W2;
        if (mb > 0) {
            const double* Pb = P + W2 * ld;
            GemmArgs tb{};
            tb.A = Pb; tb.lda = ld; tb.B = Pb; tb.ldb = ld; tb.C = A + (K0 + W + W2) * ld + (K0 + W + W2); tb.ldc = ld;
            tb.M = (int)mb; tb.N = (int)mb; tb.K = (int)W; tb.alpha = -1.0; tb.beta = 1.0; tb.lower = 1;
            cudaEvent_t e0 = nullptr, e1 = nullptr;
            if (c->opt_profile) {
                cudaEventCreate(&e0);
                cudaEventCreate(&e1);
                cudaEventRecord(e0, sm);
            }
            rc = gemm(c, GEMM_NT, tb, sm);
            if (c->opt_profile) {
                cudaEventRecord(e1, sm);
                pev.push_back(e0);
                pev.push_back(e1);
                pflops.push_back((double)mb * (double)(mb + 1) * (double)W);  // 2 * (mb (mb+1) / 2) * W: lower triangle only
            }
        }
    }
    if (la) {
        cudaEventDestroy(ev_panel);
        cudaEventDestroy(ev_ta);
    }
    if (c->opt_profile) {
        c->prof_tb_ms = c->prof_tb_flops = c->prof_tb_launches = 0.0;
        for (size_t i = 0; i + 1 < pev.size(); i += 2) {
            float ms = 0.f;
            cudaEventSynchronize(pev[i + 1]);
            if (cudaEventElapsedTime(&ms, pev[i], pev[i + 1]) == cudaSuccess) {
                c->prof_tb_ms += ms;
                c->prof_tb_flops += pflops[i / 2];
                c->prof_tb_launches += 1.0;
            }
            cudaEventDestroy(pev[i]);
            cudaEventDestroy(pev[i + 1]);
        }
    }
    if (rc) return rc;
    if (info_host) {
        GP_CUDA(c, cudaMemcpyAsync(info_host, c->dInfo, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
        GP_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return 0;
}

// Tall panel (rows x w, leading dimension ld): Cholesky of the top w x w block and the rows below solved against
// it, i.e. one block column of L. invL receives ceil(w / NB) inverted diagonal blocks. Used by the block-cyclic
// multi-GPU factorisation, where the owner of a block column factors it before broadcasting it.
int potrf_panel(Ctx* c, double* P, int64_t ld, int64_t rows, int64_t w, double* invL, int col0, int* info_host) {
    const size_t smem = sizeof(double) * (NB * PT_LD + 2 * NB);
    GP_CUDA(c, cudaFuncSetAttribute(potrf_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    GP_CUDA(c, cudaFuncSetAttribute(potrf_tile2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P2_SMEM));
    GP_CUDA(c, cudaFuncSetAttribute(potrf_tile3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P3_SMEM));
    GP_CUDA(c, cudaMemsetAsync(c->dInfo, 0, sizeof(int), c->stream));
    (void)col0;
    GP_TRY(factor_panel(c, P, ld, rows, 0, w, invL, smem, c->stream));
    if (info_host) {
        GP_CUDA(c, cudaMemcpyAsync(info_host, c->dInfo, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
        GP_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return 0;
}

// The same without any host synchronisation: INFO stays on the device. dinfo is a caller-owned device int that is NOT
// reset here; the first failing leading minor (col0 + local index, 1-based) wins (atomicCAS from 0), so one int can
// collect the status of a whole multi-step, multi-GPU factorisation and be read once at the end.
int potrf_panel_async(Ctx* c, double* P, int64_t ld, int64_t rows, int64_t w, double* invL, int col0, int* dinfo) {
    const size_t smem = sizeof(double) * (NB * PT_LD + 2 * NB);
    static bool attr_set[64] = {false};
    if (c->device >= 0 && c->device < 64 && !attr_set[c->device]) {
        GP_CUDA(c, cudaFuncSetAttribute(potrf_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        GP_CUDA(c, cudaFuncSetAttribute(potrf_tile2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P2_SMEM));
        GP_CUDA(c, cudaFuncSetAttribute(potrf_tile3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P3_SMEM));
        attr_set[c->device] = true;
    }
    return factor_panel(c, P, ld, rows, 0, w, invL, smem, c->stream, dinfo, col0);
}

int logdet_from_factor(Ctx* c, const double* L, int64_t ld, int64_t N, double* out_host) {
    logdet_kernel<<<1, 1024, 0, c->stream>>>(L, ld, N, c->dScratch);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    GP_CUDA(c, cudaMemcpyAsync(out_host, c->dScratch, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

}  // namespace gprb200
