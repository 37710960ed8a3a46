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

int trsm_lower_fwd(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* B, int64_t ldb,
                   int64_t M) {
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
    return (M >= c->opt_solve_left_min_m) ? bwd_left(c, L, ldl, invL, N, B, ldb, M)
                                          : bwd_right(c, L, ldl, invL, N, B, ldb, M);
}

int inverse_from_factor(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* C, int64_t ldc) {
    GP_TRY(set_identity(c, C, ldc, N));
    GP_TRY(trsm_lower_fwd(c, L, ldl, invL, N, C, ldc, N));
    GP_TRY(trsm_lower_bwd(c, L, ldl, invL, N, C, ldc, N));
    // the two sweeps give a numerically (not bitwise) symmetric result; the reference's LU inverse is
    // not exactly symmetric either. Mirror the lower triangle so C == C^T holds exactly.
    GP_TRY(symmetrize_from_lower(c, C, ldc, N));
    return 0;
}

}  // namespace gprb200
