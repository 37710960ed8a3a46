// Sparse (inducing-point) Gaussian process on the device: the regression part of the
// reference's SparseGaussianProcess<T> (include/SparseGaussianProcess.h:30-420) from the building blocks of the dense
// path (gram.cu, dgemm.cu, potrf.cu, solve.cu). n dense samples (the handle's training set), m <= n inducing points.
//
//   PreComputeRegression (:274-312)    K   = k(Xm, Xm) + jitter I                     m x m   gram_build + add_diag
//                                      Knm = k(X, Xm)                                 n x m   gram_build (cross)
//                                      S   = K + sigma^-2 Knm^T Knm                   m x m   DMMA GEMM (TN, depth n)
//                                      alpha = sigma^-2 S^-1 Knm^T Y                  m x M   GEMM + Cholesky solve
//   Predict (:84-90)                   Kx^T alpha                                            gram_build + GEMM
//   operator()(x, y) (:92-106)         k(x,y) - Kx^T K^-1 Ky + Kx^T S^-1 Ky                  two Cholesky solves
//
// The reference forms K^-1 and Sigma = S^-1 explicitly and computes alpha = K^-1 (sigma^-2 K Sigma Kmn Y) and the
// regression matrix K^-1 (K Sigma K) K^-1; both collapse to the expressions above (K^-1 K = I), which is what is
// evaluated here through the two Cholesky factors L_K and L_S.
//
// SparseGaussianLogLikelihood (include/SparseLikelihood.h:111-551), gprb200_sparse_loglik. The reference materialises
// C_inv = (sigma^2 I + Knm K^-1 Kmn)^-1 (n x n) and P blocks grad(C)_p (n x n each) and takes traces of their products.
// With alpha as above, beta = C_inv Y = sigma^-2 (Y - Knm alpha) and the identities Kmn C_inv = sigma^-2 K S^-1 Kmn,
// K^-1 Kmn C_inv Knm K^-1 = K^-1 - S^-1, the same numbers are
//   value_i = -1/2 y_i . beta_i - 1/2 (n log sigma^2 - log det K + log det S) - n/2 log 2 pi
//   grad_p  = sum_ab dKnm_p[a,b] (beta alpha^T - sigma^-2 Knm S^-1)[a,b] + sum_bc dKmm_p[b,c] 1/2 (K^-1 - S^-1 - alpha alpha^T)[b,c]
// i.e. two m x m inverses, one n x m x m GEMM and two fused derivative-Gram contractions (gram.cu): O(n m^2) work and
// O(n m) memory where the reference needs O(P n^3) and O(P n^2). The n x n core matrix Knm K^-1 Kmn (:360-378) is
// never formed.
#include <float.h>
#include <math.h>
#include <vector>
#include "../../include/gprb200.h"
#include "common.cuh"

namespace gprb200 {

struct SparseModel {
    int64_t m = 0, ldm = 0;
    double jitter = 0.0;
    double* dXm = nullptr;      // m x d inducing points
    double* dLK = nullptr;      // m x ldm  Cholesky factor of K
    double* dInvLK = nullptr;
    double* dLS = nullptr;      // m x ldm  Cholesky factor of S
    double* dInvLS = nullptr;
    double* dAlpha = nullptr;   // m x ldy
    int64_t alpha_ldy = 0;      // leading dimension dAlpha was allocated for (set_train may change M)
    int d = 0;                  // dimension of the inducing points
    bool learned = false;
};

namespace {

SparseModel* sparse_of(Ctx* c) { return static_cast<SparseModel*>(c->sparse); }

void sparse_free(SparseModel* s) {
    cudaFree(s->dXm); cudaFree(s->dLK); cudaFree(s->dInvLK); cudaFree(s->dLS); cudaFree(s->dInvLS); cudaFree(s->dAlpha);
    s->dXm = s->dLK = s->dInvLK = s->dLS = s->dInvLS = s->dAlpha = nullptr;
    s->learned = false;
}

template <class T>
int sp_alloc(Ctx* c, T** p, size_t count) {
    GP_CUDA(c, cudaMalloc((void**)p, sizeof(T) * (count ? count : 1)));
    return 0;
}

struct Tmp {
    Ctx* c;
    double* p = nullptr;
    explicit Tmp(Ctx* c_) : c(c_) {}
    int alloc(size_t count) {
        GP_CUDA(c, cudaMallocAsync((void**)&p, sizeof(double) * (count ? count : 1), c->stream));
        return 0;
    }
    ~Tmp() {
        if (p) cudaFreeAsync(p, c->stream);
    }
};

// knm: left allocated (n x ldm) for the caller (the likelihood needs it again)
int sparse_learn_core(Ctx* c, Tmp& knm) {
    SparseModel* s = sparse_of(c);
    if (!s || !s->dXm) return fail(c, -2, "SparseGaussianProcess::Initialize: no inducing samples defined during initialization");
    if (!c->have_train || !c->dY) return fail(c, -2, "SparseGaussianProcess::Initialize: no dense samples defined during initialization");
    if (!c->have_kernel) return fail(c, -2, "SparseGaussianProcess: no kernel");
    if (!(c->sigma > 0.0)) return fail(c, -2, "SparseGaussianProcess::ComputeCoreMatrices: sigma must be positive.");
    const int64_t n = c->N, m = s->m, M = c->M;
    if (m > n)
        return fail(c, -2, "SparseGaussianProcess::ComputeKernelVectorMatrix: number of dense samples must be higher than the "
                           "number of sparse samples");
    if (s->d != c->d)
        return fail(c, -2, "SparseGaussianProcess: the inducing points have dimension " + std::to_string(s->d) +
                               " but the dense samples have dimension " + std::to_string(c->d));
    s->learned = false;
    const int64_t ldm = s->ldm, nblk = (m + NB - 1) / NB;
    if (s->dAlpha && s->alpha_ldy != c->ldy) {   // a later set_train changed the number of label columns
        cudaFree(s->dAlpha);
        s->dAlpha = nullptr;
    }
    if (!s->dAlpha) {
        GP_TRY(sp_alloc(c, &s->dAlpha, (size_t)m * c->ldy));
        s->alpha_ldy = c->ldy;
    }
    if (!s->dLK) {
        GP_TRY(sp_alloc(c, &s->dLK, (size_t)m * ldm));
        GP_TRY(sp_alloc(c, &s->dLS, (size_t)m * ldm));
        GP_TRY(sp_alloc(c, &s->dInvLK, (size_t)nblk * NB * NB));
        GP_TRY(sp_alloc(c, &s->dInvLS, (size_t)nblk * NB * NB));
    }
    // K = k(Xm, Xm) + jitter I (ComputeKernelMatrixWithJitter, :175-180); S starts as a copy of it
    GP_TRY(gram_build(c, s->dXm, m, s->dXm, m, s->dLK, ldm, true, false, s->jitter));
    GP_TRY(copy2d(c, s->dLS, ldm, s->dLK, ldm, m, m));
    GP_TRY(knm.alloc((size_t)n * ldm));
    GP_TRY(gram_build(c, c->dX, n, s->dXm, m, knm.p, ldm, false, false, 0.0));   // ComputeKernelVectorMatrix (:214-232)
    const double is2 = 1.0 / (c->sigma * c->sigma);
    GemmArgs g{};  // S += sigma^-2 Knm^T Knm
    g.A = knm.p; g.lda = ldm; g.B = knm.p; g.ldb = ldm; g.C = s->dLS; g.ldc = ldm;
    g.M = (int)m; g.N = (int)m; g.K = (int)n; g.alpha = is2; g.beta = 1.0; g.lower = 1;
    GP_TRY(gemm(c, GEMM_TN, g));
    // B = sigma^-2 Knm^T Y  (m x M), solved in place into alpha
    GemmArgs r{};
    r.A = knm.p; r.lda = ldm; r.B = c->dY; r.ldb = c->ldy; r.C = s->dAlpha; r.ldc = c->ldy;
    r.M = (int)m; r.N = (int)M; r.K = (int)n; r.alpha = is2; r.beta = 0.0;
    GP_TRY(gemm(c, GEMM_TN, r));
    int info = 0;
    GP_TRY(potrf_blocked(c, s->dLK, ldm, m, s->dInvLK, &info));
    if (info != 0) {
        fail(c, info, "SparseGaussianProcess: the inducing kernel matrix K + jitter I is not positive definite (leading minor of order " +
                          std::to_string(info) + ")");
        return info;
    }
    GP_TRY(potrf_blocked(c, s->dLS, ldm, m, s->dInvLS, &info));
    if (info != 0) {
        fail(c, info, "SparseGaussianProcess: K + sigma^-2 Kmn Knm is not positive definite (leading minor of order " +
                          std::to_string(info) + ")");
        return info;
    }
    GP_TRY(trsm_lower_fwd(c, s->dLS, ldm, s->dInvLS, m, s->dAlpha, c->ldy, M));
    GP_TRY(trsm_lower_bwd(c, s->dLS, ldm, s->dInvLS, m, s->dAlpha, c->ldy, M));
    s->learned = true;
    return 0;
}

int sparse_learn_impl(Ctx* c) {
    Tmp knm(c);
    return sparse_learn_core(c, knm);
}

int ensure_sparse(Ctx* c) {
    SparseModel* s = sparse_of(c);
    if (s && s->learned) return 0;
    return sparse_learn_impl(c);
}

// q[t] = (L^-1 Ka_t) . (L^-1 Kb_t) with Ka, Kb generated as m x T (one column per test point)
int sparse_quad(Ctx* c, const double* L, const double* invL, const double* dXa, const double* dXb, int64_t T, double* dq) {
    SparseModel* s = sparse_of(c);
    const int64_t m = s->m, ldv = round_up(T, 16);
    Tmp va(c), vb(c);
    GP_TRY(va.alloc((size_t)m * ldv));
    GP_TRY(gram_build(c, s->dXm, m, dXa, T, va.p, ldv, false, false, 0.0));
    GP_TRY(trsm_lower_fwd(c, L, s->ldm, invL, m, va.p, ldv, T));
    const double* pb = va.p;
    if (dXb != dXa) {
        GP_TRY(vb.alloc((size_t)m * ldv));
        GP_TRY(gram_build(c, s->dXm, m, dXb, T, vb.p, ldv, false, false, 0.0));
        GP_TRY(trsm_lower_fwd(c, L, s->ldm, invL, m, vb.p, ldv, T));
        pb = vb.p;
    }
    return coldot(c, va.p, ldv, pb, ldv, m, T, dq);
}

// One factor of EfficientDeterminant (include/SparseLikelihood.h:142-150) as a long double would hold it: zero below
// the denormal range, +inf above LDBL_MAX, NaN from 0 x inf; finite values are carried as logarithms.
struct LdVal {
    enum Kind { ZERO, FINITE, INF, NOTANUMBER } kind;
    long double lg;
    static LdVal from_log(long double lg) {
        static const long double lg_max = logl(LDBL_MAX), lg_den = logl(LDBL_TRUE_MIN);
        if (isnan((double)lg)) return {NOTANUMBER, 0.0L};
        if (lg > lg_max) return {INF, 0.0L};
        if (lg < lg_den) return {ZERO, 0.0L};
        return {FINITE, lg};
    }
    LdVal times(const LdVal& o) const {
        if (kind == NOTANUMBER || o.kind == NOTANUMBER) return {NOTANUMBER, 0.0L};
        if ((kind == ZERO && o.kind == INF) || (kind == INF && o.kind == ZERO)) return {NOTANUMBER, 0.0L};
        if (kind == ZERO || o.kind == ZERO) return {ZERO, 0.0L};
        if (kind == INF || o.kind == INF) return {INF, 0.0L};
        return from_log(lg + o.lg);
    }
};

}  // namespace

int dgram_contract_xy(Ctx* c, const double* Xa, int64_t Na, const double* Xb, int64_t Nb, const double* W, int64_t ldw,
                      double* out_host);  // gram.cu

namespace {

int sparse_loglik_impl(Ctx* c, int mode, double* value, double* grad, double* jac) {
    SparseModel* s = sparse_of(c);
    if (!s || !s->dXm)
        return fail(c, -2, "SparseLikelihood::GetValueAndParameterDerivative: there are no inducing samples specified");
    const int P = c->prog.nparams;
    if (grad && c->M != 1)
        return fail(c, -2, "SparseLikelihood::GetParameterDerivatives: dimension missmatch in calculating derivative of data fit term");
    // the reference rebuilds every matrix on each call (ComputeCoreMatrices, include/SparseGaussianProcess.h:323-350)
    Tmp knm(c);
    GP_TRY(sparse_learn_core(c, knm));
    const int64_t n = c->N, m = s->m, M = c->M, ldm = s->ldm, ldy = c->ldy;
    const double s2 = c->sigma * c->sigma, is2 = 1.0 / s2;

    // beta = C_inv Y = sigma^-2 (Y - Knm alpha)                                      (C_inv of :132-138 applied to Y)
    Tmp beta(c), qd(c);
    GP_TRY(beta.alloc((size_t)n * ldy));
    GP_TRY(copy2d(c, beta.p, ldy, c->dY, ldy, n, M));
    GemmArgs b{};
    b.A = knm.p; b.lda = ldm; b.B = s->dAlpha; b.ldb = ldy; b.C = beta.p; b.ldc = ldy;
    b.M = (int)n; b.N = (int)M; b.K = (int)m; b.alpha = -is2; b.beta = is2;
    GP_TRY(gemm(c, GEMM_NN, b));

    if (value) {
        std::vector<double> q((size_t)M);
        GP_TRY(qd.alloc((size_t)M));
        GP_TRY(coldot(c, c->dY, ldy, beta.p, ldy, n, M, qd.p));
        GP_CUDA(c, cudaMemcpyAsync(q.data(), qd.p, sizeof(double) * M, cudaMemcpyDeviceToHost, c->stream));
        double ldK = 0.0, ldS = 0.0;
        GP_TRY(logdet_from_factor(c, s->dLK, ldm, m, &ldK));
        GP_TRY(logdet_from_factor(c, s->dLS, ldm, m, &ldS));
        GP_CUDA(c, cudaStreamSynchronize(c->stream));
        const long double lg_s2n = (long double)n * logl((long double)s2);
        long double cp;
        if (mode == GPRB200_LOGLIK_TRUE) {
            cp = -0.5L * (-(long double)ldK + lg_s2n + (long double)ldS);
        } else {
            // determinant = det(K^-1) * prod(sigma^2) * det(S), left to right in long double (:145-149); an infinite
            // det(K^-1) is replaced by LDBL_MAX first (:146-148)
            LdVal detB = LdVal::from_log(-(long double)ldK);
            if (detB.kind == LdVal::INF) detB = {LdVal::FINITE, logl(LDBL_MAX)};
            const LdVal det = detB.times(LdVal::from_log(lg_s2n)).times(LdVal::from_log((long double)ldS));
            const bool nan_is_min = (grad != nullptr) || (jac != nullptr);   // :351, :477 test isnan, :177 does not
            if (det.kind == LdVal::NOTANUMBER && !nan_is_min)
                return fail(c, -4, "SparseLikelihood::GetValueAndParameterDerivative: likelihood value is not a number.");
            if (det.kind == LdVal::ZERO || det.kind == LdVal::NOTANUMBER || det.lg <= logl(LDBL_MIN)) cp = -0.5L * logl(LDBL_MIN);
            else if (det.kind == LdVal::INF) cp = -0.5L * logl(LDBL_MAX);
            else cp = -0.5L * det.lg;
        }
        const double ct = -((double)n / 2.0) * log(2.0 * M_PI);   // :198
        double sum = 0.0;
        for (int64_t k = 0; k < M; ++k) {
            value[k] = (double)((long double)(-0.5 * q[(size_t)k]) + (cp + (long double)ct));
            sum += value[k];
        }
        if (isnan(sum))
            return fail(c, -4, "SparseLikelihood::GetValueAndParameterDerivative: likelihood value is not a number.");
    }
    if (!grad && !jac) return 0;

    // explicit m x m inverses from the two factors
    Tmp kinv(c), sinv(c), wnm(c);
    GP_TRY(kinv.alloc((size_t)m * ldm));
    GP_TRY(sinv.alloc((size_t)m * ldm));
    GP_TRY(inverse_from_factor(c, s->dLK, ldm, s->dInvLK, m, kinv.p, ldm));
    GP_TRY(inverse_from_factor(c, s->dLS, ldm, s->dInvLS, m, sinv.p, ldm));
    // complexity part, shared by all output columns:  Wnm = -sigma^-2 Knm S^-1,  Wmm = 1/2 (K^-1 - S^-1)
    GP_TRY(wnm.alloc((size_t)n * ldm));
    GemmArgs g{};
    g.A = knm.p; g.lda = ldm; g.B = sinv.p; g.ldb = ldm; g.C = wnm.p; g.ldc = ldm;
    g.M = (int)n; g.N = (int)m; g.K = (int)m; g.alpha = -is2; g.beta = 0.0;
    GP_TRY(gemm(c, GEMM_NN, g));
    GP_TRY(axpby2d(c, kinv.p, ldm, 0.5, kinv.p, ldm, -0.5, sinv.p, ldm, m, m));   // kinv := Wmm

    std::vector<double> t1((size_t)P), t2((size_t)P);
    if (grad) {
        // M == 1: fold the data-fit part into the same two passes: Wnm += beta alpha^T, Wmm -= 1/2 alpha alpha^T
        GemmArgs r1{};
        r1.A = beta.p; r1.lda = ldy; r1.B = s->dAlpha; r1.ldb = ldy; r1.C = wnm.p; r1.ldc = ldm;
        r1.M = (int)n; r1.N = (int)m; r1.K = (int)M; r1.alpha = 1.0; r1.beta = 1.0;
        GP_TRY(gemm(c, GEMM_NT, r1));
        GemmArgs r2{};
        r2.A = s->dAlpha; r2.lda = ldy; r2.B = s->dAlpha; r2.ldb = ldy; r2.C = kinv.p; r2.ldc = ldm;
        r2.M = (int)m; r2.N = (int)m; r2.K = (int)M; r2.alpha = -0.5; r2.beta = 1.0;
        GP_TRY(gemm(c, GEMM_NT, r2));
        GP_TRY(dgram_contract_xy(c, c->dX, n, s->dXm, m, wnm.p, ldm, t1.data()));
        GP_TRY(dgram_contract_xy(c, s->dXm, m, s->dXm, m, kinv.p, ldm, t2.data()));
        for (int p = 0; p < P; ++p) grad[p] = t1[(size_t)p] + t2[(size_t)p];
        if (!jac) return 0;
        for (int p = 0; p < P; ++p) jac[p] = grad[p];   // M == 1: the Jacobian is the gradient
        return 0;
    }
    // Jacobian, M columns: cp_grad[p] from the two contractions, then per parameter
    //   df(i,p) = beta_i^T dKnm_p alpha_i - 1/2 alpha_i^T dKmm_p alpha_i                  (:462-469 with beta, alpha)
    GP_TRY(dgram_contract_xy(c, c->dX, n, s->dXm, m, wnm.p, ldm, t1.data()));
    GP_TRY(dgram_contract_xy(c, s->dXm, m, s->dXm, m, kinv.p, ldm, t2.data()));
    Tmp dmm(c), tn(c), tm(c), q1(c), q2(c);
    GP_TRY(dmm.alloc((size_t)m * ldm));
    GP_TRY(tn.alloc((size_t)n * ldy));
    GP_TRY(tm.alloc((size_t)m * ldy));
    GP_TRY(q1.alloc((size_t)M));
    GP_TRY(q2.alloc((size_t)M));
    std::vector<double> h1((size_t)M), h2((size_t)M);
    for (int p = 0; p < P; ++p) {
        GP_TRY(dgram_build_xy(c, c->dX, n, s->dXm, m, wnm.p, ldm, p));      // wnm reused as dKnm_p
        GemmArgs a1{};
        a1.A = wnm.p; a1.lda = ldm; a1.B = s->dAlpha; a1.ldb = ldy; a1.C = tn.p; a1.ldc = ldy;
        a1.M = (int)n; a1.N = (int)M; a1.K = (int)m; a1.alpha = 1.0; a1.beta = 0.0;
        GP_TRY(gemm(c, GEMM_NN, a1));
        GP_TRY(coldot(c, beta.p, ldy, tn.p, ldy, n, M, q1.p));
        GP_TRY(dgram_build_xy(c, s->dXm, m, s->dXm, m, dmm.p, ldm, p));
        GemmArgs a2{};
        a2.A = dmm.p; a2.lda = ldm; a2.B = s->dAlpha; a2.ldb = ldy; a2.C = tm.p; a2.ldc = ldy;
        a2.M = (int)m; a2.N = (int)M; a2.K = (int)m; a2.alpha = 1.0; a2.beta = 0.0;
        GP_TRY(gemm(c, GEMM_NN, a2));
        GP_TRY(coldot(c, s->dAlpha, ldy, tm.p, ldy, m, M, q2.p));
        GP_CUDA(c, cudaMemcpyAsync(h1.data(), q1.p, sizeof(double) * M, cudaMemcpyDeviceToHost, c->stream));
        GP_CUDA(c, cudaMemcpyAsync(h2.data(), q2.p, sizeof(double) * M, cudaMemcpyDeviceToHost, c->stream));
        GP_CUDA(c, cudaStreamSynchronize(c->stream));
        for (int64_t i = 0; i < M; ++i)
            jac[i * P + p] = h1[(size_t)i] - 0.5 * h2[(size_t)i] + t1[(size_t)p] + t2[(size_t)p];
    }
    return 0;
}

__global__ void sparse_cov_finish_kernel(const double* __restrict__ kxy, const double* __restrict__ qk,
                                         const double* __restrict__ qs, int64_t T, double* __restrict__ out) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < T) out[t] = kxy[t] - qk[t] + qs[t];  // k(x,y) - Kx^T K^-1 Ky + Kx^T Sigma Ky (:103-105)
}

}  // namespace

int kpair_values(Ctx* c, const double* Xa, const double* Xb, int64_t T, double* d_out);  // gram.cu

void sparse_invalidate(Ctx* c) {
    if (c->sparse) sparse_of(c)->learned = false;
}

void sparse_destroy(Ctx* c) {
    if (!c->sparse) return;
    SparseModel* s = sparse_of(c);
    sparse_free(s);
    delete s;
    c->sparse = nullptr;
}

}  // namespace gprb200

using namespace gprb200;

extern "C" {

int gprb200_sparse_set_inducing(gprb200_handle h, const double* Xm, int64_t m, int d, double jitter) {
    if (!h) return -2;
    Ctx* c = h;
    GP_CUDA(c, cudaSetDevice(c->device));
    if (!Xm || m <= 0 || d <= 0) return fail(c, -2, "SparseGaussianProcess: need m > 0 inducing samples of dimension d > 0");
    if (c->have_train && c->d != d) return fail(c, -2, "SparseGaussianProcess::AddInducingSample: dimension of input vector does not correspond to the input dimension");
    if (!c->sparse) c->sparse = new (std::nothrow) SparseModel();
    if (!c->sparse) return fail(c, -1, "out of host memory");
    SparseModel* s = static_cast<SparseModel*>(c->sparse);
    cudaStreamSynchronize(c->stream);
    sparse_free(s);
    s->m = m;
    s->d = d;
    s->ldm = round_up(m, 16);
    s->jitter = jitter;
    GP_TRY(sp_alloc(c, &s->dXm, (size_t)m * d));
    GP_CUDA(c, cudaMemcpyAsync(s->dXm, Xm, sizeof(double) * m * d, cudaMemcpyHostToDevice, c->stream));
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

int gprb200_sparse_set_jitter(gprb200_handle h, double jitter) {
    if (!h) return -2;
    if (!h->sparse) return fail(h, -2, "SparseGaussianProcess: no inducing samples");
    SparseModel* s = static_cast<SparseModel*>(h->sparse);
    s->jitter = jitter;
    s->learned = false;
    return 0;
}

int gprb200_sparse_learn(gprb200_handle h) {
    if (!h) return -2;
    GP_CUDA(h, cudaSetDevice(h->device));
    return sparse_learn_impl(h);
}

int gprb200_sparse_loglik(gprb200_handle h, int mode, double* value, double* grad, double* jac) {
    if (!h) return -2;
    Ctx* c = h;
    GP_CUDA(c, cudaSetDevice(c->device));
    return sparse_loglik_impl(c, mode, value, grad, jac);
}

int gprb200_sparse_dgram_cross(gprb200_handle h, double* D_out) {
    if (!h || !D_out) return -2;
    Ctx* c = h;
    GP_CUDA(c, cudaSetDevice(c->device));
    SparseModel* s = sparse_of(c);
    if (!s || !s->dXm) return fail(c, -2, "SparseGaussianProcess: no inducing samples");
    if (!c->have_train || !c->have_kernel) return fail(c, -2, "SparseGaussianProcess: need dense samples and a kernel");
    if (s->m > c->N)
        return fail(c, -2, "SparseGaussianProcess::ComputeDerivativeKernelVectorMatrix: number of dense samples must be higher "
                           "than the number of sparse samples");
    const int P = c->prog.nparams;
    Tmp D(c);
    GP_TRY(D.alloc((size_t)P * c->N * s->ldm));
    GP_TRY(dgram_build_xy(c, c->dX, c->N, s->dXm, s->m, D.p, s->ldm, -1));
    GP_CUDA(c, cudaMemcpy2DAsync(D_out, sizeof(double) * s->m, D.p, sizeof(double) * s->ldm, sizeof(double) * s->m,
                                 (size_t)P * c->N, cudaMemcpyDeviceToHost, c->stream));
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

int gprb200_sparse_get_alpha(gprb200_handle h, double* alpha) {
    if (!h || !alpha) return -2;
    Ctx* c = h;
    GP_CUDA(c, cudaSetDevice(c->device));
    GP_TRY(ensure_sparse(c));
    SparseModel* s = static_cast<SparseModel*>(c->sparse);
    GP_CUDA(c, cudaMemcpy2DAsync(alpha, sizeof(double) * c->M, s->dAlpha, sizeof(double) * c->ldy, sizeof(double) * c->M,
                                 (size_t)s->m, cudaMemcpyDeviceToHost, c->stream));
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

int gprb200_sparse_predict(gprb200_handle h, const double* Xs, int64_t T, double* mean, double* var) {
    if (!h || !Xs || T <= 0) return -2;
    Ctx* c = h;
    GP_CUDA(c, cudaSetDevice(c->device));
    GP_TRY(ensure_sparse(c));
    SparseModel* s = static_cast<SparseModel*>(c->sparse);
    const int64_t m = s->m, M = c->M, ldo = round_up(M, 16);
    Tmp xs(c), kx(c), out(c), q1(c), q2(c), kk(c), v(c);
    GP_TRY(xs.alloc((size_t)T * c->d));
    GP_CUDA(c, cudaMemcpyAsync(xs.p, Xs, sizeof(double) * T * c->d, cudaMemcpyHostToDevice, c->stream));
    if (mean) {
        GP_TRY(kx.alloc((size_t)T * s->ldm));
        GP_TRY(out.alloc((size_t)T * ldo));
        GP_TRY(gram_build(c, xs.p, T, s->dXm, m, kx.p, s->ldm, false, false, 0.0));
        GemmArgs g{};
        g.A = kx.p; g.lda = s->ldm; g.B = s->dAlpha; g.ldb = c->ldy; g.C = out.p; g.ldc = ldo;
        g.M = (int)T; g.N = (int)M; g.K = (int)m; g.alpha = 1.0; g.beta = 0.0;
        GP_TRY(gemm(c, GEMM_NN, g));
        GP_CUDA(c, cudaMemcpy2DAsync(mean, sizeof(double) * M, out.p, sizeof(double) * ldo, sizeof(double) * M, (size_t)T,
                                     cudaMemcpyDeviceToHost, c->stream));
    }
    if (var) {
        GP_TRY(q1.alloc((size_t)T)); GP_TRY(q2.alloc((size_t)T)); GP_TRY(kk.alloc((size_t)T)); GP_TRY(v.alloc((size_t)T));
        GP_TRY(sparse_quad(c, s->dLK, s->dInvLK, xs.p, xs.p, T, q1.p));
        GP_TRY(sparse_quad(c, s->dLS, s->dInvLS, xs.p, xs.p, T, q2.p));
        GP_TRY(kpair_values(c, xs.p, xs.p, T, kk.p));
        sparse_cov_finish_kernel<<<(unsigned)((T + 255) / 256), 256, 0, c->stream>>>(kk.p, q1.p, q2.p, T, v.p);
        ++c->launches;
        GP_CUDA(c, cudaMemcpyAsync(var, v.p, sizeof(double) * T, cudaMemcpyDeviceToHost, c->stream));
    }
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

int gprb200_sparse_cov(gprb200_handle h, const double* Xa, const double* Xb, int64_t T, double* cov) {
    if (!h || !Xa || !Xb || !cov || T <= 0) return -2;
    Ctx* c = h;
    GP_CUDA(c, cudaSetDevice(c->device));
    GP_TRY(ensure_sparse(c));
    SparseModel* s = static_cast<SparseModel*>(c->sparse);
    Tmp xa(c), xb(c), q1(c), q2(c), kk(c), v(c);
    GP_TRY(xa.alloc((size_t)T * c->d)); GP_TRY(xb.alloc((size_t)T * c->d));
    GP_TRY(q1.alloc((size_t)T)); GP_TRY(q2.alloc((size_t)T)); GP_TRY(kk.alloc((size_t)T)); GP_TRY(v.alloc((size_t)T));
    GP_CUDA(c, cudaMemcpyAsync(xa.p, Xa, sizeof(double) * T * c->d, cudaMemcpyHostToDevice, c->stream));
    GP_CUDA(c, cudaMemcpyAsync(xb.p, Xb, sizeof(double) * T * c->d, cudaMemcpyHostToDevice, c->stream));
    GP_TRY(sparse_quad(c, s->dLK, s->dInvLK, xa.p, xb.p, T, q1.p));
    GP_TRY(sparse_quad(c, s->dLS, s->dInvLS, xa.p, xb.p, T, q2.p));
    GP_TRY(kpair_values(c, xa.p, xb.p, T, kk.p));
    sparse_cov_finish_kernel<<<(unsigned)((T + 255) / 256), 256, 0, c->stream>>>(kk.p, q1.p, q2.p, T, v.p);
    ++c->launches;
    GP_CUDA(c, cudaMemcpyAsync(cov, v.p, sizeof(double) * T, cudaMemcpyDeviceToHost, c->stream));
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

}  // extern "C"
