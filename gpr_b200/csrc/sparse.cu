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
// evaluated here through the two Cholesky factors L_K and L_S. The n x n "core matrix" Knm K^-1 Kmn (:360-378) is only
// used by the sparse likelihood and is not formed.
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

int sparse_learn_impl(Ctx* c) {
    SparseModel* s = sparse_of(c);
    if (!s || !s->dXm) return fail(c, -2, "SparseGaussianProcess::Initialize: no inducing samples defined during initialization");
    if (!c->have_train || !c->dY) return fail(c, -2, "SparseGaussianProcess::Initialize: no dense samples defined during initialization");
    if (!c->have_kernel) return fail(c, -2, "SparseGaussianProcess: no kernel");
    if (!(c->sigma > 0.0)) return fail(c, -2, "SparseGaussianProcess::ComputeCoreMatrices: sigma must be positive.");
    const int64_t n = c->N, m = s->m, M = c->M;
    if (m > n)
        return fail(c, -2, "SparseGaussianProcess::ComputeKernelVectorMatrix: number of dense samples must be higher than the "
                           "number of sparse samples");
    s->learned = false;
    const int64_t ldm = s->ldm, nblk = (m + NB - 1) / NB;
    if (!s->dLK) {
        GP_TRY(sp_alloc(c, &s->dLK, (size_t)m * ldm));
        GP_TRY(sp_alloc(c, &s->dLS, (size_t)m * ldm));
        GP_TRY(sp_alloc(c, &s->dInvLK, (size_t)nblk * NB * NB));
        GP_TRY(sp_alloc(c, &s->dInvLS, (size_t)nblk * NB * NB));
        GP_TRY(sp_alloc(c, &s->dAlpha, (size_t)m * c->ldy));
    }
    // K = k(Xm, Xm) + jitter I (ComputeKernelMatrixWithJitter, :175-180); S starts as a copy of it
    GP_TRY(gram_build(c, s->dXm, m, s->dXm, m, s->dLK, ldm, true, false, s->jitter));
    GP_TRY(copy2d(c, s->dLS, ldm, s->dLK, ldm, m, m));
    Tmp knm(c);
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
