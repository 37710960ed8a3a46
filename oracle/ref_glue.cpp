// oracle/_ref: the reference's OWN GaussianProcess / Kernel / Likelihood code behind a C interface.
//
// TEST INFRASTRUCTURE ONLY. This translation unit #includes the reference's sources from where they
// lie under /root/reference (never copied into this repository):
//     lib/GaussianProcess.cpp  (which pulls include/GaussianProcess.h, Kernel.h, KernelFactory.h,
//                               Likelihood.h, LAPACKUtils.h, MatrixIO.h)
// compiled against third_party/eigen_lite (Eigen / boost::filesystem stand-ins) and linked with the same
// LAPACK routines the reference declares (include/LAPACKUtils.h:13-27), taken from scipy's OpenBLAS.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference use the result.
#include <cstring>
#include <memory>
#include <sstream>
#include <string>

#include "lib/GaussianProcess.cpp"  // -I/root/reference
#include "include/GaussianProcessInference.h"  // the optimiser loop; boost stand-ins in eigen_lite
#include "include/PCA.h"          // the reference's PCA, BDCSVD -> LAPACK dgesdd in eigen_lite
#include "include/SparseLikelihood.h"  // SparseGaussianLogLikelihood

// ---- LAPACK: the reference declares the un-prefixed Fortran symbols; scipy's OpenBLAS exports them
//      with a scipy_ prefix (LP64).
extern "C" {
void scipy_dgetrf_(int*, int*, double*, int*, int*, int*);
void scipy_dgetri_(int*, double*, int*, int*, double*, int*, int*);
void scipy_dpotrf_(char*, int*, double*, int*, int*);
void scipy_dpotri_(char*, int*, double*, int*, int*);
int scipy_ilaenv_(int*, char*, char*, int*, int*, int*, int*, int, int);
void scipy_openblas_set_num_threads(int);
int scipy_openblas_get_num_threads(void);

void dgetrf_(int* M, int* N, double* A, int* lda, int* IPIV, int* INFO) { scipy_dgetrf_(M, N, A, lda, IPIV, INFO); }
void dgetri_(int* N, double* A, int* lda, int* IPIV, double* WORK, int* lwork, int* INFO) {
    scipy_dgetri_(N, A, lda, IPIV, WORK, lwork, INFO);
}
void dpotrf_(char* UPLO, int* N, double* A, int* LDA, int* INFO) { scipy_dpotrf_(UPLO, N, A, LDA, INFO); }
void dpotri_(char* UPLO, int* N, double* A, int* LDA, int* INFO) { scipy_dpotri_(UPLO, N, A, LDA, INFO); }
int ilaenv_(int* ispec, char* name, char* opts, int* n1, int* n2, int* n3, int* n4) {
    return scipy_ilaenv_(ispec, name, opts, n1, n2, n3, n4, (int)std::strlen(name), (int)std::strlen(opts));
}
}

namespace {

typedef gpr::GaussianProcess<double> GP;
typedef GP::VectorType Vec;
typedef GP::MatrixType Mat;

// protected members of GaussianProcess<double> are reachable from a derived class
struct Access : GP {
    explicit Access(GP::KernelTypePointer k) : GP(k) {}
    using GP::ComputeKernelMatrix;
    using GP::ComputeDerivativeKernelMatrix;
    using GP::ComputeKernelVector;
    using GP::ComputeCoreMatrix;
    using GP::m_RegressionVectors;
    using GP::m_CoreMatrix;
    using GP::m_OutputDimension;
    using GP::m_InputDimension;
};

thread_local std::string g_err;

Vec to_vec(const double* p, long n) {
    Vec v(n);
    for (long i = 0; i < n; ++i) v[i] = p[i];
    return v;
}
void from_mat(const Eigen::MatBase<double>& m, double* out) { std::memcpy(out, m.data(), sizeof(double) * m.size()); }

template <class F>
int guard(F f) {
    try {
        f();
        return 0;
    } catch (std::string& s) {
        g_err = s;
    } catch (std::exception& e) {
        g_err = e.what();
    } catch (...) {
        g_err = "unknown exception";
    }
    return -1;
}

}  // namespace

extern "C" {

const char* ref_last_error() { return g_err.c_str(); }
void ref_set_blas_threads(int n) { scipy_openblas_set_num_threads(n); }
int ref_get_blas_threads() { return scipy_openblas_get_num_threads(); }

// ---- Kernel / KernelFactory ---------------------------------------------------------------------
int ref_kernel_eval(const char* ks, const double* x, const double* y, int d, double* out) {
    return guard([&] {
        std::string s(ks);
        auto k = gpr::KernelFactory<double>::GetKernel(s);
        *out = (*k)(to_vec(x, d), to_vec(y, d));
    });
}
int ref_kernel_deriv(const char* ks, const double* x, const double* y, int d, double* out, int* nparams) {
    return guard([&] {
        std::string s(ks);
        auto k = gpr::KernelFactory<double>::GetKernel(s);
        Vec D = k->GetDerivative(to_vec(x, d), to_vec(y, d));
        *nparams = (int)D.rows();
        for (long i = 0; i < D.rows(); ++i) out[i] = D[i];
    });
}
int ref_kernel_tostring(const char* ks, char* out, int cap, double* params, int* nparams) {
    return guard([&] {
        std::string s(ks);
        auto k = gpr::KernelFactory<double>::GetKernel(s);
        std::string t = k->ToString();
        std::strncpy(out, t.c_str(), cap - 1);
        out[cap - 1] = 0;
        auto p = k->GetParameters();
        *nparams = (int)p.size();
        for (size_t i = 0; i < p.size(); ++i) params[i] = p[i];
    });
}
int ref_general_kernel_tostring(const double* p13, char* out, int cap);

// ---- GaussianProcess ------------------------------------------------------------------------------
void* ref_gp_new(const char* ks, double sigma) {
    Access* a = nullptr;
    guard([&] {
        std::string s(ks);
        a = new Access(gpr::KernelFactory<double>::GetKernel(s));
        a->SetSigma(sigma);
    });
    return a;
}
void ref_gp_free(void* h) { delete static_cast<Access*>(h); }
int ref_gp_add_samples(void* h, const double* X, long N, int d, const double* Y, int M) {
    return guard([&] {
        for (long i = 0; i < N; ++i) static_cast<Access*>(h)->AddSample(to_vec(X + i * d, d), to_vec(Y + i * M, M));
    });
}
int ref_gp_set_sigma(void* h, double s) { return guard([&] { static_cast<Access*>(h)->SetSigma(s); }); }
int ref_gp_set_kernel(void* h, const char* ks) {
    return guard([&] {
        std::string s(ks);
        static_cast<Access*>(h)->SetKernel(gpr::KernelFactory<double>::GetKernel(s));
    });
}
int ref_gp_set_inversion(void* h, int m) {
    return guard([&] { static_cast<Access*>(h)->SetInversionMethod((GP::InversionMethod)m); });
}
int ref_gp_set_efficient_storage(void* h, int s) {
    return guard([&] { static_cast<Access*>(h)->SetEfficientStorage(s != 0); });
}
int ref_gp_initialize(void* h) { return guard([&] { static_cast<Access*>(h)->Initialize(); }); }
int ref_gp_dims(void* h, long* N, int* d, int* M, long* core_size) {
    return guard([&] {
        Access* a = static_cast<Access*>(h);
        *N = a->GetNumberOfSamples();
        *d = a->m_InputDimension;
        *M = a->m_OutputDimension;
        *core_size = a->m_CoreMatrix.diagonalSize();
    });
}
int ref_gp_predict(void* h, const double* x, int d, double* out) {
    return guard([&] { from_mat(static_cast<Access*>(h)->Predict(to_vec(x, d)), out); });
}
int ref_gp_predict_derivative(void* h, const double* x, int d, double* mean, double* D) {
    return guard([&] {
        Mat Dm;
        Vec m = static_cast<Access*>(h)->PredictDerivative(to_vec(x, d), Dm);
        from_mat(m, mean);
        from_mat(Dm, D);
    });
}
int ref_gp_cov(void* h, const double* x, const double* y, int d, double* out) {
    return guard([&] { *out = (*static_cast<Access*>(h))(to_vec(x, d), to_vec(y, d)); });
}
int ref_gp_ci(void* h, const double* x, int d, double* out) {
    return guard([&] { *out = static_cast<Access*>(h)->GetCredibleInterval(to_vec(x, d)); });
}
int ref_gp_kernel_matrix(void* h, double* K) {
    return guard([&] {
        Mat M;
        static_cast<Access*>(h)->ComputeKernelMatrix(M);
        from_mat(M, K);
    });
}
int ref_gp_dkernel_matrix(void* h, double* D) {
    return guard([&] {
        Mat M;
        static_cast<Access*>(h)->ComputeDerivativeKernelMatrix(M);
        from_mat(M, D);
    });
}
int ref_gp_kernel_vector(void* h, const double* x, int d, double* out) {
    return guard([&] {
        Vec k;
        static_cast<Access*>(h)->ComputeKernelVector(to_vec(x, d), k);
        from_mat(k, out);
    });
}
int ref_gp_alpha(void* h, double* out) {
    return guard([&] { from_mat(static_cast<Access*>(h)->m_RegressionVectors, out); });
}
int ref_gp_core(void* h, double* out) { return guard([&] { from_mat(static_cast<Access*>(h)->m_CoreMatrix, out); }); }
int ref_gp_save(void* h, const char* prefix) { return guard([&] { static_cast<Access*>(h)->Save(prefix); }); }
int ref_gp_load(void* h, const char* prefix) { return guard([&] { static_cast<Access*>(h)->Load(prefix); }); }
int ref_gp_equal(void* a, void* b) { return (*static_cast<Access*>(a) == *static_cast<Access*>(b)) ? 1 : 0; }
int ref_gp_kernel_string(void* h, char* out, int cap) {
    return guard([&] {
        std::string t = static_cast<Access*>(h)->GetKernel()->ToString();
        std::strncpy(out, t.c_str(), cap - 1);
        out[cap - 1] = 0;
    });
}

// ---- GaussianLogLikelihood (include/Likelihood.h) --------------------------------------------------
// the Likelihood API takes a shared_ptr; wrap without taking ownership
static std::shared_ptr<GP> borrow(void* h) { return std::shared_ptr<GP>(static_cast<Access*>(h), [](GP*) {}); }

int ref_loglik_value(void* h, double* value) {
    return guard([&] {
        gpr::GaussianLogLikelihood<double> lik;
        from_mat(lik(borrow(h)), value);
    });
}
int ref_loglik_grad(void* h, double* grad) {
    return guard([&] {
        gpr::GaussianLogLikelihood<double> lik;
        from_mat(lik.GetParameterDerivatives(borrow(h)), grad);
    });
}
int ref_loglik_value_grad(void* h, double* value, double* grad) {
    return guard([&] {
        gpr::GaussianLogLikelihood<double> lik;
        auto r = lik.GetValueAndParameterDerivatives(borrow(h));
        from_mat(r.first, value);
        from_mat(r.second, grad);
    });
}
int ref_loglik_value_jac(void* h, double* value, double* jac) {
    return guard([&] {
        gpr::GaussianLogLikelihood<double> lik;
        auto r = lik.GetValueAndJacobian(borrow(h));
        from_mat(r.first, value);
        from_mat(r.second, jac);
    });
}
int ref_lik_value(void* h, double* value) {
    return guard([&] {
        gpr::GaussianLikelihood<double> lik;
        from_mat(lik(borrow(h)), value);
    });
}

// ---- LAPACKUtils -----------------------------------------------------------------------------------
int ref_lu_invert(double* A, int N) {
    return guard([&] {
        Mat M(N, N);
        std::memcpy(M.data(), A, sizeof(double) * N * N);
        from_mat(gpr::lapack::lu_invert<double>(M), A);
    });
}
int ref_chol_invert(double* A, int N) {
    return guard([&] {
        Mat M(N, N);
        std::memcpy(M.data(), A, sizeof(double) * N * N);
        from_mat(gpr::lapack::chol_invert<double>(M), A);
    });
}


// ---- PCA (include/PCA.h) ----------------------------------------------------------------------------
// X is D x n row-major (one sample per column, as DataParser builds it). Outputs: mean (D), sigma (n),
// U (D x n row-major).
void* ref_pca_fit(const double* X, long D, long n, int nFeatures) {
    void* out = nullptr;
    guard([&] {
        Mat M(D, n);
        std::memcpy(M.data(), X, sizeof(double) * D * n);
        out = new PCA<double>(M, nFeatures);
    });
    return out;
}
void ref_pca_free(void* p) { delete static_cast<PCA<double>*>(p); }
int ref_pca_get(void* p, double* mean, double* sigma, double* U) {
    return guard([&] {
        PCA<double>* pca = static_cast<PCA<double>*>(p);
        if (mean) from_mat(pca->GetMean(), mean);
        if (sigma) from_mat(pca->GetEigenvalues(), sigma);
        if (U) from_mat(pca->GetMatrixU(), U);
    });
}
int ref_pca_write(void* p, const char* prefix) {
    return guard([&] { static_cast<PCA<double>*>(p)->WriteMatricesToFile(prefix); });
}
void* ref_pca_load(const char* prefix, int nFeatures) {
    void* out = nullptr;
    guard([&] { out = new PCA<double>(std::string(prefix), nFeatures); });
    return out;
}
// features = basis^T (X - mean): rows_out x ncols, rows_out = n (all modes), or nFeatures with fast != 0
int ref_pca_project(void* p, const double* X, long D, long ncols, int fast, double* out, long* rows_out) {
    return guard([&] {
        PCA<double>* pca = static_cast<PCA<double>*>(p);
        Mat M(D, ncols);
        std::memcpy(M.data(), X, sizeof(double) * D * ncols);
        Mat F;
        if (fast) {
            pca->PrecomputeTranspose();
            F = pca->DimensionalityReductionFast(M);
        } else {
            F = pca->DimensionalityReduction(M);
        }
        *rows_out = F.rows();
        from_mat(F, out);
    });
}
// X = basisInverse[:, :nFeatures] W + mean: W is nFeatures x ncols, out D x ncols
int ref_pca_reconstruct(void* p, const double* W, long nF, long ncols, double* out) {
    return guard([&] {
        PCA<double>* pca = static_cast<PCA<double>*>(p);
        Mat Wm(nF, ncols);
        std::memcpy(Wm.data(), W, sizeof(double) * nF * ncols);
        Mat R = pca->GetReconstruction(Wm);
        from_mat(R, out);
    });
}


// ---- GaussianProcessInference (include/GaussianProcessInference.h:84-229) -------------------------------------
// which = 1: Optimize, 2: Optimize2. mask (may be NULL): SetParametersToOptimize. params_out: the P parameters after
// `iterations` iterations (GetParameters); the GP's kernel is left at the last parameters the loop evaluated.
int ref_infer(void* h, int which, double step, int iterations, const int* mask, int nmask, double* params_out,
              int* nparams) {
    return guard([&] {
        typedef gpr::GaussianProcessInference<double> Inference;
        Inference::LikelihoodTypePointer lh(new gpr::GaussianLogLikelihood<double>());
        Inference gpi(lh, borrow(h), step, (unsigned)iterations);
        if (mask) {
            Inference::BooleanVectorType bv;
            for (int i = 0; i < nmask; ++i) bv.push_back(mask[i] != 0);
            gpi.SetParametersToOptimize(bv);
        }
        if (which == 2) gpi.Optimize2(false, false);
        else gpi.Optimize(false, false);
        Inference::ParameterVectorType p = gpi.GetParameters();
        *nparams = (int)p.size();
        for (size_t i = 0; i < p.size(); ++i) params_out[i] = p[i];
    });
}
// pinv of include/Prior.h:38-57 on an r x c row-major matrix -> c x r
int ref_pinv(const double* A, int r, int c, double* out) {
    return guard([&] {
        Mat M(r, c);
        std::memcpy(M.data(), A, sizeof(double) * r * c);
        Mat P = gpr::pinv<Mat>(M);
        from_mat(P, out);
    });
}


// kernel->SetParameters(p) on the GP's kernel, as the optimiser loop does every iteration (GaussianProcessInference.h:90)
int ref_gp_set_params(void* h, const double* p, int n) {
    return guard([&] {
        std::vector<double> v(p, p + n);
        static_cast<Access*>(h)->GetKernel()->SetParameters(v);
    });
}


// ---- SparseGaussianProcess (include/SparseGaussianProcess.h), regression part -----------------------------------
void* ref_sgp_new(const char* ks, double sigma, double jitter) {
    gpr::SparseGaussianProcess<double>* g = nullptr;
    guard([&] {
        std::string s(ks);
        g = new gpr::SparseGaussianProcess<double>(gpr::KernelFactory<double>::GetKernel(s), jitter);
        g->SetSigma(sigma);
    });
    return g;
}
void ref_sgp_free(void* h) { delete static_cast<gpr::SparseGaussianProcess<double>*>(h); }
int ref_sgp_add(void* h, const double* X, long N, int d, const double* Y, int M, int inducing) {
    return guard([&] {
        auto* g = static_cast<gpr::SparseGaussianProcess<double>*>(h);
        for (long i = 0; i < N; ++i) {
            if (inducing) g->AddInducingSample(to_vec(X + i * d, d), to_vec(Y + i * M, M));
            else g->AddSample(to_vec(X + i * d, d), to_vec(Y + i * M, M));
        }
    });
}
int ref_sgp_predict(void* h, const double* x, int d, double* out) {
    return guard([&] { from_mat(static_cast<gpr::SparseGaussianProcess<double>*>(h)->Predict(to_vec(x, d)), out); });
}
int ref_sgp_cov(void* h, const double* x, const double* y, int d, double* out) {
    return guard([&] { *out = (*static_cast<gpr::SparseGaussianProcess<double>*>(h))(to_vec(x, d), to_vec(y, d)); });
}

// ---- SparseGaussianLogLikelihood (include/SparseLikelihood.h:111-551) -------------------------------------------
// what = 0: operator() -> value (M); 1: GetParameterDerivatives -> grad (P); 2: GetValueAndParameterDerivatives ->
// value, grad; 3: GetValueAndJacobian -> value, jac (M x P row-major)
int ref_sgp_loglik(void* h, int what, double* value, double* grad, double* jac) {
    return guard([&] {
        auto* g = static_cast<gpr::SparseGaussianProcess<double>*>(h);
        std::shared_ptr<GP> gp(static_cast<GP*>(g), [](GP*) {});
        gpr::SparseGaussianLogLikelihood<double> lik;
        if (what == 0) {
            from_mat(lik(gp), value);
        } else if (what == 1) {
            from_mat(lik.GetParameterDerivatives(gp), grad);
        } else if (what == 2) {
            auto vd = lik.GetValueAndParameterDerivatives(gp);
            from_mat(vd.first, value);
            from_mat(vd.second, grad);
        } else {
            auto vj = lik.GetValueAndJacobian(gp);
            from_mat(vj.first, value);
            from_mat(vj.second, jac);
        }
    });
}
int ref_sgp_set_params(void* h, const double* p, int n) {
    return guard([&] {
        std::vector<double> v(p, p + n);
        static_cast<gpr::SparseGaussianProcess<double>*>(h)->GetKernel()->SetParameters(v);
    });
}

}  // extern "C"
