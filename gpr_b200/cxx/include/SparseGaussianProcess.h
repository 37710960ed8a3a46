// gpr::SparseGaussianProcess<T> — the regression part of the reference's inducing-point GP
// (include/SparseGaussianProcess.h:30-420) over the B200 library (gpr_b200/csrc/sparse.cu): same constructors and
// public methods (AddInducingSample, ClearInducingSamples, Predict, operator(), Get/SetJitter, Initialize,
// GetNumberOfInducingSamples, ComputeDenseKernelMatrix). The dense samples are those of the GaussianProcess base
// class (AddSample); m inducing samples, m <= n.
//
//   Initialize -> PreComputeRegression   K = k(Xm,Xm) + jitter I, Knm, S = K + sigma^-2 Kmn Knm (DMMA GEMM of depth n),
//                                        alpha = sigma^-2 S^-1 Kmn Y through two Cholesky factors  gprb200_sparse_learn
//   Predict                              Kx^T alpha                                                gprb200_sparse_predict
//   operator()(x, y)                     k(x,y) - Kx^T K^-1 Ky + Kx^T S^-1 Ky                      gprb200_sparse_cov
//
//   SparseGaussianLogLikelihood          value / gradient / Jacobian (SparseLikelihood.h)           gprb200_sparse_loglik
//
// The protected matrix helpers the reference's sparse likelihood is built from (ComputeCoreMatrices,
// ComputeKernelVectorMatrix, ComputeDerivativeKernelVectorMatrix, ComputeCoreMatrix, ... :155-397) keep their names,
// signatures and results (materialised host matrices, generated on the device) for code that calls them; the
// likelihood itself does not go through them.
#pragma once

#include <limits>
#include <string>
#include <vector>

#include "GaussianProcess.h"
#include "gprb200.h"

namespace gpr {

template <class TScalarType>
class SparseLikelihood;

template <class TScalarType>
class SparseGaussianProcess : public GaussianProcess<TScalarType> {
public:
    typedef SparseGaussianProcess Self;
    typedef std::shared_ptr<Self> Pointer;

    typedef GaussianProcess<TScalarType> Superclass;
    typedef typename Superclass::VectorType VectorType;
    typedef typename Superclass::MatrixType MatrixType;
    typedef typename Superclass::DiagMatrixType DiagMatrixType;
    typedef typename Superclass::VectorListType VectorListType;
    typedef typename Superclass::MatrixListType MatrixListType;
    typedef typename Superclass::KernelType KernelType;
    typedef typename Superclass::KernelTypePointer KernelTypePointer;

    SparseGaussianProcess(KernelTypePointer kernel) : Superclass(kernel), m_Jitter(0), m_Initialized(false), m_InducingDirty(true) {}
    SparseGaussianProcess(KernelTypePointer kernel, TScalarType jitter)
        : Superclass(kernel), m_Jitter(jitter), m_Initialized(false), m_InducingDirty(true) {}
    virtual ~SparseGaussianProcess() {}

    void AddInducingSample(const VectorType& x, const VectorType& y) {
        if (m_InducingSampleVectors.size() == 0) this->m_InputDimension = x.size();
        if (m_InducingLabelVectors.size() == 0) this->m_OutputDimension = y.size();
        this->CheckInputDimension(x, "SparseGaussianProcess::AddInducingSample: ");
        this->CheckOutputDimension(y, "SparseGaussianProcess::AddInducingSample: ");
        m_InducingSampleVectors.push_back(x);
        m_InducingLabelVectors.push_back(y);
        m_Initialized = false;
        m_InducingDirty = true;
    }

    void ClearInducingSamples() {
        m_InducingSampleVectors.clear();
        m_InducingLabelVectors.clear();
        m_Initialized = false;
        m_InducingDirty = true;
    }

    VectorType Predict(const VectorType& x) {
        Initialize();
        this->CheckInputDimension(x, "GaussianProcess::Predict: ");
        std::vector<double> xs(x.size()), out(this->m_OutputDimension);
        for (long i = 0; i < (long)x.size(); ++i) xs[(size_t)i] = (double)x[i];
        Check(gprb200_sparse_predict(Handle(), xs.data(), 1, out.data(), nullptr));
        VectorType y(this->m_OutputDimension);
        for (unsigned k = 0; k < this->m_OutputDimension; ++k) y[k] = (TScalarType)out[k];
        return y;
    }

    TScalarType operator()(const VectorType& x, const VectorType& y) {
        Initialize();
        this->CheckInputDimension(x, "SparseGaussianProcess::(): ");
        this->CheckInputDimension(y, "SparseGaussianProcess::(): ");
        std::vector<double> xa(x.size()), xb(y.size());
        for (long i = 0; i < (long)x.size(); ++i) {
            xa[(size_t)i] = (double)x[i];
            xb[(size_t)i] = (double)y[i];
        }
        double cov = 0;
        Check(gprb200_sparse_cov(Handle(), xa.data(), xb.data(), 1, &cov));
        return (TScalarType)cov;
    }

    unsigned GetNumberOfInducingSamples() const { return m_InducingSampleVectors.size(); }
    TScalarType GetJitter() const { return m_Jitter; }
    void SetJitter(TScalarType jitter) {
        m_Jitter = jitter;
        m_Initialized = false;
        m_InducingDirty = true;
    }

    virtual void Initialize() {
        if (m_Initialized) return;
        if (!(m_InducingSampleVectors.size() > 0))
            throw std::string("SparseGaussianProcess::Initialize: no inducing samples defined during initialization");
        if (!(m_InducingLabelVectors.size() > 0))
            throw std::string("SparseGaussianProcess::Initialize: no inducing labels defined during initialization");
        if (!(this->m_SampleVectors.size() > 0))
            throw std::string("SparseGaussianProcess::Initialize: no dense samples defined during initialization");
        if (!(this->m_LabelVectors.size() > 0))
            throw std::string("SparseGaussianProcess::Initialize: no dense labels defined during initialization");
        PreComputeRegression();
        m_Initialized = true;
    }

    virtual void ComputeDenseKernelMatrix(MatrixType& M) const { Superclass::ComputeKernelMatrixInternal(M, this->m_SampleVectors); }

protected:
    // ---- matrix helpers of the reference (:155-397), host matrices generated on the device -------------------------
    virtual void ComputeKernelMatrix(MatrixType& M) const { Superclass::ComputeKernelMatrixInternal(M, m_InducingSampleVectors); }
    virtual void ComputeKernelMatrixWithJitter(MatrixType& M) const {
        ComputeKernelMatrix(M);
        for (unsigned i = 0; i < (unsigned)M.rows(); ++i) M(i, i) += m_Jitter;
    }
    virtual void ComputeLabelMatrix(MatrixType& Y) const { Superclass::ComputeLabelMatrixInternal(Y, m_InducingLabelVectors); }
    virtual void ComputeDenseLabelMatrix(MatrixType& Y) const { Superclass::ComputeLabelMatrix(Y); }
    virtual void ComputeKernelVector(const VectorType& x, VectorType& Kx) const {
        Superclass::ComputeKernelVectorInternal(x, Kx, m_InducingSampleVectors);
    }
    // Knm = k(x_i, z_j), n x m (:218-235)
    virtual void ComputeKernelVectorMatrix(MatrixType& Knm) const {
        const size_t n = this->m_SampleVectors.size(), m = m_InducingSampleVectors.size(), d = this->m_InputDimension;
        if (!(m <= n))
            throw std::string("SparseGaussianProcess::ComputeKernelVectorMatrix: number of dense samples must be higher than the "
                              "number of sparse samples");
        std::vector<double> xm(m * d), kmn(m * n);
        for (size_t i = 0; i < m; ++i)
            for (size_t k = 0; k < d; ++k) xm[i * d + k] = (double)m_InducingSampleVectors[i][k];
        Check(gprb200_kernel_vector(Handle(), xm.data(), (int64_t)m, kmn.data()));   // m x n block of kernel values
        Knm.resize(n, m);
        for (size_t i = 0; i < n; ++i)
            for (size_t j = 0; j < m; ++j) Knm(i, j) = (TScalarType)kmn[j * n + i];
    }
    // stack of d Knm / d theta_p, (P n) x m (:237-264)
    virtual void ComputeDerivativeKernelVectorMatrix(MatrixType& M) const {
        const size_t n = this->m_SampleVectors.size(), m = m_InducingSampleVectors.size();
        const size_t P = this->m_Kernel->GetNumberOfParameters();
        if (!(m <= n))
            throw std::string("SparseGaussianProcess::ComputeDerivativeKernelVectorMatrix: number of dense samples must be higher "
                              "than the number of sparse samples");
        gprb200_handle h = InducingHandle();
        std::vector<double> D(P * n * m);
        Check(gprb200_sparse_dgram_cross(h, D.data()));
        M.resize(n * P, m);
        for (size_t e = 0; e < D.size(); ++e) M.data()[e] = (TScalarType)D[e];
    }
    // K + jitter I, its inverse, Knm and sigma^2 I (:323-350)
    virtual void ComputeCoreMatrices(MatrixType& K, MatrixType& K_inv, MatrixType& Kmn, DiagMatrixType& I_sigma) {
        const bool stable = m_Jitter < std::numeric_limits<TScalarType>::min();
        ComputeKernelMatrixWithJitter(K);
        K_inv = this->InvertKernelMatrix(K, this->m_InvMethod, stable);
        ComputeKernelVectorMatrix(Kmn);
        if (this->GetSigma() <= 0) throw std::string("SparseGaussianProcess::ComputeCoreMatrices: sigma must be positive.");
        if (Kmn.rows() == 0) throw std::string("SparseGaussianProcess::ComputeCoreMatrices: empty sample set.");
        I_sigma.resize(Kmn.rows());
        I_sigma.setIdentity();
        I_sigma = (I_sigma.diagonal().array() * this->GetSigmaSquared()).matrix().asDiagonal();
    }
    // the n x n core matrix Knm K^-1 Kmn (:355-378)
    virtual void ComputeCoreMatrix(MatrixType& C, MatrixType& K_inv) const {
        const bool stable = m_Jitter < std::numeric_limits<TScalarType>::min();
        MatrixType K, Knm;
        ComputeKernelMatrixWithJitter(K);
        ComputeKernelVectorMatrix(Knm);
        K_inv = this->InvertKernelMatrix(K, this->m_InvMethod, stable);
        ComputeCoreMatrix(C, K_inv, Knm);
    }
    virtual void ComputeCoreMatrix(MatrixType& C) const {
        MatrixType K_inv;
        ComputeCoreMatrix(C, K_inv);
    }
    virtual void ComputeCoreMatrix(MatrixType& C, const MatrixType& K_inv, const MatrixType& Knm) const {
        C = Knm * K_inv * Knm.adjoint();
    }
    virtual void ComputeDerivativeKernelMatrix(MatrixType& M) const {
        Superclass::ComputeDerivativeKernelMatrixInternal(M, m_InducingSampleVectors);
    }

    // SparseGaussianLogLikelihood on the device (gprb200_sparse_loglik); any of value (M), grad (P), jacobian (M x P)
    // may be null. Every matrix is rebuilt from the current kernel parameters, sigma and jitter.
    void DeviceSparseLogLikelihood(VectorType* value, VectorType* grad, MatrixType* jacobian) const {
        gprb200_handle h = InducingHandle();
        const size_t M = this->m_OutputDimension, P = this->m_Kernel->GetNumberOfParameters();
        std::vector<double> v(M), g(P), j(M * P);
        const int mode = (this->m_DeterminantMode == Superclass::TrueLogDeterminant) ? GPRB200_LOGLIK_TRUE
                                                                                     : GPRB200_LOGLIK_REFERENCE_CLAMP;
        Check(gprb200_sparse_loglik(h, mode, value ? v.data() : nullptr, grad ? g.data() : nullptr,
                                    jacobian ? j.data() : nullptr));
        if (value) {
            value->resize(M);
            for (size_t i = 0; i < M; ++i) (*value)[i] = (TScalarType)v[i];
        }
        if (grad) {
            grad->resize(P);
            for (size_t i = 0; i < P; ++i) (*grad)[i] = (TScalarType)g[i];
        }
        if (jacobian) {
            jacobian->resize(M, P);
            for (size_t i = 0; i < M; ++i)
                for (size_t p = 0; p < P; ++p) (*jacobian)(i, p) = (TScalarType)j[i * P + p];
        }
        // the likelihood left the regression of the CURRENT parameters on the device, which is what Predict wants
    }

    virtual void PreComputeRegression() {
        Check(gprb200_sparse_learn(InducingHandle()));
    }

    TScalarType m_Jitter;
    bool m_Initialized;
    VectorListType m_InducingSampleVectors;
    VectorListType m_InducingLabelVectors;

private:
    gprb200_handle Handle() const { return (gprb200_handle)this->SyncedDeviceHandle(); }
    // kernel, dense samples and sigma in sync (nothing dense is factored) plus the inducing points and jitter
    gprb200_handle InducingHandle() const {
        gprb200_handle h = Handle();
        if (m_InducingDirty) {
            const size_t m = m_InducingSampleVectors.size(), d = this->m_InputDimension;
            if (m == 0) throw std::string("SparseGaussianProcess: no inducing samples");
            std::vector<double> xm(m * d);
            for (size_t i = 0; i < m; ++i)
                for (size_t k = 0; k < d; ++k) xm[i * d + k] = (double)m_InducingSampleVectors[i][k];
            Check(gprb200_sparse_set_inducing(h, xm.data(), (int64_t)m, (int)d, (double)m_Jitter));
            m_InducingDirty = false;
        }
        return h;
    }
    void Check(int rc) const {
        if (rc != 0) throw std::string(gprb200_last_error((gprb200_handle)this->SyncedDeviceHandle()));
    }
    mutable bool m_InducingDirty;

    SparseGaussianProcess(const Self&);
    void operator=(const Self&);

    friend class SparseLikelihood<TScalarType>;
};

}  // namespace gpr

#include "SparseLikelihood.h"
