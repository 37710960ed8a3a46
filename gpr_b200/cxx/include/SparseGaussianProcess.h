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
// Not provided: the protected matrix helpers that only the sparse likelihood uses (ComputeCoreMatrices,
// ComputeDerivativeKernelVectorMatrix, ... :316-420) and SparseLikelihood.h itself.
#pragma once

#include <string>
#include <vector>

#include "GaussianProcess.h"
#include "gprb200.h"

namespace gpr {

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
    virtual void PreComputeRegression() {
        gprb200_handle h = Handle();  // kernel, dense samples and sigma in sync; nothing dense is factored
        if (m_InducingDirty) {
            const size_t m = m_InducingSampleVectors.size(), d = this->m_InputDimension;
            std::vector<double> xm(m * d);
            for (size_t i = 0; i < m; ++i)
                for (size_t k = 0; k < d; ++k) xm[i * d + k] = (double)m_InducingSampleVectors[i][k];
            Check(gprb200_sparse_set_inducing(h, xm.data(), (int64_t)m, (int)d, (double)m_Jitter));
            m_InducingDirty = false;
        }
        Check(gprb200_sparse_learn(h));
    }

    TScalarType m_Jitter;
    bool m_Initialized;
    VectorListType m_InducingSampleVectors;
    VectorListType m_InducingLabelVectors;

private:
    gprb200_handle Handle() const { return (gprb200_handle)this->SyncedDeviceHandle(); }
    void Check(int rc) const {
        if (rc != 0) throw std::string(gprb200_last_error((gprb200_handle)this->SyncedDeviceHandle()));
    }
    bool m_InducingDirty;

    SparseGaussianProcess(const Self&);
    void operator=(const Self&);
};

}  // namespace gpr
