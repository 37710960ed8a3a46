// gpr::SparseLikelihood<T> / gpr::SparseGaussianLogLikelihood<T> of the reference (include/SparseLikelihood.h:34-551)
// over the B200 library: same class names, public methods and error strings.
//
//   operator()                         value per output column                        reference :152-217
//   GetParameterDerivatives            gradient, single output only (throws else)               :219-285
//   GetValueAndParameterDerivatives                                                             :287-409
//   GetValueAndJacobian                                                                         :411-535
//
// The reference builds C_inv = (sigma^2 I + Knm K^-1 Kmn)^-1 (n x n) through EfficientInversion, P blocks grad(C)_p
// (n x n each) and takes products and traces of them on the host: O(P n^3) work, O(P n^2) memory. Here ONE device call
// (gprb200_sparse_loglik, gpr_b200/csrc/sparse.cu) evaluates the same numbers from two m x m Cholesky factors, one
// n x m x m GEMM and two fused derivative-Gram contractions; nothing n x n exists. The public helpers the reference
// exposes "for testing" (GetCoreMatrices, GetCoreMatrix, GetKernelMatrixTrace, EfficientInversion,
// EfficientDeterminant) keep their meaning on host matrices.
// GaussianProcess::SetDeterminantMode(TrueLogDeterminant) replaces the long-double-range clamp of the complexity term
// (:177-185, :351-359) by the true log determinant.
#pragma once

#include <cmath>
#include <limits>
#include <memory>
#include <string>
#include <utility>

#include "GaussianProcess.h"
#include "Likelihood.h"
#include "SparseGaussianProcess.h"

namespace gpr {

template <class TScalarType>
class SparseLikelihood : public Likelihood<TScalarType> {
public:
    typedef SparseLikelihood Self;
    typedef std::shared_ptr<Self> Pointer;
    typedef Likelihood<TScalarType> Superclass;
    typedef typename Superclass::ValueDerivativePair ValueDerivativePair;
    typedef typename Superclass::ValueJacobianPair ValueJacobianPair;

    typedef SparseGaussianProcess<TScalarType> SparseGaussianProcessType;
    typedef typename SparseGaussianProcessType::Pointer SparseGaussianProcessTypePointer;
    typedef typename SparseGaussianProcessType::VectorType VectorType;
    typedef typename SparseGaussianProcessType::MatrixType MatrixType;
    typedef typename SparseGaussianProcessType::DiagMatrixType DiagMatrixType;
    typedef GaussianProcess<TScalarType> GaussianProcessType;
    typedef typename GaussianProcess<TScalarType>::Pointer GaussianProcessTypePointer;

    virtual std::string ToString() const = 0;
    virtual void DebugOn() { debug = true; }

    SparseLikelihood() : debug(false) {}
    virtual ~SparseLikelihood() {}

    // public in the reference "for the use in testing" (:62-84)
    virtual void GetCoreMatrices(const SparseGaussianProcessTypePointer gp, MatrixType& K, MatrixType& K_inv, MatrixType& Kmn,
                                 DiagMatrixType& I_sigma) const {
        gp->ComputeCoreMatrices(K, K_inv, Kmn, I_sigma);
    }
    virtual void GetCoreMatrix(const SparseGaussianProcessTypePointer gp, MatrixType& C, MatrixType& K_inv, MatrixType& Kmn) const {
        gp->ComputeCoreMatrix(C, K_inv, Kmn);
    }
    virtual TScalarType GetKernelMatrixTrace(const SparseGaussianProcessTypePointer gp) const {
        return gp->ComputeKernelMatrixTrace();
    }
    virtual VectorType GetDerivativeKernelMatrixTrace(const SparseGaussianProcessTypePointer gp) const {
        return gp->ComputeDerivativeKernelMatrixTrace();
    }

protected:
    virtual void GetLabelMatrix(const SparseGaussianProcessTypePointer gp, MatrixType& Y) const { gp->ComputeDenseLabelMatrix(Y); }
    virtual void GetDerivativeKernelMatrix(const SparseGaussianProcessTypePointer gp, MatrixType& D) const {
        gp->ComputeDerivativeKernelMatrix(D);
    }
    virtual void GetDerivativeKernelVectorMatrix(const SparseGaussianProcessTypePointer gp, MatrixType& D) const {
        gp->ComputeDerivativeKernelVectorMatrix(D);
    }
    virtual MatrixType GetInverseMatrix(const SparseGaussianProcessTypePointer gp, MatrixType& K) const {
        return gp->InvertKernelMatrix(K, gp->GetInversionMethod());
    }
    // the fused device evaluation (null = not wanted)
    void EvaluateSparseOnDevice(const SparseGaussianProcessTypePointer gp, VectorType* value, VectorType* grad,
                                MatrixType* jacobian) const {
        gp->DeviceSparseLogLikelihood(value, grad, jacobian);
    }

    bool debug;

private:
    SparseLikelihood(const Self&);
    void operator=(const Self&);
};

template <class TScalarType>
class SparseGaussianLogLikelihood : public SparseLikelihood<TScalarType> {
public:
    typedef SparseLikelihood<TScalarType> Superclass;
    typedef SparseGaussianLogLikelihood Self;
    typedef std::shared_ptr<Self> Pointer;
    typedef typename Superclass::VectorType VectorType;
    typedef typename Superclass::MatrixType MatrixType;
    typedef typename Superclass::DiagMatrixType DiagMatrixType;
    typedef typename Superclass::SparseGaussianProcessType SparseGaussianProcessType;
    typedef typename Superclass::SparseGaussianProcessTypePointer SparseGaussianProcessTypePointer;
    typedef typename Superclass::GaussianProcessTypePointer GaussianProcessTypePointer;
    typedef typename Superclass::ValueDerivativePair ValueDerivativePair;
    typedef typename Superclass::ValueJacobianPair ValueJacobianPair;
    typedef long double HighPrecisionType;

    // Woodbury form of (A + X B X')^-1 with A diagonal (:129-138); host matrices, small problems / tests only
    virtual inline void EfficientInversion(const SparseGaussianProcessTypePointer gp, MatrixType& D, const DiagMatrixType& A,
                                           const MatrixType& B, const MatrixType& B_inv, const MatrixType& X) const {
        (void)B;
        const VectorType a = A.diagonal();
        MatrixType XtAi = X.adjoint();                       // X' A^-1: scale column i by 1 / a_i
        for (long i = 0; i < (long)XtAi.cols(); ++i)
            for (long r = 0; r < (long)XtAi.rows(); ++r) XtAi(r, i) /= a[i];
        MatrixType inner = B_inv + XtAi * X;
        inner = this->GetInverseMatrix(gp, inner);
        D = XtAi.adjoint() * inner * XtAi;
        D = -D;
        for (long i = 0; i < (long)D.rows(); ++i) D(i, i) += TScalarType(1) / a[i];
    }

    // |A + X B X'| = |B| |A| |B^-1 + X' A^-1 X| in long double (:140-150)
    virtual inline HighPrecisionType EfficientDeterminant(const DiagMatrixType& A, const MatrixType& B, const MatrixType& B_inv,
                                                          const MatrixType& X) const {
        const VectorType a = A.diagonal();
        HighPrecisionType det_B = B.template cast<HighPrecisionType>().determinant();
        if (std::isinf(det_B)) det_B = std::numeric_limits<HighPrecisionType>::max();
        HighPrecisionType det_A = 1;
        MatrixType XtAi = X.adjoint();
        for (long i = 0; i < (long)XtAi.cols(); ++i) {
            det_A *= (HighPrecisionType)a[i];
            for (long r = 0; r < (long)XtAi.rows(); ++r) XtAi(r, i) /= a[i];
        }
        const MatrixType inner = B_inv + XtAi * X;
        return det_B * det_A * inner.template cast<HighPrecisionType>().determinant();
    }

    virtual inline VectorType operator()(const GaussianProcessTypePointer gp) const {
        SparseGaussianProcessTypePointer sgp = Cast(gp);
        VectorType value;
        this->EvaluateSparseOnDevice(sgp, &value, nullptr, nullptr);
        return value;
    }

    virtual inline VectorType GetParameterDerivatives(const GaussianProcessTypePointer gp) const {
        SparseGaussianProcessTypePointer sgp = Cast(gp);
        VectorType grad;
        this->EvaluateSparseOnDevice(sgp, nullptr, &grad, nullptr);
        return grad;
    }

    virtual inline ValueDerivativePair GetValueAndParameterDerivatives(const GaussianProcessTypePointer gp) const {
        SparseGaussianProcessTypePointer sgp = Cast(gp);
        VectorType value, grad;
        this->EvaluateSparseOnDevice(sgp, &value, &grad, nullptr);
        return std::make_pair(value, grad);
    }

    // Row i is the gradient of the single-output problem on label column i. (The reference's own last line, :530, adds a
    // P-vector to the columns of an M x P matrix and is only defined for M == P == 1.)
    virtual inline ValueJacobianPair GetValueAndJacobian(const GaussianProcessTypePointer gp) const {
        SparseGaussianProcessTypePointer sgp = Cast(gp);
        VectorType value;
        MatrixType jacobian;
        this->EvaluateSparseOnDevice(sgp, &value, nullptr, &jacobian);
        return std::make_pair(value, jacobian);
    }

    SparseGaussianLogLikelihood() : Superclass() {}
    virtual ~SparseGaussianLogLikelihood() {}
    virtual std::string ToString() const { return "SparseGaussianLogLikelihood"; }

private:
    SparseGaussianProcessTypePointer Cast(const GaussianProcessTypePointer gp) const {
        SparseGaussianProcessTypePointer sgp(std::dynamic_pointer_cast<SparseGaussianProcessType>(gp));
        if (sgp.get() == NULL) throw std::string("SparseGaussianLogLikelihood: cannot cast to SparseGaussianProcess");
        if (sgp->GetNumberOfInducingSamples() == 0)
            throw std::string("SparseLikelihood::GetValueAndParameterDerivative: there are no inducing samples specified");
        return sgp;
    }
    SparseGaussianLogLikelihood(const Self&);
    void operator=(const Self&);
};

}  // namespace gpr
