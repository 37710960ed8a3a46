// Likelihood classes of the reference (include/Likelihood.h) over the B200 library.
//
//   GaussianLogLikelihood::operator()                       reference :166-202
//   ::GetParameterDerivatives / ::GetValueAndParameterDerivatives     :204-285
//   ::GetValueAndJacobian                                             :287-344
//   GaussianLikelihood::operator()                                    :106-139
//
// The reference materialises C = inv(K + sigma^2 I), a long-double determinant and the whole
// (P N) x N derivative stack on the host and takes P full N^3 products for the traces. Here one device
// call re-factorises (Cholesky), takes log det from the factor's diagonal and contracts
// (alpha alpha^T - C) with dK/dtheta_p tile by tile without storing dK. The reference's clamped
// complexity term is reproduced by default (GaussianProcess::SetDeterminantMode selects the true value).
// The protected accessors of Likelihood<T> keep their reference meaning (materialised matrices) for
// subclasses that use them.
#ifndef GPRB200_CXX_LIKELIHOOD_H
#define GPRB200_CXX_LIKELIHOOD_H

#include <cmath>
#include <limits>
#include <memory>
#include <sstream>
#include <string>
#include <utility>

#include "GaussianProcess.h"

namespace gpr {

template <class TScalarType>
class Likelihood {
public:
    typedef Likelihood Self;
    typedef std::shared_ptr<Self> Pointer;
    typedef GaussianProcess<TScalarType> GaussianProcessType;
    typedef std::shared_ptr<GaussianProcessType> GaussianProcessTypePointer;
    typedef typename GaussianProcessType::VectorType VectorType;
    typedef typename GaussianProcessType::MatrixType MatrixType;
    typedef typename GaussianProcessType::DiagMatrixType DiagMatrixType;
    typedef typename std::pair<VectorType, VectorType> ValueDerivativePair;
    typedef typename std::pair<VectorType, MatrixType> ValueJacobianPair;
    typedef long double HighPrecisionType;

    virtual VectorType operator()(const GaussianProcessTypePointer) const {
        throw std::string("Likelihood: operator() is not implemented.");
    }
    virtual VectorType GetParameterDerivatives(const GaussianProcessTypePointer) const {
        throw std::string("Likelihood: GetParameterDerivatives is not implemented.");
    }
    virtual ValueDerivativePair GetValueAndParameterDerivatives(const GaussianProcessTypePointer) const {
        throw std::string("Likelihood: GetValueAndParameterDerivatives is not implemented.");
    }
    virtual ValueJacobianPair GetValueAndJacobian(const GaussianProcessTypePointer) const {
        throw std::string("Likelihood: GetValueAndJacobian is not implemented.");
    }
    virtual std::string ToString() const = 0;

    Likelihood() {}
    virtual ~Likelihood() {}

protected:
    // friendship with GaussianProcess (reference :73-87)
    virtual void GetLabelMatrix(const GaussianProcessTypePointer gp, MatrixType& Y) const { gp->ComputeLabelMatrix(Y); }
    virtual TScalarType GetCoreMatrix(const GaussianProcessTypePointer gp, MatrixType& C) const {
        return gp->ComputeCoreMatrixWithDeterminant(C);  // long double -> TScalarType, as the reference returns it
    }
    virtual void GetDerivativeKernelMatrix(const GaussianProcessTypePointer gp, MatrixType& D) const {
        gp->ComputeDerivativeKernelMatrix(D);
    }
    virtual TScalarType GetSigma(const GaussianProcessTypePointer gp) const { return gp->GetSigma(); }

    // the fused device evaluation (value / gradient / Jacobian; null = not wanted)
    void EvaluateOnDevice(const GaussianProcessTypePointer gp, VectorType* value, VectorType* grad,
                          MatrixType* jacobian) const {
        gp->DeviceLogLikelihood(value, grad, jacobian);
    }

private:
    Likelihood(const Self&);
    void operator=(const Self&);
};

template <class TScalarType>
class GaussianLikelihood : public Likelihood<TScalarType> {
public:
    typedef Likelihood<TScalarType> Superclass;
    typedef GaussianLikelihood Self;
    typedef std::shared_ptr<Self> SelfPointer;
    typedef typename Superclass::VectorType VectorType;
    typedef typename Superclass::MatrixType MatrixType;
    typedef typename Superclass::GaussianProcessTypePointer GaussianProcessTypePointer;
    typedef typename Superclass::HighPrecisionType HighPrecisionType;

    // exp(-0.5 y^T C y) * det^-1/2 * (2 pi)^-N/2, from the log-likelihood pieces of the device call
    virtual VectorType operator()(const GaussianProcessTypePointer gp) const {
        typename GaussianProcess<TScalarType>::DeterminantMode keep = gp->GetDeterminantMode();
        gp->SetDeterminantMode(GaussianProcess<TScalarType>::TrueLogDeterminant);
        VectorType logv;
        try {
            this->EvaluateOnDevice(gp, &logv, 0, 0);
        } catch (...) {
            gp->SetDeterminantMode(keep);
            throw;
        }
        gp->SetDeterminantMode(keep);
        // logv = df - 0.5 logdet - N/2 log(2 pi); the reference multiplies exp(df), det^-1/2 and the
        // constant separately (with det already truncated to TScalarType), which is the same number
        // whenever none of the three factors leaves the scalar range.
        VectorType out = logv;
        for (unsigned i = 0; i < static_cast<unsigned>(out.rows()); ++i) out[i] = std::exp(logv[i]);
        return out;
    }

    GaussianLikelihood() : Superclass() {}
    virtual ~GaussianLikelihood() {}
    virtual std::string ToString() const { return "GaussianLikelihood"; }

private:
    GaussianLikelihood(const Self&);
    void operator=(const Self&);
};

template <class TScalarType>
class GaussianLogLikelihood : public Likelihood<TScalarType> {
public:
    typedef Likelihood<TScalarType> Superclass;
    typedef GaussianLogLikelihood Self;
    typedef std::shared_ptr<Self> SelfPointer;
    typedef typename Superclass::VectorType VectorType;
    typedef typename Superclass::MatrixType MatrixType;
    typedef typename Superclass::GaussianProcessTypePointer GaussianProcessTypePointer;
    typedef typename Superclass::ValueDerivativePair ValueDerivativePair;
    typedef typename Superclass::ValueJacobianPair ValueJacobianPair;
    typedef typename Superclass::HighPrecisionType HighPrecisionType;

    virtual VectorType operator()(const GaussianProcessTypePointer gp) const {
        VectorType value;
        this->EvaluateOnDevice(gp, &value, 0, 0);
        return value;
    }
    virtual VectorType GetParameterDerivatives(const GaussianProcessTypePointer gp) const {
        VectorType delta;
        this->EvaluateOnDevice(gp, 0, &delta, 0);
        return delta;
    }
    virtual ValueDerivativePair GetValueAndParameterDerivatives(const GaussianProcessTypePointer gp) const {
        VectorType value, delta;
        this->EvaluateOnDevice(gp, &value, &delta, 0);
        return std::make_pair(value, delta);
    }
    virtual ValueJacobianPair GetValueAndJacobian(const GaussianProcessTypePointer gp) const {
        VectorType value;
        MatrixType jacobian;
        this->EvaluateOnDevice(gp, &value, 0, &jacobian);
        return std::make_pair(value, jacobian);
    }

    GaussianLogLikelihood() : Superclass() {}
    virtual ~GaussianLogLikelihood() {}
    virtual std::string ToString() const { return "GaussianLogLikelihood"; }

private:
    GaussianLogLikelihood(const Self&);
    void operator=(const Self&);
};

}  // namespace gpr

#endif
