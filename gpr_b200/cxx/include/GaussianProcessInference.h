// gpr::GaussianProcessInference<T> — the reference's hyper-parameter optimiser (include/GaussianProcessInference.h:40-243),
// SURVEY.md section 8(f) row 2. Same constructor, Optimize / Optimize2 / GetParameters / SetParametersToOptimize and the
// same update rules, over the device-resident likelihood:
//
//   every iteration: kernel->SetParameters(theta)           P scalars cross the host/device boundary
//                    likelihood value + gradient/Jacobian    Gram, Cholesky, inverse and the trace contraction run
//                                                            on the GPU against the samples already resident there
//                                                            (GaussianProcess::DeviceLogLikelihood -> gprb200_loglik)
//                    P-dimensional update on the host        pinv of a P x P matrix (Prior.h)
//
// Update rules (reference :100-138 and :195-203):
//   Optimize   u = pinv(g g^T) g. Where u_p == 0: theta_p += sign * w^3 * slog(g_p); otherwise
//              theta_p -= sign * w * slog(u_p * L_0), with slog(v) = sgn(v) log(1 + |v|), sign = -1 if L_0 > 0 else +1.
//   Optimize2  per-output signs folded into L, u = pinv(J^T J) J^T L, theta_p -= w * slog(u_p) for the selected
//              parameters; stops early when the likelihood vector stops changing.
// With SetDeterminantMode(TrueLogDeterminant) on the GaussianProcess the optimiser sees the true log-likelihood
// instead of the reference's clamped determinant term (include/Likelihood.h:180-188), which makes N >= 2000 meaningful.
#pragma once

#include <cmath>
#include <iostream>
#include <memory>
#include <string>
#include <vector>

#include <Eigen/Dense>

#include "GaussianProcess.h"
#include "Kernel.h"
#include "Likelihood.h"
#include "Prior.h"

namespace gpr {

template <class TScalarType>
class GaussianProcessInference {
public:
    typedef GaussianProcessInference Self;
    typedef std::shared_ptr<Self> Pointer;

    typedef GaussianProcess<TScalarType> GaussianProcessType;
    typedef typename GaussianProcessType::Pointer GaussianProcessTypePointer;
    typedef typename GaussianProcessType::VectorListType VectorListType;
    typedef typename GaussianProcessType::VectorType VectorType;
    typedef typename GaussianProcessType::MatrixType MatrixType;
    typedef std::vector<TScalarType> ParameterVectorType;
    typedef std::vector<bool> BooleanVectorType;

    typedef Likelihood<TScalarType> LikelihoodType;
    typedef typename LikelihoodType::Pointer LikelihoodTypePointer;
    typedef typename LikelihoodType::ValueDerivativePair ValueDerivativePair;
    typedef typename LikelihoodType::ValueJacobianPair ValueJacobianPair;

    GaussianProcessInference(LikelihoodTypePointer lh, GaussianProcessTypePointer gp, double stepwidth, unsigned iterations)
        : m_StepWidth(stepwidth),
          m_StepWidth3(stepwidth * stepwidth * stepwidth),
          m_NumberOfIterations(iterations),
          m_Likelihood(lh),
          m_GaussianProcess(gp) {
        m_Parameters = m_GaussianProcess->GetKernel()->GetParameters();
        m_ParametersToOptimize.assign(m_Parameters.size(), true);
    }
    ~GaussianProcessInference() {}

    ParameterVectorType GetParameters() { return m_Parameters; }

    void SetParametersToOptimize(const BooleanVectorType& v) {
        for (unsigned i = 0; i < std::min(m_ParametersToOptimize.size(), v.size()); i++) m_ParametersToOptimize[i] = v[i];
    }

    // additive: likelihood value(s) seen at the start of each iteration of the last Optimize / Optimize2 call
    const std::vector<VectorType>& GetLikelihoodTrace() const { return m_Trace; }

    void Optimize(bool output = true, bool exp_output = false) {
        m_Trace.clear();
        const unsigned P = m_Parameters.size();
        for (unsigned it = 0; it < m_NumberOfIterations; it++) {
            try {
                m_GaussianProcess->GetKernel()->SetParameters(m_Parameters);
                ValueDerivativePair vd = m_Likelihood->GetValueAndParameterDerivatives(m_GaussianProcess);
                const VectorType& value = vd.first;
                const VectorType& grad = vd.second;
                m_Trace.push_back(value);
                const double sign = (value[0] > 0) ? -1 : 1;
                MatrixType outer = grad * grad.adjoint();
                VectorType update = pinv<MatrixType>(outer) * grad;
                if (output) {
                    std::cout << "Likelihood " << value << ", : ";
                    PrintParameters(exp_output);
                    std::cout << ", Gradients: " << grad.adjoint() << ", inf(J'J)J': " << update.adjoint() << ", update: " << std::flush;
                }
                for (unsigned p = 0; p < P; p++) {
                    double u;
                    if (update[p] == 0) {  // no curvature information: damped gradient step
                        u = m_StepWidth3 * SignedLog(grad[p]);
                        m_Parameters[p] += u * sign;
                    } else {  // Gauss-Newton step, damped
                        u = m_StepWidth * SignedLog(update[p] * value[0]);
                        m_Parameters[p] -= u * sign;
                    }
                    if (output) std::cout << u << " " << std::flush;
                }
                if (output) {
                    std::cout << ", new parameters: " << std::flush;
                    PrintParameters(exp_output);
                    std::cout << std::endl;
                }
            } catch (std::string& s) {
                std::cout << "[failed] " << s << std::endl;
                return;
            }
        }
    }

    void Optimize2(bool output = true, bool exp_output = false) {
        m_Trace.clear();
        VectorType previous;
        for (unsigned it = 0; it < m_NumberOfIterations; it++) {
            try {
                m_GaussianProcess->GetKernel()->SetParameters(m_Parameters);
                ValueJacobianPair vj = m_Likelihood->GetValueAndJacobian(m_GaussianProcess);
                const MatrixType& J = vj.second;
                VectorType value = vj.first;
                m_Trace.push_back(value);
                if (it > 0 && (previous - value).norm() == 0) break;
                for (unsigned l = 0; l < value.rows(); l++)
                    if (value[l] > 0) value[l] = -value[l];  // every output contributes -|L_l|
                MatrixType JtJ = J.adjoint() * J;
                MatrixType Jt = J.adjoint();
                VectorType update = (pinv<MatrixType>(JtJ) * Jt) * value;
                if (output) {
                    std::cout << "Likelihood " << vj.first << ", : ";
                    PrintParameters(exp_output);
                    std::cout << ", update: " << update.adjoint() << std::flush;
                }
                for (unsigned p = 0; p < m_Parameters.size(); p++) {
                    if (!m_ParametersToOptimize[p]) continue;
                    m_Parameters[p] -= m_StepWidth * SignedLog(update[p]);
                }
                if (output) {
                    std::cout << ", new parameters: " << std::flush;
                    PrintParameters(exp_output);
                    std::cout << std::endl;
                }
                previous = value;
            } catch (std::string& s) {
                std::cout << "[failed] " << s << std::endl;
                return;
            }
        }
    }

private:
    static double SignedLog(double v) { return v >= 0 ? std::log(1 + v) : -std::log(1 + std::fabs(v)); }
    void PrintParameters(bool exp_output) const {
        for (unsigned p = 0; p < m_Parameters.size(); p++)
            std::cout << (exp_output ? std::exp(m_Parameters[p]) : m_Parameters[p]) << ", " << std::flush;
    }

    TScalarType m_StepWidth;
    TScalarType m_StepWidth3;
    unsigned m_NumberOfIterations;
    BooleanVectorType m_ParametersToOptimize;

    LikelihoodTypePointer m_Likelihood;
    GaussianProcessTypePointer m_GaussianProcess;
    ParameterVectorType m_Parameters;
    std::vector<VectorType> m_Trace;

    GaussianProcessInference(const Self&);  // purposely not implemented
    void operator=(const Self&);            // purposely not implemented
};

}  // namespace gpr
