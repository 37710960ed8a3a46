// The slice of the reference's include/Prior.h that the GP hot path and its callers use: the SVD pseudo-inverse
// `pinv` (include/Prior.h:38-57, used by GaussianProcessInference::Optimize/Optimize2) and `sgn` (:60-62).
// The prior density classes (Gaussian / LogGaussian / InverseGaussian / Gamma, include/Prior.h:66-753) are scalar
// statistics outside the data-parallel path (out of scope) and are not provided.
#pragma once

#include <cmath>
#include <limits>

#include <Eigen/Dense>
#include <Eigen/SVD>

namespace gpr {

// Moore-Penrose pseudo-inverse through the SVD; singular values <= epsilon are dropped (not inverted).
template <class TMatrixType>
TMatrixType pinv(const TMatrixType& m, double epsilon = std::numeric_limits<double>::epsilon()) {
    Eigen::JacobiSVD<TMatrixType> svd(m, Eigen::ComputeFullU | Eigen::ComputeFullV);
    typedef typename TMatrixType::Scalar Scalar;
    Eigen::Matrix<Scalar, Eigen::Dynamic, 1> inv = svd.singularValues();
    for (int i = 0; i < inv.rows(); ++i) inv(i) = (inv(i) <= epsilon) ? Scalar(0) : Scalar(1) / inv(i);
    return TMatrixType(svd.matrixV() * inv.asDiagonal() * svd.matrixU().transpose());
}

template <typename T>
int sgn(T val) {
    return (T(0) < val) - (val < T(0));
}

}  // namespace gpr
