// GetLocalPeriodLength: rough period estimate of a 1-D signal from its amplitude spectrum, the helper of
// reference include/LikelihoodUtils.h:33-75 used to seed the PeriodicKernel's period hyper-parameter.
// Host-only scalar work (not on the Learn/Predict path); kept so code including this header compiles.
// Returns (samples per period, dominant amplitude, "sinus likeness" = integral / (integral - max) - 1).
#ifndef GPRB200_CXX_LIKELIHOOD_UTILS_H
#define GPRB200_CXX_LIKELIHOOD_UTILS_H

#include <cmath>
#include <complex>
#include <iostream>
#include <limits>
#include <string>
#include <tuple>
#include <vector>

#include "GaussianProcess.h"

namespace gpr {

template <class TScalarType>
std::tuple<TScalarType, TScalarType, TScalarType> GetLocalPeriodLength(
    const typename GaussianProcess<TScalarType>::VectorType& vec, unsigned ommit = 1) {
    if (vec.rows() < 4 + ommit)
        throw std::string("GetLocalPeriodLength: longer signal required. Check if a column vector is provided!");
    const unsigned n = vec.rows();
    // single-precision spectrum, like the reference's Eigen::FFT<float>
    unsigned best = 0;
    float amp_best = std::numeric_limits<float>::lowest(), amp_sum = 0;
    for (unsigned k = ommit; k < n / 2; ++k) {
        std::complex<double> acc(0.0, 0.0);
        for (unsigned t = 0; t < n; ++t) {
            const double a = -2.0 * M_PI * (double)((unsigned long long)k * t % n) / (double)n;
            acc += (double)(float)vec[t] * std::complex<double>(std::cos(a), std::sin(a));
        }
        const float amp = 2 * (float)std::abs(acc) / n;
        if (amp > amp_best) {
            amp_best = amp;
            best = k;
        }
        amp_sum += amp;
    }
    std::cout << "amp max: " << amp_best << ", amp integral: " << amp_sum << std::endl;
    const double period = static_cast<TScalarType>(n) / static_cast<TScalarType>(best);
    double likeness;
    if (amp_sum - amp_best < std::numeric_limits<TScalarType>::min())
        likeness = std::numeric_limits<TScalarType>::max();
    else
        likeness = amp_sum / (amp_sum - amp_best) - 1;
    return std::make_tuple((TScalarType)period, (TScalarType)amp_best, (TScalarType)likeness);
}

}  // namespace gpr

#endif
