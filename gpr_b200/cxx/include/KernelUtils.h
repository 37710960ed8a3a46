// GetGeneralKernel: the 13-parameter composite kernel of reference include/KernelUtils.h:45-89,
//   ((G(p1,p0) + G(p3,p2) * P(p4,p5,p6)) + RQ(p7,p8,p9)) + (G(p11,p10) + W(p12))
// Pinned by tests/HighlyGeneralKernelTest.cpp:73-150 (nested string round trip).
#ifndef GPRB200_CXX_KERNEL_UTILS_H
#define GPRB200_CXX_KERNEL_UTILS_H

#include <memory>
#include <string>
#include <vector>

#include "Kernel.h"

namespace gpr {

template <class TScalarType>
std::shared_ptr<Kernel<TScalarType> > GetGeneralKernel(const std::vector<TScalarType>& params) {
    typedef std::shared_ptr<Kernel<TScalarType> > K;
    if (params.size() != 13) throw std::string("Wrong number of arguments.");
    const std::vector<TScalarType>& p = params;
    K smooth(new GaussianKernel<TScalarType>(p[1], p[0]));
    K envelope(new GaussianKernel<TScalarType>(p[3], p[2]));
    K wave(new PeriodicKernel<TScalarType>(p[4], p[5], p[6]));
    K decaying_wave(new ProductKernel<TScalarType>(envelope, wave));
    K multiscale(new RationalQuadraticKernel<TScalarType>(p[7], p[8], p[9]));
    K local(new GaussianKernel<TScalarType>(p[11], p[10]));
    K nugget(new WhiteKernel<TScalarType>(p[12]));
    K noise(new SumKernel<TScalarType>(local, nugget));
    K a(new SumKernel<TScalarType>(smooth, decaying_wave));
    K b(new SumKernel<TScalarType>(a, multiscale));
    return K(new SumKernel<TScalarType>(b, noise));
}

}  // namespace gpr

#endif
