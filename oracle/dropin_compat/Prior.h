// TEST INFRASTRUCTURE ONLY. The reference's tests/GaussianLikelihoodTest.cpp includes "Prior.h" without using
// anything from it; the priors (include/Prior.h: scalar densities over Boost.Math) are outside the GP hot path
// and are not part of the product. This empty header only lets that test program compile unmodified.
#ifndef GPRB200_DROPIN_COMPAT_PRIOR_H
#define GPRB200_DROPIN_COMPAT_PRIOR_H
#endif
