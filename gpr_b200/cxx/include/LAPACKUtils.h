// lapack::lu_invert / lapack::chol_invert over the device library (reference include/LAPACKUtils.h:76-111).
// The reference binds dgetrf/dgetri and dpotrf/dpotri; here both run on the B200 through the C-ABI:
//   lu_invert   -> gprb200_lu_inverse   (Gauss-Jordan with partial pivoting, general matrices; csrc/lu.cu)
//   chol_invert -> gprb200_spd_inverse  (the blocked Cholesky of the GP path + TRTRI/LAUUM; csrc/potrf.cu, solve.cu)
// Same names, template parameters, return types and exception type, so tests/InversionMethodsTest.cpp compiles
// unchanged. As in the reference, float instantiations cast to double around the inversion (:87, :96).
// do not replace ifndef with pragma once
#ifndef LAPACK_UTILS_H
#define LAPACK_UTILS_H

#include <cstdlib>
#include <exception>
#include <string>
#include "GaussianProcess.h"
#include "gprb200.h"

namespace gpr {
namespace lapack {

struct LAPACKException : public std::exception {
    std::string s;
    LAPACKException(std::string ss) : s(ss) {}
    ~LAPACKException() throw() {}
    const char* what() const throw() { return s.c_str(); }
};

namespace detail {
// one device handle per process for the free functions (no GaussianProcess object to borrow one from)
inline gprb200_handle Handle() {
    static gprb200_handle h = [] {
        gprb200_handle p = nullptr;
        const char* env = std::getenv("GPRB200_DEVICE");
        if (gprb200_create(&p, env ? std::atoi(env) : 0) != 0)
            throw LAPACKException(std::string("lapack: ") + gprb200_last_error(nullptr));
        return p;
    }();
    return h;
}
}  // namespace detail

// wrapper around the LU inversion (general square matrix)
template <class T>
typename GaussianProcess<T>::MatrixType lu_invert(const typename GaussianProcess<T>::MatrixType& matrix) {
    GaussianProcess<double>::MatrixType inv_matrix(matrix.template cast<double>());
    if (inv_matrix.rows() != inv_matrix.cols()) throw LAPACKException("lu_invert: matrix is not square.");
    GaussianProcess<double>::MatrixType out(inv_matrix.rows(), inv_matrix.cols());
    const int info = gprb200_lu_inverse(detail::Handle(), inv_matrix.data(), inv_matrix.cols(), out.data());
    if (info != 0) throw LAPACKException("lu_invert: error in inversion. Matrix is singular or contains an illegal value.");
    return out.template cast<T>();
}

// wrapper around the Cholesky inversion (symmetric positive definite matrix; never use float as T)
// The reference ignores dpotrf's INFO (:59-73) and returns garbage for a matrix that is not positive definite; here that
// case throws, like lu_invert does for a singular one.
template <class T>
typename GaussianProcess<T>::MatrixType chol_invert(const typename GaussianProcess<T>::MatrixType& matrix) {
    GaussianProcess<double>::MatrixType inv_matrix(matrix.template cast<double>());
    if (inv_matrix.rows() != inv_matrix.cols()) throw LAPACKException("chol_invert: matrix is not square.");
    GaussianProcess<double>::MatrixType out(inv_matrix.rows(), inv_matrix.cols());
    const int info = gprb200_spd_inverse(detail::Handle(), inv_matrix.data(), inv_matrix.cols(), out.data(), nullptr);
    if (info != 0) throw LAPACKException(std::string("chol_invert: ") + gprb200_last_error(detail::Handle()));
    return out.template cast<T>();
}

}  // namespace lapack
}  // namespace gpr

#endif
