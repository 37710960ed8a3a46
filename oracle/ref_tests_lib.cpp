// Library half of the reference's test build: lib/GaussianProcess.cpp compiled once (it carries the
// explicit instantiations the tests link against). TEST INFRASTRUCTURE ONLY.
#include "lib/GaussianProcess.cpp"
extern "C" {
void scipy_dgetrf_(int*, int*, double*, int*, int*, int*);
void scipy_dgetri_(int*, double*, int*, int*, double*, int*, int*);
void scipy_dpotrf_(char*, int*, double*, int*, int*);
void scipy_dpotri_(char*, int*, double*, int*, int*);
int scipy_ilaenv_(int*, char*, char*, int*, int*, int*, int*, int, int);
void dgetrf_(int* M, int* N, double* A, int* lda, int* IPIV, int* INFO) { scipy_dgetrf_(M, N, A, lda, IPIV, INFO); }
void dgetri_(int* N, double* A, int* lda, int* IPIV, double* WORK, int* lwork, int* INFO) {
    scipy_dgetri_(N, A, lda, IPIV, WORK, lwork, INFO);
}
void dpotrf_(char* UPLO, int* N, double* A, int* LDA, int* INFO) { scipy_dpotrf_(UPLO, N, A, LDA, INFO); }
void dpotri_(char* UPLO, int* N, double* A, int* LDA, int* INFO) { scipy_dpotri_(UPLO, N, A, LDA, INFO); }
int ilaenv_(int* ispec, char* name, char* opts, int* n1, int* n2, int* n3, int* n4) {
    return scipy_ilaenv_(ispec, name, opts, n1, n2, n3, n4, 6, 1);
}
}
