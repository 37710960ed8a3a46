// On-disk matrix format of the reference (lib/MatrixIO.cpp:38-99): one ASCII header line
// "<rows> <cols>\n" followed by rows*cols raw scalars in ROW-MAJOR order. GaussianProcess::Save/Load
// model files stay byte-compatible with agiger/GPR. Explicitly instantiated for the row-major float
// and double matrix types in src/MatrixIO.cpp, like the reference (:121-124).
#ifndef GPRB200_CXX_MATRIX_IO_H
#define GPRB200_CXX_MATRIX_IO_H

#include <string>

namespace gpr {

template <typename _Matrix_Type_>
_Matrix_Type_ ReadMatrix(std::string filename);

template <typename _Matrix_Type_>
void WriteMatrix(const _Matrix_Type_& matrix, std::string filename);

bool MatrixIOTest();

}  // namespace gpr

#endif
