// MatrixIO: "<rows> <cols>\n" + raw row-major scalars (format of reference lib/MatrixIO.cpp:38-99).
// Streams the body in bounded chunks straight from / into the matrix storage, so a 25 GB
// -RegressionVectors.txt (BASELINE config C2) needs no second host copy.
#include <cstdio>
#include <fstream>
#include <sstream>
#include <string>

#include <Eigen/Dense>

#include "MatrixIO.h"

namespace gpr {

namespace {
const size_t kChunkBytes = size_t(64) << 20;

template <class Scalar>
void StreamBytes(std::istream& in, Scalar* dst, size_t count) {
    char* p = reinterpret_cast<char*>(dst);
    size_t left = count * sizeof(Scalar);
    while (left > 0 && in) {
        const size_t n = left < kChunkBytes ? left : kChunkBytes;
        in.read(p, static_cast<std::streamsize>(n));
        p += n;
        left -= n;
    }
}
}  // namespace

template <typename _Matrix_Type_>
_Matrix_Type_ ReadMatrix(std::string filename) {
    typedef typename _Matrix_Type_::Scalar Scalar;
    std::ifstream in(filename.c_str(), std::ios::in | std::ios::binary);
    std::string header;
    unsigned long long rows = 0, cols = 0;
    if (std::getline(in, header)) {
        std::istringstream hs(header);
        if (!(hs >> rows && hs >> cols)) {
            std::ostringstream msg;
            msg << "ReadMatrix: header is corrupt (filename " << filename << ")." << std::endl;
            throw msg.str();
        }
    }
    _Matrix_Type_ m(rows, cols);
    if (rows * cols > 0) StreamBytes<Scalar>(in, m.data(), static_cast<size_t>(rows * cols));
    return m;
}

template <typename _Matrix_Type_>
void WriteMatrix(const _Matrix_Type_& matrix, std::string filename) {
    typedef typename _Matrix_Type_::Scalar Scalar;
    std::ofstream out(filename.c_str(), std::ios::out | std::ios::binary);
    const unsigned long long rows = matrix.rows(), cols = matrix.cols();
    std::ostringstream hs;
    hs << rows << " " << cols << std::endl;
    out.write(hs.str().data(), static_cast<std::streamsize>(hs.str().size()));
    const char* p = reinterpret_cast<const char*>(matrix.data());  // row-major storage == file order
    size_t left = static_cast<size_t>(rows * cols) * sizeof(Scalar);
    while (left > 0) {
        const size_t n = left < kChunkBytes ? left : kChunkBytes;
        out.write(p, static_cast<std::streamsize>(n));
        p += n;
        left -= n;
    }
}

bool MatrixIOTest() {
    typedef Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor> M;
    M a(10, 3);
    for (int i = 0; i < 10; ++i)
        for (int j = 0; j < 3; ++j) a(i, j) = 0.25 * i - 1.5 * j + 1.0 / (1 + i + j);
    const std::string fn = "/tmp/gprb200_matrixio_selftest.txt";
    WriteMatrix<M>(a, fn);
    M b = ReadMatrix<M>(fn);
    std::remove(fn.c_str());
    if (a.rows() != b.rows() || a.cols() != b.cols()) return false;
    for (int i = 0; i < 10; ++i)
        for (int j = 0; j < 3; ++j)
            if (a(i, j) != b(i, j)) return false;
    return true;
}

typedef Eigen::Matrix<float, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor> MatrixTypeF;
typedef Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor> MatrixTypeD;
template MatrixTypeF ReadMatrix<MatrixTypeF>(std::string);
template MatrixTypeD ReadMatrix<MatrixTypeD>(std::string);
template void WriteMatrix<MatrixTypeF>(const MatrixTypeF&, std::string);
template void WriteMatrix<MatrixTypeD>(const MatrixTypeD&, std::string);

}  // namespace gpr
