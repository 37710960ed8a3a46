// Small data-movement / reduction kernels shared by the drivers (all HBM-bound, coalesced).
#include "common.cuh"

namespace gprb200 {

int fail(Ctx* c, int code, const std::string& msg) {
    if (c) c->err = msg;
    return code;
}

__global__ void copy2d_kernel(double* dst, int64_t ldd, const double* src, int64_t lds, int64_t rows, int64_t cols) {
    const int64_t total = rows * cols;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / cols, c = e - r * cols;
        dst[r * ldd + c] = src[r * lds + c];
    }
}
__global__ void fill2d_kernel(double* dst, int64_t ldd, int64_t rows, int64_t cols, double v, int identity) {
    const int64_t total = rows * cols;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / cols, c = e - r * cols;
        dst[r * ldd + c] = (identity && r == c) ? 1.0 : v;
    }
}
// 32x32 tiles below the diagonal are transposed through shared memory into the upper triangle
__global__ void symmetrize_kernel(double* A, int64_t ld, int64_t n, int zero_only) {
    __shared__ double t[32][33];
    const int64_t bi = blockIdx.y, bj = blockIdx.x;
    if (bj > bi) return;
    const int64_t r0 = bi * 32, c0 = bj * 32;
    for (int y = threadIdx.y; y < 32; y += blockDim.y) {
        const int64_t r = r0 + y, c = c0 + threadIdx.x;
        t[y][threadIdx.x] = (r < n && c < n) ? A[r * ld + c] : 0.0;
    }
    __syncthreads();
    for (int y = threadIdx.y; y < 32; y += blockDim.y) {
        // destination (c0 + y, r0 + x) = source (r0 + x, c0 + y)
        const int64_t dr = c0 + y, dc = r0 + threadIdx.x;
        if (dr < n && dc < n && dc > dr) A[dr * ld + dc] = zero_only ? 0.0 : t[threadIdx.x][y];
    }
}
// out[c] = sum_r A(r,c) B(r,c): one thread per column, coalesced across columns; fixed order.
__global__ void coldot_kernel(const double* A, int64_t lda, const double* B, int64_t ldb, int64_t rows, int64_t cols,
                              double* out) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= cols) return;
    double s = 0.0;
    for (int64_t r = 0; r < rows; ++r) s = fma(A[r * lda + c], B[r * ldb + c], s);
    out[c] = s;
}
// Few columns, many rows (the quadratic form y^T alpha at M = 1: one thread would walk all N rows): rows split into
// chunks over blockIdx.y and over the 8 row lanes of a 32 x 8 block; partial sums [chunk][col] are then added in chunk
// order by coldot_finish_kernel, so the result does not depend on scheduling.
constexpr int CD_ROWS_PER_CHUNK = 512;
__global__ void coldot_part_kernel(const double* A, int64_t lda, const double* B, int64_t ldb, int64_t rows, int64_t cols,
                                   double* part) {
    __shared__ double red[8][33];
    const int64_t c = (int64_t)blockIdx.x * 32 + threadIdx.x;
    const int64_t r0 = (int64_t)blockIdx.y * CD_ROWS_PER_CHUNK;
    const int64_t r1 = (r0 + CD_ROWS_PER_CHUNK < rows) ? r0 + CD_ROWS_PER_CHUNK : rows;
    double s0 = 0.0, s1 = 0.0;
    if (c < cols) {
        int64_t r = r0 + threadIdx.y;
        for (; r + 8 < r1; r += 16) {
            s0 = fma(A[r * lda + c], B[r * ldb + c], s0);
            s1 = fma(A[(r + 8) * lda + c], B[(r + 8) * ldb + c], s1);
        }
        if (r < r1) s0 = fma(A[r * lda + c], B[r * ldb + c], s0);
    }
    red[threadIdx.y][threadIdx.x] = s0 + s1;
    __syncthreads();
    if (threadIdx.y == 0 && c < cols) {
        double s = 0.0;
#pragma unroll
        for (int y = 0; y < 8; ++y) s += red[y][threadIdx.x];
        part[(int64_t)blockIdx.y * cols + c] = s;
    }
}
__global__ void coldot_finish_kernel(const double* part, int64_t nchunks, int64_t cols, double* out) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= cols) return;
    double s = 0.0;
    for (int64_t k = 0; k < nchunks; ++k) s += part[k * cols + c];
    out[c] = s;
}
// out[r] = sum_c A(r,c) B(r,c): one warp per row
__global__ void rowdot_kernel(const double* A, int64_t lda, const double* B, int64_t ldb, int64_t rows, int64_t cols,
                              double* out) {
    const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= rows) return;
    double s = 0.0;
    for (int64_t c = threadIdx.x & 31; c < cols; c += 32) s = fma(A[r * lda + c], B[r * ldb + c], s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) out[r] = s;
}

static unsigned grid_for(int64_t total, int threads) {
    int64_t b = (total + threads - 1) / threads;
    const int64_t cap = 148 * 16;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (unsigned)b;
}

int copy2d(Ctx* c, double* dst, int64_t ldd, const double* src, int64_t lds, int64_t rows, int64_t cols) {
    if (rows <= 0 || cols <= 0) return 0;
    copy2d_kernel<<<grid_for(rows * cols, 256), 256, 0, c->stream>>>(dst, ldd, src, lds, rows, cols);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}
__global__ void axpby2d_kernel(double* dst, int64_t ldd, double a, const double* A, int64_t lda, double b, const double* B,
                               int64_t ldb, int64_t rows, int64_t cols) {
    const int64_t total = rows * cols;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / cols, cc = e - r * cols;
        dst[r * ldd + cc] = a * A[r * lda + cc] + b * B[r * ldb + cc];
    }
}
int axpby2d(Ctx* c, double* dst, int64_t ldd, double a, const double* A, int64_t lda, double b, const double* B,
            int64_t ldb, int64_t rows, int64_t cols) {
    if (rows <= 0 || cols <= 0) return 0;
    axpby2d_kernel<<<grid_for(rows * cols, 256), 256, 0, c->stream>>>(dst, ldd, a, A, lda, b, B, ldb, rows, cols);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}
int fill2d(Ctx* c, double* dst, int64_t ldd, int64_t rows, int64_t cols, double v) {
    if (rows <= 0 || cols <= 0) return 0;
    fill2d_kernel<<<grid_for(rows * cols, 256), 256, 0, c->stream>>>(dst, ldd, rows, cols, v, 0);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}
__global__ void add_diag_kernel(double* A, int64_t ld, int64_t n, double v) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) A[i * ld + i] += v;
}
int add_diag(Ctx* c, double* A, int64_t ld, int64_t n, double v) {
    if (n <= 0) return 0;
    add_diag_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(A, ld, n, v);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}

int set_identity(Ctx* c, double* dst, int64_t ldd, int64_t n) {
    if (n <= 0) return 0;
    fill2d_kernel<<<grid_for(n * n, 256), 256, 0, c->stream>>>(dst, ldd, n, n, 0.0, 1);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}
static int sym_launch(Ctx* c, double* A, int64_t ld, int64_t n, int zero_only) {
    if (n <= 0) return 0;
    const unsigned t = (unsigned)((n + 31) / 32);
    symmetrize_kernel<<<dim3(t, t), dim3(32, 8), 0, c->stream>>>(A, ld, n, zero_only);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}
int symmetrize_from_lower(Ctx* c, double* A, int64_t ld, int64_t n) { return sym_launch(c, A, ld, n, 0); }
int zero_upper(Ctx* c, double* A, int64_t ld, int64_t n) { return sym_launch(c, A, ld, n, 1); }

int coldot(Ctx* c, const double* A, int64_t lda, const double* B, int64_t ldb, int64_t rows, int64_t cols,
           double* d_out) {
    if (cols <= 0) return 0;
    const int64_t nchunks = (rows + CD_ROWS_PER_CHUNK - 1) / CD_ROWS_PER_CHUNK;
    if (rows >= 4 * CD_ROWS_PER_CHUNK && cols < 148 * 128 && nchunks <= 65535) {
        // not enough columns to fill the GPU with one thread per column: split the rows as well
        double* part = nullptr;
        GP_CUDA(c, cudaMallocAsync((void**)&part, sizeof(double) * nchunks * cols, c->stream));
        dim3 grid((unsigned)((cols + 31) / 32), (unsigned)nchunks), block(32, 8);
        coldot_part_kernel<<<grid, block, 0, c->stream>>>(A, lda, B, ldb, rows, cols, part);
        coldot_finish_kernel<<<(unsigned)((cols + 127) / 128), 128, 0, c->stream>>>(part, nchunks, cols, d_out);
        c->launches += 2;
        GP_CUDA(c, cudaGetLastError());
        GP_CUDA(c, cudaFreeAsync(part, c->stream));
        return 0;
    }
    coldot_kernel<<<(unsigned)((cols + 127) / 128), 128, 0, c->stream>>>(A, lda, B, ldb, rows, cols, d_out);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}
int rowdot(Ctx* c, const double* A, int64_t lda, const double* B, int64_t ldb, int64_t rows, int64_t cols,
           double* d_out) {
    if (rows <= 0) return 0;
    rowdot_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, c->stream>>>(A, lda, B, ldb, rows, cols, d_out);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}

}  // namespace gprb200
