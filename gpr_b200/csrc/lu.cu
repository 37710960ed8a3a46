// General (non-symmetric) matrix inverse on the device: what lapack::lu_invert (reference include/LAPACKUtils.h:38-56,
// :85-97: dgetrf + dgetri) returns, for the public LAPACKUtils interface and tests/InversionMethodsTest.cpp Test5.
// NOT on the GP hot path: every GaussianProcess inversion method goes through the Cholesky factorisation (potrf.cu).
// Algorithm: in-place Gauss-Jordan elimination with partial (row) pivoting - the same pivot choice as dgetrf - two
// launches per column, then the recorded row exchanges are undone as column exchanges in reverse order. 2 n^3 flop on
// the FP64 FMA pipe, O(n) launches: a utility sized for the reference's use (a few hundred to a few thousand rows).
#include "common.cuh"
#include <string>

namespace gprb200 {

constexpr int LU_THREADS = 1024;

// step k, one CTA: pivot search in column k (rows >= k), row exchange, column k saved to colk, pivot row scaled.
__global__ void __launch_bounds__(LU_THREADS) gj_pivot_kernel(double* A, int64_t ld, int n, int k, int* piv, double* colk,
                                                              int* info) {
    __shared__ double sval[LU_THREADS / 32];
    __shared__ int sidx[LU_THREADS / 32];
    __shared__ int sp;
    if (*info != 0) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double best = -1.0;
    int bi = n;
    for (int i = k + tid; i < n; i += LU_THREADS) {
        const double v = fabs(A[(int64_t)i * ld + k]);
        if (v > best || (v == best && i < bi)) {   // NaN never wins: a column of NaNs reports singular
            best = v;
            bi = i;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) {
            best = ov;
            bi = oi;
        }
    }
    if (lane == 0) {
        sval[warp] = best;
        sidx[warp] = bi;
    }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < LU_THREADS / 32; ++w)
            if (sval[w] > best || (sval[w] == best && sidx[w] < bi)) {
                best = sval[w];
                bi = sidx[w];
            }
        if (!(best > 0.0) || !isfinite(best)) {   // exactly singular (or not finite): dgetrf's INFO = k + 1
            *info = k + 1;
            bi = -1;
        } else {
            piv[k] = bi;
        }
        sp = bi;
    }
    __syncthreads();
    const int p = sp;
    if (p < 0) return;
    double* rk = A + (int64_t)k * ld;
    if (p != k) {
        double* rp = A + (int64_t)p * ld;
        for (int j = tid; j < n; j += LU_THREADS) {
            const double t = rk[j];
            rk[j] = rp[j];
            rp[j] = t;
        }
    }
    __syncthreads();
    const double d = rk[k];
    __syncthreads();
    // column k leaves the matrix (it becomes column k of the inverse being built in place)
    for (int i = tid; i < n; i += LU_THREADS) {
        colk[i] = (i == k) ? 0.0 : A[(int64_t)i * ld + k];
        A[(int64_t)i * ld + k] = (i == k) ? 1.0 : 0.0;
    }
    __syncthreads();
    for (int j = tid; j < n; j += LU_THREADS) rk[j] = rk[j] / d;
}

// row i -= colk[i] * row k for every i != k (colk[k] = 0)
__global__ void gj_eliminate_kernel(double* A, int64_t ld, int n, int k, const double* __restrict__ colk, const int* info) {
    if (*info != 0) return;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i0 = blockIdx.y * 8;
    if (j >= n) return;
    const double rkj = A[(int64_t)k * ld + j];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        const int i = i0 + u;
        if (i < n && i != k) A[(int64_t)i * ld + j] = fma(-colk[i], rkj, A[(int64_t)i * ld + j]);
    }
}

// undo the row exchanges: columns k and piv[k] swapped for k = n-1 .. 0; a thread owns one row
__global__ void gj_unpivot_kernel(double* A, int64_t ld, int n, const int* __restrict__ piv, const int* info) {
    if (*info != 0) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double* r = A + (int64_t)i * ld;
    for (int k = n - 1; k >= 0; --k) {
        const int p = piv[k];
        if (p != k) {
            const double t = r[k];
            r[k] = r[p];
            r[p] = t;
        }
    }
}

// A (n x n, ld) <- inv(A) in place on c->stream. piv: n ints, colk: n doubles, info: device int (0 on entry).
int lu_inverse_inplace(Ctx* c, double* A, int64_t ld, int64_t n, int* piv, double* colk, int* info) {
    const dim3 grid((unsigned)((n + 255) / 256), (unsigned)((n + 7) / 8));
    for (int k = 0; k < (int)n; ++k) {
        gj_pivot_kernel<<<1, LU_THREADS, 0, c->stream>>>(A, ld, (int)n, k, piv, colk, info);
        gj_eliminate_kernel<<<grid, 256, 0, c->stream>>>(A, ld, (int)n, k, colk, info);
        c->launches += 2;
    }
    gj_unpivot_kernel<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(A, ld, (int)n, piv, info);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}

}  // namespace gprb200

extern "C" int gprb200_lu_inverse(gprb200_handle h, const double* A, int64_t n, double* Ainv) {
    using namespace gprb200;
    if (!h || !A || !Ainv || n <= 0) return -2;
    Ctx* c = h;
    if (n > 32768) return fail(c, -2, "lu_inverse: n <= 32768 (a utility; the GP path factors by Cholesky)");
    GP_CUDA(c, cudaSetDevice(c->device));
    const int64_t ld = round_up(n, 16);
    double *dA = nullptr, *colk = nullptr;
    int* ip = nullptr;   // [0] = info, [1..n] = pivots
    GP_CUDA(c, cudaMallocAsync((void**)&dA, sizeof(double) * n * ld, c->stream));
    GP_CUDA(c, cudaMallocAsync((void**)&colk, sizeof(double) * n, c->stream));
    GP_CUDA(c, cudaMallocAsync((void**)&ip, sizeof(int) * (n + 1), c->stream));
    int rc = 0, info = 0;
    auto ck = [&](cudaError_t e, const char* what) {
        if (e != cudaSuccess && rc == 0) rc = fail(c, -100 - (int)e, std::string("lu_inverse: ") + what + ": " + cudaGetErrorString(e));
    };
    ck(cudaMemsetAsync(ip, 0, sizeof(int), c->stream), "memset");
    ck(cudaMemcpy2DAsync(dA, sizeof(double) * ld, A, sizeof(double) * n, sizeof(double) * n, (size_t)n, cudaMemcpyHostToDevice,
                         c->stream), "upload");
    if (rc == 0) rc = lu_inverse_inplace(c, dA, ld, n, ip + 1, colk, ip);
    ck(cudaMemcpyAsync(&info, ip, sizeof(int), cudaMemcpyDeviceToHost, c->stream), "info");
    ck(cudaStreamSynchronize(c->stream), "synchronize");
    if (rc == 0 && info == 0)
        ck(cudaMemcpy2DAsync(Ainv, sizeof(double) * n, dA, sizeof(double) * ld, sizeof(double) * n, (size_t)n,
                             cudaMemcpyDeviceToHost, c->stream), "download");
    ck(cudaStreamSynchronize(c->stream), "synchronize");
    cudaFreeAsync(dA, c->stream);
    cudaFreeAsync(colk, c->stream);
    cudaFreeAsync(ip, c->stream);
    if (rc != 0) return rc;
    if (info != 0)
        return fail(c, info, "lu_invert: error in inversion. Matrix is singular or contains an illegal value.");
    return 0;
}
