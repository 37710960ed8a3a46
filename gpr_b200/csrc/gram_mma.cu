// K1 for a LARGE input dimension: the Gram matrix of kernels that depend on the pair only through r^2 = ||x - y||^2
// (Gaussian, GaussianExp, RationalQuadratic, White and Sum / Product trees of them; reference include/Kernel.h:465-808)
// with the cross term of  r^2 = ||x||^2 + ||y||^2 - 2 x.y  on the FP64 tensor pipe.
//
// When it applies. The direct form (gram.cu) spends 2 d FP64 instructions per pair on differences and squares; at
// d = 8 (BASELINE configs C3-C5) that is nothing next to the exponential, at d = 30 (C2) it is half of the pair, and at
// the reference's production setting d = 238 (scripts/params/pmb/config_ZL0001_vol01.yaml:49-52, n_inputModes) it IS
// the pair: 2 N^2 d flop, a real contraction, which DMMA.8x8x4 runs at twice the FMA-pipe rate with none of the issue
// slots. Used for d >= 64 (option "gram_mma_min_d"; measured crossover, tools/gram_mma_probe.py) when the tree has no PERIODIC leaf.
//
// Steps (all on c->stream):
//   1. centre: xc = x - mean(Xb) (r^2 is translation invariant; smaller norms = less cancellation), row norms n_i;
//   2. G = Xc_a Xc_b^T with the DMMA GEMM of dgemm.cu straight into the output buffer (lower tiles only when the
//      result is symmetric);
//   3. in place: r^2 = n_i + n_j - 2 G_ij  ->  k(r^2) by the kernel-tree evaluator (kprog_value), + sigma^2 on the
//      diagonal, mirrored into the upper triangle when the full matrix is wanted.
// Cancellation. The expansion carries an ABSOLUTE error of a few eps (n_i + n_j) in r^2, whatever r^2 is, so the
// error of K is bounded by max |dk / dr^2| * c eps max ||xc||^2. The bound is evaluated on the host from the tree
// (gram_mma_bound) and the call FALLS BACK to the direct kernels (returns 1) unless it is below 2e-14 of the largest
// kernel value, which keeps the parity bar of K (1e-13 max |K|). Pairs whose expanded r^2 is below 1e-9 (n_i + n_j)
// (near-coincident points, where the RELATIVE error of r^2 would matter, and exact duplicates, which WhiteKernel's
// ||x - y|| == 0 test must see as exactly 0, include/Kernel.h:695-702) are recomputed by direct differences.
#include "common.cuh"
#include "kfunc.cuh"
#include <math.h>
#include <string.h>

namespace gprb200 {

// column means of X (N x d) -> mu[d]; one CTA per column chunk, fixed order
__global__ void colmean_kernel(const double* __restrict__ X, int64_t N, int d, double* __restrict__ mu) {
    const int j = blockIdx.x;
    __shared__ double red[256];
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < N; i += blockDim.x) s += X[i * d + j];
    red[threadIdx.x] = s;
    __syncthreads();
    for (int o = blockDim.x >> 1; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) mu[j] = red[0] / (double)N;
}

// Xc (N x ldc, zero-padded) = X - mu, nrm[i] = ||xc_i||^2, and the maximum norm as an ordered-int atomic max
__global__ void centre_norm_kernel(const double* __restrict__ X, int64_t N, int d, const double* __restrict__ mu,
                                   double* __restrict__ Xc, int ldc, double* __restrict__ nrm, unsigned long long* maxbits) {
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= N) return;
    double s = 0.0;
    for (int k = lane; k < ldc; k += 32) {
        const double v = (k < d) ? X[i * d + k] - mu[k] : 0.0;
        Xc[i * ldc + k] = v;
        s = fma(v, v, s);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) {
        nrm[i] = s;
        atomicMax(maxbits, (unsigned long long)__double_as_longlong(s));   // s >= 0: bit order = value order
    }
}

// in place: out[i][j] = k(r^2), r^2 from the expansion (or from direct differences for near-coincident pairs)
__global__ void __launch_bounds__(256) gram_mma_transform_kernel(double* __restrict__ out, int64_t ldo, int64_t Na, int64_t Nb,
                                                                 const double* __restrict__ na, const double* __restrict__ nb,
                                                                 const double* __restrict__ Xca, const double* __restrict__ Xcb,
                                                                 int ldc, int d, DevProg P, int symmetric, int lower_only,
                                                                 double diag_add, int* nonfinite) {
    // 32 x 32 tiles, 256 threads: thread (tx, ty) owns column tx and rows ty, ty + 8, ...
    __shared__ double tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int64_t i0 = (int64_t)blockIdx.y * 32, j0 = (int64_t)blockIdx.x * 32;
    if (symmetric && j0 > i0) return;                   // lower tiles only; the mirror is written from here
    const int64_t j = j0 + tx;
    const double nj = (j < Nb) ? nb[j] : 0.0;
    int bad = 0;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int64_t i = i0 + ty + 8 * r;
        double v = 0.0;
        if (i < Na && j < Nb && !(symmetric && j > i)) {
            const double ni = na[i];
            double r2 = fma(-2.0, out[i * ldo + j], ni + nj);
            if (symmetric && i == j) {
                r2 = 0.0;
            } else if (r2 < 1e-9 * (ni + nj)) {
                const double* xa = Xca + i * ldc;
                const double* xb = Xcb + j * ldc;
                double s = 0.0;
                for (int k = 0; k < d; ++k) {
                    const double dx = xa[k] - xb[k];
                    s = fma(dx, dx, s);
                }
                r2 = s;
            }
            Geo<false> g;
            geo_init(g);
            g.r2 = r2;
            v = kprog_value(P, g);
            if (symmetric && i == j) v += diag_add;
            bad |= (((unsigned)__double2hiint(v) & 0x7ff00000u) == 0x7ff00000u) ? 1 : 0;
            out[i * ldo + j] = v;
        }
        tile[ty + 8 * r][tx] = v;
    }
    if (symmetric && !lower_only && i0 != j0) {   // transposed store of the tile into the upper triangle
        __syncthreads();
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int64_t jj = j0 + ty + 8 * r, ii = i0 + tx;
            if (jj < Nb && ii < Na) out[jj * ldo + ii] = tile[tx][ty + 8 * r];
        }
    } else if (symmetric && !lower_only) {        // diagonal tile: mirror its strict lower part
        __syncthreads();
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int a = ty + 8 * r, b = tx;     // (row a, col b) of the upper part takes (b, a) of the lower
            if (b > a && i0 + a < Na && j0 + b < Nb) out[(i0 + a) * ldo + j0 + b] = tile[b][a];
        }
    }
    if (nonfinite && __syncthreads_or(bad) && threadIdx.x == 0) atomicAdd(nonfinite, 1);
}

// (max |k|, max |dk / dr^2|) of the tree over r^2 >= 0; false if a leaf is not a function of r^2 alone
static bool tree_bounds(const DevProg& P, double* maxval, double* maxder) {
    double val[GPRB200_MAX_OPS], der[GPRB200_MAX_OPS];
    for (int i = 0; i < P.nops; ++i) {
        const double* c = P.cst[i];
        switch (P.op[i]) {
            case GPRB200_K_GAUSSIAN:
            case GPRB200_K_GAUSSIAN_EXP: val[i] = fabs(c[0]); der[i] = fabs(c[0] * c[1]); break;       // c0 exp(c1 r2), c1 < 0
            case GPRB200_K_WHITE: val[i] = fabs(c[0]); der[i] = 0.0; break;                              // handled by the fix-up
            case GPRB200_K_RQ: val[i] = fabs(c[0]); der[i] = fabs(c[0] * c[1] * c[2]); break;            // c0 (1 + c1 r2)^c2
            case GPRB200_K_SUM:
                val[i] = val[P.lchild[i]] + val[P.rchild[i]];
                der[i] = der[P.lchild[i]] + der[P.rchild[i]];
                break;
            case GPRB200_K_PRODUCT:
                val[i] = val[P.lchild[i]] * val[P.rchild[i]];
                der[i] = der[P.lchild[i]] * val[P.rchild[i]] + val[P.lchild[i]] * der[P.rchild[i]];
                break;
            default: return false;   // PERIODIC: needs its own feature products
        }
    }
    *maxval = val[P.nops - 1];
    *maxder = der[P.nops - 1];
    return true;
}

// Returns 0 if the Gram was built here, 1 if the path does not apply (the caller uses the direct kernels), < 0 on error.
int gram_build_mma(Ctx* c, const DevProg& P, const double* Xa, int64_t Na, const double* Xb, int64_t Nb, double* out,
                   int64_t ldo, bool symmetric, bool lower_only, double diag_add, int* d_nonfinite) {
    const int d = c->d;
    if (!c->opt_gram_mma || d < c->opt_gram_mma_min_d || P.nper > 0) return 1;
    if (Na * Nb < (1ll << 16) || Na < 32 || Nb < 32) return 1;                       // kernel vectors, tiny blocks
    if ((ldo & 1) || (reinterpret_cast<uintptr_t>(out) & 15)) return 1;
    double maxval = 0.0, maxder = 0.0;
    if (!tree_bounds(P, &maxval, &maxder) || !(maxval > 0.0)) return 1;
    const int ldc = (int)round_up(d, 2);
    const bool same = (Xa == Xb && Na == Nb);
    double* buf = nullptr;   // [mu: ldc][na: Na][nb: Nb][Xca: Na ldc][Xcb: Nb ldc][max: 1]
    const size_t need = (size_t)ldc + Na + (same ? 0 : Nb) + (size_t)Na * ldc + (same ? 0 : (size_t)Nb * ldc) + 2;
    GP_CUDA(c, cudaMallocAsync((void**)&buf, sizeof(double) * need, c->stream));
    double* mu = buf;
    double* na = mu + ldc;
    double* nb = same ? na : na + Na;
    double* Xca = nb + Nb;
    if ((reinterpret_cast<uintptr_t>(Xca) & 15)) ++Xca;   // 16-byte rows for the GEMM loader (one spare double allocated)
    double* Xcb = same ? Xca : Xca + (size_t)Na * ldc;
    unsigned long long* maxbits = reinterpret_cast<unsigned long long*>(buf + need - 1);
    int rc = 0;
    auto body = [&]() -> int {
        GP_CUDA(c, cudaMemsetAsync(maxbits, 0, sizeof(unsigned long long), c->stream));
        colmean_kernel<<<d, 256, 0, c->stream>>>(Xb, Nb, d, mu);
        centre_norm_kernel<<<(unsigned)((Nb + 7) / 8), 256, 0, c->stream>>>(Xb, Nb, d, mu, Xcb, ldc, nb, maxbits);
        c->launches += 2;
        if (!same) {
            centre_norm_kernel<<<(unsigned)((Na + 7) / 8), 256, 0, c->stream>>>(Xa, Na, d, mu, Xca, ldc, na, maxbits);
            ++c->launches;
        }
        GP_CUDA(c, cudaGetLastError());
        unsigned long long bits = 0;
        GP_CUDA(c, cudaMemcpyAsync(&bits, maxbits, sizeof(bits), cudaMemcpyDeviceToHost, c->stream));
        GP_CUDA(c, cudaStreamSynchronize(c->stream));
        double maxn;
        memcpy(&maxn, &bits, sizeof(double));
        // absolute error of the expanded r^2: the GEMM's d-term sum and the two additions, ~ (4 + sqrt(d)) eps max ||xc||^2
        const double err_r2 = (4.0 + sqrt((double)d)) * 2.220446049250313e-16 * 2.0 * maxn;
        if (!(maxder * err_r2 <= 2e-14 * maxval)) return 1;   // cancellation guard: use the direct kernels
        GemmArgs g{};
        g.A = Xca; g.lda = ldc; g.B = Xcb; g.ldb = ldc; g.C = out; g.ldc = ldo;
        g.M = (int)Na; g.N = (int)Nb; g.K = ldc; g.alpha = 1.0; g.beta = 0.0; g.lower = (symmetric && same) ? 1 : 0;
        GP_TRY(gemm(c, GEMM_NT, g));
        const dim3 grid((unsigned)((Nb + 31) / 32), (unsigned)((Na + 31) / 32));
        gram_mma_transform_kernel<<<grid, 256, 0, c->stream>>>(out, ldo, Na, Nb, na, nb, Xca, Xcb, ldc, d, P,
                                                               (symmetric && same) ? 1 : 0, lower_only ? 1 : 0, diag_add,
                                                               d_nonfinite);
        ++c->launches;
        GP_CUDA(c, cudaGetLastError());
        return 0;
    };
    rc = body();
    cudaFreeAsync(buf, c->stream);
    return rc;
}

}  // namespace gprb200
