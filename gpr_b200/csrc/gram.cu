// K1 / K2 / K6: Gram-matrix kernels.
//   gram_kernel<false>  K(i,j) = k(xa_i, xb_j)       replaces ComputeKernelMatrixInternal (reference
//                       lib/GaussianProcess.cpp:384-402; + sigma^2 on the diagonal, :375-381) and
//                       ComputeKernelVectorInternal (:684-693) for a batch of test points.
//   gram_kernel<true>   the stack dK/dtheta_p        replaces ComputeDerivativeKernelMatrixInternal (:472-495)
//                       either stored ((P*N) x N) or contracted on the fly with a symmetric S
//                       (the trace of include/Likelihood.h:280-282 without materialising dK).
// Bound: HBM write (8 B per entry) for the value kernel once the FP64 ALU work per entry
// (~2d + exp) is below ~11 DFMA-equivalents/byte; PERIODIC leaves are ALU-bound (sin per dimension).
#include "kfunc.cuh"
#include <math.h>
#include <vector>

namespace gprb200 {

int gram_build_mma(Ctx* c, const DevProg& P, const double* Xa, int64_t Na, const double* Xb, int64_t Nb, double* out,
                   int64_t ldo, bool symmetric, bool lower_only, double diag_add, int* d_nonfinite);

constexpr int GT = 64;        // CTA tile is GT x GT entries
constexpr int GDC = 8;        // dimensions staged per shared-memory chunk
constexpr int GTHREADS = 256;

struct GramArgs {
    const double* Xa;  // Na x d
    const double* Xb;  // Nb x d
    int64_t Na, Nb;
    int d;
    double* out;       // value: Na x Nb (ldo); deriv store: (P*Na) x Nb
    int64_t ldo;
    int symmetric;     // Xa == Xb: only tiles ti >= tj are visited
    int lower_only;    // symmetric: do not mirror into the upper triangle
    double diag_add;   // symmetric: added to out(i,i)
    const double* S;   // deriv contraction: symmetric N x N weights (lds) or nullptr
    int64_t lds;
    double* partial;   // deriv contraction: [grid][P] block partials
    int tiles_b;       // number of tile columns
    int p_select;      // deriv store: -1 = whole stack, else only parameter p (out is then Na x Nb)
    const double* Fa;  // value kernel: per-point features [row][leaf][dim][cos,sin] of the PERIODIC leaves (or nullptr)
    const double* Fb;
    int* nonfinite;    // value kernel: incremented once per CTA that produced a non-finite entry (or nullptr)
};

__device__ __forceinline__ void tile_from_linear(int lin, int symmetric, int tiles_b, int& ti, int& tj) {
    if (symmetric) {
        // lin = ti (ti + 1) / 2 + tj, tj <= ti
        int t = (int)((sqrt(8.0 * (double)lin + 1.0) - 1.0) * 0.5);
        while ((int64_t)(t + 1) * (t + 2) / 2 <= lin) ++t;
        while ((int64_t)t * (t + 1) / 2 > lin) --t;
        ti = t;
        tj = lin - (int)((int64_t)t * (t + 1) / 2);
    } else {
        ti = lin / tiles_b;
        tj = lin - ti * tiles_b;
    }
}

template <bool DERIV>
__global__ void __launch_bounds__(GTHREADS, DERIV ? 1 : 2) gram_kernel(GramArgs a, DevProg P) {
    __shared__ double xa[GDC][GT + 1];
    __shared__ double xb[GDC][GT + 1];
    __shared__ double tr[GT][GT / 2 + 1];        // transpose staging for the mirrored write, one j-half per pass
    __shared__ double feat[2][2][GDC][GT + 1];   // value kernel: [cos|sin][a|b][dim][row] features of one PERIODIC leaf

    int ti, tj;
    tile_from_linear(blockIdx.x, a.symmetric, a.tiles_b, ti, tj);
    const int64_t i0 = (int64_t)ti * GT, j0 = (int64_t)tj * GT;
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;  // entries i = ty + 16 a, j = tx + 16 b

    double acc[DERIV ? GPRB200_MAX_PARAMS : 1];
    if constexpr (DERIV) {
        for (int p = 0; p < P.nparams; ++p) acc[p] = 0.0;
    }
    int bad = 0;
    const bool diag_tile = a.symmetric && (ti == tj);
    const bool mirror = a.symmetric && !a.lower_only && !diag_tile;

    // two passes over the j-direction (b in {0,1} then {2,3}) keep the per-thread geometry at 8 pairs
    for (int pass = 0; pass < 2; ++pass) {
        Geo<DERIV> g[4][2];
#pragma unroll
        for (int ia = 0; ia < 4; ++ia)
#pragma unroll
            for (int ib = 0; ib < 2; ++ib) geo_init(g[ia][ib]);

        for (int d0 = 0; d0 < a.d; d0 += GDC) {
            const int dc = min(GDC, a.d - d0);
            __syncthreads();
            // stage xa / xb chunk, transposed to [dim][row]
            for (int e = tid; e < GT * GDC; e += GTHREADS) {
                const int r = e / GDC, dd = e - r * GDC;
                double va = 0.0, vb = 0.0;
                if (dd < dc) {
                    if (i0 + r < a.Na) va = a.Xa[(i0 + r) * a.d + d0 + dd];
                    if (j0 + r < a.Nb) vb = a.Xb[(j0 + r) * a.d + d0 + dd];
                }
                xa[dd][r] = va;
                xb[dd][r] = vb;
            }
            __syncthreads();
            for (int dd = 0; dd < dc; ++dd) {
                double va[4], vb[2];
#pragma unroll
                for (int ia = 0; ia < 4; ++ia) va[ia] = xa[dd][ty + 16 * ia];
#pragma unroll
                for (int ib = 0; ib < 2; ++ib) vb[ib] = xb[dd][tx + 16 * (pass * 2 + ib)];
#pragma unroll
                for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                    for (int ib = 0; ib < 2; ++ib) {
                        if constexpr (DERIV) geo_acc(g[ia][ib], P, va[ia] - vb[ib]);
                        else geo_acc_r2(g[ia][ib], va[ia] - vb[ib]);
                    }
            }
            if constexpr (!DERIV) {
                // PERIODIC leaves: sum_d sin^2(b (x_d - y_d)) = d/2 - 1/2 sum_d [cos(2b x_d) cos(2b y_d) + sin(2b x_d) sin(2b y_d)].
                // The per-point features (cos, sin of 2 b x_d) are computed once per Gram build by
                // periodic_features_kernel (N d sincos instead of N^2 d / 2 sin); the pair work is a 2 dc-long dot product. Absolute error of the sum ~ d eps, i.e. relative 1e-15 on k (it enters
                // through exp); checked against the oracle at 1e-13 of max|K|.
#pragma unroll
                for (int q = 0; q < MAXPER; ++q) {
                    if (q >= P.nper) break;  // uniform across the CTA
                    __syncthreads();
                    for (int e = tid; e < GT * GDC; e += GTHREADS) {
                        const int dd = e % GDC, r = e / GDC;
                        double2 fa = make_double2(1.0, 0.0), fb = make_double2(1.0, 0.0);
                        if (dd < dc) {
                            if (i0 + r < a.Na)
                                fa = *reinterpret_cast<const double2*>(a.Fa + ((((i0 + r) * P.nper + q) * a.d) + d0 + dd) * 2);
                            if (j0 + r < a.Nb)
                                fb = *reinterpret_cast<const double2*>(a.Fb + ((((j0 + r) * P.nper + q) * a.d) + d0 + dd) * 2);
                        }
                        feat[0][0][dd][r] = fa.x; feat[1][0][dd][r] = fa.y;
                        feat[0][1][dd][r] = fb.x; feat[1][1][dd][r] = fb.y;
                    }
                    __syncthreads();
                    for (int dd = 0; dd < dc; ++dd) {
                        double ca[4], sa[4], cb[2], sb[2];
#pragma unroll
                        for (int ia = 0; ia < 4; ++ia) {
                            ca[ia] = feat[0][0][dd][ty + 16 * ia];
                            sa[ia] = feat[1][0][dd][ty + 16 * ia];
                        }
#pragma unroll
                        for (int ib = 0; ib < 2; ++ib) {
                            cb[ib] = feat[0][1][dd][tx + 16 * (pass * 2 + ib)];
                            sb[ib] = feat[1][1][dd][tx + 16 * (pass * 2 + ib)];
                        }
#pragma unroll
                        for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                            for (int ib = 0; ib < 2; ++ib)
                                g[ia][ib].S[q] = fma(ca[ia], cb[ib], fma(sa[ia], sb[ib], g[ia][ib].S[q]));
                    }
                }
            }
        }
        if constexpr (!DERIV) {
            // S currently holds the feature dot product; turn it into sum_d sin^2
#pragma unroll
            for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                for (int ib = 0; ib < 2; ++ib)
#pragma unroll
                    for (int q = 0; q < MAXPER; ++q)
                        if (q < P.nper) g[ia][ib].S[q] = fmax(0.0, 0.5 * ((double)a.d - g[ia][ib].S[q]));
        }

#pragma unroll
        for (int ia = 0; ia < 4; ++ia) {
#pragma unroll
            for (int ib = 0; ib < 2; ++ib) {
                const int li = ty + 16 * ia, lj = tx + 16 * (pass * 2 + ib);
                const int64_t i = i0 + li, j = j0 + lj;
                const bool inb = (i < a.Na) && (j < a.Nb);
                if constexpr (!DERIV) {
                    double v = kprog_value(P, g[ia][ib]);
                    if (a.symmetric && i == j) v += a.diag_add;
                    if (inb && !isfinite(v)) bad = 1;
                    if (inb && !(a.symmetric && a.lower_only && diag_tile && j > i)) a.out[i * a.ldo + j] = v;
                    if (mirror) tr[li][lj - pass * (GT / 2)] = v;
                } else {
                    double dk[GPRB200_MAX_PARAMS];
                    if (inb) {
                        kprog_deriv(P, g[ia][ib], dk);
                        if (a.S) {
                            // symmetric S and dK: off-diagonal tiles count twice; a rectangular (cross) contraction
                            // visits every tile once
                            const double w = a.S[i * a.lds + j] * ((a.symmetric && !diag_tile) ? 2.0 : 1.0);
                            for (int p = 0; p < P.nparams; ++p) acc[p] = fma(w, dk[p], acc[p]);
                        } else if (a.p_select >= 0) {
                            a.out[i * a.ldo + j] = dk[a.p_select];
                            if (a.symmetric && !diag_tile) a.out[j * a.ldo + i] = dk[a.p_select];
                        } else {
                            for (int p = 0; p < P.nparams; ++p) {
                                a.out[((int64_t)p * a.Na + i) * a.ldo + j] = dk[p];
                                if (a.symmetric && !diag_tile) a.out[((int64_t)p * a.Na + j) * a.ldo + i] = dk[p];
                            }
                        }
                    }
                }
            }
        }
        if constexpr (!DERIV) {
            if (mirror) {
                __syncthreads();
                // out(j0 + pass 32 + r, i0 + c) = tile(c, r): consecutive threads walk the contiguous direction
                for (int e = tid; e < GT * (GT / 2); e += GTHREADS) {
                    const int r = e / GT, cc = e - r * GT;
                    const int64_t j = j0 + pass * (GT / 2) + r, i = i0 + cc;
                    if (j < a.Nb && i < a.Na) a.out[j * a.ldo + i] = tr[cc][r];
                }
                __syncthreads();
            }
        }
    }

    if constexpr (!DERIV) {
        // finite check of ComputeKernelMatrixInternal (lib/GaussianProcess.cpp:399), fused into the build
        if (a.nonfinite && __syncthreads_or(bad) && tid == 0) atomicAdd(a.nonfinite, 1);
    }

    if constexpr (DERIV) {
      if (a.S) {
        // deterministic block reduction of acc[p] -> partial[block][p]
        double* red = &tr[0][0];  // GTHREADS doubles are enough
        for (int p = 0; p < P.nparams; ++p) {
            double v = acc[p];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            __syncthreads();
            if ((tid & 31) == 0) red[tid >> 5] = v;
            __syncthreads();
            if (tid == 0) {
                double s = 0.0;
                for (int w = 0; w < GTHREADS / 32; ++w) s += red[w];
                a.partial[(int64_t)blockIdx.x * P.nparams + p] = s;
            }
        }
      }
    }
}

// ---------------------------------------------------------------------------------------------
// exp(x) for x <= 0 (every exponential of the kernel family has a non-positive argument: -r^2/(2 sigma^2),
// -sum sin^2/(2 sigma^2)). Branch-free: n = rint(x log2 e) by the magic-constant trick, r = x - n ln2 (two-term ln2),
// degree-13 Taylor polynomial on |r| <= ln2/2 (truncation 4e-18 relative), 2^n folded into the exponent field.
// Below -708 the result is 0 (the true value is < 3.3e-308; no denormal path). Relative error <= 2 ulp.
__device__ __forceinline__ double exp_nonpos(double x) {
    const double MAGIC = 6755399441055744.0;   // 1.5 * 2^52
    const double t = fma(x, 1.4426950408889634074, MAGIC);
    const int n = __double2loint(t);
    const double nd = t - MAGIC;
    double r = fma(nd, -6.93147180559945286227e-01, x);
    r = fma(nd, -2.31904681384629955842e-17, r);
    double p = 1.6059043836821613e-10;                 // 1/13!
    p = fma(p, r, 2.08767569878680990e-09);            // 1/12!
    p = fma(p, r, 2.50521083854417188e-08);            // 1/11!
    p = fma(p, r, 2.75573192239858907e-07);            // 1/10!
    p = fma(p, r, 2.75573192239858907e-06);            // 1/9!
    p = fma(p, r, 2.48015873015873016e-05);            // 1/8!
    p = fma(p, r, 1.98412698412698413e-04);            // 1/7!
    p = fma(p, r, 1.38888888888888889e-03);            // 1/6!
    p = fma(p, r, 8.33333333333333333e-03);            // 1/5!
    p = fma(p, r, 4.16666666666666667e-02);            // 1/4!
    p = fma(p, r, 1.66666666666666667e-01);            // 1/3!
    p = fma(p, r, 0.5);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    const double res = __hiloint2double(__double2hiint(p) + (n << 20), __double2loint(p));
    return (x < -708.0) ? 0.0 : res;
}

// ---------------------------------------------------------------------------------------------
// Value Gram in term form (DevProg::nterms > 0): the default K1 kernel.
//   * a thread owns 4 consecutive rows x 2 consecutive columns per pass (two passes over the 64 tile columns), so the
//     per-dimension operands come from shared memory as 16-byte loads (3 LDS per 8 pairs) and K leaves as 16-byte stores;
//   * the kernel program is walked ONCE per thread and pass, term by term and leaf by leaf, each leaf evaluated for the
//     8 pairs in a row with its opcode branch outside the loop - no per-entry interpreter, no evaluation stack;
//   * the points and the trig features of ALL PERIODIC leaves are staged together, once per CTA when d <= GDC
//     (one barrier per tile instead of two per pass and leaf);
//   * NPER (number of PERIODIC leaves) is a template parameter so the trig-feature accumulators are exactly sized.
// lib/GaussianProcess.cpp:384-402 (+ :375-381 for the diagonal), ComputeKernelVectorInternal :684-693 for Xa != Xb.
template <int NPER>
__global__ void __launch_bounds__(GTHREADS, 2) gram_value_kernel(GramArgs a, DevProg P, int vec_ok) {
    constexpr int NS = NPER > 0 ? NPER : 1;
    extern __shared__ __align__(16) unsigned char gv_smem[];
    typedef double Row[GT];
    typedef double2 Row2[GT];
    Row* xa = reinterpret_cast<Row*>(gv_smem);                                   // [GDC][GT]
    Row* xb = xa + GDC;                                                          // [GDC][GT]
    Row2* fa = reinterpret_cast<Row2*>(xb + GDC);                                // [NPER][GDC][GT] (cos, sin)(2 b x_d)
    Row2* fb = fa + NPER * GDC;                                                  // [NPER][GDC][GT]
    double (*tr)[GT / 2 + 1] = reinterpret_cast<double (*)[GT / 2 + 1]>(fb + NPER * GDC);   // [GT][GT/2+1] mirror staging

    int ti, tj;
    tile_from_linear(blockIdx.x, a.symmetric, a.tiles_b, ti, tj);
    const int64_t i0 = (int64_t)ti * GT, j0 = (int64_t)tj * GT;
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const bool diag_tile = a.symmetric && (ti == tj);
    const bool mirror = a.symmetric && !a.lower_only && !diag_tile;
    const bool one_chunk = a.d <= GDC;
    int bad = 0;

    auto stage = [&](int d0, int dc) {
        for (int e = tid; e < GT * GDC; e += GTHREADS) {
            const int r = e & (GT - 1), dd = e / GT;
            const bool ina = (dd < dc) && (i0 + r < a.Na), inb = (dd < dc) && (j0 + r < a.Nb);
            xa[dd][r] = ina ? a.Xa[(i0 + r) * a.d + d0 + dd] : 0.0;
            xb[dd][r] = inb ? a.Xb[(j0 + r) * a.d + d0 + dd] : 0.0;
#pragma unroll
            for (int q = 0; q < NPER; ++q) {
                fa[q * GDC + dd][r] = ina ? *reinterpret_cast<const double2*>(a.Fa + ((((i0 + r) * NPER + q) * a.d) + d0 + dd) * 2)
                                          : make_double2(1.0, 0.0);
                fb[q * GDC + dd][r] = inb ? *reinterpret_cast<const double2*>(a.Fb + ((((j0 + r) * NPER + q) * a.d) + d0 + dd) * 2)
                                          : make_double2(1.0, 0.0);
            }
        }
    };
    if (one_chunk) {
        stage(0, a.d);
        __syncthreads();
    }

    for (int pass = 0; pass < 2; ++pass) {
        const int lj0 = pass * (GT / 2) + 2 * tx;   // first of this thread's two tile columns
        double r2[4][2], S[NS][4][2];
#pragma unroll
        for (int ia = 0; ia < 4; ++ia)
#pragma unroll
            for (int ib = 0; ib < 2; ++ib) {
                r2[ia][ib] = 0.0;
#pragma unroll
                for (int q = 0; q < NS; ++q) S[q][ia][ib] = 0.0;
            }

        for (int d0 = 0; d0 < a.d; d0 += GDC) {
            const int dc = min(GDC, a.d - d0);
            if (!one_chunk) {
                __syncthreads();
                stage(d0, dc);
                __syncthreads();
            }
            for (int dd = 0; dd < dc; ++dd) {
                const double2 a01 = *reinterpret_cast<const double2*>(&xa[dd][4 * ty]);
                const double2 a23 = *reinterpret_cast<const double2*>(&xa[dd][4 * ty + 2]);
                const double2 b01 = *reinterpret_cast<const double2*>(&xb[dd][lj0]);
                const double va[4] = {a01.x, a01.y, a23.x, a23.y};
                const double vb[2] = {b01.x, b01.y};
#pragma unroll
                for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                    for (int ib = 0; ib < 2; ++ib) {
                        const double dx = va[ia] - vb[ib];
                        r2[ia][ib] = fma(dx, dx, r2[ia][ib]);
                    }
            }
            // sum_d sin^2(b (x_d - y_d)) = d/2 - 1/2 sum_d [cos(2b x_d) cos(2b y_d) + sin(2b x_d) sin(2b y_d)]
            // with the per-point features of periodic_features_kernel: a 2 dc-long dot product per pair
#pragma unroll
            for (int q = 0; q < NPER; ++q) {
                for (int dd = 0; dd < dc; ++dd) {
                    double2 ca[4], cb[2];
#pragma unroll
                    for (int ia = 0; ia < 4; ++ia) ca[ia] = fa[q * GDC + dd][4 * ty + ia];
#pragma unroll
                    for (int ib = 0; ib < 2; ++ib) cb[ib] = fb[q * GDC + dd][lj0 + ib];
#pragma unroll
                    for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                        for (int ib = 0; ib < 2; ++ib)
                            S[q][ia][ib] = fma(ca[ia].x, cb[ib].x, fma(ca[ia].y, cb[ib].y, S[q][ia][ib]));
                }
            }
        }
#pragma unroll
        for (int q = 0; q < NPER; ++q)
#pragma unroll
            for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                for (int ib = 0; ib < 2; ++ib) S[q][ia][ib] = fmax(0.0, 0.5 * ((double)a.d - S[q][ia][ib]));

        // k = sum_t prod_l leaf_l, leaf by leaf over the 8 pairs
        double acc[4][2];
#pragma unroll
        for (int ia = 0; ia < 4; ++ia)
#pragma unroll
            for (int ib = 0; ib < 2; ++ib) acc[ia][ib] = 0.0;
        for (int t = 0; t < P.nterms; ++t) {
            double prod[4][2];
            const int k0 = P.term_start[t], k1 = P.term_start[t + 1];
            for (int k = k0; k < k1; ++k) {
                const int node = P.term_leaf[k];
                const int op = P.op[node];
                const double c0 = P.cst[node][0], c1 = P.cst[node][1], c2 = P.cst[node][2];
                double v[4][2];
                switch (op) {
                    case GPRB200_K_GAUSSIAN:
                    case GPRB200_K_GAUSSIAN_EXP:
#pragma unroll
                        for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                            for (int ib = 0; ib < 2; ++ib) v[ia][ib] = c0 * exp_nonpos(c1 * r2[ia][ib]);
                        break;
                    case GPRB200_K_WHITE:
#pragma unroll
                        for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                            for (int ib = 0; ib < 2; ++ib) v[ia][ib] = (r2[ia][ib] == 0.0) ? c0 : 0.0;
                        break;
                    case GPRB200_K_RQ:
#pragma unroll
                        for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                            for (int ib = 0; ib < 2; ++ib) v[ia][ib] = c0 * pow(fma(c1, r2[ia][ib], 1.0), c2);
                        break;
                    default: {  // PERIODIC
                        const int slot = P.perslot[node];
#pragma unroll
                        for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                            for (int ib = 0; ib < 2; ++ib) v[ia][ib] = 0.0;
#pragma unroll
                        for (int q = 0; q < NS; ++q)
                            if (slot == q) {   // uniform
#pragma unroll
                                for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                                    for (int ib = 0; ib < 2; ++ib) v[ia][ib] = c0 * exp_nonpos(c1 * S[q][ia][ib]);
                            }
                    } break;
                }
                if (k == k0) {
#pragma unroll
                    for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                        for (int ib = 0; ib < 2; ++ib) prod[ia][ib] = v[ia][ib];
                } else {
#pragma unroll
                    for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                        for (int ib = 0; ib < 2; ++ib) prod[ia][ib] *= v[ia][ib];
                }
            }
#pragma unroll
            for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                for (int ib = 0; ib < 2; ++ib) acc[ia][ib] += prod[ia][ib];
        }

#pragma unroll
        for (int ia = 0; ia < 4; ++ia) {
            const int li = 4 * ty + ia;
            const int64_t i = i0 + li, j = j0 + lj0;
            if (a.symmetric) {
                if (i == j) acc[ia][0] += a.diag_add;
                if (i == j + 1) acc[ia][1] += a.diag_add;
            }
            const bool in0 = (i < a.Na) && (j < a.Nb), in1 = (i < a.Na) && (j + 1 < a.Nb);
            if ((in0 && !isfinite(acc[ia][0])) || (in1 && !isfinite(acc[ia][1]))) bad = 1;
            const bool tri = a.symmetric && a.lower_only && diag_tile;   // keep only j <= i
            if (vec_ok && in1 && !(tri && j + 1 > i)) {
                *reinterpret_cast<double2*>(&a.out[i * a.ldo + j]) = make_double2(acc[ia][0], acc[ia][1]);
            } else {
                if (in0 && !(tri && j > i)) a.out[i * a.ldo + j] = acc[ia][0];
                if (in1 && !(tri && j + 1 > i)) a.out[i * a.ldo + j + 1] = acc[ia][1];
            }
            if (mirror) {
                tr[li][2 * tx] = acc[ia][0];
                tr[li][2 * tx + 1] = acc[ia][1];
            }
        }
        if (mirror) {
            __syncthreads();
            // out(j0 + pass 32 + r, i0 + c) = tile(c, r): consecutive threads walk the contiguous direction
            for (int e = tid; e < GT * (GT / 2); e += GTHREADS) {
                const int r = e / GT, cc = e - r * GT;
                const int64_t j = j0 + pass * (GT / 2) + r, i = i0 + cc;
                if (j < a.Nb && i < a.Na) a.out[j * a.ldo + i] = tr[cc][r];
            }
            __syncthreads();
        }
    }
    // finite check of ComputeKernelMatrixInternal (lib/GaussianProcess.cpp:399), fused into the build
    if (a.nonfinite && __syncthreads_or(bad) && tid == 0) atomicAdd(a.nonfinite, 1);
}

// ---------------------------------------------------------------------------------------------
// Value Gram, row-group form: the kernel gram_build launches whenever all d dimensions of a 64-point tile fit in
// shared memory (always for the BASELINE configs). On top of gram_value_kernel:
//   * a CTA owns tile row ti and GROUP consecutive tile columns: the row side (points + trig features) is staged once,
//     the column side is double-buffered with cp.async - the next tile's points land while this tile is computed, so no
//     global-memory latency is exposed after the first tile;
//   * FP64 DFMA has a ~40-cycle dependent latency on sm_100 (tools/fp64_probe.cu: one warp with 16 independent chains
//     reaches 79 % of the pipe), so the exponentials of a thread's 8 pairs are evaluated in lock step (8 independent
//     Horner chains per warp, 32 per scheduler) instead of one call after another;
//   * single-leaf terms accumulate straight into the result (no product registers live).
template <int MODE>   // 0: dst = v, 1: dst *= v, 2: dst += v   with v = c0 exp(c1 src), src >= 0 >= c1
__device__ __forceinline__ void exp_leaf8(const double (&src)[4][2], double c0, double c1, double (&dst)[4][2]) {
    const double MAGIC = 6755399441055744.0;
    double r[4][2], p[4][2];
    int n[4][2];
#pragma unroll
    for (int ia = 0; ia < 4; ++ia)
#pragma unroll
        for (int ib = 0; ib < 2; ++ib) {
            const double x = __dmul_rn(c1, src[ia][ib]);
            const double t = fma(x, 1.4426950408889634074, MAGIC);
            n[ia][ib] = __double2loint(t);
            const double nd = t - MAGIC;
            r[ia][ib] = fma(nd, -2.31904681384629955842e-17, fma(nd, -6.93147180559945286227e-01, x));
            p[ia][ib] = fma(1.6059043836821613e-10, r[ia][ib], 2.08767569878680990e-09);
        }
    const double cf[12] = {2.50521083854417188e-08, 2.75573192239858907e-07, 2.75573192239858907e-06,
                           2.48015873015873016e-05, 1.98412698412698413e-04, 1.38888888888888889e-03,
                           8.33333333333333333e-03, 4.16666666666666667e-02, 1.66666666666666667e-01, 0.5, 1.0, 1.0};
#pragma unroll
    for (int k = 0; k < 12; ++k)
#pragma unroll
        for (int ia = 0; ia < 4; ++ia)
#pragma unroll
            for (int ib = 0; ib < 2; ++ib) p[ia][ib] = fma(p[ia][ib], r[ia][ib], cf[k]);
#pragma unroll
    for (int ia = 0; ia < 4; ++ia)
#pragma unroll
        for (int ib = 0; ib < 2; ++ib) {
            double v = __hiloint2double(__double2hiint(p[ia][ib]) + (n[ia][ib] << 20), __double2loint(p[ia][ib]));
            v = (n[ia][ib] < -1021) ? 0.0 : v;     // below 2^-1022: no denormal path
            // explicit roundings: K(i,j) and K(j,i) of a diagonal tile sit in different register slots of different
            // threads and must come out bit-identical, so the compiler may not choose the mul/add contractions
            if (MODE == 0) dst[ia][ib] = __dmul_rn(c0, v);
            else if (MODE == 1) dst[ia][ib] = __dmul_rn(dst[ia][ib], __dmul_rn(c0, v));
            else dst[ia][ib] = __fma_rn(c0, v, dst[ia][ib]);
        }
}

template <int MODE, int NS>
__device__ __forceinline__ void leaf8(const DevProg& P, int node, const double (&r2)[4][2], const double (&S)[NS][4][2],
                                      double (&dst)[4][2]) {
    const int op = P.op[node];
    const double c0 = P.cst[node][0], c1 = P.cst[node][1];
    switch (op) {
        case GPRB200_K_GAUSSIAN:
        case GPRB200_K_GAUSSIAN_EXP:
            exp_leaf8<MODE>(r2, c0, c1, dst);
            break;
        case GPRB200_K_PERIODIC: {
            const int slot = P.perslot[node];
#pragma unroll
            for (int q = 0; q < NS; ++q)
                if (slot == q) exp_leaf8<MODE>(S[q], c0, c1, dst);   // uniform branch, static register indexing
        } break;
        default: {
            const double c2 = P.cst[node][2];
#pragma unroll
            for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                for (int ib = 0; ib < 2; ++ib) {
                    // RQ: f^-alpha = exp(-alpha ln f), f >= 1 (relative error <= (1 + alpha ln f) ulp)
                    const double v = (op == GPRB200_K_WHITE) ? ((r2[ia][ib] == 0.0) ? c0 : 0.0)
                                     : (c2 <= 0.0) ? __dmul_rn(c0, exp_nonpos(__dmul_rn(c2, log(__fma_rn(c1, r2[ia][ib], 1.0)))))
                                                   : __dmul_rn(c0, pow(__fma_rn(c1, r2[ia][ib], 1.0), c2));
                    if (MODE == 0) dst[ia][ib] = v;
                    else if (MODE == 1) dst[ia][ib] = __dmul_rn(dst[ia][ib], v);
                    else dst[ia][ib] = __dadd_rn(dst[ia][ib], v);
                }
        } break;
    }
}

__device__ __forceinline__ void cp_async8(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;\n" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem));
}

constexpr int GGROUP = 4;   // tile columns per CTA

template <int NPER>
__global__ void __launch_bounds__(GTHREADS, 2) gram_rows_kernel(GramArgs a, DevProg P, int vec_ok, int use_tr) {
    constexpr int NS = NPER > 0 ? NPER : 1;
    extern __shared__ __align__(16) unsigned char gv_smem[];
    const int d = a.d;
    // layout: xa [d][GT] | xb [2][d][GT] | fa [NPER][d][GT] double2 | fb [2][NPER][d][GT] double2 | tr [GT][GT/2+1]
    double* xa = reinterpret_cast<double*>(gv_smem);
    double* xb = xa + d * GT;
    double2* fa = reinterpret_cast<double2*>(xb + 2 * d * GT);
    double2* fb = fa + NPER * d * GT;
    double (*tr)[GT / 2 + 1] = reinterpret_cast<double (*)[GT / 2 + 1]>(fb + 2 * NPER * d * GT);

    const int ti = blockIdx.y;
    const int tj_begin = blockIdx.x * GGROUP;
    int tj_end = min(tj_begin + GGROUP, a.tiles_b);
    if (a.symmetric) tj_end = min(tj_end, ti + 1);
    if (tj_begin >= tj_end) return;
    const int64_t i0 = (int64_t)ti * GT;
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    int bad = 0, expmax = 0;

    // out-of-range points are staged as zeros / (1, 0): their pairs are computed and never stored
    auto stage_side = [&](const double* X, const double* F, int64_t row0, int64_t nrows, double* xs, double2* fs) {
        for (int e = tid; e < GT * d; e += GTHREADS) {
            const int r = e & (GT - 1), dd = e / GT;
            if (row0 + r < nrows) {
                cp_async8(&xs[dd * GT + r], X + (row0 + r) * d + dd);
#pragma unroll
                for (int q = 0; q < NPER; ++q)
                    cp_async16(&fs[(q * d + dd) * GT + r], F + (((row0 + r) * NPER + q) * d + dd) * 2);
            } else {
                xs[dd * GT + r] = 0.0;
#pragma unroll
                for (int q = 0; q < NPER; ++q) fs[(q * d + dd) * GT + r] = make_double2(1.0, 0.0);
            }
        }
    };
    stage_side(a.Xa, a.Fa, i0, a.Na, xa, fa);
    stage_side(a.Xb, a.Fb, (int64_t)tj_begin * GT, a.Nb, xb, fb);
    asm volatile("cp.async.commit_group;\n" ::);

    for (int tj = tj_begin; tj < tj_end; ++tj) {
        const int buf = (tj - tj_begin) & 1;
        const double* xbc = xb + buf * d * GT;
        const double2* fbc = fb + buf * NPER * d * GT;
        if (tj + 1 < tj_end) {
            stage_side(a.Xb, a.Fb, (int64_t)(tj + 1) * GT, a.Nb, xb + (buf ^ 1) * d * GT, fb + (buf ^ 1) * NPER * d * GT);
            asm volatile("cp.async.commit_group;\n" ::);
            asm volatile("cp.async.wait_group 1;\n" ::);
        } else {
            asm volatile("cp.async.wait_group 0;\n" ::);
        }
        __syncthreads();

        const int64_t j0 = (int64_t)tj * GT;
        const bool diag_tile = a.symmetric && (ti == tj);
        const bool mirror = a.symmetric && !a.lower_only && !diag_tile;
        const bool full_tile = vec_ok && !diag_tile && (i0 + GT <= a.Na) && (j0 + GT <= a.Nb);   // uniform

        for (int pass = 0; pass < 2; ++pass) {
            const int lj0 = pass * (GT / 2) + 2 * tx;
            double r2[4][2], S[NS][4][2];
#pragma unroll
            for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                for (int ib = 0; ib < 2; ++ib) {
                    r2[ia][ib] = 0.0;
#pragma unroll
                    for (int q = 0; q < NS; ++q) S[q][ia][ib] = 0.0;
                }
            for (int dd = 0; dd < d; ++dd) {
                const double2 a01 = *reinterpret_cast<const double2*>(&xa[dd * GT + 4 * ty]);
                const double2 a23 = *reinterpret_cast<const double2*>(&xa[dd * GT + 4 * ty + 2]);
                const double2 b01 = *reinterpret_cast<const double2*>(&xbc[dd * GT + lj0]);
                const double va[4] = {a01.x, a01.y, a23.x, a23.y};
                const double vb[2] = {b01.x, b01.y};
#pragma unroll
                for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                    for (int ib = 0; ib < 2; ++ib) {
                        const double dx = va[ia] - vb[ib];
                        r2[ia][ib] = fma(dx, dx, r2[ia][ib]);
                    }
            }
#pragma unroll
            for (int q = 0; q < NPER; ++q) {
                for (int dd = 0; dd < d; ++dd) {
                    double2 ca[4], cb[2];
#pragma unroll
                    for (int ia = 0; ia < 4; ++ia) ca[ia] = fa[(q * d + dd) * GT + 4 * ty + ia];
#pragma unroll
                    for (int ib = 0; ib < 2; ++ib) cb[ib] = fbc[(q * d + dd) * GT + lj0 + ib];
#pragma unroll
                    for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                        for (int ib = 0; ib < 2; ++ib)
                            S[q][ia][ib] = fma(ca[ia].x, cb[ib].x, fma(ca[ia].y, cb[ib].y, S[q][ia][ib]));
                }
#pragma unroll
                for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                    for (int ib = 0; ib < 2; ++ib) S[q][ia][ib] = fmax(0.0, __dmul_rn(0.5, __dsub_rn((double)d, S[q][ia][ib])));
            }

            double acc[4][2];
#pragma unroll
            for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                for (int ib = 0; ib < 2; ++ib) acc[ia][ib] = 0.0;
            for (int t = 0; t < P.nterms; ++t) {
                const int k0 = P.term_start[t], k1 = P.term_start[t + 1];
                if (k1 - k0 == 1) {
                    leaf8<2, NS>(P, P.term_leaf[k0], r2, S, acc);
                } else {
                    double prod[4][2];
                    leaf8<0, NS>(P, P.term_leaf[k0], r2, S, prod);
                    for (int k = k0 + 1; k < k1; ++k) leaf8<1, NS>(P, P.term_leaf[k], r2, S, prod);
#pragma unroll
                    for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                        for (int ib = 0; ib < 2; ++ib) acc[ia][ib] = __dadd_rn(acc[ia][ib], prod[ia][ib]);
                }
            }

            if (full_tile) {
                // interior tile: no bounds, diagonal or triangle logic; 16-byte stores
                double* o = a.out + (i0 + 4 * ty) * a.ldo + j0 + lj0;
#pragma unroll
                for (int ia = 0; ia < 4; ++ia) {
                    *reinterpret_cast<double2*>(o + ia * a.ldo) = make_double2(acc[ia][0], acc[ia][1]);
                    // largest |high word|: >= 0x7ff00000 means an Inf or NaN was produced
                    expmax = max(expmax, max(__double2hiint(acc[ia][0]) & 0x7fffffff, __double2hiint(acc[ia][1]) & 0x7fffffff));
                    if (use_tr) {
                        tr[4 * ty + ia][2 * tx] = acc[ia][0];
                        tr[4 * ty + ia][2 * tx + 1] = acc[ia][1];
                    }
                }
            } else {
#pragma unroll
                for (int ia = 0; ia < 4; ++ia) {
                    const int li = 4 * ty + ia;
                    const int64_t i = i0 + li, j = j0 + lj0;
                    if (a.symmetric) {
                        if (i == j) acc[ia][0] += a.diag_add;
                        if (i == j + 1) acc[ia][1] += a.diag_add;
                    }
                    const bool in0 = (i < a.Na) && (j < a.Nb), in1 = (i < a.Na) && (j + 1 < a.Nb);
                    if ((in0 && !isfinite(acc[ia][0])) || (in1 && !isfinite(acc[ia][1]))) bad = 1;
                    const bool tri = a.symmetric && a.lower_only && diag_tile;   // keep only j <= i
                    if (vec_ok && in1 && !(tri && j + 1 > i)) {
                        *reinterpret_cast<double2*>(&a.out[i * a.ldo + j]) = make_double2(acc[ia][0], acc[ia][1]);
                    } else {
                        if (in0 && !(tri && j > i)) a.out[i * a.ldo + j] = acc[ia][0];
                        if (in1 && !(tri && j + 1 > i)) a.out[i * a.ldo + j + 1] = acc[ia][1];
                    }
                    if (mirror) {
                        tr[li][2 * tx] = acc[ia][0];
                        tr[li][2 * tx + 1] = acc[ia][1];
                    }
                }
            }
            if (use_tr) {   // uniform over the grid: the barriers below are reached by every thread
                __syncthreads();
                if (mirror) {
                    for (int e = tid; e < GT * (GT / 2); e += GTHREADS) {
                        const int r = e / GT, cc = e - r * GT;
                        const int64_t j = j0 + pass * (GT / 2) + r, i = i0 + cc;
                        if (j < a.Nb && i < a.Na) a.out[j * a.ldo + i] = tr[cc][r];
                    }
                }
                __syncthreads();
            }
        }
        __syncthreads();   // every warp is done with this column buffer before the next prefetch overwrites it
    }
    if (expmax >= 0x7ff00000) bad = 1;
    if (a.nonfinite && __syncthreads_or(bad) && tid == 0) atomicAdd(a.nonfinite, 1);
}

// ---------------------------------------------------------------------------------------------
// Fused derivative contraction out[p] = sum_ij W_ij dK_p(xa_i, xb_j) (K2: the trace 1/2 tr((alpha alpha^T - C) dK_p)
// of include/Likelihood.h:280-282 and the two sums of the sparse likelihood) in the same row-group / term form as
// gram_rows_kernel. dK is never stored. Every leaf derivative of the family is leaf value x constant x a factor of
// the pair geometry (r^2; S, F for PERIODIC; r^2/f, ln f for RQ), so for a term t = prod_l leaf_l
//     d t / d theta_p = t x dcst_p x factor_p          (p a parameter of a leaf of t)
// and one weighted product g = W t per pair serves every parameter of the term. Per-thread sums live in shared memory
// ([P][256], the thread's own column: no synchronisation), reduced once per CTA in a fixed order (deterministic).
template <int NPER>
__global__ void __launch_bounds__(GTHREADS, 2) gram_dcontract_kernel(GramArgs a, DevProg P, int vec_ok) {
    constexpr int NS = NPER > 0 ? NPER : 1;
    extern __shared__ __align__(16) unsigned char gv_smem[];
    const int d = a.d;
    double* xa = reinterpret_cast<double*>(gv_smem);
    double* xb = xa + d * GT;
    double2* fa = reinterpret_cast<double2*>(xb + 2 * d * GT);
    double2* fb = fa + NPER * d * GT;
    double* accs = reinterpret_cast<double*>(fb + 2 * NPER * d * GT);   // [nparams][GTHREADS]

    const int ti = blockIdx.y;
    const int tj_begin = blockIdx.x * GGROUP;
    int tj_end = min(tj_begin + GGROUP, a.tiles_b);
    if (a.symmetric) tj_end = min(tj_end, ti + 1);
    if (tj_begin >= tj_end) return;   // its partial row stays at the memset zero
    const int64_t i0 = (int64_t)ti * GT;
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    for (int p = 0; p < P.nparams; ++p) accs[p * GTHREADS + tid] = 0.0;

    auto stage_side = [&](const double* X, const double* F, int64_t row0, int64_t nrows, double* xs, double2* fs) {
        for (int e = tid; e < GT * d; e += GTHREADS) {
            const int r = e & (GT - 1), dd = e / GT;
            if (row0 + r < nrows) {
                cp_async8(&xs[dd * GT + r], X + (row0 + r) * d + dd);
#pragma unroll
                for (int q = 0; q < NPER; ++q)
                    cp_async16(&fs[(q * d + dd) * GT + r], F + (((row0 + r) * NPER + q) * d + dd) * 2);
            } else {
                xs[dd * GT + r] = 0.0;
#pragma unroll
                for (int q = 0; q < NPER; ++q) fs[(q * d + dd) * GT + r] = make_double2(1.0, 0.0);
            }
        }
    };
    stage_side(a.Xa, a.Fa, i0, a.Na, xa, fa);
    stage_side(a.Xb, a.Fb, (int64_t)tj_begin * GT, a.Nb, xb, fb);
    asm volatile("cp.async.commit_group;\n" ::);

    for (int tj = tj_begin; tj < tj_end; ++tj) {
        const int buf = (tj - tj_begin) & 1;
        const double* xbc = xb + buf * d * GT;
        const double2* fbc = fb + buf * NPER * d * GT;
        if (tj + 1 < tj_end) {
            stage_side(a.Xb, a.Fb, (int64_t)(tj + 1) * GT, a.Nb, xb + (buf ^ 1) * d * GT, fb + (buf ^ 1) * NPER * d * GT);
            asm volatile("cp.async.commit_group;\n" ::);
            asm volatile("cp.async.wait_group 1;\n" ::);
        } else {
            asm volatile("cp.async.wait_group 0;\n" ::);
        }
        __syncthreads();

        const int64_t j0 = (int64_t)tj * GT;
        const bool diag_tile = a.symmetric && (ti == tj);
        const double wscale = (a.symmetric && !diag_tile) ? 2.0 : 1.0;   // symmetric W and dK: off-diagonal tiles count twice
        const bool full_tile = vec_ok && (i0 + GT <= a.Na) && (j0 + GT <= a.Nb);

        for (int pass = 0; pass < 2; ++pass) {
            const int lj0 = pass * (GT / 2) + 2 * tx;
            double w[4][2];
            if (full_tile) {
                const double* wp = a.S + (i0 + 4 * ty) * a.lds + j0 + lj0;
#pragma unroll
                for (int ia = 0; ia < 4; ++ia) {
                    const double2 t = __ldg(reinterpret_cast<const double2*>(wp + ia * a.lds));
                    w[ia][0] = t.x * wscale;
                    w[ia][1] = t.y * wscale;
                }
            } else {
#pragma unroll
                for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                    for (int ib = 0; ib < 2; ++ib) {
                        const int64_t i = i0 + 4 * ty + ia, j = j0 + lj0 + ib;
                        w[ia][ib] = (i < a.Na && j < a.Nb) ? a.S[i * a.lds + j] * wscale : 0.0;
                    }
            }
            double r2[4][2], S[NS][4][2], F[NS][4][2];
#pragma unroll
            for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                for (int ib = 0; ib < 2; ++ib) {
                    r2[ia][ib] = 0.0;
#pragma unroll
                    for (int q = 0; q < NS; ++q) S[q][ia][ib] = F[q][ia][ib] = 0.0;
                }
            for (int dd = 0; dd < d; ++dd) {
                const double2 a01 = *reinterpret_cast<const double2*>(&xa[dd * GT + 4 * ty]);
                const double2 a23 = *reinterpret_cast<const double2*>(&xa[dd * GT + 4 * ty + 2]);
                const double2 b01 = *reinterpret_cast<const double2*>(&xbc[dd * GT + lj0]);
                const double va[4] = {a01.x, a01.y, a23.x, a23.y};
                const double vb[2] = {b01.x, b01.y};
                double dx[4][2];
#pragma unroll
                for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                    for (int ib = 0; ib < 2; ++ib) {
                        dx[ia][ib] = va[ia] - vb[ib];
                        r2[ia][ib] = fma(dx[ia][ib], dx[ia][ib], r2[ia][ib]);
                    }
#pragma unroll
                for (int q = 0; q < NPER; ++q) {
                    // cos / sin of 2 b dx from the per-point features: S accumulates cos (turned into sum sin^2 below),
                    // F = sum_d dx sin(2 b dx) = sum_d 2 dx cos(b dx) sin(b dx)   (include/Kernel.h:930-941)
                    double2 ca[4], cb[2];
#pragma unroll
                    for (int ia = 0; ia < 4; ++ia) ca[ia] = fa[(q * d + dd) * GT + 4 * ty + ia];
#pragma unroll
                    for (int ib = 0; ib < 2; ++ib) cb[ib] = fbc[(q * d + dd) * GT + lj0 + ib];
#pragma unroll
                    for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                        for (int ib = 0; ib < 2; ++ib) {
                            S[q][ia][ib] = fma(ca[ia].x, cb[ib].x, fma(ca[ia].y, cb[ib].y, S[q][ia][ib]));
                            const double sn = fma(ca[ia].y, cb[ib].x, -(ca[ia].x * cb[ib].y));
                            F[q][ia][ib] = fma(dx[ia][ib], sn, F[q][ia][ib]);
                        }
                }
            }
#pragma unroll
            for (int q = 0; q < NPER; ++q)
#pragma unroll
                for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                    for (int ib = 0; ib < 2; ++ib) S[q][ia][ib] = fmax(0.0, 0.5 * ((double)d - S[q][ia][ib]));

            for (int t = 0; t < P.nterms; ++t) {
                const int k0 = P.term_start[t], k1 = P.term_start[t + 1];
                double g[4][2];
                leaf8<0, NS>(P, P.term_leaf[k0], r2, S, g);
                for (int k = k0 + 1; k < k1; ++k) leaf8<1, NS>(P, P.term_leaf[k], r2, S, g);
                double sg = 0.0;
#pragma unroll
                for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                    for (int ib = 0; ib < 2; ++ib) {
                        g[ia][ib] *= w[ia][ib];
                        sg += g[ia][ib];
                    }
                for (int k = k0; k < k1; ++k) {
                    const int node = P.term_leaf[k];
                    const int op = P.op[node];
                    double* o = accs + (int)P.poff[node] * GTHREADS + tid;
                    const double dc0 = P.dcst[node][0], dc1 = P.dcst[node][1], dc2 = P.dcst[node][2];
                    switch (op) {
                        case GPRB200_K_GAUSSIAN:
                        case GPRB200_K_GAUSSIAN_EXP: {
                            double s0 = 0.0;
#pragma unroll
                            for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                                for (int ib = 0; ib < 2; ++ib) s0 = fma(g[ia][ib], r2[ia][ib], s0);
                            o[0] = fma(dc0, s0, o[0]);
                            o[GTHREADS] = fma(dc1, sg, o[GTHREADS]);
                        } break;
                        case GPRB200_K_WHITE:
                            o[0] = fma(dc0, sg, o[0]);
                            break;
                        case GPRB200_K_RQ: {
                            const double c1 = P.cst[node][1];
                            double s1 = 0.0, s2 = 0.0;
#pragma unroll
                            for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                                for (int ib = 0; ib < 2; ++ib) {
                                    const double f = fma(c1, r2[ia][ib], 1.0);
                                    const double q1 = r2[ia][ib] / f;
                                    s1 = fma(g[ia][ib], q1, s1);
                                    s2 = fma(g[ia][ib], fma(c1, q1, -log(f)), s2);
                                }
                            o[0] = fma(dc0, sg, o[0]);
                            o[GTHREADS] = fma(dc1, s1, o[GTHREADS]);
                            o[2 * GTHREADS] = fma(dc2, s2, o[2 * GTHREADS]);
                        } break;
                        default: {  // PERIODIC [scale, b, sigma]
                            const int slot = P.perslot[node];
                            double s1 = 0.0, s2 = 0.0;
#pragma unroll
                            for (int q = 0; q < NS; ++q)
                                if (slot == q) {
#pragma unroll
                                    for (int ia = 0; ia < 4; ++ia)
#pragma unroll
                                        for (int ib = 0; ib < 2; ++ib) {
                                            s1 = fma(g[ia][ib], F[q][ia][ib], s1);
                                            s2 = fma(g[ia][ib], S[q][ia][ib], s2);
                                        }
                                }
                            o[0] = fma(dc0, sg, o[0]);
                            o[GTHREADS] = fma(dc1, s1, o[GTHREADS]);
                            o[2 * GTHREADS] = fma(dc2, s2, o[2 * GTHREADS]);
                        } break;
                    }
                }
            }
        }
        __syncthreads();   // every warp is done with this column buffer before the next prefetch overwrites it
    }

    // deterministic CTA reduction: warp shuffle, then the 8 warp sums in order
    __shared__ double red[GTHREADS / 32];
    double* out = a.partial + ((int64_t)blockIdx.y * gridDim.x + blockIdx.x) * P.nparams;
    for (int p = 0; p < P.nparams; ++p) {
        double v = accs[p * GTHREADS + tid];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        __syncthreads();
        if ((tid & 31) == 0) red[tid >> 5] = v;
        __syncthreads();
        if (tid == 0) {
            double s = 0.0;
            for (int wv = 0; wv < GTHREADS / 32; ++wv) s += red[wv];
            out[p] = s;
        }
    }
}

template <int NPER>
int launch_gram_dcontract(Ctx* c, const GramArgs& a, const DevProg& P, int vec_ok, dim3 grid, bool* launched) {
    const size_t smem = sizeof(double) * 3 * a.d * GT + sizeof(double2) * 3 * NPER * a.d * GT +
                        sizeof(double) * (size_t)P.nparams * GTHREADS;
    *launched = false;
    if (smem > 200 * 1024) return 0;
    static size_t configured = 0;
    if (smem > configured) {
        GP_CUDA(c, cudaFuncSetAttribute(gram_dcontract_kernel<NPER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    gram_dcontract_kernel<NPER><<<grid, GTHREADS, smem, c->stream>>>(a, P, vec_ok);
    *launched = true;
    return 0;
}

// shared memory of gram_rows_kernel<NPER> for d resident dimensions
template <int NPER>
size_t gram_rows_smem(int d, bool use_tr) {
    return sizeof(double) * 3 * d * GT + sizeof(double2) * 3 * NPER * d * GT + (use_tr ? sizeof(double) * GT * (GT / 2 + 1) : 0);
}

template <int NPER>
int launch_gram_rows(Ctx* c, const GramArgs& a, const DevProg& P, int vec_ok, bool* launched) {
    const bool use_tr = a.symmetric && !a.lower_only;
    const size_t smem = gram_rows_smem<NPER>(a.d, use_tr);
    *launched = false;
    if (smem > 200 * 1024) return 0;   // too many dimensions to keep resident: chunked kernel
    const int64_t ta = (a.Na + GT - 1) / GT;
    if (ta > 65535) return 0;
    static size_t configured = 0;   // per instantiation
    if (smem > configured) {
        GP_CUDA(c, cudaFuncSetAttribute(gram_rows_kernel<NPER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    dim3 grid((unsigned)((a.tiles_b + GGROUP - 1) / GGROUP), (unsigned)ta);
    gram_rows_kernel<NPER><<<grid, GTHREADS, smem, c->stream>>>(a, P, vec_ok, use_tr ? 1 : 0);
    *launched = true;
    return 0;
}

template <int NPER>
int launch_gram_value(Ctx* c, const GramArgs& a, const DevProg& P, int vec_ok, int64_t grid) {
    const size_t smem = sizeof(double) * 2 * GDC * GT + sizeof(double2) * 2 * NPER * GDC * GT + sizeof(double) * GT * (GT / 2 + 1);
    static bool configured = false;   // per instantiation
    if (!configured) {
        GP_CUDA(c, cudaFuncSetAttribute(gram_value_kernel<NPER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = true;
    }
    gram_value_kernel<NPER><<<(unsigned)grid, GTHREADS, smem, c->stream>>>(a, P, vec_ok);
    return 0;
}

// F[row][leaf][dim] = (cos, sin)(2 b_leaf x[row][dim]) for the PERIODIC leaves of the kernel tree
__global__ void periodic_features_kernel(const double* X, int64_t n, int d, DevProg P, double* F) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t total = n * P.nper * d;
    if (e >= total) return;
    const int dd = (int)(e % d);
    const int q = (int)((e / d) % P.nper);
    const int64_t row = e / ((int64_t)d * P.nper);
    double sn, cs;
    sincos(2.0 * P.perb[q] * X[row * d + dd], &sn, &cs);
    F[e * 2] = cs;
    F[e * 2 + 1] = sn;
}

// out[t] = k(xa_t, xb_t): diagonal terms of the predictive variance (lib/GaussianProcess.cpp:98)
__global__ void kpair_kernel(const double* Xa, const double* Xb, int64_t T, int d, double* out, DevProg P) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    Geo<false> g;
    geo_init(g);
    for (int dd = 0; dd < d; ++dd) geo_acc(g, P, Xa[t * d + dd] - Xb[t * d + dd]);
    out[t] = kprog_value(P, g);
}

__global__ void reduce_partials_kernel(const double* partial, int nblocks, int P, double* out) {
    // one warp per parameter; fixed summation order -> bitwise reproducible
    const int p = blockIdx.x;
    double s = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += 32) s += partial[(int64_t)b * P + p];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (threadIdx.x == 0) out[p] = s;
}

__global__ void count_nonfinite_kernel(const double* A, int64_t rows, int64_t cols, int64_t ld, int lower_only,
                                       int* count) {
    int local = 0;
    const int64_t total = rows * cols;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / cols, cidx = e - r * cols;
        if (lower_only && cidx > r) continue;
        if (!isfinite(A[r * ld + cidx])) ++local;
    }
    local = __reduce_add_sync(0xffffffffu, local);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(count, local);
}

// ---------------------------------------------------------------------------------------------
int lower_program(const KProg& in, DevProg* out, std::string* err) {
    DevProg& D = *out;
    D = DevProg{};
    D.nops = in.nops;
    D.nparams = in.nparams;
    int stack[GPRB200_MAX_OPS];
    int sp = 0, poff = 0, nper = 0, max_sp = 0;
    for (int i = 0; i < in.nops; ++i) {
        const int op = in.op[i];
        D.op[i] = (int8_t)op;
        D.poff[i] = (int8_t)poff;
        const double* th = in.par + poff;
        double* c = D.cst[i];
        auto need = [&](int n) -> bool {
            if (poff + n > in.nparams) {
                *err = "set_kernel: parameter vector too short for the kernel program";
                return false;
            }
            return true;
        };
        switch (op) {
            case GPRB200_K_GAUSSIAN: {
                if (!need(2)) return -2;
                const double sg = th[0], s = th[1];
                c[0] = s * s;
                c[1] = -0.5 / (sg * sg);
                c[2] = (s * s) / (sg * sg * sg);
                c[3] = 2.0 * s;
                D.dcst[i][0] = 1.0 / (sg * sg * sg);            // x r^2
                D.dcst[i][1] = (s != 0.0) ? 2.0 / s : 0.0;      // x 1
                poff += 2;
            } break;
            case GPRB200_K_GAUSSIAN_EXP: {
                if (!need(2)) return -2;
                const double sg = th[0], s = th[1];
                const double es = exp(s), eg = exp(sg);
                c[0] = es * es;
                c[1] = -0.5 / (eg * eg);
                c[2] = exp(-2.0 * sg);
                c[3] = (4.0 * sg - 4.0 * s) * exp(2.0 * sg);
                c[4] = 4.0 * exp(2.0 * sg) * s;
                // include/Kernel.h:591-596: d/dsigma = r^2 exp(-2 sigma) k, d/dscale = 2 k
                D.dcst[i][0] = exp(-2.0 * sg);
                D.dcst[i][1] = 2.0;
                poff += 2;
            } break;
            case GPRB200_K_WHITE: {
                if (!need(1)) return -2;
                c[0] = th[0] * th[0];
                c[1] = 2.0 * th[0];
                D.dcst[i][0] = (th[0] != 0.0) ? 2.0 / th[0] : 0.0;
                poff += 1;
            } break;
            case GPRB200_K_RQ: {
                if (!need(3)) return -2;
                const double s = th[0], sg = th[1], al = th[2];
                c[0] = s * s;
                c[1] = 0.5 / (sg * sg * al);
                c[2] = -al;
                c[3] = 2.0 * s;
                c[4] = (s * s) / (sg * sg * sg);
                D.dcst[i][0] = (s != 0.0) ? 2.0 / s : 0.0;      // x 1
                D.dcst[i][1] = 1.0 / (sg * sg * sg);            // x r^2 / f
                D.dcst[i][2] = 1.0;                             // x (r^2 / (2 sg^2 a f) - ln f)
                poff += 3;
            } break;
            case GPRB200_K_PERIODIC: {
                if (!need(3)) return -2;
                if (nper >= MAXPER) {
                    *err = "set_kernel: more than " + std::to_string(MAXPER) + " PeriodicKernel leaves are not supported";
                    return -2;
                }
                const double s = th[0], b = th[1], sg = th[2];
                c[0] = s * s;
                c[1] = -0.5 / (sg * sg);
                c[2] = 2.0 * s;
                c[3] = (s * s) / (sg * sg * sg);
                D.dcst[i][0] = (s != 0.0) ? 2.0 / s : 0.0;      // x 1
                D.dcst[i][1] = -0.5 / (sg * sg);                // x F = sum_d dx sin(2 b dx)
                D.dcst[i][2] = 1.0 / (sg * sg * sg);            // x S = sum_d sin^2(b dx)
                D.perb[nper] = b;
                D.perslot[i] = (int8_t)nper;
                ++nper;
                poff += 3;
            } break;
            case GPRB200_K_SUM:
            case GPRB200_K_PRODUCT: {
                if (sp < 2) {
                    *err = "set_kernel: malformed kernel program (binary op with < 2 operands)";
                    return -2;
                }
                D.rchild[i] = (int8_t)stack[--sp];
                D.lchild[i] = (int8_t)stack[--sp];
            } break;
            default:
                *err = "set_kernel: unknown kernel op " + std::to_string(op);
                return -2;
        }
        stack[sp++] = i;
        if (sp > max_sp) max_sp = sp;
    }
    if (sp != 1) {
        *err = "set_kernel: malformed kernel program (does not reduce to one kernel)";
        return -2;
    }
    if (poff != in.nparams) {
        *err = "set_kernel: wrong number of parameters for the kernel program";
        return -2;
    }
    if (max_sp > GPRB200_MAX_STACK) {
        *err = "set_kernel: kernel tree needs an evaluation stack deeper than " + std::to_string(GPRB200_MAX_STACK);
        return -2;
    }
    D.nper = nper;
    // expand the tree into a sum of products of leaves (distributing products over sums)
    {
        typedef std::vector<std::vector<int>> Poly;   // list of terms, each a list of leaf op indices
        std::vector<Poly> st;
        bool ok = true;
        for (int i = 0; i < in.nops && ok; ++i) {
            const int op = in.op[i];
            if (op == GPRB200_K_SUM || op == GPRB200_K_PRODUCT) {
                Poly b = st.back(); st.pop_back();
                Poly a = st.back(); st.pop_back();
                Poly r;
                if (op == GPRB200_K_SUM) {
                    r = a;
                    r.insert(r.end(), b.begin(), b.end());
                } else {
                    for (const auto& ta : a)
                        for (const auto& tb : b) {
                            std::vector<int> t = ta;
                            t.insert(t.end(), tb.begin(), tb.end());
                            r.push_back(t);
                        }
                }
                size_t leaves = 0;
                for (const auto& t : r) leaves += t.size();
                if (r.size() > (size_t)MAX_TERMS || leaves > (size_t)MAX_TERM_LEAVES) ok = false;
                st.push_back(r);
            } else {
                st.push_back(Poly{{i}});
            }
        }
        D.nterms = 0;
        if (ok && st.size() == 1) {
            int k = 0, t = 0;
            for (const auto& term : st[0]) {
                D.term_start[t++] = (int8_t)k;
                for (int leaf : term) D.term_leaf[k++] = (int8_t)leaf;
            }
            D.term_start[t] = (int8_t)k;
            D.nterms = t;
        }
    }
    return 0;
}

static int64_t sym_tiles(int64_t t) { return t * (t + 1) / 2; }

int gram_build(Ctx* c, const double* Xa, int64_t Na, const double* Xb, int64_t Nb, double* out, int64_t ldo,
               bool symmetric, bool lower_only, double diag_add, int* d_nonfinite) {
    DevProg P;
    if (int r = lower_program(c->prog, &P, &c->err)) return r;
    {   // large input dimension: cross term of r^2 on the FP64 tensor pipe (gram_mma.cu); 1 = does not apply
        const int r = gram_build_mma(c, P, Xa, Na, Xb, Nb, out, ldo, symmetric, lower_only, diag_add, d_nonfinite);
        if (r <= 0) return r;
    }
    GramArgs a{};
    a.Xa = Xa; a.Xb = Xb; a.Na = Na; a.Nb = Nb; a.d = c->d;
    a.out = out; a.ldo = ldo;
    a.symmetric = symmetric; a.lower_only = lower_only; a.diag_add = diag_add; a.p_select = -1;
    a.nonfinite = d_nonfinite;
    const int64_t ta = (Na + GT - 1) / GT, tb = (Nb + GT - 1) / GT;
    a.tiles_b = (int)tb;
    const int64_t grid = symmetric ? sym_tiles(ta) : ta * tb;
    if (grid <= 0) return 0;
    double* feat = nullptr;
    if (P.nper > 0) {
        const int64_t per_row = (int64_t)P.nper * c->d * 2;
        const bool same = (Xa == Xb && Na == Nb);
        GP_CUDA(c, cudaMallocAsync((void**)&feat, sizeof(double) * per_row * (Na + (same ? 0 : Nb)), c->stream));
        const int64_t ea = Na * P.nper * c->d;
        periodic_features_kernel<<<(unsigned)((ea + 255) / 256), 256, 0, c->stream>>>(Xa, Na, c->d, P, feat);
        ++c->launches;
        a.Fa = feat;
        a.Fb = feat;
        if (!same) {
            const int64_t eb = Nb * P.nper * c->d;
            a.Fb = feat + per_row * Na;
            periodic_features_kernel<<<(unsigned)((eb + 255) / 256), 256, 0, c->stream>>>(Xb, Nb, c->d, P,
                                                                                          feat + per_row * Na);
            ++c->launches;
        }
        GP_CUDA(c, cudaGetLastError());
    }
    const int vec_ok = ((ldo & 1) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) ? 1 : 0;
    bool launched = false;
    if (P.nterms > 0 && c->opt_gram_terms >= 2) {
        switch (P.nper) {
            case 0: GP_TRY(launch_gram_rows<0>(c, a, P, vec_ok, &launched)); break;
            case 1: GP_TRY(launch_gram_rows<1>(c, a, P, vec_ok, &launched)); break;
            case 2: GP_TRY(launch_gram_rows<2>(c, a, P, vec_ok, &launched)); break;
            case 3: GP_TRY(launch_gram_rows<3>(c, a, P, vec_ok, &launched)); break;
            default: GP_TRY(launch_gram_rows<4>(c, a, P, vec_ok, &launched)); break;
        }
    }
    if (launched) {
    } else if (P.nterms > 0 && c->opt_gram_terms) {
        switch (P.nper) {
            case 0: GP_TRY(launch_gram_value<0>(c, a, P, vec_ok, grid)); break;
            case 1: GP_TRY(launch_gram_value<1>(c, a, P, vec_ok, grid)); break;
            case 2: GP_TRY(launch_gram_value<2>(c, a, P, vec_ok, grid)); break;
            case 3: GP_TRY(launch_gram_value<3>(c, a, P, vec_ok, grid)); break;
            default: GP_TRY(launch_gram_value<4>(c, a, P, vec_ok, grid)); break;
        }
    } else {
        gram_kernel<false><<<(unsigned)grid, GTHREADS, 0, c->stream>>>(a, P);   // interpreter: any tree
    }
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    if (feat) GP_CUDA(c, cudaFreeAsync(feat, c->stream));
    return 0;
}

int kpair_values(Ctx* c, const double* Xa, const double* Xb, int64_t T, double* d_out) {
    DevProg P;
    if (int r = lower_program(c->prog, &P, &c->err)) return r;
    kpair_kernel<<<(unsigned)((T + 127) / 128), 128, 0, c->stream>>>(Xa, Xb, T, c->d, d_out, P);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}

int dgram_build_one(Ctx* c, double* out, int64_t ldo, int p_select);
int dgram_build(Ctx* c, double* out, int64_t ldo) { return dgram_build_one(c, out, ldo, -1); }

// dK_p between two point sets (Xa == Xb, Na == Nb: symmetric, mirrored). p_select < 0 stores the whole stack
// ((P*Na) x Nb, block p at rows [p Na, (p+1) Na)), else parameter p only (Na x Nb).
int dgram_build_xy(Ctx* c, const double* Xa, int64_t Na, const double* Xb, int64_t Nb, double* out, int64_t ldo,
                   int p_select) {
    DevProg P;
    if (int r = lower_program(c->prog, &P, &c->err)) return r;
    if (p_select >= P.nparams) return fail(c, -2, "dgram: parameter index out of range");
    const bool symmetric = (Xa == Xb && Na == Nb);
    GramArgs a{};
    a.Xa = Xa; a.Xb = Xb; a.Na = Na; a.Nb = Nb; a.d = c->d;
    a.out = out; a.ldo = ldo; a.symmetric = symmetric ? 1 : 0; a.p_select = p_select;
    const int64_t ta = (Na + GT - 1) / GT, tb = (Nb + GT - 1) / GT;
    a.tiles_b = (int)tb;
    const int64_t grid = symmetric ? sym_tiles(ta) : ta * tb;
    if (grid <= 0) return 0;
    gram_kernel<true><<<(unsigned)grid, GTHREADS, 0, c->stream>>>(a, P);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}

int dgram_build_one(Ctx* c, double* out, int64_t ldo, int p_select) {
    return dgram_build_xy(c, c->dX, c->N, c->dX, c->N, out, ldo, p_select);
}

// out[p] = sum_ij W_ij dK_p(xa_i, xb_j); Xa == Xb (Na == Nb) takes W symmetric and visits the lower tiles only.
int dgram_contract_xy(Ctx* c, const double* Xa, int64_t Na, const double* Xb, int64_t Nb, const double* W, int64_t ldw,
                      double* out_host) {
    DevProg P;
    if (int r = lower_program(c->prog, &P, &c->err)) return r;
    const bool symmetric = (Xa == Xb && Na == Nb);
    const int64_t ta = (Na + GT - 1) / GT, tb = (Nb + GT - 1) / GT;
    GramArgs a{};
    a.Xa = Xa; a.Xb = Xb; a.Na = Na; a.Nb = Nb; a.d = c->d;
    a.symmetric = symmetric ? 1 : 0; a.S = W; a.lds = ldw; a.tiles_b = (int)tb; a.p_select = -1;

    if (P.nterms > 0 && c->opt_gram_terms >= 2 && ta <= 65535) {
        // row-group term-form kernel (gram_dcontract_kernel)
        dim3 grid((unsigned)((tb + GGROUP - 1) / GGROUP), (unsigned)ta);
        const int64_t nblocks = (int64_t)grid.x * grid.y;
        double* partial = nullptr;
        double* feat = nullptr;
        GP_CUDA(c, cudaMallocAsync(&partial, sizeof(double) * (nblocks + 1) * P.nparams, c->stream));
        GP_CUDA(c, cudaMemsetAsync(partial, 0, sizeof(double) * nblocks * P.nparams, c->stream));
        if (P.nper > 0) {
            const int64_t per_row = (int64_t)P.nper * c->d * 2;
            GP_CUDA(c, cudaMallocAsync((void**)&feat, sizeof(double) * per_row * (Na + (symmetric ? 0 : Nb)), c->stream));
            periodic_features_kernel<<<(unsigned)((Na * P.nper * c->d + 255) / 256), 256, 0, c->stream>>>(Xa, Na, c->d, P, feat);
            ++c->launches;
            a.Fa = a.Fb = feat;
            if (!symmetric) {
                a.Fb = feat + per_row * Na;
                periodic_features_kernel<<<(unsigned)((Nb * P.nper * c->d + 255) / 256), 256, 0, c->stream>>>(
                    Xb, Nb, c->d, P, feat + per_row * Na);
                ++c->launches;
            }
        }
        a.partial = partial;
        const int vec_ok = ((ldw & 1) == 0 && (reinterpret_cast<uintptr_t>(W) & 15) == 0) ? 1 : 0;
        bool launched = false;
        switch (P.nper) {
            case 0: GP_TRY(launch_gram_dcontract<0>(c, a, P, vec_ok, grid, &launched)); break;
            case 1: GP_TRY(launch_gram_dcontract<1>(c, a, P, vec_ok, grid, &launched)); break;
            case 2: GP_TRY(launch_gram_dcontract<2>(c, a, P, vec_ok, grid, &launched)); break;
            case 3: GP_TRY(launch_gram_dcontract<3>(c, a, P, vec_ok, grid, &launched)); break;
            default: GP_TRY(launch_gram_dcontract<4>(c, a, P, vec_ok, grid, &launched)); break;
        }
        if (launched) {
            ++c->launches;
            GP_CUDA(c, cudaGetLastError());
            double* dres = partial + nblocks * P.nparams;
            reduce_partials_kernel<<<P.nparams, 32, 0, c->stream>>>(partial, (int)nblocks, P.nparams, dres);
            ++c->launches;
            GP_CUDA(c, cudaGetLastError());
            GP_CUDA(c, cudaMemcpyAsync(out_host, dres, sizeof(double) * P.nparams, cudaMemcpyDeviceToHost, c->stream));
            GP_CUDA(c, cudaStreamSynchronize(c->stream));
        }
        if (feat) GP_CUDA(c, cudaFreeAsync(feat, c->stream));
        GP_CUDA(c, cudaFreeAsync(partial, c->stream));
        if (launched) return 0;
    }

    // per-entry interpreter (any tree, any d)
    const int64_t grid = symmetric ? sym_tiles(ta) : ta * tb;
    double* partial = nullptr;
    GP_CUDA(c, cudaMallocAsync(&partial, sizeof(double) * (grid + 1) * P.nparams, c->stream));
    a.partial = partial;
    gram_kernel<true><<<(unsigned)grid, GTHREADS, 0, c->stream>>>(a, P);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    double* dres = partial + grid * P.nparams;
    reduce_partials_kernel<<<P.nparams, 32, 0, c->stream>>>(partial, (int)grid, P.nparams, dres);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    GP_CUDA(c, cudaMemcpyAsync(out_host, dres, sizeof(double) * P.nparams, cudaMemcpyDeviceToHost, c->stream));
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    GP_CUDA(c, cudaFreeAsync(partial, c->stream));
    return 0;
}

int dgram_contract(Ctx* c, const double* S, int64_t lds, double* out_host) {
    return dgram_contract_xy(c, c->dX, c->N, c->dX, c->N, S, lds, out_host);
}

int check_finite(Ctx* c, const double* A, int64_t rows, int64_t cols, int64_t ld, bool lower_only, int* bad_host) {
    GP_CUDA(c, cudaMemsetAsync(c->dInfo + 1, 0, sizeof(int), c->stream));
    count_nonfinite_kernel<<<148 * 8, 256, 0, c->stream>>>(A, rows, cols, ld, lower_only ? 1 : 0, c->dInfo + 1);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    GP_CUDA(c, cudaMemcpyAsync(bad_host, c->dInfo + 1, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    GP_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

}  // namespace gprb200
