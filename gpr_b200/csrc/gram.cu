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
                poff += 2;
            } break;
            case GPRB200_K_WHITE: {
                if (!need(1)) return -2;
                c[0] = th[0] * th[0];
                c[1] = 2.0 * th[0];
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
    return 0;
}

static int64_t sym_tiles(int64_t t) { return t * (t + 1) / 2; }

int gram_build(Ctx* c, const double* Xa, int64_t Na, const double* Xb, int64_t Nb, double* out, int64_t ldo,
               bool symmetric, bool lower_only, double diag_add, int* d_nonfinite) {
    DevProg P;
    if (int r = lower_program(c->prog, &P, &c->err)) return r;
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
    gram_kernel<false><<<(unsigned)grid, GTHREADS, 0, c->stream>>>(a, P);
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
    const int64_t grid = symmetric ? sym_tiles(ta) : ta * tb;
    double* partial = nullptr;
    GP_CUDA(c, cudaMallocAsync(&partial, sizeof(double) * (grid + 1) * P.nparams, c->stream));
    GramArgs a{};
    a.Xa = Xa; a.Xb = Xb; a.Na = Na; a.Nb = Nb; a.d = c->d;
    a.symmetric = symmetric ? 1 : 0; a.S = W; a.lds = ldw; a.partial = partial; a.tiles_b = (int)tb; a.p_select = -1;
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
