// Device-side evaluation of the reference's kernel tree (include/Kernel.h) as a flat post-order
// program.  One evaluator for Gram (K1), cross-Gram (K6) and the parameter derivatives (K2).
//
// Arithmetic restated from the reference (file:line in /root/reference):
//   GaussianKernel     k = s^2 exp(-0.5 r^2 / sg^2)                        include/Kernel.h:465-468
//                      dk = [ s^2 r^2 / sg^3 * f , 2 s f ]                  include/Kernel.h:471-479
//   GaussianExpKernel  (log-space parameters)                               include/Kernel.h:580-598
//   WhiteKernel        k = s^2 [ ||x-y|| == 0 ],  dk = 2 s [..]             include/Kernel.h:695-713
//   RationalQuadratic  k = s^2 (1 + r^2/(2 a sg^2))^-a                      include/Kernel.h:794-808
//   PeriodicKernel     k = s^2 exp(-0.5 sum_d sin^2(b dx_d) / sg^2)         include/Kernel.h:912-948
//   Sum / Product      k1+k2, [dk1; dk2]  /  k1*k2, [dk1*k2; dk2*k1]        include/Kernel.h:165-178, 314-327
// r^2 is accumulated as sum_d dx_d^2 (the reference squares Eigen's norm(); the two differ by
// rounding only, and Eigen's own summation order is not reproducible anyway).
#pragma once
#include "common.cuh"

namespace gprb200 {

constexpr int MAXPER = 4;  // PERIODIC leaves per kernel tree handled in one pass
constexpr int MAX_TERMS = 12;        // expanded sum-of-products form of the kernel value
constexpr int MAX_TERM_LEAVES = 24;  // leaf evaluations over all terms

// Device form of the program: per-op derived constants are precomputed on the host in fp64.
struct DevProg {
    int nops;
    int nparams;
    int nper;
    int pad_;
    int8_t op[GPRB200_MAX_OPS];
    int8_t poff[GPRB200_MAX_OPS];    // first parameter index of a leaf
    int8_t lchild[GPRB200_MAX_OPS];  // binary ops: node index of the left / right operand
    int8_t rchild[GPRB200_MAX_OPS];
    int8_t perslot[GPRB200_MAX_OPS];  // PERIODIC leaf -> slot in the per-pair S/F accumulators
    double cst[GPRB200_MAX_OPS][6];
    double perb[MAXPER];  // b of each periodic leaf
    // Term form of the value: k = sum_t prod_{l in term t} leaf_l, the Sum/Product tree expanded on the host
    // (lower_program). nterms == 0: the expansion was too large, the value kernels fall back to the interpreter.
    int nterms;
    int8_t term_start[MAX_TERMS + 1];   // term t = term_leaf[term_start[t] .. term_start[t+1])
    int8_t term_leaf[MAX_TERM_LEAVES];  // op index of the leaf (its constants are cst[that index])
    // d leaf / d theta_p = leaf value x dcst[p] x factor_p(r^2, S, F): every leaf derivative of the family is
    // proportional to the leaf value (gram_dcontract_kernel); 0 where the scale parameter is 0
    double dcst[GPRB200_MAX_OPS][3];
};

// geometry of one (x, y) pair
template <bool DERIV>
struct Geo {
    double r2;
    double S[MAXPER];                   // sum_d sin^2(b dx_d)
    double F[DERIV ? MAXPER : 1];       // sum_d 2 dx_d cos(b dx_d) sin(b dx_d)
};

template <bool DERIV>
__device__ __forceinline__ void geo_init(Geo<DERIV>& g) {
    g.r2 = 0.0;
#pragma unroll
    for (int q = 0; q < MAXPER; ++q) g.S[q] = 0.0;
    if (DERIV) {
#pragma unroll
        for (int q = 0; q < MAXPER; ++q) g.F[DERIV ? q : 0] = 0.0;
    }
}

template <bool DERIV>
__device__ __forceinline__ void geo_acc_r2(Geo<DERIV>& g, double dx) {
    g.r2 = fma(dx, dx, g.r2);
}

template <bool DERIV>
__device__ __forceinline__ void geo_acc(Geo<DERIV>& g, const DevProg& P, double dx) {
    g.r2 = fma(dx, dx, g.r2);
#pragma unroll
    for (int q = 0; q < MAXPER; ++q) {
        if (q < P.nper) {
            if (DERIV) {
                double sn, cs;
                sincos(P.perb[q] * dx, &sn, &cs);
                g.S[q] = fma(sn, sn, g.S[q]);
                g.F[DERIV ? q : 0] += 2.0 * dx * cs * sn;
            } else {
                const double sn = sin(P.perb[q] * dx);
                g.S[q] = fma(sn, sn, g.S[q]);
            }
        }
    }
}

__device__ __forceinline__ double sel_per(const double* S, int slot) {
    double v = S[0];
#pragma unroll
    for (int q = 1; q < MAXPER; ++q) v = (slot == q) ? S[q] : v;
    return v;
}

// value of one leaf
__device__ __forceinline__ double leaf_value(int op, const double* c, double r2, double S) {
    switch (op) {
        case GPRB200_K_GAUSSIAN:
        case GPRB200_K_GAUSSIAN_EXP:
            return c[0] * exp(c[1] * r2);
        case GPRB200_K_WHITE:
            return (r2 == 0.0) ? c[0] : 0.0;
        case GPRB200_K_RQ:
            return c[0] * pow(fma(c[1], r2, 1.0), c[2]);
        default:  // PERIODIC
            return c[0] * exp(c[1] * S);
    }
}

// k(x,y) from the pair geometry; shift-register stack keeps everything in registers.
template <bool DERIV>
__device__ __forceinline__ double kprog_value(const DevProg& P, const Geo<DERIV>& g) {
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0, s5 = 0, s6 = 0, s7 = 0;
    for (int i = 0; i < P.nops; ++i) {
        const int op = P.op[i];
        if (op >= GPRB200_K_SUM) {
            const double v = (op == GPRB200_K_SUM) ? (s1 + s0) : (s1 * s0);
            s0 = v; s1 = s2; s2 = s3; s3 = s4; s4 = s5; s5 = s6; s6 = s7;
        } else {
            const double Sq = (op == GPRB200_K_PERIODIC) ? sel_per(g.S, P.perslot[i]) : 0.0;
            const double v = leaf_value(op, P.cst[i], g.r2, Sq);
            s7 = s6; s6 = s5; s5 = s4; s4 = s3; s3 = s2; s2 = s1; s1 = s0; s0 = v;
        }
    }
    return s0;
}

// k(x,y) and dk/dtheta_p (p < P.nparams) into out[]; returns k.
__device__ __forceinline__ double kprog_deriv(const DevProg& P, const Geo<true>& g, double* out) {
    double val[GPRB200_MAX_OPS];
    double cof[GPRB200_MAX_OPS];
    for (int i = 0; i < P.nops; ++i) {
        const int op = P.op[i];
        if (op >= GPRB200_K_SUM) {
            const double a = val[P.lchild[i]], b = val[P.rchild[i]];
            val[i] = (op == GPRB200_K_SUM) ? (a + b) : (a * b);
        } else {
            const double Sq = (op == GPRB200_K_PERIODIC) ? sel_per(g.S, P.perslot[i]) : 0.0;
            val[i] = leaf_value(op, P.cst[i], g.r2, Sq);
        }
    }
    cof[P.nops - 1] = 1.0;
    for (int i = P.nops - 1; i >= 0; --i) {
        const int op = P.op[i];
        const double ci = cof[i];
        const double* c = P.cst[i];
        double* o = out + P.poff[i];
        switch (op) {
            case GPRB200_K_SUM:
                cof[P.lchild[i]] = ci;
                cof[P.rchild[i]] = ci;
                break;
            case GPRB200_K_PRODUCT:
                cof[P.lchild[i]] = ci * val[P.rchild[i]];
                cof[P.rchild[i]] = ci * val[P.lchild[i]];
                break;
            case GPRB200_K_GAUSSIAN: {
                const double f = exp(c[1] * g.r2);
                o[0] = ci * (c[2] * g.r2 * f);  // d/dsigma = s^2 r^2 / sigma^3 f
                o[1] = ci * (c[3] * f);         // d/dscale = 2 s f
            } break;
            case GPRB200_K_GAUSSIAN_EXP: {
                // include/Kernel.h:591-596 with f1 = exp(-2 sigma) = c[2], (4 sigma - 4 scale) f2 = c[3], 4 f2 scale = c[4]
                o[0] = ci * (g.r2 * exp(-0.5 * c[2] * (c[3] + g.r2)));
                o[1] = ci * (2.0 * exp(0.5 * c[2] * (c[4] - g.r2)));
            } break;
            case GPRB200_K_WHITE:
                o[0] = ci * ((g.r2 == 0.0) ? c[1] : 0.0);
                break;
            case GPRB200_K_RQ: {
                const double f = fma(c[1], g.r2, 1.0);
                const double pw = pow(f, c[2]);
                o[0] = ci * (c[3] * pw);                                   // 2 s f^-a
                o[1] = ci * (c[4] * g.r2 * (pw / f));                      // s^2 r^2 f^(-a-1) / sigma^3
                o[2] = ci * (c[0] * ((c[1] * g.r2 / f) - log(f)) * pw);    // s^2 (r^2/(2 sg^2 f a) - ln f) f^-a
            } break;
            default: {  // PERIODIC [scale, b, sigma]
                const int q = P.perslot[i];
                const double S = sel_per(g.S, q), F = sel_per(g.F, q);
                const double e = exp(c[1] * S);
                o[0] = ci * (c[2] * e);            // 2 s e
                o[1] = ci * (c[1] * c[0] * e * F); // -0.5 s^2 e F / sigma^2
                o[2] = ci * (c[3] * e * S);        // s^2 e S / sigma^3
            } break;
        }
    }
    return val[P.nops - 1];
}

// host: lower the public KProg into the device form (derived constants in fp64)
int lower_program(const KProg& in, DevProg* out, std::string* err);

}  // namespace gprb200
