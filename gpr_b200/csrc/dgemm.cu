// fp64 tensor-core GEMM for sm_100a: C = alpha * op(A) * op(B) + beta * C, everything row-major.
//
// This is the work-horse behind the blocked Cholesky (SYRK trailing update and the TRSM-as-GEMM
// panel solve), the triangular solves for alpha = (K + sigma^2 I)^-1 Y and batched prediction.
// It replaces the reference's LAPACK / Eigen products (include/LAPACKUtils.h:46-50,
// lib/GaussianProcess.cpp:661, :60).
//
// Hardware mapping (measured on B200, profiles/r01_fp64_probe.txt): every f64 mma.sync shape
// lowers to DMMA.8x8x4, which issues at 37.1 TFLOP/s chip-wide (128 flop/clk/SM @ 1.965 GHz);
// tcgen05 has no f64 kind, so this warp-level path IS the FP64 tensor path on sm_100a.
// One warp owns a 64x32 accumulator tile (8x4 DMMA atoms, 64 doubles/thread); per k4 step it
// loads 8 A + 4 B fragments (12 LDS.64) for 32 DMMAs, i.e. shared-memory traffic is ~20 % of the
// crossbar while the tensor pipe is saturated.
//
// Operand forms (K = contraction index):
//   NT: A is M x K (k contiguous), B is N x K (k contiguous)        C = A B^T   (SYRK, panel solve)
//   NN: A is M x K (k contiguous), B is K x N (n contiguous)        C = A B     (forward solve, predict)
//   TN: A is K x M (m contiguous), B is K x N (n contiguous)        C = A^T B   (backward solve)
// Global->shared: 16-byte cp.async (LDGSTS) with zero-fill for ragged edges, STAGES-deep ring.
// A 64-bit LDS is served per HALF-warp (16 lanes x 8 B = one 128-byte wavefront if conflict-free); a half-warp
// of the 8x4 DMMA fragment is 4 rows x 4 consecutive k = 4 x 32 B. k-contiguous tiles are stored [row][16] with
// the 32-byte group index XOR-swizzled by (row & 3), m/n-contiguous tiles [16][rows + 4] (k-rows 32 B apart
// mod 128), so those four 32-byte pieces always fall in four different bank groups: 2 wavefronts per LDS, the
// minimum. (Round-1 ncu showed 4 per LDS with a (row & 7) 16-byte swizzle and the shared-memory pipe at 90 %.)
#include "common.cuh"
#include "gemm_tile_map.cuh"

namespace gprb200 {

constexpr int BK = 16;

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, int src_bytes) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

// tile geometry helpers ------------------------------------------------------------------------
template <int ROWS, bool KC>
struct OperandTile {
    static constexpr int LD = KC ? BK : (ROWS + 4);
    static constexpr int ELEMS = KC ? ROWS * BK : BK * (ROWS + 4);
    // element (r = row in tile, k = 0..15)
    __device__ static __forceinline__ int idx(int r, int k) {
        if (KC) return r * BK + ((((k >> 1) ^ ((r & 3) << 1)) << 1) | (k & 1));
        return k * LD + r;
    }
    // cooperative async load of one k-tile. base points at the operand's (row 0, k 0) element.
    // rows_valid / k_valid: remaining extent from the tile origin (may be <= 0 or > tile size).
    template <int THREADS>
    __device__ static __forceinline__ void load(double* tile, const double* base, int64_t ld, int rows_valid,
                                                int k_valid, int tid) {
        if (KC) {
            constexpr int CHUNKS = ROWS * (BK / 2);
#pragma unroll
            for (int c = tid; c < CHUNKS; c += THREADS) {
                const int r = c >> 3, ch = c & 7;
                const int k = ch * 2;
                int bytes = 0;
                if (r < rows_valid) bytes = min(max((k_valid - k) * 8, 0), 16);
                const double* src = bytes ? (base + (int64_t)r * ld + k) : base;
                cp_async16(tile + r * BK + ((ch ^ ((r & 3) << 1)) << 1), src, bytes);
            }
        } else {
            constexpr int CPR = ROWS / 2;  // chunks per k-row
            constexpr int CHUNKS = BK * CPR;
#pragma unroll
            for (int c = tid; c < CHUNKS; c += THREADS) {
                const int k = c / CPR, ch = c - k * CPR;
                const int r = ch * 2;
                int bytes = 0;
                if (k < k_valid) bytes = min(max((rows_valid - r) * 8, 0), 16);
                const double* src = bytes ? (base + (int64_t)k * ld + r) : base;
                cp_async16(tile + k * LD + r, src, bytes);
            }
        }
    }
};

__device__ __forceinline__ void cp_async16_full(unsigned smem_addr, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(smem_addr), "l"(gmem));
}

// Branch-free loader for interior tiles (all rows valid, k-tile fully inside K): each thread's chunks lie on an
// arithmetic progression in global memory (stride gstride doubles) and in shared memory (stride sstride bytes),
// so a k-tile is NI back-to-back LDGSTS with no predicates. The zero-filling OperandTile::load above handles edges.
// (Round-1 SASS: the predicated path compiled to one branch diamond per 16-byte chunk, 12 per thread and k-tile,
// fencing the DMMA stream at every k-tile.)
template <int ROWS, bool KC, int THREADS>
struct FastLoader {
    static constexpr int CPR = KC ? (BK / 2) : (ROWS / 2);          // 16-byte chunks per tile row (KC) / per k-row
    static constexpr int LINES = KC ? ROWS : BK;                      // rows (KC) or k-rows walked by the progression
    static constexpr int STEP = THREADS / CPR;                        // lines advanced per chunk index i
    static constexpr int NI = LINES / STEP;
    static_assert(THREADS % CPR == 0 && LINES % STEP == 0, "tile does not split evenly over the CTA");
    const double* src;   // chunk 0 of k-tile 0
    int64_t gstride;     // doubles between consecutive chunks of this thread
    unsigned dst;        // shared address of chunk 0 in stage 0
    __device__ __forceinline__ void init(const double* base, int64_t ld, double* tile0, int tid) {
        using T = OperandTile<ROWS, KC>;
        const int line = tid / CPR, ch = tid - line * CPR;
        if (KC) {
            src = base + (int64_t)line * ld + ch * 2;
            dst = (unsigned)__cvta_generic_to_shared(tile0 + line * BK + ((ch ^ ((line & 3) << 1)) << 1));
            static_assert(!KC || (STEP % 4 == 0), "swizzle phase must be constant along the progression");
        } else {
            src = base + (int64_t)line * ld + ch * 2;
            dst = (unsigned)__cvta_generic_to_shared(tile0 + line * T::LD + ch * 2);
        }
        gstride = (int64_t)STEP * ld;
    }
    // kofs: element offset of this k-tile from k-tile 0 (k0 for KC, k0 * ld otherwise); stage_bytes: ring offset
    __device__ __forceinline__ void issue(int64_t kofs, unsigned stage_bytes) const {
        using T = OperandTile<ROWS, KC>;
        constexpr unsigned SSTRIDE = (KC ? STEP * BK : STEP * T::LD) * sizeof(double);
        const double* p = src + kofs;
#pragma unroll
        for (int i = 0; i < NI; ++i) cp_async16_full(dst + stage_bytes + i * SSTRIDE, p + i * gstride);
    }
};

// STAGES k-tiles of BK live in the shared-memory ring; KSTEP of them are consumed per CTA barrier
// (KSTEP = 2 halves the barrier rate at the same BK = 16 tile geometry).
// EXT = false is the Cholesky / solve hot path: no batch index, no triangular k range (the 64x16 variant sits exactly at
// its 128-register cap; carrying the extra strides and offsets cost 3 % of the factorisation when it was unconditional).
template <int BM, int BN, int WM, int WN, bool AKC, bool BKC, int STAGES, int MINB, int KSTEP = 1, bool EXT = false>
__global__ void __launch_bounds__((BM / WM) * (BN / WN) * 32, MINB) dgemm_kernel(GemmArgs g) {
    // (the tile counts are recomputed here rather than passed: the parameter block stays at sizeof(GemmArgs) = 128 bytes)
    const int tiles_m = (g.M + BM - 1) / BM, tiles_n = (g.N + BN - 1) / BN;
    constexpr int THREADS = (BM / WM) * (BN / WN) * 32;
    constexpr int WARPS_N = BN / WN;
    constexpr int FM = WM / 8, FN = WN / 8;
    using TA = OperandTile<BM, AKC>;
    using TB = OperandTile<BN, BKC>;
    extern __shared__ __align__(128) double smem[];
    double* sA = smem;
    double* sB = smem + STAGES * TA::ELEMS;

    int ti, tj;
    if (!map_tile<BM, BN>(blockIdx.x, g.lower, tiles_m, tiles_n, ti, tj)) return;
    const int m0 = ti * BM, n0 = tj * BN;
    if (m0 >= g.M || n0 >= g.N) return;
    if (g.ktri == 5) {   // staircase C: nothing of this tile is below the stair of its columns
        const int* stair = g.stair();
        int first = stair[n0 >> 6];
        if (BN > 64 && n0 + 64 < g.N) first = min(first, stair[(n0 >> 6) + 1]);
        if (g.stair_row0() + m0 + BM <= first) return;
    }

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm0 = (warp / WARPS_N) * WM, wn0 = (warp % WARPS_N) * WN;
    const int grp = lane >> 2, t4 = lane & 3;

    const int64_t z = EXT ? blockIdx.y : 0;  // batch index
    // triangular operands: skip the leading k range that is known to be zero (whole BK steps only)
    const int kskip = EXT ? (g.ktri == 1 ? n0
                                         : (g.ktri == 2 ? max(m0, n0)
                                                        : (g.ktri == 3 ? (g.tri_first + (n0 / g.tri_cb) * g.tri_stride) * g.tri_cb + n0 % g.tri_cb
                                                                       : 0))) / BK * BK
                          : 0;
    // ktri == 4: B (N x K, NT) is lower triangular, B[n][k] = 0 for k > n: the k loop ENDS after the tile's last column
    const int Kt = EXT ? (g.ktri == 4 ? min(g.K, n0 + BN) : g.K) - kskip : g.K;  // this tile's contraction depth
    const double* Abase = AKC ? (g.A + (int64_t)m0 * g.lda) : (g.A + m0);
    const double* Bbase = BKC ? (g.B + (int64_t)n0 * g.ldb) : (g.B + n0);
    if (EXT) {
        Abase += (AKC ? (int64_t)kskip : (int64_t)kskip * g.lda) + z * g.sA;
        Bbase += (BKC ? (int64_t)kskip : (int64_t)kskip * g.ldb) + z * g.sB;
    }
    const int mrem = g.M - m0, nrem = g.N - n0;
    const int nk = (Kt + BK - 1) / BK;  // Kt > 0: callers keep kskip < K

    FastLoader<BM, AKC, THREADS> fa;
    FastLoader<BN, BKC, THREADS> fb;
    fa.init(Abase, g.lda, sA, tid);
    fb.init(Bbase, g.ldb, sB, tid);
    const bool interior = (mrem >= BM) && (nrem >= BN);  // uniform across the CTA

    auto issue = [&](int kt) {
        const int buf = kt % STAGES;
        const int k0 = kt * BK;
        if (interior && k0 + BK <= Kt) {
            fa.issue(AKC ? (int64_t)k0 : (int64_t)k0 * g.lda, (unsigned)(buf * TA::ELEMS * sizeof(double)));
            fb.issue(BKC ? (int64_t)k0 : (int64_t)k0 * g.ldb, (unsigned)(buf * TB::ELEMS * sizeof(double)));
            return;
        }
        const double* a = AKC ? (Abase + k0) : (Abase + (int64_t)k0 * g.lda);
        const double* b = BKC ? (Bbase + k0) : (Bbase + (int64_t)k0 * g.ldb);
        TA::template load<THREADS>(sA + buf * TA::ELEMS, a, g.lda, mrem, Kt - k0, tid);
        TB::template load<THREADS>(sB + buf * TB::ELEMS, b, g.ldb, nrem, Kt - k0, tid);
    };

    double acc[FM][FN][2];
#pragma unroll
    for (int i = 0; i < FM; ++i)
#pragma unroll
        for (int j = 0; j < FN; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    static_assert(STAGES % KSTEP == 0 && STAGES / KSTEP >= 2, "ring must hold at least two barrier groups");
    constexpr int GROUPS = STAGES / KSTEP;
    const int ng = (nk + KSTEP - 1) / KSTEP;
    auto issue_group = [&](int gi) {
#pragma unroll
        for (int u = 0; u < KSTEP; ++u)
            if (gi * KSTEP + u < nk) issue(gi * KSTEP + u);
    };
#pragma unroll
    for (int s = 0; s < GROUPS - 1; ++s) {
        if (s < ng) issue_group(s);
        cp_async_commit();
    }

    for (int gi = 0; gi < ng; ++gi) {
        cp_async_wait<GROUPS - 2>();
        __syncthreads();
        if (gi + GROUPS - 1 < ng) issue_group(gi + GROUPS - 1);
        cp_async_commit();

#pragma unroll
        for (int u = 0; u < KSTEP; ++u) {
            const int kt = gi * KSTEP + u;
            if (KSTEP > 1 && kt >= nk) break;
            const double* a_t = sA + (kt % STAGES) * TA::ELEMS;
            const double* b_t = sB + (kt % STAGES) * TB::ELEMS;
#pragma unroll
            for (int kk = 0; kk < BK / 4; ++kk) {
                double af[FM], bf[FN];
#pragma unroll
                for (int i = 0; i < FM; ++i) af[i] = a_t[TA::idx(wm0 + i * 8 + grp, kk * 4 + t4)];
#pragma unroll
                for (int j = 0; j < FN; ++j) bf[j] = b_t[TB::idx(wn0 + j * 8 + grp, kk * 4 + t4)];
#pragma unroll
                for (int i = 0; i < FM; ++i)
#pragma unroll
                    for (int j = 0; j < FN; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
            }
        }
    }
    cp_async_wait<0>();

    // epilogue: each thread owns (row, col..col+1) pairs -> 16-byte accesses (ldc and col are even)
    const bool has_beta = (g.beta != 0.0);
#pragma unroll
    for (int i = 0; i < FM; ++i) {
        const int row = m0 + wm0 + i * 8 + grp;
        if (row >= g.M) continue;
        double* crow = g.C + (EXT ? z * g.sC : 0) + (int64_t)row * g.ldc;
#pragma unroll
        for (int j = 0; j < FN; ++j) {
            const int col = n0 + wn0 + j * 8 + t4 * 2;
            if (col + 1 < g.N) {
                double2 v = make_double2(g.alpha * acc[i][j][0], g.alpha * acc[i][j][1]);
                double2* p = reinterpret_cast<double2*>(crow + col);
                if (has_beta) {
                    const double2 o = *p;
                    v.x = fma(g.beta, o.x, v.x);
                    v.y = fma(g.beta, o.y, v.y);
                }
                *p = v;
            } else if (col < g.N) {
                double v = g.alpha * acc[i][j][0];
                if (has_beta) v = fma(g.beta, crow[col], v);
                crow[col] = v;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
template <int BM, int BN, int WM, int WN, bool AKC, bool BKC, int STAGES, int MINB, int KSTEP, bool EXT>
static int launch_impl(Ctx* c, const GemmArgs& g, cudaStream_t stream) {
    using TA = OperandTile<BM, AKC>;
    using TB = OperandTile<BN, BKC>;
    constexpr int THREADS = (BM / WM) * (BN / WN) * 32;
    constexpr size_t SMEM = sizeof(double) * STAGES * (TA::ELEMS + TB::ELEMS);
    auto kern = dgemm_kernel<BM, BN, WM, WN, AKC, BKC, STAGES, MINB, KSTEP, EXT>;
    GP_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
    const int64_t tm = (g.M + BM - 1) / BM, tn = (g.N + BN - 1) / BN;
    int64_t grid;
    if (g.lower) {
        constexpr int R = BM / BN;
        const int64_t ngroups = (tm + GM - 1) / GM;
        grid = (int64_t)GM * GM * R * ngroups * (ngroups + 1) / 2;  // includes the early-exit tiles right of the diagonal
    } else {
        grid = tm * tn;
    }
    if (grid <= 0) return 0;
    kern<<<dim3((unsigned)grid, (unsigned)(g.batch > 1 ? g.batch : 1)), THREADS, SMEM, stream>>>(g);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}

// the extended kernel (batch / triangular k range) exists only for the default tile shapes
template <int BM, int BN, int WM, int WN, bool AKC, bool BKC, int STAGES, int MINB, int KSTEP = 1>
static int launch(Ctx* c, const GemmArgs& g, cudaStream_t stream) {
    if (g.batch > 1 || (g.ktri != 0 && g.ktri != 5)) {
        if (KSTEP == 1 && STAGES == 3) return launch_impl<BM, BN, WM, WN, AKC, BKC, 3, MINB, 1, true>(c, g, stream);
        return fail(c, -2, "gemm: batch / triangular k range need the default pipeline (gemm_stages = 3, variant 0 or 1)");
    }
    return launch_impl<BM, BN, WM, WN, AKC, BKC, STAGES, MINB, KSTEP, false>(c, g, stream);
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

int gemm(Ctx* c, GemmForm form, const GemmArgs& g, cudaStream_t stream) {
    if (g.M <= 0 || g.N <= 0) return 0;
    if (!stream) stream = c->stream;
    // few right-hand sides: HBM-bound streaming kernels instead of a mostly empty tensor-core tile (skinny.cu).
    // In-place use (C aliasing B: the diagonal-block step Z_k = inv(L_kk) B_k) keeps the tile kernel, whose CTA
    // reads all of its B columns before it writes.
    if (!g.lower && g.N <= 8 && g.batch <= 1 && (form == GEMM_NN || form == GEMM_TN) && g.C != g.B && g.C != g.A &&
        c->opt_skinny)
        return gemm_skinny(c, form, g, stream);
    if (!aligned16(g.A) || !aligned16(g.B) || !aligned16(g.C) || (g.lda & 1) || (g.ldb & 1) || (g.ldc & 1))
        return fail(c, -2, "gemm: operands must be 16-byte aligned with even leading dimensions");
    if (gemm_tma_eligible(c, form, g)) return gemm_tma_nt(c, g, stream);
    if (g.lower && g.M != g.N) return fail(c, -2, "gemm: lower-only update needs a square C");
    // wide tile (128x128, 8 warps) when C has a single column tile that must be owned by one CTA
    // (in-place panel solves), otherwise 128x64 tiles with two CTAs per SM so one CTA's epilogue
    // overlaps the other's main loop.
    const bool wide = (g.N > 64 && g.N <= 128 && !g.lower);
    const bool s3 = (c->opt_gemm_stages == 3);
    // auto (-1): the 16-warp variant wins where the epilogue is heavy (beta != 0 and a short K: the Cholesky updates)
    const int variant = c->opt_gemm_variant >= 0 ? (int)c->opt_gemm_variant : ((g.beta != 0.0 && g.K <= 2048) ? 1 : 0);
    // variant 1: 64x16 warp tiles, 8 warps per CTA, 16 warps per SM (<= 128 registers): four warps per scheduler
    // keep the DMMA pipe fed while another CTA sits in its epilogue or at a barrier.
    // short-K products with one column tile (the K = 128 panel solves and next-block-column updates of potrf.cu):
    // a 128 x 128 x 128 product is 17 us of DMMA on ONE SM, and a tall panel has only m / 128 such CTAs; 32-row tiles
    // give four times the CTAs at a quarter of the latency each
    if (wide && form == GEMM_NT && g.K <= 256 && c->opt_gemm_small && g.batch <= 1 && g.ktri == 0)
        return launch<32, 128, 16, 32, true, true, 3, 2>(c, g, stream);
    switch (form) {
        case GEMM_NT:
            if (wide) return launch<128, 128, 64, 32, true, true, 3, 1>(c, g, stream);
            if (variant == 1) return launch<128, 64, 64, 16, true, true, 3, 2>(c, g, stream);
            if (variant == 2) return launch<128, 64, 64, 32, true, true, 4, 2, 2>(c, g, stream);
            return s3 ? launch<128, 64, 64, 32, true, true, 3, 2>(c, g, stream)
                      : launch<128, 64, 64, 32, true, true, 4, 2>(c, g, stream);
        case GEMM_NN:
            if (wide) return launch<128, 128, 64, 32, true, false, 3, 1>(c, g, stream);
            if (variant == 1) return launch<128, 64, 64, 16, true, false, 3, 2>(c, g, stream);
            if (variant == 2) return launch<128, 64, 64, 32, true, false, 4, 2, 2>(c, g, stream);
            return s3 ? launch<128, 64, 64, 32, true, false, 3, 2>(c, g, stream)
                      : launch<128, 64, 64, 32, true, false, 4, 2>(c, g, stream);
        case GEMM_TN:
            if (wide) return launch<128, 128, 64, 32, false, false, 3, 1>(c, g, stream);
            if (variant == 1) return launch<128, 64, 64, 16, false, false, 3, 2>(c, g, stream);
            if (variant == 2) return launch<128, 64, 64, 32, false, false, 4, 2, 2>(c, g, stream);
            return s3 ? launch<128, 64, 64, 32, false, false, 3, 2>(c, g, stream)
                      : launch<128, 64, 64, 32, false, false, 4, 2>(c, g, stream);
    }
    return fail(c, -2, "gemm: unknown form");
}

}  // namespace gprb200
