// Persistent, TMA-fed fp64 tensor-core GEMM (NT form) for sm_100a: C = alpha A B^T + beta C, A M x K and B N x K with
// k contiguous, everything row-major. An A/B alternative to the cp.async kernel of dgemm.cu for the SYRK trailing updates
// of the blocked Cholesky (potrf.cu, the replacement of the reference's dgetrf/dgetri and dpotrf/dpotri,
// include/LAPACKUtils.h:38-73). OFF by default (option "gemm_tma"): measured equal to the cp.async kernel, see below.
//
// Design. The cp.async kernel measured 91 % of the DMMA pipe in its largest launch and 83 % averaged over a
// factorisation (profiles/r01b_ncu_summary.md): every k-tile ends in a CTA-wide barrier, every thread spends issue slots
// and registers on LDGSTS addressing, and every tile pays a pipeline fill. Here
//   * one CTA per SM lives for the whole launch and walks the (lower-triangle) tile list; the list is the grouped
//     rasterisation of dgemm.cu without the tiles right of the diagonal, so co-resident CTAs share A / B panels in L2;
//   * lane 0 of warp 0 is the PRODUCER (between its own k-steps): it issues
//     cp.async.bulk.tensor.2d (TMA, SASS UTMALDG) for the 128 x 16 A and B boxes of a stage and arms the stage's mbarrier
//     with the byte count; out-of-range rows / k are zero-filled by the tensor map, so there is no edge code on the load
//     side. It runs up to STAGES k-tiles ahead, ACROSS tile boundaries: the next tile's operands land while the
//     consumers are in the epilogue;
//   * all 16 warps are CONSUMERS (64 x 16 accumulators each, DMMA.8x8x4): they wait on the stage's "full" mbarrier,
//     read fragments, and release the stage with one arrive per warp on its "empty" mbarrier. No __syncthreads in the
//     main loop: a warp is only ever held by data that has not landed.
// Shared-memory layout: the TMA writes dense 128-byte rows (16 doubles) with SWIZZLE_128B (16-byte chunk index XOR
// row mod 8). A 64-bit LDS is served per half-warp (4 rows x 4 k of the DMMA fragment); with the natural k order the
// rows r and r+1 would collide. The k order INSIDE a k-tile is therefore permuted - at step kk lanes t4 = 0,1 read
// k = 2kk, 2kk+1 and lanes t4 = 2,3 read k = 8+2kk, 9+2kk (chunks kk and kk+4) - for A and B alike, which leaves the
// product unchanged (a different summation order) and makes every fragment load hit 16 distinct 8-byte bank slots.
#include "common.cuh"
#include "gemm_tile_map.cuh"
#include <cuda.h>
#include <cudaTypedefs.h>

namespace gprb200 {

constexpr int TBK = 16;
constexpr int TBM = 128, TBN = 128;
constexpr int TSTAGES = 6;
constexpr int TCONS = 16;                       // consumer warps: 2 (M) x 8 (N), 64 x 16 each
constexpr int TTHREADS = TCONS * 32;            // 512 threads -> 128 registers per thread; no dedicated producer warp
// Who issues the TMA. A first version had a producer WARPGROUP (640 threads, 96 registers per thread) that handed
// registers to the consumers with setmaxnreg (24 / 112 and 56 / 104 were tried). MEASURED wrong: a few tiles per launch
// differ once a launch has ~1800+ tiles, with either split, and the results are exact without the hand-over
// (profiles/r02_gemm_tma_probe.md). Without it the consumers were capped at 96 registers and the kernel lost 8 %
// (30.97 vs 28.64 ms on the 30720^2 update); one extra producer warp does not help either (warps are allocated four at a
// time: 17 warps are budgeted as 20, 96 registers again). So lane 0 of warp 0 is the producer: at k-step `it` it refills
// the stage that all warps released at step it - 1 with k-tile it + STAGES - 1 (which may belong to the NEXT tile).
constexpr int TA_BYTES = TBM * TBK * 8;
constexpr int TB_BYTES = TBN * TBK * 8;
constexpr int TSTAGE_BYTES = TA_BYTES + TB_BYTES;
constexpr size_t TSMEM = 1024 + (size_t)TSTAGES * TSTAGE_BYTES + 2 * TSTAGES * sizeof(uint64_t);

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}

// tile `lin` of this launch: compact lower-triangle list, or the grouped order of dgemm.cu
__device__ __forceinline__ bool next_tile(int lin, int lower, int tiles_m, int tiles_n, int& ti, int& tj) {
    if (lower) return map_tile_lower_compact(lin, tiles_m, ti, tj);
    return map_tile<TBM, TBN>(lin, 0, tiles_m, tiles_n, ti, tj);
}

__global__ void __launch_bounds__(TTHREADS, 1)
dgemm_tma_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, GemmArgs g, int tiles_m,
                 int tiles_n, int total) {
    extern __shared__ unsigned char tsm_raw[];
    // SWIZZLE_128B: the pattern is a function of address bits 7..9, so stage bases must be 1024-byte aligned
    unsigned char* tsm = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(tsm_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(tsm);
    const uint32_t bar_full = sbase + TSTAGES * TSTAGE_BYTES;
    const uint32_t bar_empty = bar_full + TSTAGES * 8;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < TSTAGES; ++s) {
            mbar_init(bar_full + s * 8, 1);
            mbar_init(bar_empty + s * 8, TCONS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    const int nk = (g.K + TBK - 1) / TBK;

    // ---- producer cursor (used by thread 0 only): the next k-tile to load, over ALL tiles of this CTA
    int p_lin = blockIdx.x, p_kt = 0, p_m0 = 0, p_n0 = 0;
    uint32_t p_it = 0;
    bool p_live = false;
    auto p_seek = [&]() {   // first valid tile at or after p_lin
        p_live = false;
        for (; p_lin < total; p_lin += gridDim.x) {
            int ti, tj;
            if (next_tile(p_lin, g.lower, tiles_m, tiles_n, ti, tj)) {
                p_m0 = ti * TBM;
                p_n0 = tj * TBN;
                p_live = true;
                return;
            }
        }
    };
    auto p_issue = [&]() {
        if (!p_live) return;
        const uint32_t s = p_it % TSTAGES, use = p_it / TSTAGES;
        if (use > 0) mbar_wait(bar_empty + s * 8, (use - 1) & 1);   // every warp released the previous use of this stage
        mbar_expect_tx(bar_full + s * 8, TSTAGE_BYTES);
        tma_load_2d(sbase + s * TSTAGE_BYTES, &mapA, bar_full + s * 8, p_kt * TBK, p_m0);
        tma_load_2d(sbase + s * TSTAGE_BYTES + TA_BYTES, &mapB, bar_full + s * 8, p_kt * TBK, p_n0);
        ++p_it;
        if (++p_kt == nk) {
            p_kt = 0;
            p_lin += gridDim.x;
            p_seek();
        }
    };
    if (tid == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&mapA)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&mapB)) : "memory");
        p_seek();
#pragma unroll 1
        for (int i = 0; i < TSTAGES - 1; ++i) p_issue();
    }

    const int grp = lane >> 2, t4 = lane & 3;
    const int wm0 = (warp >> 3) * 64, wn0 = (warp & 7) * 16;
    // byte offset of this lane's element inside a 128-byte row at k-step kk (see the header comment)
    uint32_t koff[4];
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) koff[kk] = ((uint32_t)((kk + 4 * (t4 >> 1)) ^ grp) << 4) + (uint32_t)(t4 & 1) * 8;
    const uint32_t arow = (uint32_t)(wm0 + grp) * 128, brow = (uint32_t)TA_BYTES + (uint32_t)(wn0 + grp) * 128;
    const bool has_beta = (g.beta != 0.0);

    uint32_t it = 0;
    for (int lin = blockIdx.x; lin < total; lin += gridDim.x) {
        int ti, tj;
        if (!next_tile(lin, g.lower, tiles_m, tiles_n, ti, tj)) continue;
        const int m0 = ti * TBM, n0 = tj * TBN;
        double acc[8][2][2];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

        for (int kt = 0; kt < nk; ++kt, ++it) {
            if (tid == 0) p_issue();   // k-tile it + STAGES - 1 into the stage released at step it - 1
            const uint32_t s = it % TSTAGES, use = it / TSTAGES;
            mbar_wait(bar_full + s * 8, use & 1);
            const unsigned char* st = tsm + s * TSTAGE_BYTES;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                double af[8], bf[2];
#pragma unroll
                for (int i = 0; i < 8; ++i) af[i] = *reinterpret_cast<const double*>(st + arow + i * 1024 + koff[kk]);
#pragma unroll
                for (int j = 0; j < 2; ++j) bf[j] = *reinterpret_cast<const double*>(st + brow + j * 1024 + koff[kk]);
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int j = 0; j < 2; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_empty + s * 8);
        }
        // epilogue: (row, col..col+1) pairs -> 16-byte accesses (ldc and col are even); the producer is already
        // filling the ring with the next tile's operands
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int row = m0 + wm0 + i * 8 + grp;
            if (row >= g.M) continue;
            double* crow = g.C + (int64_t)row * g.ldc;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
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
}

// ---------------------------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled encode_fn() {
    static PFN_cuTensorMapEncodeTiled fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<PFN_cuTensorMapEncodeTiled>(p);
    }();
    return fn;
}

// rows x K operand, k contiguous, leading dimension ld (doubles): boxes of box_rows x 16 doubles, SWIZZLE_128B
static int make_map(Ctx* c, CUtensorMap* m, const double* base, int64_t rows, int64_t K, int64_t ld, int box_rows) {
    PFN_cuTensorMapEncodeTiled enc = encode_fn();
    if (!enc) return fail(c, -1, "gemm_tma: cuTensorMapEncodeTiled is not available from this driver");
    const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(double)};
    const cuuint32_t box[2] = {(cuuint32_t)TBK, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double*>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(c, -1, "gemm_tma: cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
    return 0;
}

bool gemm_tma_eligible(const Ctx* c, GemmForm form, const GemmArgs& g) {
    if (!c->opt_gemm_tma || form != GEMM_NT || g.batch > 1 || g.ktri != 0) return false;
    if (g.C == g.A || g.C == g.B) return false;                       // in-place panel solves keep the CTA-owned kernel
    if (g.K < c->opt_gemm_tma_min_k) return false;                    // short K: the pipeline fill dominates
    const int64_t tm = (g.M + TBM - 1) / TBM, tn = (g.N + TBN - 1) / TBN;
    const int64_t tiles = g.lower ? tm * (tm + 1) / 2 : tm * tn;
    if (tiles < c->opt_gemm_tma_min_tiles) return false;              // too few 128 x 128 tiles for 148 persistent CTAs
    auto a16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    return a16(g.A) && a16(g.B) && a16(g.C) && !(g.lda & 1) && !(g.ldb & 1) && !(g.ldc & 1);
}

int gemm_tma_nt(Ctx* c, const GemmArgs& g, cudaStream_t stream) {
    if (g.M <= 0 || g.N <= 0) return 0;
    if (g.lower && g.M != g.N) return fail(c, -2, "gemm: lower-only update needs a square C");
    int sm_count = 0;
    GP_CUDA(c, cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, c->device));
    static bool attr_set[64] = {false};
    if (c->device >= 0 && c->device < 64 && !attr_set[c->device]) {
        GP_CUDA(c, cudaFuncSetAttribute(dgemm_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TSMEM));
        attr_set[c->device] = true;
    }
    CUtensorMap ma, mb;
    GP_TRY(make_map(c, &ma, g.A, g.M, g.K, g.lda, TBM));
    GP_TRY(make_map(c, &mb, g.B, g.N, g.K, g.ldb, TBN));
    const int64_t tm = (g.M + TBM - 1) / TBM, tn = (g.N + TBN - 1) / TBN;
    const int64_t total = g.lower ? lower_compact_count(tm) : tm * tn;
    const int64_t valid = g.lower ? tm * (tm + 1) / 2 : tm * tn;
    const int grid = (int)((valid < sm_count) ? valid : sm_count);
    if (total > 0x7fffffff) return fail(c, -2, "gemm_tma: tile space exceeds 31 bits");
    dgemm_tma_kernel<<<grid, TTHREADS, TSMEM, stream>>>(ma, mb, g, (int)tm, (int)tn, (int)total);
    ++c->launches;
    GP_CUDA(c, cudaGetLastError());
    return 0;
}

}  // namespace gprb200

