// Shared declarations for the gprb200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include "../../include/gprb200.h"

namespace gprb200 {

constexpr int NB = 128;  // Cholesky diagonal-block size (one CTA factors NB x NB in shared memory)

// Flat kernel program, passed by value to kernels (lives in constant bank / param space).
struct KProg {
    int nops;
    int nparams;
    int nper;  // number of PERIODIC leaves
    int max_stack;
    int8_t op[GPRB200_MAX_OPS];
    int8_t poff[GPRB200_MAX_OPS];  // offset of the leaf's first parameter
    double par[GPRB200_MAX_PARAMS];
};

struct GemmArgs {
    const double* A;
    const double* B;
    double* C;
    int64_t lda, ldb, ldc;
    int M, N, K;
    double alpha, beta;
    int lower;  // 1: only tiles that intersect {row >= col} are computed (SYRK trailing update)
    // batch > 1: blockIdx.y selects operand set z with A + z*sA, B + z*sB, C + z*sC (split-K partial products
    // of the PCA Gram, pca.cu). 0 and 1 both mean a single GEMM.
    int batch;
    int64_t sA, sB, sC;
    // ktri: triangular operand structure known to the caller; the k loop of a tile starts later (rounded down to BK):
    //   1: B (K x N) is lower triangular, B[k][n] = 0 for k < n  ->  k starts at the tile's first column (TRMM, NN)
    //   2: A and B (K x M, K x N) are lower triangular            ->  k starts at max(first row, first column) (LAUUM, TN)
    //   3: as 1, but B's columns are a COMPACT selection of cyclic column blocks (trtri_cols, solve.cu): local column n
    //      is global column (tri_first + (n / tri_cb) * tri_stride) * tri_cb + n % tri_cb, and B[k][n] = 0 for k < that
    //   4: NT form, B (N x K) is lower triangular, B[n][k] = 0 for k > n  ->  k ENDS after the tile's last column (the
    //      product with an inverted Cholesky block: X = A inv(L)^T)
    //   5: "staircase" C, the trailing update of a block-cyclic Cholesky in ONE launch (no k skip). stair() is a device
    //      table: entry u is the first row (in the caller's row numbering) that columns [64 u, 64 u + 64) of C need; row r
    //      of C carries the number stair_row0() + r. A tile whose rows all lie before the stair of all of its columns exits
    //      at once. The table pointer and the row number travel in sC and tri_first, which mode 5 does not use otherwise:
    //      the struct stays at 128 bytes - growing the kernel's parameter block by 16 bytes cost the single-GPU
    //      Cholesky 4 % (364 -> 380 ms at N = 32768, same SASS instruction mix; profiles/r02_gemmargs_ab.md).
    int ktri;
    int tri_cb, tri_first, tri_stride;
    __host__ __device__ const int* stair() const { return ktri == 5 ? reinterpret_cast<const int*>(sC) : nullptr; }
    __host__ __device__ int stair_row0() const { return tri_first; }
    void set_stair(const int* table, int row0) { ktri = 5; sC = reinterpret_cast<int64_t>(table); tri_first = row0; }
};
static_assert(sizeof(GemmArgs) == 128, "GemmArgs is a kernel parameter: keep its size");

enum GemmForm { GEMM_NT = 0, GEMM_NN = 1, GEMM_TN = 2 };

struct Ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    int64_t launches = 0;

    KProg prog{};
    bool have_kernel = false;

    // training data (device)
    int64_t N = 0, M = 0;
    int d = 0;
    double sigma = 0.0;
    double* dX = nullptr;   // N x d (ld = d)
    double* dY = nullptr;   // N x ldy
    double* dA = nullptr;   // alpha, N x ldy
    int64_t ldy = 0;
    double* dL = nullptr;   // N x ldk: lower Cholesky factor of K + sigma^2 I (upper part: stale K)
    int64_t ldk = 0;
    double* dInvL = nullptr;  // ceil(N/NB) blocks of NB x NB: inverses of the diagonal blocks of L
    double* dC = nullptr;     // N x ldk explicit core matrix (lazy)
    double* dW = nullptr;     // N x ldk inverse factor inv(L) (lazy; single-point credible intervals, api.cu: quad_forms)
    bool have_winv = false;
    int64_t ci_small_calls = 0;   // single-point interval calls since the last factorisation
    int64_t opt_ci_winv = 1;      // 0: never form inv(L); 1: from the 3rd single-point call; 2: at the first
    int* dInfo = nullptr;
    double* dScratch = nullptr;  // small reductions
    size_t scratch_bytes = 0;

    bool have_train = false, have_factor = false, have_alpha = false, have_core = false;
    bool core_is_loaded = false;  // C installed by set_state (Load), no factor behind it
    double logdet = 0.0;

    int64_t opt_gram_lower_only = 1;
    int64_t opt_gram_mma = 1;         // r^2 = |x|^2 + |y|^2 - 2 x.y with the cross term on DMMA for large d (gram_mma.cu)
    int64_t opt_gram_mma_min_d = 64;  // ... from this input dimension (measured: d = 30 6.5 vs 4.0 ms direct, d = 64 7.4 vs 8.4, d = 238 13.0 vs 41.7)
    int64_t opt_gram_terms = 2;   // value Gram: 2 row-group kernel (gram_rows_kernel), 1 per-tile term form, 0 interpreter
    int64_t opt_nbo = 0;          // outer (trailing-update) block of the two-level Cholesky; 0 = auto (512 / 1024)
    int64_t opt_chol_left = 0;    // left-looking outer loop (long-K column updates): measured slower, see potrf.cu
    int64_t opt_lookahead = 1;    // factor panel k+1 on a second stream while update k runs
    int64_t opt_potrf_tile = 3;    // 3: DMMA block products + fused diagonal factor/inverse (potrf_tile3_kernel), 2: round-1 scalar blocked kernel, 1: unblocked
    int64_t opt_skinny = 1;        // route n <= 8 NN/TN products to the streaming kernels of skinny.cu
    int64_t opt_gemm_variant = -1; // -1: auto; 0: 64x32 warp tiles (8 warps/SM), 1: 64x16 warp tiles (16 warps/SM)
    int64_t opt_gemm_stages = 3;  // cp.async ring depth of the 128x64 GEMM
    int64_t opt_gemm_tma = 0;            // 1: persistent TMA-fed kernel (dgemm_tma.cu) for large NT products (A/B: no gain)
    int64_t opt_gemm_tma_min_k = 256;    // ... from this contraction depth
    int64_t opt_gemm_tma_min_tiles = 64; // ... and this many 128 x 128 output tiles
    int64_t opt_trsv_chain = 1;           // M <= 2 triangular solves as one chained launch (solve.cu)
    int64_t opt_inverse_long_k = 1;       // explicit inverse as TRTRI + LAUUM with long-K GEMMs (solve.cu)
    int64_t opt_solve_left_min_m = 8192;  // left-looking (long-K) triangular solves from this many RHS
    cudaStream_t panel_stream = nullptr;  // high-priority stream for the look-ahead panel
    cudaStream_t aux_stream = nullptr;    // off-critical-path part of a panel step (potrf.cu: factor_panel)
    cudaEvent_t ev_pcrit = nullptr, ev_paux = nullptr;
    int64_t opt_panel_split = 1;          // 1: only the next block column is updated on the panel's critical stream
    int64_t opt_gemm_small = 1;           // 1: 32 x 128 CTA tiles for N <= 128, K <= 256 products (panel solves)
    int64_t opt_profile = 0;      // time every trailing-update GEMM launch with CUDA events (bench.py roofline)
    double prof_tb_ms = 0.0, prof_tb_flops = 0.0, prof_tb_launches = 0.0;  // last factorisation
    double prof_gemv_ms = 0.0, prof_gemv_bytes = 0.0, prof_gemv_launches = 0.0;  // accumulated since reset

    cudaEvent_t ev0 = nullptr, ev1 = nullptr;  // Predict phase timing (resolved lazily by gprb200_phase_ms)
    bool predict_time_pending = false;

    // pinned staging ring for large host <-> device matrices (hostio.cu)
    static constexpr int NPIN = 3;
    void* pin[2][NPIN] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};   // ring 0: to device, ring 1: to host
    cudaEvent_t pin_ev[2][NPIN] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};
    cudaStream_t io_stream[2] = {nullptr, nullptr};   // copy streams of the pipelined Learn (gprb200_learn_rows)
    int64_t opt_stream_io = 1;   // 1: label / regression matrices >= 64 MB cross PCIe through the pinned ring

    void* sparse = nullptr;  // SparseModel (sparse.cu): inducing points and the two m x m factors
    double phase_ms[GPRB200_T_COUNT] = {0};
};

// error helpers -----------------------------------------------------------------------------
int fail(Ctx* c, int code, const std::string& msg);
#define GP_CUDA(c, expr)                                                                              \
    do {                                                                                              \
        cudaError_t _e = (expr);                                                                      \
        if (_e != cudaSuccess)                                                                        \
            return gprb200::fail((c), -100 - (int)_e,                                                 \
                                 std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ + \
                                     ":" + std::to_string(__LINE__) + ")");                           \
    } while (0)
#define GP_TRY(expr)                \
    do {                            \
        int _r = (expr);            \
        if (_r != 0) return _r;     \
    } while (0)

static inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

#ifdef __CUDACC__
// FP64 tensor-core atom (SASS DMMA.8x8x4): C(8x8) += A(8x4) B(4x8). Lane (grp = lane >> 2, t4 = lane & 3) holds
// a = A[grp][t4], b = B[t4][grp], c0 = C[grp][2 t4], c1 = C[grp][2 t4 + 1]. tcgen05 has no f64 kind, so this
// warp-level instruction is the FP64 tensor path on sm_100a.
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}
#endif

// kernels / drivers (each returns 0 or an error code; all async on c->stream) -----------------
// gram.cu
int gram_build(Ctx* c, const double* Xa, int64_t Na, const double* Xb, int64_t Nb, double* out, int64_t ldo,
               bool symmetric, bool lower_only, double diag_add, int* d_nonfinite = nullptr);
int dgram_build(Ctx* c, double* out /* (P*N) x N */, int64_t ldo);
int check_finite(Ctx* c, const double* A, int64_t rows, int64_t cols, int64_t ld, bool lower_only, int* bad_host);
// fused likelihood-gradient contraction: out[p] = sum_ij S_ij * dK_p(i,j), S symmetric N x N (lds)
int dgram_contract(Ctx* c, const double* S, int64_t lds, double* out_host /* P */);
// the same between two point sets (sparse likelihood: n x m cross-Gram weights); Xa == Xb means symmetric
int dgram_contract_xy(Ctx* c, const double* Xa, int64_t Na, const double* Xb, int64_t Nb, const double* W, int64_t ldw,
                      double* out_host /* P */);
int dgram_build_xy(Ctx* c, const double* Xa, int64_t Na, const double* Xb, int64_t Nb, double* out, int64_t ldo,
                   int p_select);
// dgemm.cu
int gemm(Ctx* c, GemmForm form, const GemmArgs& g, cudaStream_t stream = nullptr);
int gemm_skinny(Ctx* c, GemmForm form, const GemmArgs& g, cudaStream_t stream);  // skinny.cu, n <= 8
// dgemm_tma.cu: persistent TMA + mbarrier producer / DMMA consumers, NT form
bool gemm_tma_eligible(const Ctx* c, GemmForm form, const GemmArgs& g);
int gemm_tma_nt(Ctx* c, const GemmArgs& g, cudaStream_t stream);
// potrf.cu
int potrf_blocked(Ctx* c, double* A, int64_t ld, int64_t N, double* invL, int* info_host);
int potrf_panel(Ctx* c, double* P, int64_t ld, int64_t rows, int64_t w, double* invL, int col0, int* info_host);
int potrf_panel_async(Ctx* c, double* P, int64_t ld, int64_t rows, int64_t w, double* invL, int col0, int* dinfo);
int logdet_from_factor(Ctx* c, const double* L, int64_t ld, int64_t N, double* out_host);
// solve.cu
int trsm_lower_fwd(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* B, int64_t ldb,
                   int64_t M);
int trsm_lower_bwd(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* B, int64_t ldb,
                   int64_t M);
int inverse_from_factor(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, double* C, int64_t ldc);
int trtri_cols(Ctx* c, const double* L, int64_t ldl, const double* invL, int64_t N, int64_t cb, int64_t first,
               int64_t stride, double* W, int64_t ldw);
// util kernels (util.cu)
int copy2d(Ctx* c, double* dst, int64_t ldd, const double* src, int64_t lds, int64_t rows, int64_t cols);
int fill2d(Ctx* c, double* dst, int64_t ldd, int64_t rows, int64_t cols, double v);
int add_diag(Ctx* c, double* A, int64_t ld, int64_t n, double v);  // A(i,i) += v, i < n
// dst = a A + b B element-wise on rows x cols (dst may alias A or B)
int axpby2d(Ctx* c, double* dst, int64_t ldd, double a, const double* A, int64_t lda, double b, const double* B,
            int64_t ldb, int64_t rows, int64_t cols);
int set_identity(Ctx* c, double* dst, int64_t ldd, int64_t n);
int symmetrize_from_lower(Ctx* c, double* A, int64_t ld, int64_t n);
int zero_upper(Ctx* c, double* A, int64_t ld, int64_t n);
int coldot(Ctx* c, const double* A, int64_t lda, const double* B, int64_t ldb, int64_t rows, int64_t cols,
           double* d_out /* cols */);
int rowdot(Ctx* c, const double* A, int64_t lda, const double* B, int64_t ldb, int64_t rows, int64_t cols,
           double* d_out /* rows */);
// sparse.cu
void sparse_invalidate(Ctx* c);
void sparse_destroy(Ctx* c);
// hostio.cu: pinned-ring transfers of large matrices; rowptr (one pointer per row) or flat (leading dimension ldf)
// st == nullptr: c->stream. hoff: column offset applied to every host row (a column block of a wider host matrix).
int stream_h2d(Ctx* c, double* dst, int64_t ldd, const double* const* rowptr, const double* flat, int64_t ldf, int64_t rows,
               int64_t cols, cudaStream_t st = nullptr, int64_t hoff = 0);
int stream_d2h(Ctx* c, double* flat, int64_t ldf, const double* src, int64_t lds, int64_t rows, int64_t cols,
               cudaStream_t st = nullptr);
void hostio_destroy(Ctx* c);
// predict.cu
int gemv_t(Ctx* c, const double* kx, const double* A, int64_t lda, int64_t N, int64_t M, double* out);
int trmv_lower(Ctx* c, const double* W, int64_t ldw, int64_t N, const double* kx, int64_t ldk, int T, double* V, int64_t ldv);

}  // namespace gprb200

// the opaque handle of include/gprb200.h
struct gprb200_ctx : gprb200::Ctx {};
