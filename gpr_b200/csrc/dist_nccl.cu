// libgprb200_dist.so: 2-D block-cyclic GP Learn over the GPUs of one node, NCCL from C++ (include/gprb200_dist.h).
//
// Built ON TOP of the public C-ABI of libgprb200.so (gram_block / potrf_panel_async / gemm / trsm / logdet), so the
// multi-GPU schedule uses exactly the kernels the single-GPU path is tested with; this file adds the ownership maps,
// the stream / event choreography and the NCCL exchanges. Same schedule as gpr_b200/dist.py:BlockCyclicGP (which stays
// as the torch.distributed form and as the one the CPU gloo tests exercise). What it reproduces of the reference:
// the result of InvertKernelMatrix + C * Y (lib/GaussianProcess.cpp:531-560, :642-672).
//
// Layout. Block (i, j), i >= j, of nb x nb lives on rank rowof(i) * Pc + colof(j) (cyclic maps, the row map mirrored
// by default: see cyc() below). A rank keeps ONE local matrix: its
// owned block rows stacked (H rows) by its owned block columns side by side (Wtot columns); block column j uses the rows
// from its first owned block row i >= j on (the part above is never touched: twice the memory of a packed lower
// triangle, which buys a trailing update that is ONE tensor-core launch per step - gprb200_gemm_stair_dev - instead of
// one per owned block column, whose partial last waves added up to several per cent). Per step k:
//   grid column k mod Pc, panel stream: L_kk on the diagonal owner, W = inv(L_kk) broadcast down the column, each rank
//     solves its rows X_i = A_ik W^T with one tensor-core GEMM;
//   everybody, panel stream: X rows broadcast along the grid rows; the blocks that are a column operand L_jk broadcast
//     down the grid columns (Pr roots, one NCCL group);
//   everybody, main stream: A_ij -= X_i L_jk^T for all owned block columns j > k in one staircase GEMM; block column
//     k + 1 is updated and factored first, on the panel stream (look-ahead), so its broadcasts overlap the update.
// INFO is a device int for the whole factorisation, agreed on by one all-reduce(min) at the end.
#include <cuda_runtime.h>
#include <nccl.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/gprb200.h"
#include "../../include/gprb200_dist.h"

namespace {

thread_local std::string g_create_error;

struct Problem {
    int64_t N = 0, nb = 0, nblk = 0, H = 0;
    int dim = 0;
    double* dX = nullptr;                    // training inputs (N x dim)
    double* dXr = nullptr;                   // the owned block rows of dX, stacked (H x dim)
    std::vector<int64_t> hgt, roff;          // block heights; offset of block row i in a stacked panel (-1: not mine)
    std::vector<int64_t> myrows, mine;       // owned block rows / block columns
    double* A = nullptr;                     // the local matrix, H x Wtot (ld LDA)
    int64_t LDA = 0, Wtot = 0;
    std::vector<int64_t> coff;               // local column offset of owned block column j (-1: not mine)
    std::vector<double*> cols, invl;         // per block column: view into A from its first row on; inverted diagonal blocks
    std::vector<int64_t> ld, off;            // leading dimension (LDA) and first stacked row (offset of first i >= j)
    std::vector<int64_t> first;              // first owned block row >= j (-1: none)
    int* stair = nullptr;                    // per 64 local columns: first local row they need (gprb200_gemm_stair_dev)
    double* Bop[2] = {nullptr, nullptr};     // column operands L_jk of step k for the owned j > k, in local column order
    double* row[2] = {nullptr, nullptr};     // solved panel rows of step k (H x nb)
    double* W[2] = {nullptr, nullptr};       // inv(L_kk)
    double* eye = nullptr;
    std::vector<double*> colbuf[2];          // per source grid row: column operands of step k
    int* dinfo = nullptr;
    long long* dflag = nullptr;
    double* dsum = nullptr;
    std::vector<void*> allocs;
    std::vector<void*> ws;                   // workspace of solve() for right-hand sides of leading dimension ws_ld
    int64_t ws_ld = 0;
    std::vector<cudaEvent_t> ev_main, ev_panel, ev_start;
    bool factored = false;
    double logdet = 0.0;
};

}  // namespace

struct gprb200_dist {
    int dev = 0, rank = 0, world = 1, Pr = 1, Pc = 1, pr = 0, pc = 0;
    ncclComm_t comm = nullptr, row_comm = nullptr, col_comm = nullptr;
    gprb200_handle hm = nullptr, hp = nullptr;
    cudaStream_t sm = nullptr, sp = nullptr;
    cudaEvent_t t0 = nullptr, t1 = nullptr, t2 = nullptr, t3 = nullptr, ev_gram = nullptr;
    std::string err;
    bool have_kernel = false;
    Problem P;
    double ms[3] = {0, 0, 0};
    double n_coll = 0, n_bytes = 0, n_flop = 0, launches0 = 0, n_launch = 0;
    int mirror_rows = 1, mirror_cols = 0;    // options: boustrophedon block-cyclic maps (all ranks must agree)
    int profile = 0;                         // option "profile": timed per-step events on the main stream
    double n_flop_main = 0;                  // flop of the main-stream trailing updates (last learn)
    double stall_ms = 0, busy_ms = 0;        // main stream: waiting for a panel / inside trailing updates (last learn)
};

namespace {

int fail(gprb200_dist* d, int code, const std::string& msg) {
    if (d) d->err = msg; else g_create_error = msg;
    return code;
}

#define CK(call)                                                                                                   \
    do {                                                                                                           \
        cudaError_t e_ = (call);                                                                                   \
        if (e_ != cudaSuccess)                                                                                     \
            return fail(d, -3, std::string(#call) + ": " + cudaGetErrorString(e_));                                \
    } while (0)
#define NK(call)                                                                                                   \
    do {                                                                                                           \
        ncclResult_t r_ = (call);                                                                                  \
        if (r_ != ncclSuccess) return fail(d, -4, std::string(#call) + ": " + ncclGetErrorString(r_));             \
    } while (0)
#define GK(h, call)                                                                                                \
    do {                                                                                                           \
        int r_ = (call);                                                                                           \
        if (r_ != 0) return fail(d, r_ < 0 ? r_ : -5, std::string(#call) + ": " + gprb200_last_error(h));          \
    } while (0)
#define RK(call)                                                                                                   \
    do {                                                                                                           \
        int r_ = (call);                                                                                           \
        if (r_ != 0) return r_;                                                                                    \
    } while (0)

__global__ void sub_kernel(double* __restrict__ out, const double* __restrict__ a, const double* __restrict__ b, int64_t n) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) out[i] = a[i] - b[i];
}

// rows of M doubles <-> rows of ld doubles (right-hand sides are kept with an even leading dimension)
__global__ void repitch_kernel(double* __restrict__ dst, int64_t ldd, const double* __restrict__ src, int64_t lds, int64_t rows,
                               int64_t M) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < rows * M) dst[(i / M) * ldd + i % M] = src[(i / M) * lds + i % M];
}

__global__ void eye_kernel(double* __restrict__ out, int64_t n) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n * n) out[i] = (i / n == i % n) ? 1.0 : 0.0;
}

int dmalloc(gprb200_dist* d, void** p, size_t bytes) {
    CK(cudaMalloc(p, std::max<size_t>(bytes, 256)));
    d->P.allocs.push_back(*p);
    return 0;
}

int release_problem(gprb200_dist* d) {
    Problem& P = d->P;
    cudaDeviceSynchronize();
    for (void* p : P.allocs) cudaFree(p);
    for (void* p : P.ws) cudaFree(p);
    for (cudaEvent_t e : P.ev_main) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : P.ev_panel) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : P.ev_start) if (e) cudaEventDestroy(e);
    P = Problem();
    return 0;
}

// Grid row of block row i / grid column of block column j. Plain cyclic, or (option "mirror_rows" / "mirror_cols")
// boustrophedon: 0 1 .. P-1 P-1 .. 1 0, which evens out the lower triangle's bias towards high grid rows and low grid
// columns (with plain cyclic maps the busiest rank of a 4 x 2 grid at 128 block columns does 7 % more than the mean).
inline int cyc(int64_t i, int P, int mirror) {
    if (!mirror || P == 1) return (int)(i % P);
    const int t = (int)(i % (2 * P));
    return t < P ? t : 2 * P - 1 - t;
}
inline int rowof(const gprb200_dist* d, int64_t i) { return cyc(i, d->Pr, d->mirror_rows); }
inline int colof(const gprb200_dist* d, int64_t j) { return cyc(j, d->Pc, d->mirror_cols); }

// offset (in rows of a stacked panel) of the first owned block row i >= j, and that i (-1: none)
inline void first_at_or_after(const gprb200_dist* d, int64_t j, int64_t* offp, int64_t* ip) {
    const Problem& P = d->P;
    auto it = std::lower_bound(P.myrows.begin(), P.myrows.end(), j);
    if (it == P.myrows.end()) { *offp = P.H; *ip = -1; return; }
    *ip = *it;
    *offp = P.roff[*it];
}

inline int bcast(gprb200_dist* d, double* buf, int64_t count, int root, ncclComm_t comm, cudaStream_t st) {
    NK(ncclBroadcast(buf, buf, (size_t)count, ncclDouble, root, comm, st));
    d->n_coll += 1;
    d->n_bytes += 8.0 * (double)count;
    return 0;
}

}  // namespace

namespace {

// Host-only part of the set-up: who owns what, where it sits in the local matrix, and the staircase table. No CUDA call
// (gprb200_dist_plan runs it without a GPU; tests/test_dist_plan.py).
void plan_problem(gprb200_dist* d, int64_t N, int dim, int64_t nb, std::vector<int>* stair_out) {
    Problem& P = d->P;
    P.N = N; P.nb = nb; P.dim = dim; P.nblk = (N + nb - 1) / nb;
    P.hgt.resize(P.nblk); P.roff.assign(P.nblk, -1);
    for (int64_t i = 0; i < P.nblk; ++i) P.hgt[i] = std::min(nb, N - i * nb);
    for (int64_t i = 0; i < P.nblk; ++i)
        if (rowof(d, i) == d->pr) { P.roff[i] = P.H; P.H += P.hgt[i]; P.myrows.push_back(i); }
    for (int64_t j = 0; j < P.nblk; ++j)
        if (colof(d, j) == d->pc) P.mine.push_back(j);
    P.cols.assign(P.nblk, nullptr); P.invl.assign(P.nblk, nullptr);
    P.ld.assign(P.nblk, 0); P.off.assign(P.nblk, 0); P.first.assign(P.nblk, -1);
    P.coff.assign(P.nblk, -1);
    for (int64_t j : P.mine) { P.coff[j] = P.Wtot; P.Wtot += P.hgt[j]; }
    P.LDA = (P.Wtot + 1) / 2 * 2;
    std::vector<int>& stair = *stair_out;
    stair.assign((size_t)(P.Wtot + 63) / 64 + 1, (int)P.H);
    for (int64_t j : P.mine) {
        int64_t off, first;
        first_at_or_after(d, j, &off, &first);
        P.off[j] = off; P.first[j] = first;
        if (first < 0) continue;
        P.ld[j] = P.LDA;
        // columns [c, c + 64) of the diagonal block need the rows from c on (lower triangle), the others the whole block row
        for (int64_t c = 0; c < P.hgt[j]; c += 64) stair[(size_t)(P.coff[j] + c) / 64] = (int)(off + (first == j ? c : 0));
    }
}

// flop this rank's GEMMs execute in one factorisation (the same accounting as the n_flop counter of a real run)
double plan_flops(gprb200_dist* d) {
    Problem& P = d->P;
    double f = 0.0;
    for (int64_t k = 0; k + 1 < P.nblk; ++k) {
        int64_t o1, first;
        first_at_or_after(d, k + 1, &o1, &first);
        if (colof(d, k) == d->pc && first >= 0) f += 1.0 * (P.H - o1) * P.hgt[k] * (P.hgt[k] + 64);
        for (int64_t j : P.mine)
            if (j > k && P.first[j] >= 0) f += 2.0 * (P.H - P.off[j]) * P.hgt[j] * P.nb;
    }
    return f;
}

// ---- ownership maps and device buffers ----------------------------------------------------------------------------
int setup_problem(gprb200_dist* d, int64_t N, int dim, int64_t nb) {
    if (d->P.N == N && d->P.nb == nb && d->P.dim == dim && d->P.dX) {   // same shape as the last learn: keep the buffers
        d->P.factored = false;
        return 0;
    }
    release_problem(d);
    Problem& P = d->P;
    std::vector<int> stair;
    plan_problem(d, N, dim, nb, &stair);
    RK(dmalloc(d, (void**)&P.A, (size_t)std::max<int64_t>(P.H, 1) * std::max<int64_t>(P.LDA, 2) * 8));
    for (int64_t j : P.mine) {
        if (P.first[j] < 0) continue;
        P.cols[j] = P.A + P.off[j] * P.LDA + P.coff[j];
        if (P.first[j] == j) RK(dmalloc(d, (void**)&P.invl[j], (size_t)((P.hgt[j] + 127) / 128) * 128 * 128 * 8));
    }
    RK(dmalloc(d, (void**)&P.stair, stair.size() * sizeof(int)));
    CK(cudaMemcpyAsync(P.stair, stair.data(), stair.size() * sizeof(int), cudaMemcpyHostToDevice, d->sm));
    CK(cudaStreamSynchronize(d->sm));   // `stair` is a local
    for (int s = 0; s < 2; ++s) {
        RK(dmalloc(d, (void**)&P.Bop[s], (size_t)std::max<int64_t>(P.Wtot, 1) * nb * 8));
        RK(dmalloc(d, (void**)&P.row[s], (size_t)std::max<int64_t>(P.H, 1) * nb * 8));
        RK(dmalloc(d, (void**)&P.W[s], (size_t)nb * nb * 8));
        if (d->Pr > 1) {
            P.colbuf[s].assign(d->Pr, nullptr);
            for (int r2 = 0; r2 < d->Pr; ++r2) {
                int64_t cnt = 0;
                for (int64_t j : P.mine) if (rowof(d, j) == r2) cnt += P.hgt[j];
                RK(dmalloc(d, (void**)&P.colbuf[s][r2], (size_t)std::max<int64_t>(cnt, 1) * nb * 8));
            }
        }
    }
    RK(dmalloc(d, (void**)&P.eye, (size_t)nb * nb * 8));
    RK(dmalloc(d, (void**)&P.dX, (size_t)N * dim * 8));
    RK(dmalloc(d, (void**)&P.dXr, (size_t)std::max<int64_t>(P.H, 1) * dim * 8));
    RK(dmalloc(d, (void**)&P.dinfo, 256));
    RK(dmalloc(d, (void**)&P.dflag, 256));
    RK(dmalloc(d, (void**)&P.dsum, 256));
    eye_kernel<<<(unsigned)((nb * nb + 255) / 256), 256, 0, d->sm>>>(P.eye, nb);
    CK(cudaGetLastError());
    P.ev_main.assign(P.nblk, nullptr); P.ev_panel.assign(P.nblk, nullptr);
    if (d->profile) P.ev_start.assign(P.nblk, nullptr);
    for (int64_t k = 0; k < P.nblk; ++k) {
        CK(cudaEventCreateWithFlags(&P.ev_main[k], d->profile ? cudaEventDefault : cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&P.ev_panel[k], cudaEventDisableTiming));
        if (d->profile) CK(cudaEventCreate(&P.ev_start[k]));
    }
    return 0;
}

// ---- Gram: every rank builds exactly the blocks it owns, no exchange ------------------------------------------------
int build_gram(gprb200_dist* d) {
    Problem& P = d->P;
    CK(cudaMemsetAsync(P.dinfo, 0, sizeof(int), d->sm));
    for (int64_t j : P.mine) {
        if (!P.cols[j]) continue;
        if (d->Pr == 1) {   // the rows of a block column are contiguous in the matrix: one launch
            GK(d->hm, gprb200_gram_block_dev(d->hm, j * P.nb, P.H - P.off[j], j * P.nb, P.hgt[j], P.cols[j], P.ld[j]));
            continue;
        }
        // the owned rows below are not contiguous in the sample list: one launch against the gathered rows, then the
        // diagonal block again through gram_block (which adds sigma^2 on the global diagonal)
        GK(d->hm, gprb200_gram_pair_dev(d->hm, P.dXr + P.off[j] * P.dim, P.H - P.off[j], P.dX + j * P.nb * P.dim, P.hgt[j],
                                        P.cols[j], P.ld[j]));
        if (P.first[j] == j) GK(d->hm, gprb200_gram_block_dev(d->hm, j * P.nb, P.hgt[j], j * P.nb, P.hgt[j], P.cols[j], P.ld[j]));
    }
    P.factored = false;
    return 0;
}

// ---- one step of the factorisation ------------------------------------------------------------------------------------
// Ranks of grid column k mod Pc, panel stream: L_kk, W = inv(L_kk) down the column, X_i = A_ik W^T for the owned rows
// below; X is left in the row buffer k % 2 AND in cols[k] (the factor, needed by the solve).
int panel_factor(gprb200_dist* d, int64_t k) {
    Problem& P = d->P;
    if (colof(d, k) != d->pc) return 0;
    const int64_t w = P.hgt[k], nb = P.nb;
    const bool diag = (rowof(d, k) == d->pr);
    if (diag) GK(d->hp, gprb200_potrf_panel_async_dev(d->hp, P.cols[k], P.ld[k], w, w, P.invl[k], k * nb, P.dinfo));
    if (k == P.nblk - 1) return 0;          // nothing below the last diagonal block (and w == nb everywhere else)
    double* W = P.W[k % 2];
    if (diag) {
        CK(cudaMemcpyAsync(W, P.eye, (size_t)nb * nb * 8, cudaMemcpyDeviceToDevice, d->sp));
        GK(d->hp, gprb200_trsm_dev(d->hp, P.cols[k], P.ld[k], P.invl[k], w, W, nb, nb, 0));
    }
    if (d->Pr > 1) RK(bcast(d, W, nb * nb, rowof(d, k), d->col_comm, d->sp));
    int64_t o1, first;
    first_at_or_after(d, k + 1, &o1, &first);
    if (first < 0) return 0;
    double* below = P.cols[k] + (o1 - P.off[k]) * P.ld[k];
    double* X = P.row[k % 2];
    const int64_t rows = P.H - o1;
    GK(d->hp, gprb200_gemm_dev(d->hp, 3, rows, w, w, 1.0, below, P.ld[k], W, nb, 0.0, X, nb, 0));    // W is lower triangular
    d->n_flop += 1.0 * rows * w * (w + 64);
    CK(cudaMemcpy2DAsync(below, (size_t)P.ld[k] * 8, X, (size_t)nb * 8, (size_t)w * 8, (size_t)rows, cudaMemcpyDeviceToDevice, d->sp));
    return 0;
}

// All ranks, panel stream: the solved panel rows along the grid rows, then the column operands L_jk of the owned block
// columns j > k down the grid columns, left in Bop[k % 2] in local column order (the B operand of the staircase update).
int distribute(gprb200_dist* d, int64_t k) {
    Problem& P = d->P;
    if (k == P.nblk - 1) return 0;
    const int64_t nb = P.nb;
    int64_t o1, first;
    first_at_or_after(d, k + 1, &o1, &first);
    double* X = P.row[k % 2];
    if (d->Pc > 1 && first >= 0) RK(bcast(d, X, (P.H - o1) * nb, colof(d, k), d->row_comm, d->sp));
    double* B = P.Bop[k % 2];
    auto jbeg = std::upper_bound(P.mine.begin(), P.mine.end(), k);
    if (jbeg == P.mine.end()) return 0;
    const int64_t c0 = P.coff[*jbeg];                  // B row of block j: coff[j] - c0
    auto copy_block = [&](double* dst, const double* src, int64_t rows) -> int {
        CK(cudaMemcpyAsync(dst, src, (size_t)rows * nb * 8, cudaMemcpyDeviceToDevice, d->sp));
        return 0;
    };
    if (d->Pr == 1) {   // this rank holds every block row of X: pick the owned ones
        for (auto it = jbeg; it != P.mine.end(); ++it)
            RK(copy_block(B + (P.coff[*it] - c0) * nb, X + (P.roff[*it] - o1) * nb, P.hgt[*it]));
        return 0;
    }
    std::vector<int64_t> total(d->Pr, 0);
    int groups = 0;
    for (auto it = jbeg; it != P.mine.end(); ++it) total[rowof(d, *it)] += P.hgt[*it];
    for (int r2 = 0; r2 < d->Pr; ++r2) groups += total[r2] > 0;
    if (groups == 1) {  // (Pc a multiple of Pr: all owned block columns meet the same grid row) straight into B
        const int r2 = rowof(d, *jbeg);
        if (r2 == d->pr)
            for (auto it = jbeg; it != P.mine.end(); ++it)
                RK(copy_block(B + (P.coff[*it] - c0) * nb, X + (P.roff[*it] - o1) * nb, P.hgt[*it]));
        return bcast(d, B, total[r2] * nb, r2, d->col_comm, d->sp);
    }
    std::vector<int64_t> pos(d->Pr, 0);
    for (auto it = jbeg; it != P.mine.end(); ++it) {
        const int r2 = rowof(d, *it);
        if (r2 == d->pr) RK(copy_block(P.colbuf[k % 2][r2] + pos[r2] * nb, X + (P.roff[*it] - o1) * nb, P.hgt[*it]));
        pos[r2] += P.hgt[*it];
    }
    NK(ncclGroupStart());
    int grc = 0;    // an error inside the group must not leave it open
    for (int r2 = 0; r2 < d->Pr && grc == 0; ++r2)
        if (total[r2] > 0) grc = bcast(d, P.colbuf[k % 2][r2], total[r2] * nb, r2, d->col_comm, d->sp);
    NK(ncclGroupEnd());
    if (grc) return grc;
    std::fill(pos.begin(), pos.end(), 0);
    for (auto it = jbeg; it != P.mine.end(); ++it) {
        const int r2 = rowof(d, *it);
        RK(copy_block(B + (P.coff[*it] - c0) * nb, P.colbuf[k % 2][r2] + pos[r2] * nb, P.hgt[*it]));
        pos[r2] += P.hgt[*it];
    }
    return 0;
}

// A_ij -= X_i L_jk^T for the owned block columns jfrom <= j <= jto (consecutive in `mine`), all their rows, ONE launch
int update(gprb200_dist* d, int64_t k, int64_t jfrom, int64_t jto, gprb200_handle h) {
    Problem& P = d->P;
    if (P.first[jfrom] < 0) return 0;                   // no owned rows at or below the first (hence any) of these columns
    int64_t o1, f1;
    first_at_or_after(d, k + 1, &o1, &f1);
    const int64_t c0 = P.coff[*std::upper_bound(P.mine.begin(), P.mine.end(), k)];
    const int64_t oS = P.off[jfrom], rows = P.H - oS, ncols = P.coff[jto] + P.hgt[jto] - P.coff[jfrom];
    GK(h, gprb200_gemm_stair_dev(h, rows, ncols, P.nb, -1.0, P.row[k % 2] + (oS - o1) * P.nb, P.nb,
                                 P.Bop[k % 2] + (P.coff[jfrom] - c0) * P.nb, P.nb, 1.0, P.A + oS * P.LDA + P.coff[jfrom], P.LDA,
                                 P.stair + P.coff[jfrom] / 64, oS));
    for (auto it = std::lower_bound(P.mine.begin(), P.mine.end(), jfrom); it != P.mine.end() && *it <= jto; ++it)
        if (P.first[*it] >= 0) {
            const double f = 2.0 * (P.H - P.off[*it]) * P.hgt[*it] * P.nb;
            d->n_flop += f;
            if (h == d->hm) d->n_flop_main += f;
        }
    return 0;
}

int factor(gprb200_dist* d, int* info_out) {
    Problem& P = d->P;
    const int64_t nblk = P.nblk;
    CK(cudaEventRecord(d->ev_gram, d->sm));
    CK(cudaStreamWaitEvent(d->sp, d->ev_gram, 0));
    RK(panel_factor(d, 0));
    for (int64_t k = 0; k < nblk; ++k) {
        const bool nxt_mine = k + 1 < nblk && colof(d, k + 1) == d->pc;
        if (k >= 2) CK(cudaStreamWaitEvent(d->sp, P.ev_main[k - 2], 0));     // buffers k % 2 were last read at step k - 2
        RK(distribute(d, k));
        CK(cudaEventRecord(P.ev_panel[k], d->sp));
        if (nxt_mine) {
            // look-ahead: block column k + 1 gets this panel first, off the main stream, and is factored at once so
            // that its broadcasts overlap everybody's trailing update
            if (k >= 1) CK(cudaStreamWaitEvent(d->sp, P.ev_main[k - 1], 0));
            RK(update(d, k, k + 1, k + 1, d->hp));
            RK(panel_factor(d, k + 1));
        }
        CK(cudaStreamWaitEvent(d->sm, P.ev_panel[k], 0));
        if (!P.ev_start.empty()) CK(cudaEventRecord(P.ev_start[k], d->sm));
        {
            auto it = std::upper_bound(P.mine.begin(), P.mine.end(), nxt_mine ? k + 1 : k);
            if (it != P.mine.end()) RK(update(d, k, *it, P.mine.back(), d->hm));
        }
        CK(cudaEventRecord(P.ev_main[k], d->sm));
    }
    // A rank that meets a non-SPD block keeps following the schedule (nobody hangs in a broadcast); all ranks agree on
    // INFO here: the smallest failing leading minor.
    CK(cudaStreamSynchronize(d->sp));
    CK(cudaStreamSynchronize(d->sm));
    if (!P.ev_start.empty()) {   // where the main stream's time went: inside updates, or waiting for the panel of the step
        d->stall_ms = d->busy_ms = 0.0;
        for (int64_t k = 0; k < nblk; ++k) {
            float a = 0.f, b = 0.f;
            CK(cudaEventElapsedTime(&a, P.ev_start[k], P.ev_main[k]));
            if (k > 0) CK(cudaEventElapsedTime(&b, P.ev_main[k - 1], P.ev_start[k]));
            else CK(cudaEventElapsedTime(&b, d->t1, P.ev_start[0]));
            d->busy_ms += a; d->stall_ms += b;
        }
    }
    int info = 0;
    CK(cudaMemcpy(&info, P.dinfo, sizeof(int), cudaMemcpyDeviceToHost));
    const long long big = 1LL << 62;
    long long flag = info > 0 ? info : big;
    if (d->world > 1) {
        CK(cudaMemcpyAsync(P.dflag, &flag, sizeof(flag), cudaMemcpyHostToDevice, d->sm));
        NK(ncclAllReduce(P.dflag, P.dflag, 1, ncclInt64, ncclMin, d->comm, d->sm));
        CK(cudaMemcpyAsync(&flag, P.dflag, sizeof(flag), cudaMemcpyDeviceToHost, d->sm));
        CK(cudaStreamSynchronize(d->sm));
        d->n_coll += 1;
    }
    info = flag == big ? 0 : (int)flag;
    P.factored = info == 0;
    *info_out = info;
    return 0;
}

int reduce_logdet(gprb200_dist* d) {
    Problem& P = d->P;
    double part = 0.0;
    for (int64_t j : P.mine)
        if (P.invl[j]) {
            double v = 0.0;
            GK(d->hm, gprb200_logdet_dev(d->hm, P.cols[j], P.ld[j], P.hgt[j], &v));
            part += v;
        }
    if (d->world > 1) {
        CK(cudaMemcpyAsync(P.dsum, &part, 8, cudaMemcpyHostToDevice, d->sm));
        NK(ncclAllReduce(P.dsum, P.dsum, 1, ncclDouble, ncclSum, d->comm, d->sm));
        CK(cudaMemcpyAsync(&part, P.dsum, 8, cudaMemcpyDeviceToHost, d->sm));
        CK(cudaStreamSynchronize(d->sm));
        d->n_coll += 1;
    }
    P.logdet = part;
    return 0;
}

}  // namespace

namespace {

// ---- alpha = (K + sigma^2 I)^-1 Y -----------------------------------------------------------------------------------------
// Forward L z = y, block row k: the ranks of grid row k hold the partial sums  sum_{j<k} L_kj z_j  of the blocks they own
// -> ONE small reduce (nb x M doubles) to the diagonal owner, which solves with L_kk and sends z_k down grid column k,
// whose ranks add L_ik z_k to their partial sums with one skinny GEMM over the stacked column. Backward L^T x = z mirrors
// it. Per block step: at most two collectives of nb x M doubles and a handful of small launches, no host
// synchronisation; the cost is latency (2 nblk steps), independent of the number of ranks.
int solve(gprb200_dist* d, const double* Yhost, int64_t M, double* alpha_host) {
    Problem& P = d->P;
    if (!P.factored) return fail(d, -2, "gprb200_dist_solve: no factor (learn first)");
    if (M < 1) return fail(d, -2, "gprb200_dist_solve: M < 1");
    const int64_t nb = P.nb, nblk = P.nblk, N = P.N, ld = (M + 1) / 2 * 2;
    cudaStream_t st = d->sm;
    gprb200_handle h = d->hm;
    double *y = nullptr, *acc = nullptr, *out = nullptr, *msg[2] = {nullptr, nullptr}, *zown = nullptr;
    int64_t ndiag = 0;
    std::vector<int64_t> zoff(nblk, -1);
    for (int64_t k = 0; k < nblk; ++k)
        if (rowof(d, k) == d->pr && colof(d, k) == d->pc) { zoff[k] = ndiag; ndiag += P.hgt[k]; }
    double* stage = nullptr;     // contiguous N x M image of the host array (a pitched copy of 8-byte rows crawls)
    // the workspace is kept for the next call with the same number of columns (one interval per call: GetCredibleInterval)
    const size_t sizes[7] = {(size_t)N * ld, (size_t)N * ld, (size_t)N * ld /* stage: N x M, M <= ld */, (size_t)std::max<int64_t>(P.H, 1) * ld, (size_t)nb * ld,
                             (size_t)nb * ld, (size_t)std::max<int64_t>(ndiag, 1) * ld};
    if (P.ws_ld != ld || P.ws.size() != 7) {
        for (void* p : P.ws) cudaFree(p);
        P.ws.clear();
        P.ws_ld = 0;
        for (int i = 0; i < 7; ++i) {
            void* p = nullptr;
            CK(cudaMalloc(&p, std::max<size_t>(sizes[i] * 8, 256)));
            P.ws.push_back(p);
        }
        P.ws_ld = ld;
    }
    y = (double*)P.ws[0]; out = (double*)P.ws[1]; stage = (double*)P.ws[2]; acc = (double*)P.ws[3];
    msg[0] = (double*)P.ws[4]; msg[1] = (double*)P.ws[5]; zown = (double*)P.ws[6];
    for (int i = 0; i < 7; ++i) CK(cudaMemsetAsync(P.ws[i], 0, std::max<size_t>(sizes[i] * 8, 256), st));
    auto body = [&]() -> int {
        if (ld == M) {
            CK(cudaMemcpyAsync(y, Yhost, (size_t)N * M * 8, cudaMemcpyHostToDevice, st));
        } else {
            CK(cudaMemcpyAsync(stage, Yhost, (size_t)N * M * 8, cudaMemcpyHostToDevice, st));
            repitch_kernel<<<(unsigned)((N * M + 255) / 256), 256, 0, st>>>(y, ld, stage, M, N, M);
            CK(cudaGetLastError());
        }
        // ---- forward
        for (int64_t k = 0; k < nblk; ++k) {
            const int64_t w = P.hgt[k];
            const bool in_row = (rowof(d, k) == d->pr), in_col = (colof(d, k) == d->pc), diag = in_row && in_col;
            double* buf = msg[k % 2];
            if (in_row) {
                double* part = acc + P.roff[k] * ld;
                if (d->Pc > 1) {
                    NK(ncclReduce(part, part, (size_t)(w * ld), ncclDouble, ncclSum, colof(d, k), d->row_comm, st));
                    d->n_coll += 1; d->n_bytes += 8.0 * w * ld;
                }
                if (diag) {
                    sub_kernel<<<(unsigned)((w * ld + 255) / 256), 256, 0, st>>>(buf, y + k * nb * ld, part, w * ld);
                    CK(cudaGetLastError());
                    GK(h, gprb200_trsm_dev(h, P.cols[k], P.ld[k], P.invl[k], w, buf, ld, ld, 0));
                    CK(cudaMemcpyAsync(zown + zoff[k] * ld, buf, (size_t)w * ld * 8, cudaMemcpyDeviceToDevice, st));
                }
            }
            if (!in_col || k == nblk - 1) continue;
            if (d->Pr > 1) RK(bcast(d, buf, nb * ld, rowof(d, k), d->col_comm, st));
            int64_t o1, first;
            first_at_or_after(d, k + 1, &o1, &first);
            if (first >= 0)
                GK(h, gprb200_gemm_dev(h, 1, P.H - o1, ld, w, 1.0, P.cols[k] + (o1 - P.off[k]) * P.ld[k], P.ld[k], buf, ld, 1.0,
                                       acc + o1 * ld, ld, 0));
        }
        // ---- backward: acc now collects x_i for the owned rows
        for (int64_t k = nblk - 1; k >= 0; --k) {
            const int64_t w = P.hgt[k];
            const bool in_row = (rowof(d, k) == d->pr), in_col = (colof(d, k) == d->pc), diag = in_row && in_col;
            double* buf = msg[k % 2];
            if (in_col) {
                int64_t o1, first;
                first_at_or_after(d, k + 1, &o1, &first);
                if (first >= 0)
                    GK(h, gprb200_gemm_dev(h, 2, w, ld, P.H - o1, 1.0, P.cols[k] + (o1 - P.off[k]) * P.ld[k], P.ld[k], acc + o1 * ld,
                                           ld, 0.0, buf, ld, 0));
                else
                    CK(cudaMemsetAsync(buf, 0, (size_t)w * ld * 8, st));
                if (d->Pr > 1) {
                    NK(ncclReduce(buf, buf, (size_t)(nb * ld), ncclDouble, ncclSum, rowof(d, k), d->col_comm, st));
                    d->n_coll += 1; d->n_bytes += 8.0 * nb * ld;
                }
                if (diag) {
                    sub_kernel<<<(unsigned)((w * ld + 255) / 256), 256, 0, st>>>(buf, zown + zoff[k] * ld, buf, w * ld);
                    CK(cudaGetLastError());
                    GK(h, gprb200_trsm_dev(h, P.cols[k], P.ld[k], P.invl[k], w, buf, ld, ld, 1));
                }
            }
            if (in_row) {
                if (d->Pc > 1) RK(bcast(d, buf, nb * ld, colof(d, k), d->row_comm, st));
                CK(cudaMemcpyAsync(acc + P.roff[k] * ld, buf, (size_t)w * ld * 8, cudaMemcpyDeviceToDevice, st));
            }
        }
        // ---- every rank gets the whole alpha: each block row is contributed by grid column 0 of its grid row
        if (d->pc == 0)
            for (int64_t i : P.myrows)
                CK(cudaMemcpyAsync(out + i * nb * ld, acc + P.roff[i] * ld, (size_t)P.hgt[i] * ld * 8, cudaMemcpyDeviceToDevice, st));
        if (d->Pr > 1) {
            NK(ncclAllReduce(out, out, (size_t)(N * ld), ncclDouble, ncclSum, d->comm, st));
            d->n_coll += 1; d->n_bytes += 8.0 * N * ld;
        } else if (d->pc != 0) {
            CK(cudaMemcpyAsync(out, acc, (size_t)N * ld * 8, cudaMemcpyDeviceToDevice, st));   // Pr = 1: every rank's acc already is the whole x
        }
        CK(cudaEventRecord(d->t3, st));
        if (alpha_host) {
            const double* src = out;
            if (ld != M) {
                repitch_kernel<<<(unsigned)((N * M + 255) / 256), 256, 0, st>>>(stage, M, out, ld, N, M);
                CK(cudaGetLastError());
                src = stage;
            }
            CK(cudaMemcpyAsync(alpha_host, src, (size_t)N * M * 8, cudaMemcpyDeviceToHost, st));
        }
        CK(cudaStreamSynchronize(st));
        return 0;
    };
    const int rc = body();
    cudaStreamSynchronize(st);
    return rc;
}

}  // namespace

// ---- C-ABI ------------------------------------------------------------------------------------------------------------------
extern "C" {

int gprb200_dist_get_unique_id(void* id) {
    gprb200_dist* d = nullptr;
    static_assert(sizeof(ncclUniqueId) == GPRB200_DIST_ID_BYTES, "ncclUniqueId size");
    if (!id) return fail(d, -2, "gprb200_dist_get_unique_id: null id");
    ncclUniqueId u;
    NK(ncclGetUniqueId(&u));
    std::memcpy(id, &u, sizeof(u));
    return 0;
}

const char* gprb200_dist_last_error(gprb200_dist_handle d) { return d ? d->err.c_str() : g_create_error.c_str(); }

int gprb200_dist_destroy(gprb200_dist_handle d) {
    if (!d) return 0;
    cudaSetDevice(d->dev);
    release_problem(d);
    if (d->row_comm && d->row_comm != d->comm) ncclCommDestroy(d->row_comm);
    if (d->col_comm && d->col_comm != d->comm) ncclCommDestroy(d->col_comm);
    if (d->comm) ncclCommDestroy(d->comm);
    if (d->hm) gprb200_destroy(d->hm);
    if (d->hp) gprb200_destroy(d->hp);
    for (cudaEvent_t e : {d->t0, d->t1, d->t2, d->t3, d->ev_gram}) if (e) cudaEventDestroy(e);
    if (d->sm) cudaStreamDestroy(d->sm);
    if (d->sp) cudaStreamDestroy(d->sp);
    delete d;
    return 0;
}

static int create_impl(gprb200_dist* d, const void* id) {
    CK(cudaSetDevice(d->dev));
    int lo = 0, hi = 0;
    CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CK(cudaStreamCreateWithPriority(&d->sm, cudaStreamNonBlocking, lo));
    CK(cudaStreamCreateWithPriority(&d->sp, cudaStreamNonBlocking, hi));   // the panel chain is the critical path
    for (cudaEvent_t* e : {&d->t0, &d->t1, &d->t2, &d->t3}) CK(cudaEventCreate(e));
    CK(cudaEventCreateWithFlags(&d->ev_gram, cudaEventDisableTiming));
    if (gprb200_create(&d->hm, d->dev) != 0) return fail(d, -3, std::string("gprb200_create: ") + gprb200_last_error(nullptr));
    if (gprb200_create(&d->hp, d->dev) != 0) return fail(d, -3, std::string("gprb200_create: ") + gprb200_last_error(nullptr));
    GK(d->hm, gprb200_set_stream(d->hm, d->sm));
    GK(d->hp, gprb200_set_stream(d->hp, d->sp));
    if (d->world > 1) {
        ncclUniqueId u;
        std::memcpy(&u, id, sizeof(u));
        NK(ncclCommInitRank(&d->comm, d->world, u, d->rank));
        // row communicator: the Pc ranks of one grid row, rank inside = grid column; column communicator likewise
        // (a 1 x g or g x 1 grid uses the world communicator itself)
        if (d->Pc > 1) {
            if (d->Pr == 1) d->row_comm = d->comm;
            else NK(ncclCommSplit(d->comm, d->pr, d->pc, &d->row_comm, nullptr));
        }
        if (d->Pr > 1) {
            if (d->Pc == 1) d->col_comm = d->comm;
            else NK(ncclCommSplit(d->comm, d->pc, d->pr, &d->col_comm, nullptr));
        }
    }
    return 0;
}

int gprb200_dist_create(gprb200_dist_handle* out, int device, int rank, int world, const void* id, int Pr, int Pc) {
    gprb200_dist* d = nullptr;
    if (!out) return fail(d, -2, "gprb200_dist_create: null out");
    *out = nullptr;
    if (world < 1 || rank < 0 || rank >= world) return fail(d, -2, "gprb200_dist_create: rank / world out of range");
    if (world > 1 && !id) return fail(d, -2, "gprb200_dist_create: unique id needed when world > 1");
    if (Pr == 0 && Pc == 0) {   // tall: 8 -> 4 x 2, 4 -> 2 x 2, 2 -> 2 x 1 (measured: profiles/r02_c5_grid_comparison.json)
        Pc = 1;
        while ((2 * Pc) * (2 * Pc) <= world && world % (2 * Pc) == 0) Pc *= 2;
        Pr = world / Pc;
    }
    if (Pr < 1 || Pc < 1 || Pr * Pc != world) return fail(d, -2, "gprb200_dist_create: Pr * Pc must equal world");
    gprb200_dist* n = new gprb200_dist();
    n->dev = device; n->rank = rank; n->world = world; n->Pr = Pr; n->Pc = Pc; n->pr = rank / Pc; n->pc = rank % Pc;
    int rc = create_impl(n, id);
    if (rc != 0) {
        g_create_error = n->err;
        gprb200_dist_destroy(n);
        return rc;
    }
    *out = n;
    return 0;
}

int gprb200_dist_set_kernel(gprb200_dist_handle d, const int32_t* ops, int nops, const double* params, int nparams) {
    if (!d) return -2;
    CK(cudaSetDevice(d->dev));
    GK(d->hm, gprb200_set_kernel(d->hm, ops, nops, params, nparams));
    d->have_kernel = true;
    return 0;
}

int gprb200_dist_set_option(gprb200_dist_handle d, const char* name, int64_t value) {
    if (!d) return -2;
    if (name && std::strcmp(name, "profile") == 0) {    // local: timed per-step events (takes effect at the next allocation)
        d->profile = value != 0;
        release_problem(d);
        return 0;
    }
    if (name && (std::strcmp(name, "mirror_rows") == 0 || std::strcmp(name, "mirror_cols") == 0)) {
        (name[7] == 'r' ? d->mirror_rows : d->mirror_cols) = value != 0;
        release_problem(d);
        return 0;
    }
    GK(d->hm, gprb200_set_option(d->hm, name, value));
    GK(d->hp, gprb200_set_option(d->hp, name, value));
    return 0;
}

int gprb200_dist_learn(gprb200_dist_handle d, const double* X, int64_t N, int dim, const double* Y, int64_t M, double sigma,
                       int64_t nb, double* alpha, double* logdet) {
    if (!d) return -2;
    if (!d->have_kernel) return fail(d, -2, "gprb200_dist_learn: set the kernel first");
    if (!X || N < 1 || dim < 1) return fail(d, -2, "gprb200_dist_learn: X / N / dim");
    if (alpha && (!Y || M < 1)) return fail(d, -2, "gprb200_dist_learn: alpha requested without Y");
    if (nb == 0) nb = 1024;
    if (nb < 128 || nb % 128 != 0) return fail(d, -2, "gprb200_dist_learn: nb must be a multiple of 128 (the Cholesky diagonal block)");
    CK(cudaSetDevice(d->dev));
    d->n_coll = d->n_bytes = d->n_flop = d->n_flop_main = 0.0;
    const double l0 = (double)gprb200_launch_count(d->hm) + (double)gprb200_launch_count(d->hp);
    RK(setup_problem(d, N, dim, nb));
    CK(cudaMemcpyAsync(d->P.dX, X, (size_t)N * dim * 8, cudaMemcpyHostToDevice, d->sm));
    GK(d->hm, gprb200_set_train_dev(d->hm, d->P.dX, N, dim, nullptr, 0, sigma));
    if (d->Pr > 1)
        for (int64_t i : d->P.myrows)
            CK(cudaMemcpyAsync(d->P.dXr + d->P.roff[i] * dim, d->P.dX + i * nb * dim, (size_t)d->P.hgt[i] * dim * 8,
                               cudaMemcpyDeviceToDevice, d->sm));
    CK(cudaEventRecord(d->t0, d->sm));
    RK(build_gram(d));
    CK(cudaEventRecord(d->t1, d->sm));
    int info = 0;
    RK(factor(d, &info));
    CK(cudaEventRecord(d->t2, d->sm));
    CK(cudaEventSynchronize(d->t2));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, d->t0, d->t1)); d->ms[0] = ms;
    CK(cudaEventElapsedTime(&ms, d->t1, d->t2)); d->ms[1] = ms;
    d->ms[2] = 0.0;
    if (info != 0) {
        d->n_launch = (double)gprb200_launch_count(d->hm) + (double)gprb200_launch_count(d->hp) - l0;
        char b[160];
        std::snprintf(b, sizeof b, "kernel matrix not positive definite: leading minor %d (of N = %lld)", info, (long long)N);
        d->err = b;
        return info;
    }
    RK(reduce_logdet(d));
    if (logdet) *logdet = d->P.logdet;
    if (alpha) {
        CK(cudaEventRecord(d->t2, d->sm));
        RK(solve(d, Y, M, alpha));
        CK(cudaEventElapsedTime(&ms, d->t2, d->t3)); d->ms[2] = ms;
    }
    d->n_launch = (double)gprb200_launch_count(d->hm) + (double)gprb200_launch_count(d->hp) - l0;
    return 0;
}

int gprb200_dist_solve(gprb200_dist_handle d, const double* Y, int64_t M, double* alpha) {
    if (!d) return -2;
    if (!Y || !alpha) return fail(d, -2, "gprb200_dist_solve: null Y / alpha");
    CK(cudaSetDevice(d->dev));
    CK(cudaEventRecord(d->t2, d->sm));
    RK(solve(d, Y, M, alpha));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, d->t2, d->t3)); d->ms[2] = ms;
    return 0;
}

int gprb200_dist_plan(int64_t N, int64_t nb, int Pr, int Pc, int rank, int mirror_rows, int mirror_cols, double* summary,
                      int32_t* row_owned, int32_t* col_owned, int32_t* stair, int64_t stair_len) {
    gprb200_dist* d = nullptr;
    if (nb == 0) nb = 1024;
    if (N < 1 || nb < 128 || nb % 128 != 0 || Pr < 1 || Pc < 1 || rank < 0 || rank >= Pr * Pc || !summary)
        return fail(d, -2, "gprb200_dist_plan: bad arguments");
    gprb200_dist t;     // never touches CUDA or NCCL
    t.Pr = Pr; t.Pc = Pc; t.world = Pr * Pc; t.rank = rank; t.pr = rank / Pc; t.pc = rank % Pc;
    t.mirror_rows = mirror_rows != 0; t.mirror_cols = mirror_cols != 0;
    std::vector<int> st;
    plan_problem(&t, N, 1, nb, &st);
    const Problem& P = t.P;
    summary[0] = (double)P.H; summary[1] = (double)P.Wtot; summary[2] = 8.0 * (double)std::max<int64_t>(P.H, 1) * (double)std::max<int64_t>(P.LDA, 2);
    summary[3] = plan_flops(&t); summary[4] = (double)P.nblk; summary[5] = (double)st.size();
    if (row_owned) for (int64_t i = 0; i < P.nblk; ++i) row_owned[i] = P.roff[i] >= 0 ? (int32_t)P.roff[i] : -1;
    if (col_owned) for (int64_t j = 0; j < P.nblk; ++j) col_owned[j] = P.coff[j] >= 0 ? (int32_t)P.coff[j] : -1;
    if (stair) for (int64_t u = 0; u < stair_len && u < (int64_t)st.size(); ++u) stair[u] = st[u];
    return 0;
}

double gprb200_dist_phase_ms(gprb200_dist_handle d, int which) { return (d && which >= 0 && which < 3) ? d->ms[which] : -1.0; }

double gprb200_dist_counter(gprb200_dist_handle d, int which) {
    if (!d) return -1.0;
    switch (which) {
        case 0: return d->n_coll;
        case 1: return d->n_bytes;
        case 2: return d->n_flop;
        case 3: return d->n_launch;
        case 4: return d->stall_ms;
        case 5: return d->busy_ms;
        case 6: return d->n_flop_main;
        case 7: return d->Pr;
        case 8: return d->Pc;
    }
    return -1.0;
}

}  // extern "C"
