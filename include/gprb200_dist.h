/* gprb200_dist.h - C-ABI of libgprb200_dist.so: GP Learn over the GPUs of one node, NCCL called from C++.
 *
 * The reference (agiger/GPR) has no multi-process code; what this library reproduces, for N beyond one GPU's memory or
 * patience (BASELINE config C5, N = 131072: the kernel matrix alone is 137 GB), is the RESULT of
 * GaussianProcess::ComputeRegressionVectors - InvertKernelMatrix + C * Y, lib/GaussianProcess.cpp:642-672, :531-560 -
 * i.e. alpha = (K + sigma^2 I)^-1 Y and log det(K + sigma^2 I), by a Cholesky factorisation on a Pr x Pc block-cyclic
 * process grid. It is the schedule of gpr_b200/dist.py:BlockCyclicGP moved under the C-ABI:
 * every local block operation goes through libgprb200.so (include/gprb200.h), every exchange is an NCCL broadcast /
 * reduce on the row, column or world communicator, issued on the library's own CUDA streams; the host never waits for
 * the device inside the factorisation.
 *
 * Process model: ONE RANK PER GPU. A rank is a process (torchrun, mpirun, a service's workers) or a thread; every rank
 * creates its own handle. Bootstrap: rank 0 calls gprb200_dist_get_unique_id and hands the 128 bytes to the other ranks
 * by whatever the host application has (a file, a socket, MPI_Bcast, torch.distributed); then every rank calls
 * gprb200_dist_create with the same id, world, Pr, Pc. All entry points below are COLLECTIVE: every rank of the grid
 * calls them with the same arguments (X, Y, kernel) in the same order.
 *
 * Return codes as in gprb200.h: 0 ok, > 0 LAPACK-style INFO of the factorisation (same value on every rank), < 0 argument /
 * CUDA / NCCL error with text in gprb200_dist_last_error(). */
#ifndef GPRB200_DIST_H
#define GPRB200_DIST_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct gprb200_dist* gprb200_dist_handle;

#define GPRB200_DIST_ID_BYTES 128

/* Rank 0 only: fills id[GPRB200_DIST_ID_BYTES] (ncclGetUniqueId). */
int gprb200_dist_get_unique_id(void* id);
/* Every rank: device = CUDA ordinal of this rank's GPU; Pr * Pc == world; rank r sits at grid row r / Pc, grid column
 * r % Pc. Pr = Pc = 0 picks the default grid (tall: Pc the largest power of two with Pc^2 <= world dividing world;
 * 8 -> 4 x 2, 4 -> 2 x 2, 2 -> 2 x 1). Block row i lives on grid row 0 1 .. Pr-1 Pr-1 .. 1 0 0 1 .. (mirrored cyclic,
 * which evens out the lower triangle's bias; option "mirror_rows" 0 gives the plain i mod Pr), block column j on grid
 * column j mod Pc (option "mirror_cols"). */
int gprb200_dist_create(gprb200_dist_handle* out, int device, int rank, int world, const void* id, int Pr, int Pc);
int gprb200_dist_destroy(gprb200_dist_handle d);
const char* gprb200_dist_last_error(gprb200_dist_handle d); /* d may be NULL: last create() error of this thread */

/* The kernel in the op-code form of gprb200_set_kernel (gprb200.h; KernelFactory::GetKernel, include/KernelFactory.h). */
int gprb200_dist_set_kernel(gprb200_dist_handle d, const int32_t* ops, int nops, const double* params, int nparams);
/* Options of the LOCAL single-GPU library handles (gprb200_set_option names, e.g. "gemm_small"), both streams; and
 * "profile" (0 / 1): timed events around every trailing update, for counters 4 and 5 below; "mirror_rows" /
 * "mirror_cols" (0 / 1, the same on every rank): the block-to-grid maps described at gprb200_dist_create. */
int gprb200_dist_set_option(gprb200_dist_handle d, const char* name, int64_t value);

/* Learn: X is N x dim, Y is N x M (row-major HOST arrays, the same on every rank; M small - many outputs shard by
 * column instead, see gpr_b200.dist.ShardedOutputsGP). nb = block size, a multiple of 128 (0: 1024).
 * On return every rank has alpha (N x M, host) and *logdet. alpha / logdet may be NULL (factorisation only).
 * Device memory: the owned blocks of the lower triangle (about 4 N^2 / world bytes) stay allocated until the next
 * learn / destroy so that gprb200_dist_solve can reuse the factor; a learn of the same (N, dim, nb) reuses them without
 * any allocation. */
int gprb200_dist_learn(gprb200_dist_handle d, const double* X, int64_t N, int dim, const double* Y, int64_t M,
                       double sigma, int64_t nb, double* alpha, double* logdet);
/* Another right-hand side against the factor of the last learn (collective). */
int gprb200_dist_solve(gprb200_dist_handle d, const double* Y, int64_t M, double* alpha);

/* Device time of the phases of the last learn on THIS rank, in ms (CUDA events on the library's streams):
 * which = 0 Gram blocks, 1 factorisation, 2 solve. Take the max over ranks for a job time. */
double gprb200_dist_phase_ms(gprb200_dist_handle d, int which);
/* Counters of the last learn on this rank: 0 NCCL calls issued, 1 bytes broadcast / reduced, 2 GEMM flop issued,
 * 3 kernels launched through libgprb200.so; with option "profile": 4 ms the main stream waited for a panel, 5 ms it spent
 * inside trailing updates, 6 the flop of those updates (factorisation only); 7 / 8: Pr / Pc of the grid in use. */
double gprb200_dist_counter(gprb200_dist_handle d, int which);

/* Host-only (no GPU, no NCCL call): the layout rank `rank` of a Pr x Pc grid would use for an N x N problem - for capacity
 * planning and for tests of the ownership logic. summary[6]: 0 stacked rows H, 1 local columns Wtot, 2 bytes of the local
 * matrix, 3 flop this rank's GEMMs execute in one factorisation (compare N^3 / (3 Pr Pc)), 4 number of block rows,
 * 5 entries of the staircase table. row_owned / col_owned (nblk ints each, may be NULL): offset of block row i / block
 * column j in the local matrix, -1 if another rank's. stair (may be NULL): the first stair_len entries of the table
 * gprb200_gemm_stair_dev gets (first local row needed by each 64 local columns). */
int gprb200_dist_plan(int64_t N, int64_t nb, int Pr, int Pc, int rank, int mirror_rows, int mirror_cols, double* summary,
                      int32_t* row_owned, int32_t* col_owned, int32_t* stair, int64_t stair_len);

#ifdef __cplusplus
}
#endif
#endif
