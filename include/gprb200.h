/*
 * gprb200.h — C-ABI of the B200-native GaussianProcess Learn/Predict hot path.
 *
 * This is the seam between the host C++ API (gpr::GaussianProcess<T>, gpr::Kernel<T>,
 * gpr::GaussianLogLikelihood<T>; see gpr_b200/include/) and the hand-written sm_100a CUDA
 * kernels in gpr_b200/csrc/.  Plain pointers and sizes only; no C++/torch types.
 *
 * Conventions
 *   - all matrices are dense, ROW-MAJOR, IEEE fp64 (the reference's MatrixType is a row-major
 *     Eigen matrix, reference include/GaussianProcess.h:41-46; its float instantiation casts to
 *     double around LAPACK, reference include/LAPACKUtils.h:87,96);
 *   - "host" entry points take host pointers and do the H2D/D2H copies themselves;
 *     the *_dev entry points take DEVICE pointers (inputs already resident in HBM);
 *   - every function returns 0 on success, >0 = LAPACK-style INFO ("leading minor of order k is
 *     not positive definite"), <0 = argument / CUDA / NCCL error; gprb200_last_error() gives text.
 *     The C++ wrapper rethrows as std::string, the reference's error convention
 *     (reference lib/GaussianProcess.cpp:123,136,198,400,713).
 *   - one handle == one GP model on one GPU; calls on one handle must be serialised by the
 *     caller (the reference is not thread-safe either: include/GaussianProcess.h:164-165).
 */
#ifndef GPRB200_H
#define GPRB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gprb200_ctx* gprb200_handle;

/* Kernel "program": the kernel tree in post-order.  Leaves consume their parameters from the
 * flat parameter vector in order; the parameter order is exactly Kernel::GetParameters() of the
 * reference (k1's block first: include/Kernel.h:187-191).
 *   GAUSSIAN      [sigma, scale]          include/Kernel.h:465-479
 *   GAUSSIAN_EXP  [sigma, scale] (log)    include/Kernel.h:580-598
 *   WHITE         [scale]                 include/Kernel.h:695-713
 *   RQ            [scale, sigma, alpha]   include/Kernel.h:794-808
 *   PERIODIC      [scale, b, sigma]       include/Kernel.h:912-948
 *   SUM / PRODUCT pop two, push one       include/Kernel.h:165-178, 314-327            */
enum gprb200_kernel_op {
    GPRB200_K_GAUSSIAN = 1,
    GPRB200_K_GAUSSIAN_EXP = 2,
    GPRB200_K_WHITE = 3,
    GPRB200_K_RQ = 4,
    GPRB200_K_PERIODIC = 5,
    GPRB200_K_SUM = 6,
    GPRB200_K_PRODUCT = 7
};
#define GPRB200_MAX_OPS 32
#define GPRB200_MAX_PARAMS 48
#define GPRB200_MAX_STACK 8

/* learn() flags */
#define GPRB200_WANT_CORE 1u     /* also form C = inv(K + sigma^2 I) explicitly (for -CoreMatrix.txt) */

/* loglik() modes */
#define GPRB200_LOGLIK_REFERENCE_CLAMP 0 /* reproduce include/Likelihood.h:180-188 (det truncated to double, clamped) */
#define GPRB200_LOGLIK_TRUE 1            /* cp = -0.5 * logdet from the Cholesky diagonal */

/* phase timers (milliseconds, CUDA events on the handle's stream) */
enum gprb200_phase {
    GPRB200_T_GRAM = 0,
    GPRB200_T_POTRF = 1,
    GPRB200_T_SOLVE = 2,
    GPRB200_T_INVERSE = 3,
    GPRB200_T_PREDICT = 4,
    GPRB200_T_DGRAM = 5,
    GPRB200_T_LOGLIK = 6,
    GPRB200_T_COUNT = 8
};

/* ---- lifetime -------------------------------------------------------------------------- */
int gprb200_create(gprb200_handle* h, int device);
int gprb200_destroy(gprb200_handle h);
const char* gprb200_last_error(gprb200_handle h); /* h may be NULL: last create() error */
const char* gprb200_version(void);
/* Run all work of this handle on an existing cudaStream_t (e.g. torch's current stream). */
int gprb200_set_stream(gprb200_handle h, void* cuda_stream);
int gprb200_synchronize(gprb200_handle h);
/* number of CUDA kernels this handle has launched since creation (bench.py's gpu_launches) */
int64_t gprb200_launch_count(gprb200_handle h);
int gprb200_phase_ms(gprb200_handle h, double* ms /* [GPRB200_T_COUNT] */);
/* tuning knobs: "nb" (Cholesky block), "gram_lower_only" ... returns <0 for unknown names */
int gprb200_set_option(gprb200_handle h, const char* name, int64_t value);
/* with option "profile" = 1: per-launch CUDA-event statistics of the dominant kernels.
 * "tb_ms"/"tb_flops"/"tb_launches": SYRK trailing-update GEMM launches of the last factorisation;
 * "gemv_ms"/"gemv_bytes"/"gemv_launches": Predict GEMV launches since "profile" was last set. */
int gprb200_get_stat(gprb200_handle h, const char* name, double* out);

/* ---- model definition ------------------------------------------------------------------ */
/* replaces Kernel<T> virtual dispatch per matrix element (lib/GaussianProcess.cpp:392) */
int gprb200_set_kernel(gprb200_handle h, const int32_t* ops, int nops, const double* params, int nparams);
/* parameters only (hyper-parameter loop, include/GaussianProcessInference.h:90) */
int gprb200_set_kernel_params(gprb200_handle h, const double* params, int nparams);


/* replaces the AoS sample/label lists (lib/GaussianProcess.cpp:36-51) + ComputeLabelMatrix (:621-639)
 * X: N x d, Y: N x M, sigma: observation noise (SetSigma). Copies; invalidates any factorisation. */
int gprb200_set_train(gprb200_handle h, const double* X, int64_t N, int d, const double* Y, int64_t M, double sigma);
/* The same from the reference's own storage, std::vector<VectorType> (lib/GaussianProcess.cpp:48-49): one host pointer
 * per sample (d doubles) and per label vector (M doubles). The label rows are streamed through a ring of pinned
 * staging buffers (worker threads fill piece k+1 while piece k crosses PCIe), so ComputeLabelMatrixInternal's
 * N x M gather (:621-639; 25.8 GB at N = 4096, M = 786432) never exists on the host. gprb200_set_train with a
 * contiguous Y >= 64 MB and gprb200_get_alpha use the same pipeline (option "stream_io" = 0 turns it off). */
int gprb200_set_train_rows(gprb200_handle h, const double* const* Xrows, int64_t N, int d, const double* const* Yrows,
                           int64_t M, double sigma);
/* AddSample x N + Initialize (+ the host copy of the regression vectors that Save and operator== need) in one pipelined
 * call: label vectors are uploaded, solved and downloaded in column blocks of ~512 MB on three streams, so at
 * N = 4096, M = 786432 the 25.8 GB each way overlap the solve instead of adding to it. alpha_out: N x M host buffer
 * or NULL (alpha then stays on the device only). Equivalent to set_train_rows + learn + get_alpha (to rounding: a narrow last block may take a differently
 * ordered triangular solve). */
int gprb200_learn_rows(gprb200_handle h, const double* const* Xrows, int64_t N, int d, const double* const* Yrows, int64_t M,
                       double sigma, double* alpha_out);
int gprb200_set_train_dev(gprb200_handle h, const double* dX, int64_t N, int d, const double* dY, int64_t M,
                          double sigma);
int gprb200_set_sigma(gprb200_handle h, double sigma);

/* Load path (lib/GaussianProcess.cpp:184-268): install alpha (N x M) and optionally the core
 * matrix C (N x N, may be NULL = efficient storage) WITHOUT re-factorising (m_Initialized = true, :267). */
int gprb200_set_state(gprb200_handle h, const double* X, int64_t N, int d, const double* Y /* may be NULL */,
                      const double* alpha, int64_t M, const double* core /* may be NULL */, double sigma);

/* ---- Learn ------------------------------------------------------------------------------- */
/* ComputeKernelMatrix (+AddNoiseToKernelMatrix when add_noise): lib/GaussianProcess.cpp:363-402, 375-381.
 * K_out: N x N host buffer (full symmetric). Returns -3 if K has non-finite entries (:399). */
int gprb200_gram(gprb200_handle h, double* K_out, int add_noise);
int gprb200_gram_dev(gprb200_handle h, double* dK_out, int64_t ldk, int add_noise);

/* ComputeRegressionVectors (lib/GaussianProcess.cpp:642-672): Gram + sigma^2 I, Cholesky
 * (replaces lu_invert, include/LAPACKUtils.h:38-56,85-97), alpha = (K + sigma^2 I)^-1 Y (:661). */
int gprb200_learn(gprb200_handle h, unsigned flags);
/* separately callable stages (bench / tests) */
int gprb200_factorize(gprb200_handle h);        /* Gram + potrf only */
int gprb200_solve_alpha(gprb200_handle h);      /* alpha from the current factor */

int gprb200_get_alpha(gprb200_handle h, double* alpha /* N x M */);
int gprb200_get_alpha_dev(gprb200_handle h, const double** d_alpha, int64_t* ld);
/* Device view of the factor held by the handle: L (N x N, ld, lower; the strict upper part is stale K) and, if d_invL is
 * not NULL, the ceil(N/128) inverted 128 x 128 diagonal blocks gprb200_trsm_dev takes. Valid until the next call that
 * re-factorises. Lets a multi-GPU driver share the solve / inverse work of one replicated factor (dist.py). */
int gprb200_get_factor_dev(gprb200_handle h, const double** d_L, int64_t* ld, const double** d_invL);
/* C = inv(K + sigma^2 I), formed lazily from the factor (InvertKernelMatrix, lib/GaussianProcess.cpp:531-618) */
int gprb200_get_core(gprb200_handle h, double* C /* N x N */);
/* lower Cholesky factor L (row-major, strictly-upper part zeroed) */
int gprb200_get_factor(gprb200_handle h, double* L /* N x N */);
/* log det(K + sigma^2 I) = 2 sum log L_ii (replaces the long-double determinant, lib/GaussianProcess.cpp:527) */
int gprb200_get_logdet(gprb200_handle h, double* logdet);

/* InvertKernelMatrix (lib/GaussianProcess.cpp:531-618; include/LAPACKUtils.h:38-73) for an arbitrary symmetric
 * positive-definite host matrix A (n x n): Ainv = inv(A) through the same Cholesky kernels, logdet (may be NULL)
 * = log det A. Does not touch the model held by the handle. Returns k > 0 if the leading minor k is not SPD. */
int gprb200_spd_inverse(gprb200_handle h, const double* A, int64_t n, double* Ainv, double* logdet);

/* lapack::lu_invert (include/LAPACKUtils.h:38-56, :85-97: dgetrf + dgetri) for a GENERAL square host matrix A (n x n):
 * Ainv = inv(A) by in-place Gauss-Jordan elimination with partial pivoting on the device (the pivot choice of dgetrf).
 * A utility of the public LAPACKUtils interface, not on the GP path (which always factors by Cholesky). n <= 32768.
 * Returns k > 0 (dgetrf's INFO) if the matrix is exactly singular at step k or holds non-finite values. */
int gprb200_lu_inverse(gprb200_handle h, const double* A, int64_t n, double* Ainv);

/* ---- device-pointer building blocks (multi-GPU drivers, probes) -------------------------------------- */
/* C = alpha op(A) op(B) + beta C on the FP64 tensor path, row-major, device pointers, 16-byte aligned, even lds.
 * form 0 (NT): A is M x K, B is N x K, C = A B^T; 1 (NN): B is K x N; 2 (TN): A is K x M, B is K x N, C = A^T B;
 * 3: as 0 with B LOWER TRIANGULAR (B[n][k] = 0 for k > n; entries above the diagonal are not read beyond the tile's
 * last column): the product with an inverted Cholesky block, X = A inv(L)^T, at half the flop.
 * lower != 0 (square C): only tiles meeting the lower triangle are computed (SYRK-style trailing update).
 * Replaces the Eigen / BLAS products of lib/GaussianProcess.cpp:60, :661. */
int gprb200_gemm_dev(gprb200_handle h, int form, int64_t M, int64_t N, int64_t K, double alpha, const double* dA,
                     int64_t lda, const double* dB, int64_t ldb, double beta, double* dC, int64_t ldc, int lower);

/* C = alpha A B^T + beta C (NT form as above) for a "staircase" C: the trailing update of a block-cyclic Cholesky,
 * where a rank's owned block columns start at different rows, in ONE launch instead of one per block column (whose
 * partial last waves add up). dStair (DEVICE ints, one per 64 columns of C): the first row, in the caller's own row
 * numbering, that those columns need; row r of C has the number row0 + r. 128-row tiles lying entirely before the stair of
 * all their columns are skipped, everything else is computed in full (so entries above a column's stair may or may not be
 * updated: the caller must not care). Part of the dpotrf that replaces lib/GaussianProcess.cpp:531-560 at large N. */
int gprb200_gemm_stair_dev(gprb200_handle h, int64_t M, int64_t N, int64_t K, double alpha, const double* dA, int64_t lda,
                           const double* dB, int64_t ldb, double beta, double* dC, int64_t ldc, const int* dStair, int64_t row0);

/* Block [row0, row0+nrows) x [col0, col0+ncols) of K + sigma^2 I over the training samples (ComputeKernelMatrix +
 * AddNoiseToKernelMatrix, lib/GaussianProcess.cpp:363-402, :375-381), written to a device buffer: lets every rank of
 * a multi-GPU factorisation build exactly the block columns it owns, with no exchange. */
int gprb200_gram_block_dev(gprb200_handle h, int64_t row0, int64_t nrows, int64_t col0, int64_t ncols, double* dOut,
                           int64_t ldo);
/* k(a_i, b_j) for two explicit DEVICE sample sets (na x d and nb x d, d the dimension of the installed samples), no noise
 * term: ComputeKernelMatrix (lib/GaussianProcess.cpp:363-402) for rows that are not contiguous in the sample list - a
 * rank of a 2-D block-cyclic grid gathers its block rows once and builds each owned block column in one launch. */
int gprb200_gram_pair_dev(gprb200_handle h, const double* dXa, int64_t na, const double* dXb, int64_t nb, double* dOut,
                          int64_t ldo);
/* One block column of the Cholesky factor: dP is rows x w (ld), its top w x w block is factored in place and the
 * rows below are solved against it (the dpotrf + dtrsm a blocked LAPACK dpotrf does per panel;
 * include/LAPACKUtils.h:59-73). dInvL: ceil(w/128) blocks of 128 x 128 (inverted diagonal blocks, for trsm_dev). */
int gprb200_potrf_panel_dev(gprb200_handle h, double* dP, int64_t ld, int64_t rows, int64_t w, double* dInvL, int* info);
/* The same with NO host synchronisation (multi-GPU schedules must not block the host once per panel): dInfo is a
 * caller-owned DEVICE int that is not reset by the call; the first leading minor that is not positive definite stores
 * col0 + its 1-based local index there (compare-and-swap from 0), so one int collects the INFO of a whole distributed
 * factorisation and is read once at the end. */
int gprb200_potrf_panel_async_dev(gprb200_handle h, double* dP, int64_t ld, int64_t rows, int64_t w, double* dInvL,
                                  int64_t col0, int* dInfo);
/* B <- L^-1 B (transposed = 0) or L^-T B (transposed != 0); L n x n lower (ldl), B n x M (ldb), dInvL as above. */
int gprb200_trsm_dev(gprb200_handle h, const double* dL, int64_t ldl, const double* dInvL, int64_t n, double* dB,
                     int64_t ldb, int64_t M, int transposed);
/* Columns of W = inv(L) (the TRTRI half of dpotri, include/LAPACKUtils.h:59-73) for the cyclic column blocks
 * j = first, first + stride, ... of width col_block (a multiple of 128), stored compactly in dW (n x ldw): local column
 * q * col_block + t is global column (first + q * stride) * col_block + t; rows above a column block are zero. The
 * columns are independent, so g ranks with first = rank, stride = g split the n^3 / 3 flop without any exchange. */
int gprb200_trtri_cols_dev(gprb200_handle h, const double* dL, int64_t ldl, const double* dInvL, int64_t n, int64_t col_block,
                           int64_t first, int64_t stride, double* dW, int64_t ldw);
/* 2 * sum_i log L(i,i) of an n x n device factor (replaces the determinant of lib/GaussianProcess.cpp:527) */
int gprb200_logdet_dev(gprb200_handle h, const double* dL, int64_t ld, int64_t n, double* logdet);

/* ---- Predict ----------------------------------------------------------------------------- */
/* Predict (lib/GaussianProcess.cpp:55-61) for T test points at once; T = 1 is the app's call.
 * Xs: T x d.  mean: T x M.  ci (may be NULL): T values of GetCredibleInterval (:103-114),
 * i.e. 2*sqrt(max(0, k(x,x) - Kx^T C Kx)). */
int gprb200_predict(gprb200_handle h, const double* Xs, int64_t T, double* mean, double* ci);
int gprb200_predict_dev(gprb200_handle h, const double* dXs, int64_t T, double* d_mean, int64_t ld_mean,
                        double* d_ci);
/* operator()(x,y) (lib/GaussianProcess.cpp:85-99): posterior covariance for T pairs (x_t, y_t). */
int gprb200_posterior_cov(gprb200_handle h, const double* Xa, const double* Xb, int64_t T, double* cov);
/* PredictDerivative (lib/GaussianProcess.cpp:65-81): mean (M) and D (d x M) = -Xd^T (Kx o alpha_i);
 * no 1/sigma^2 factor, as in the reference. */
int gprb200_predict_derivative(gprb200_handle h, const double* x, double* mean, double* D);
/* ComputeKernelVector (lib/GaussianProcess.cpp:675-693) for T points: Kx is T x N */
int gprb200_kernel_vector(gprb200_handle h, const double* Xs, int64_t T, double* Kx);

/* Allocation-free Predict GEMV. NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * dOut[m] = sum_n dKx[n] dA[n * lda + m] (N x M row-major dA, e.g. the regression vectors from gprb200_get_alpha_dev
 * with dKx the kernel vector of one test point): the product gprb200_predict_dev forms for T = 1, same kernels, same
 * summation order. All pointers are DEVICE pointers; dA 16-byte aligned, lda even and >= M. The row-split partial
 * sums go to a file-scope static __device__ array of 4 Mi doubles (32 MiB per GPU, held until the process exits)
 * instead of a stream-ordered allocation; the call makes no cudaMalloc / cudaFree / allocating library call.
 * Limit: splits * round_up(M, 2) <= 4194304, splits = ceil(N / rows_per_split) as chosen by the kernel's geometry
 * (e.g. N = 4096: up to M = 262144 columns); larger shapes return -2 with a message naming the limit.
 * Every call on a GPU shares that one array (it keeps the last call's contents; nothing needs to start at zero):
 * calls that OVERLAP on the same GPU - different streams or handles - race. Serialise them. */
int gprb200_gemv_t_noalloc_dev(gprb200_handle h, const double* dKx, const double* dA, int64_t lda, int64_t N, int64_t M,
                               double* dOut);

/* ---- Likelihood -------------------------------------------------------------------------- */
/* ComputeDerivativeKernelMatrix (lib/GaussianProcess.cpp:460-495): D is (P*N) x N, block p at rows [pN,(p+1)N). */
int gprb200_dgram(gprb200_handle h, double* D_out);
/* GaussianLogLikelihood (include/Likelihood.h:166-344). Re-factorises with the current kernel
 * parameters (as the reference does on every call, :237), then
 *   value[k] = -0.5 y_k^T C y_k + cp + ct                       (M values; may be NULL)
 *   grad[p]  = 0.5 tr((alpha alpha^T - C) dK/dtheta_p)           (P values; may be NULL)  (:280-282)
 *   jac[i,p] = 0.5 tr((alpha_i alpha_i^T - C) dK/dtheta_p)       (M x P;   may be NULL)  (:335-341) */
int gprb200_loglik(gprb200_handle h, int mode, double* value, double* grad, double* jac);
/* Row-block share of the likelihood-gradient contraction (include/Likelihood.h:268-282 computes P full N x N products
 * for these traces): out[p] = sum_{i < nrows, j < ncols} dS[i * lds + j] * dK_p(x_{row0+i}, x_j), dS a DEVICE
 * nrows x ncols block of a symmetric weight matrix (e.g. rows of alpha alpha^T - C), dK generated on the fly, out P
 * HOST doubles. ncols < N restricts the sum to the leading columns (the lower part of a symmetric matrix, whose
 * strictly-lower blocks the caller weights twice). Summing disjoint row blocks over ranks (ncclAllReduce of P doubles)
 * gives tr(S dK_p). */
int gprb200_dgram_contract_rows_dev(gprb200_handle h, int64_t row0, int64_t nrows, int64_t ncols, const double* dS,
                                    int64_t lds, double* out);
/* q[k] = y_k^T (K + sigma^2 I)^-1 y_k, the data-fit term of include/Likelihood.h:175 (M values) */
int gprb200_quadform(gprb200_handle h, double* q);

/* ---- PCA (reference include/PCA.h) --------------------------------------
 * The reference's PCA<T> maps the GP's predicted principal weights to the displacement field, once per frame
 * (include/DataParser.h:471-495, :211-286). Data matrices are D x n row-major, ONE SAMPLE PER COLUMN, exactly the
 * MatrixType the reference passes. A PCA object lives on the device and stream of the handle it was created from and
 * must be destroyed before that handle. Singular vectors are defined up to sign (as with Eigen's BDCSVD). */
typedef struct gprb200_pca* gprb200_pca_handle;
int gprb200_pca_create(gprb200_handle h, gprb200_pca_handle* p);
int gprb200_pca_destroy(gprb200_pca_handle p);
/* "passes" (default 2): Gram/Jacobi passes of the fit; the second restores the accuracy the first Gram squares away. */
int gprb200_pca_set_option(gprb200_pca_handle p, const char* name, int64_t value);
/* PCA(MatrixType& X, int nFeatures), include/PCA.h:33-49: mean = X.rowwise().mean(), thin SVD of the mean-free data,
 * sigma = singular values / sqrt(n), U (D x n). n <= 8192, n <= D. */
int gprb200_pca_fit(gprb200_pca_handle p, const double* X, int64_t D, int64_t n, int64_t n_features);
int gprb200_pca_fit_dev(gprb200_pca_handle p, const double* dX, int64_t ldx, int64_t D, int64_t n, int64_t n_features);
/* PCA(std::string path, int nFeatures), include/PCA.h:51-65, after the three ReadMatrix calls (Mean/Sigma/U.bin). */
int gprb200_pca_set(gprb200_pca_handle p, const double* mean, const double* sigma, const double* U, int64_t D, int64_t n,
                    int64_t n_features);
/* GetMean / GetEigenvalues / GetMatrixU (include/PCA.h:71-81); any pointer may be NULL. */
int gprb200_pca_get(gprb200_pca_handle p, double* mean, double* sigma, double* U);
int gprb200_pca_dims(gprb200_pca_handle p, int64_t* D, int64_t* n, int64_t* n_features);
int gprb200_pca_get_stat(gprb200_pca_handle p, const char* name, double* out); /* "fit_ms", "sweeps" */
/* DimensionalityReduction / DimensionalityReductionFast (include/PCA.h:92-108): F = (U diag(1/sigma))^T (X - mean),
 * first rows_out rows (rows_out <= n; rows_out <= nFeatures streams the compact basis). X: D x T, F: rows_out x T. */
int gprb200_pca_project(gprb200_pca_handle p, const double* X, int64_t T, int64_t rows_out, double* F);
int gprb200_pca_project_dev(gprb200_pca_handle p, const double* dX, int64_t ldx, int64_t T, int64_t rows_out, double* dF,
                            int64_t ldf);
/* GetReconstruction (include/PCA.h:110-115): Xout (D x T) = (U diag(sigma))[:, :nFeatures] W + mean,
 * W: nFeatures x T. _dev: transposed != 0 reads W as T x nFeatures (the layout Predict writes). */
int gprb200_pca_reconstruct(gprb200_pca_handle p, const double* W, int64_t n_weights, int64_t T, double* Xout);
int gprb200_pca_reconstruct_dev(gprb200_pca_handle p, const double* dW, int64_t ldw, int transposed, int64_t n_weights,
                                int64_t T, double* dOut, int64_t ldo);
/* gp->Predict(v) for every frame (apps/GaussianProcessPredict.cpp:188-203) followed by inversePca()
 * (include/DataParser.h:471-495) without leaving the device. Xs: T x d; fields: D x T (may be NULL);
 * weights: T x M predicted principal weights (may be NULL). The GP must have M = nFeatures outputs. */
int gprb200_predict_reconstruct(gprb200_handle h, gprb200_pca_handle p, const double* Xs, int64_t T, double* fields,
                                double* weights);
int gprb200_predict_reconstruct_dev(gprb200_handle h, gprb200_pca_handle p, const double* dXs, int64_t T, double* dFields,
                                    int64_t ldo, double* dWeights, int64_t ldw);

/* ---- Sparse (inducing-point) GP (reference include/SparseGaussianProcess.h) ----
 * The dense samples are the handle's training set (set_train), the inducing points are added here. m <= n.
 * PreComputeRegression, Predict, operator()(x, y), and the sparse log-likelihood (include/SparseLikelihood.h). */
/* AddInducingSample x m (:61-76) + SetJitter (:118-121); Xm: m x d */
int gprb200_sparse_set_inducing(gprb200_handle h, const double* Xm, int64_t m, int d, double jitter);
int gprb200_sparse_set_jitter(gprb200_handle h, double jitter);
/* Initialize -> PreComputeRegression (:124-141, :274-312): K = k(Xm,Xm) + jitter I, Knm, S = K + sigma^-2 Kmn Knm,
 * alpha = sigma^-2 S^-1 Kmn Y (the value the reference's K^-1 (sigma^-2 K Sigma Kmn Y) reduces to). */
int gprb200_sparse_learn(gprb200_handle h);
int gprb200_sparse_get_alpha(gprb200_handle h, double* alpha); /* m x M regression vectors */
/* SparseGaussianLogLikelihood (include/SparseLikelihood.h:152-535). Recomputes every matrix from the current kernel
 * parameters, sigma and jitter, as the reference does on each call. mode as gprb200_loglik: REFERENCE_CLAMP follows
 * EfficientDeterminant's long-double range and the clamps at :177-185 / :351-359, TRUE uses
 * log det(sigma^2 I + Knm K^-1 Kmn) = n log sigma^2 - log det K + log det S from the two Cholesky factors.
 *   value (M or NULL)      operator() (:152-217): -1/2 y_k^T C_inv y_k + cp + ct per output column k
 *   grad  (P or NULL)      GetParameterDerivatives (:219-285); as the reference, an error unless M == 1
 *   jac   (M x P or NULL)  GetValueAndJacobian (:411-535): row k = the gradient of the single-output problem on
 *                          column k (the reference's own last line, :530, is only defined for M == P == 1)
 * A NaN determinant is an error when only the value is asked for (:177-185 do not test for it) and clamps like an
 * underflow otherwise (:351, :477). */
int gprb200_sparse_loglik(gprb200_handle h, int mode, double* value, double* grad, double* jac);
/* ComputeDerivativeKernelVectorMatrix (include/SparseGaussianProcess.h:237-264): the stack d k(x_i, z_j) / d theta_p,
 * (P*n) x m row-major, block p at rows [p n, (p+1) n). The likelihood above never materialises it. */
int gprb200_sparse_dgram_cross(gprb200_handle h, double* D_out);
/* Predict (:84-90) for T points: mean T x M (may be NULL); var (may be NULL): operator()(x, x) per point */
int gprb200_sparse_predict(gprb200_handle h, const double* Xs, int64_t T, double* mean, double* var);
/* operator()(x, y) (:92-106) for T pairs: k(x,y) - Kx^T K^-1 Ky + Kx^T S^-1 Ky */
int gprb200_sparse_cov(gprb200_handle h, const double* Xa, const double* Xb, int64_t T, double* cov);

#ifdef __cplusplus
}
#endif
#endif /* GPRB200_H */
