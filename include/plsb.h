/* include/plsb.h — C-ABI of libplsb.so, the B200 (sm_100a) engine behind PLS::Model.
 *
 * The reference (tjhladish/PLS @ d976786) has no FFI layer: its boundary is the C++ API of
 * include/PLS/pls.h, implemented on Eigen in src/pls.cpp.  This header is what a binding for the
 * fit / predict / cross-validation path binds instead of Eigen arithmetic; every entry point cites
 * the reference lines it replaces.  The repo's own include/PLS/pls.h + pls_b200/host/model.cpp is
 * that binding for C++ (source-compatible PLS::Model), pls_b200/_capi.py the one for ctypes.
 *
 * Conventions
 *  - plain C types only: FP64 pointers + sizes; no Eigen / torch / CUDA types cross the ABI.
 *  - host matrices are COLUMN-MAJOR with an explicit leading dimension (Eigen's layout).
 *  - every call returns 0 on success, non-zero on CUDA / NCCL / argument error;
 *    plsb_last_error() then holds a message.  There is no CPU fallback: without a usable
 *    sm_100 device every compute entry point fails with PLSB_ERR_CUDA.
 *  - method: 0 = KERNEL_TYPE1, 1 = KERNEL_TYPE2 (pls.h:131).
 *  - residual blocks "E" use the reference's Ev layout (pls.cpp:475, 500-507, 520):
 *    M stacked column-major (rows x A) matrices, E[(m*A + (c-1))*rows + i].
 */
#ifndef PLSB_H
#define PLSB_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PLSB_OK 0
#define PLSB_ERR_ARG 1
#define PLSB_ERR_CUDA 2
#define PLSB_ERR_NCCL 3
#define PLSB_ERR_UNSUPPORTED 4

#define PLSB_KERNEL_TYPE1 0
#define PLSB_KERNEL_TYPE2 1

typedef struct plsb_ctx *plsb_ctx_t;         /* one device (+ optional communicator) */
typedef struct plsb_dataset *plsb_dataset_t; /* device-resident X, Y (row shard of this rank) */
typedef struct plsb_model *plsb_model_t;     /* device-resident fit results W, P, R, Q, T */

/* ---- context ------------------------------------------------------------------------------- */
const char *plsb_last_error(void);
const char *plsb_version(void);
int plsb_device_count(int *count);
int plsb_ctx_create(int device, plsb_ctx_t *ctx);
int plsb_ctx_destroy(plsb_ctx_t ctx);
int plsb_ctx_sync(plsb_ctx_t ctx);
/* Device allocations are stream-ordered and cached in the device's memory pool between calls; this returns the
 * cache to the driver (e.g. before handing the GPU to another library), and frees the pinned host staging buffers and
 * host threads that large uploads / downloads of this context left behind (they come back with the next one). */
int plsb_ctx_trim(plsb_ctx_t ctx);
/* Row-sharded single fits: one context per process/GPU.  Rank 0 calls
 * plsb_comm_unique_id, the 128 bytes are broadcast by the caller's own plumbing
 * (torch.distributed / MPI / a file), then every rank calls plsb_comm_init.  After that,
 * plsb_fit on a dataset flagged `sharded` all-reduces the K*M / K+1 / K*K partial sums. */
int plsb_comm_unique_id(void *id128);
int plsb_comm_init(plsb_ctx_t ctx, int rank, int nranks, const void *id128);
int plsb_comm_allreduce_test(plsb_ctx_t ctx, double *host_buf, int64_t n);
/* What the context's communicator is: its rank / size, and whether the per-component [g | tt] exchange of a row-sharded
 * fit runs inside the partial-sum kernel over NVLink peer memory (1) or as a separate ncclAllReduce (0). */
int plsb_comm_info(plsb_ctx_t ctx, int *rank, int *nranks, int *peer_exchange);

/* ---- datasets: Model's `_X,_Y` copy (pls.cpp:344) becomes a device-resident mirror ---------- */
/* X: N x K, Y: N x M host column-major (pinned or pageable).  Rows are this rank's shard.
 * The transfer is PCIe-bound; for tall matrices (256 <= K <= 3072, N >= max(16 K, 65536); PLSB_UPLOAD_GRAM=0/1
 * overrides) the rows go up in blocks and XX = X'X, XY = X'Y (pls.cpp:396-398) are accumulated on the FP64 tensor
 * cores while the next block is in flight; the data set then carries them and plsb_fit runs from them. */
int plsb_dataset_create(plsb_ctx_t ctx, const double *X, int64_t ldx, int64_t N, int64_t K,
                        const double *Y, int64_t ldy, int64_t M, plsb_dataset_t *ds);
/* Same, but X / Y are DEVICE pointers (column-major) on ctx's device. */
int plsb_dataset_create_device(plsb_ctx_t ctx, const double *dX, int64_t ldx, int64_t N, int64_t K,
                               const double *dY, int64_t ldy, int64_t M, plsb_dataset_t *ds);
/* Deterministic synthetic latent-factor data generated on the device; bit-identical
 * to oracle/pls_oracle.c:orc_synth.  Rows [row0, row0+N) of the global data set. */
int plsb_dataset_create_synth(plsb_ctx_t ctx, uint64_t seed, int64_t row0, int64_t N, int64_t K,
                              int64_t M, int64_t L, plsb_dataset_t *ds);
/* A new data set on ctx's device holding the rows of `parts` one after the other (device-to-device; parts on other
 * GPUs are read over NVLink).  nparts = 1 replicates a data set onto another GPU, nparts = G gathers the row shards
 * of a sharded fit into the full replica cross-validation wants — the training data are the Model's own copy
 * (`_X,_Y`, pls.cpp:344, pls.h:251) and here that copy lives in HBM only, so replicas are made from it, not from
 * the caller's host matrix.  Pageable host sources of plsb_dataset_create are packed into pinned staging buffers by
 * a pool of host threads (PLSB_PACK_THREADS, default min(16, cores / contexts)) while earlier blocks are in flight. */
int plsb_dataset_concat(plsb_ctx_t ctx, int64_t nparts, const plsb_dataset_t *parts, plsb_dataset_t *out);
int plsb_dataset_destroy(plsb_dataset_t ds);
int plsb_dataset_shape(plsb_dataset_t ds, int64_t *N, int64_t *K, int64_t *M);
/* Copy the data set back to host column-major buffers (either may be NULL). */
int plsb_dataset_download(plsb_dataset_t ds, double *X, int64_t ldx, double *Y, int64_t ldy);
/* Mark the data set as one row shard of a communicator-wide fit. */
int plsb_dataset_set_sharded(plsb_dataset_t ds, int sharded);

/* ---- fit: Model::plsr (pls.cpp:390-437) ------------------------------------------------------ */
/* Outputs (host, column-major, real parts; the reference's complex members have zero imaginary
 * part): W,P,R are K x A, Q is M x A, T is N x A (this rank's rows; KERNEL_TYPE1 only; may be NULL).
 * `method` selects the reference's kernel (pls.cpp:419-424): KERNEL_TYPE1 streams X once per component,
 * KERNEL_TYPE2 builds X'X first.  A data set that already carries X'X (see plsb_dataset_create) is fitted from it
 * whatever `method` says — the same model (both kernels sit at the same distance from the reference's result,
 * profiles/r1c_fullsize_parity.json) — and T, when asked for, comes from one extra pass T = X R;
 * PLSB_FIT_NO_GRAM=1 forces the streaming passes.  T in pinned memory is sent home while the fit is still running. */
int plsb_fit(plsb_dataset_t ds, int64_t A, int method, double *W, double *P, double *R, double *Q,
             double *T);
/* The same fit with every result FORMED AND KEPT IN HBM (W, P, R, Q and — KERNEL_TYPE1, like the reference,
 * pls.cpp:394/434 — the N x A scores T of this rank's rows): nothing crosses PCIe until plsb_model_download asks
 * for it (any output pointer may be NULL).  What a caller that only wants a few of the matrices, or repeated fits
 * (bench.py's device-resident figure), uses. */
int plsb_fit_resident(plsb_dataset_t ds, int64_t A, int method, plsb_model_t *model);
int plsb_model_download(plsb_model_t model, double *W, double *P, double *R, double *Q, double *T);
/* The N x A scores alone (column-major).  as_complex != 0: T receives N*A (re, im) pairs with im = 0 — the reference's
 * Model::T is complex-typed (pls.h:251-254, Mat2Dc), so the drop-in fills its member straight from the transfer's
 * staging buffers instead of widening a second host copy. */
int plsb_model_download_scores(plsb_model_t model, double *T, int as_complex);
int plsb_model_destroy(plsb_model_t model);
/* One-shot form taking host matrices (upload + fit), the shape Model's constructor has (pls.cpp:340-353). */
int plsb_fit_host(plsb_ctx_t ctx, const double *X, int64_t ldx, int64_t N, int64_t K, const double *Y,
                  int64_t ldy, int64_t M, int64_t A, int method, double *W, double *P, double *R,
                  double *Q, double *T);

/* ---- cross products: XY = X'Y (pls.cpp:396), XX = X'X (pls.cpp:398, FP64 DMMA) -------------- */
int plsb_xty(plsb_dataset_t ds, double *XY /* K x M col-major */);
int plsb_xtx(plsb_dataset_t ds, double *XX /* K x K col-major, may be NULL: device only */);

/* Build XX = X'X and XY = X'Y once and KEEP them on the data set (pls.cpp:396, 398; the matrices an upload from host
 * memory accumulates under the transfer): plsb_cv_folds* and fits that follow start from them instead of rebuilding them
 * per call.  across_ranks != 0 — a replicated data set on a context with a communicator, collective: rank g reads rows
 * [g N/G, (g+1) N/G) only and the partial matrices are summed over the ranks. */
int plsb_dataset_build_gram(plsb_dataset_t ds, int across_ranks);

/* ---- prediction: scores / fitted_values / residuals / cv_NEW_DATA (pls.cpp:439-459, 494-510) -- */
/* S = Xn * R[:, :comp]   (Nn x comp) */
int plsb_scores(plsb_ctx_t ctx, const double *R, int64_t K, int64_t comp, const double *Xn, int64_t ldx,
                int64_t Nn, double *S);
/* F = Xn * (R_c Q_c')    (Nn x M);  if Yn != NULL returns residuals Yn - F instead (pls.cpp:453-455) */
int plsb_fitted(plsb_ctx_t ctx, const double *R, const double *Q, int64_t K, int64_t M, int64_t comp,
                const double *Xn, int64_t ldx, int64_t Nn, const double *Yn, int64_t ldy, double *F);
/* residuals for 1..A components in one pass: E = M stacked (Nn x A) blocks (cv_NEW_DATA layout) */
int plsb_predict_residuals(plsb_ctx_t ctx, const double *R, const double *Q, int64_t K, int64_t M,
                           int64_t A, const double *Xn, int64_t ldx, int64_t Nn, const double *Yn,
                           int64_t ldy, double *E);

/* B = R[:, :comp] * Q[:, :comp]'   (K x M col-major; pls.cpp:444-447) */
int plsb_coefficients(plsb_ctx_t ctx, const double *R, const double *Q, int64_t K, int64_t M, int64_t comp,
                      double *B);
/* sse[(c-1)*M + m] = sum_i (Yn - Xn B_c)[i, m]^2 for every c = 1..A in ONE pass over Xn (M x A col-major):
 * Model::SSE (pls.cpp:457-459) for all component counts, what print_explained_variance (pls.cpp:551-562)
 * and explained_variance (461-467) consume. */
int plsb_sse_prefixes(plsb_ctx_t ctx, const double *R, const double *Q, int64_t K, int64_t M, int64_t A,
                      const double *Xn, int64_t ldx, int64_t Nn, const double *Yn, int64_t ldy, double *sse);

/* sse[m] for one component count (Model::SSE, pls.cpp:457-459) */
int plsb_sse(plsb_ctx_t ctx, const double *R, const double *Q, int64_t K, int64_t M, int64_t comp, const double *Xn,
             int64_t ldx, int64_t Nn, const double *Yn, int64_t ldy, double *sse);

/* ---- cross-validation: cv_LOO (pls.cpp:469-491), cv_LSO (pls.cpp:512-549) ------------------- */
/* nfolds folds, each leaving out test_size rows: test_idx[f*test_size + i].  The training set of a
 * fold is the complement (cross products are order-independent up to rounding, so the reference's
 * training-row order does not need to be shipped).  Folds [fold0, fold0+nfolds_local) are computed
 * by this call (fold sharding across GPUs: no communication); E has `rows_total` = total folds *
 * test_size rows per (m, c) column and only this call's rows are written.
 * `method` is validated (0 or 1) and otherwise has no effect here: every fold is refitted from the downdated cross
 * products X'X - X_H'X_H, X'Y - X_H'Y_H (the KERNEL_TYPE2 recurrence, pls.cpp:423-424), or with the KERNEL_TYPE1
 * recurrence in the one-launch small-data kernel; the two give the same residuals to rounding (tests hold both to
 * the reference's cv_LOO / cv_LSO fixtures at 1e-9). */
int plsb_cv_folds(plsb_dataset_t ds, int64_t A, int method, int64_t nfolds_total, int64_t test_size,
                  const int64_t *test_idx, int64_t fold0, int64_t nfolds_local, double *E);
/* Same folds, with validation's PRESS sums (pls.cpp:235-252: column sums of squared residuals) reduced ON THE DEVICE:
 * press[(c-1)*M + m] (M x A column-major) = sum over THIS CALL's folds and held-out rows of E[m][row][c]^2 (fixed-order
 * reduction, no atomics; fold shards add their partial matrices in shard order, MSE = sum / rows_total).  E may be
 * NULL: then the residual block never leaves the device (LOO on 1M x 2048 would ship M*A*N*8 = 2 GB per call).
 * optimal_num_components (pls.cpp:265-289) ranks individual residuals (Wilcoxon) and therefore still needs E.
 * Data sets small enough for one SM's shared memory (N*K*8 <~ 200 KB: the reference's nir / toy files) run as ONE
 * launch, each CTA fitting whole folds with the KERNEL_TYPE1 recurrence; larger ones use Gram downdates + one
 * batched DMMA GEMM per component (PLSB_CV_NO_SMALL=1 forces the latter). */
int plsb_cv_folds_press(plsb_dataset_t ds, int64_t A, int method, int64_t nfolds_total, int64_t test_size,
                        const int64_t *test_idx, int64_t fold0, int64_t nfolds_local, double *E, double *press);

/* ---- column statistics and z-scoring on the device: SST (pls.cpp:69-77), colwise_stdev (79-87),
 * colwise_z_scores (93-111) — the CLI's pre-processing (main.cpp:24-25).  Two-pass formulas like the reference;
 * a zero-variance column becomes NaN exactly as there. ------------------------------------------------------ */
/* mean / sst / stdev (each K, any may be NULL) of host column-major X; means_in (K or NULL): use these means */
int plsb_colstats(plsb_ctx_t ctx, const double *X, int64_t ldx, int64_t N, int64_t K, const double *means_in,
                  double *mean, double *sst, double *stdev);
/* The same column statistics for callers that must not allocate device memory.  No counterpart in the reference.
 * plsb_colstats needs a context (streams, pools) and allocates a staging copy of X plus scratch for its partial sums.
 * A plug-in running inside an application that has already claimed the GPU's memory for its own arena, a call made from
 * a latency-critical loop where cudaMalloc/cudaFree would stall every stream of the device, or code held to a
 * "no allocations after start-up" budget cannot afford that.  This variant takes no context and calls no cudaMalloc,
 * cudaFree or pool routine: the caller hands in DEVICE pointers it already owns — dX, the N x K matrix, column-major
 * with leading dimension ldx >= N, and d_mean / d_sst / d_stdev, K doubles each, any of them NULL — plus the device
 * ordinal and a cudaStream_t (as void *, NULL = default stream).  The kernels are queued on that stream and the call
 * returns without synchronising; results are valid in stream order.  N < 2 gives sst = 0 like the reference.
 * Limit: K <= 8192 (PLSB_ERR_ARG beyond it, nothing is launched; N is unbounded).  The partial sums live in a 4.06 MiB
 * __device__ array compiled into the library, which every GPU that loads libplsb.so holds until the process exits.
 * All calls on one GPU share that array: calls that may overlap on the GPU (two streams, two threads) race and
 * must be ordered by the caller; back-to-back calls on one stream need nothing cleared between them.
 * The first call on each device resolves the array's device address (cudaGetSymbolAddress) and caches it: make that
 * call during initialisation, before starting any stream capture.  The caller's current device is restored. */
int plsb_colstats_static(int device, void *stream, const double *dX, int64_t ldx, int64_t N, int64_t K,
                         double *d_mean, double *d_sst, double *d_stdev);
/* Z = (X - mean) / stdev by column; mean_in/stdev_in both NULL: computed from X (N-1 variance) */
int plsb_colwise_z_scores(plsb_ctx_t ctx, const double *X, int64_t ldx, int64_t N, int64_t K, const double *mean_in,
                          const double *stdev_in, double *Z, int64_t ldz);
/* standardise a device-resident data set in place (X and Y); outputs (K / M values) may be NULL */
int plsb_dataset_zscore(plsb_dataset_t ds, double *x_mean, double *x_stdev, double *y_mean, double *y_stdev);

/* ---- host-side validation statistics (stay on the CPU: exact integer results, libstdc++ semantics) -- */
double plsb_normalcdf(double z);                                        /* pls.cpp:152-160 */
double plsb_wilcoxon(const double *e1, const double *e2, int64_t n);    /* pls.cpp:190-211 */
/* out: M x A column-major; out_type 0 = RESS, 1 = MSE (pls.cpp:235-261) */
int plsb_validation(const double *E, int64_t M, int64_t rows, int64_t A, int out_type, double *out);
int plsb_optimal_num_components(const double *E, int64_t M, int64_t rows, int64_t A, double alpha,
                                int64_t *out);                          /* pls.cpp:265-289 */
/* fold lists of cv_LSO (pls.cpp:515-538, rand_nchoosek 218-227) for std::mt19937(seed);
 * test_idx: num_trials x test_size, train_idx (may be NULL): num_trials x train_size, row-major */
int plsb_lso_folds(int64_t N, double test_fraction, int64_t num_trials, uint32_t seed, int64_t *test_idx,
                   int64_t *train_idx);

/* ---- measurement hooks (bench.py roofline: per-kernel CUDA-event times on the launch stream) -- */
#define PLSB_PROF_PASS_FUSED 0   /* g = X'(X r), tt, t : one streaming read of X */
#define PLSB_PROF_PASS_XTY 1     /* S = X'Y           : one streaming read of X */
#define PLSB_PROF_STEP 2         /* per-component single-CTA step */
#define PLSB_PROF_XTX 3          /* DMMA X'X */
#define PLSB_PROF_CV_GEMM 4      /* batched-fold DMMA GEMM */
#define PLSB_PROF_CV_FOLD 5      /* per-fold small kernels */
#define PLSB_PROF_FIT_TOTAL 6    /* whole fit, device side */
#define PLSB_PROF_UPLOAD 7       /* H2D + layout transform */
#define PLSB_PROF_REDUCE 8      /* cross-CTA partial-row sums after a pass */
#define PLSB_PROF_SCORES 9      /* DMMA scores S = X R */
#define PLSB_PROF_NSLOTS 10
/* on: 0 off; 1 a CUDA event pair around every launch group; 2 only around whole fits (device time of a fit without
 * an event between its kernels, which would switch off their programmatic dependent launch); 3 like 1, but the
 * two-lane cross-validation engine keeps both lanes in flight and its kernels are timed inside the overlap (1 runs it
 * as one lane, so that kernel times add up to the call) */
int plsb_profile_enable(plsb_ctx_t ctx, int on);
int plsb_profile_reset(plsb_ctx_t ctx);
/* total milliseconds and launch count accumulated in a slot since the last reset */
int plsb_profile_get(plsb_ctx_t ctx, int slot, double *ms_total, int64_t *launches);
/* number of kernels this library launched on ctx since creation (bench.py "gpu_launches") */
int plsb_launch_count(plsb_ctx_t ctx, int64_t *count);

#ifdef __cplusplus
}
#endif
#endif /* PLSB_H */
