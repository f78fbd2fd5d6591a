// pls_b200/csrc/internal.h — private structures and kernel launchers of libplsb.so.
#pragma once
#include <atomic>
#include <vector>

#include "common.cuh"

namespace plsb { class HostPool; }

#define PLSB_MAX_RANKS 16
#define PLSB_XCHG_LEN 8192          /* doubles per exchange slot */

struct plsb_ctx {
    // Lifetime: the creator's handle is one reference, every data set / resident model holds another.  plsb_ctx_destroy
    // drops the creator's; streams, communicator and scratch are torn down when the LAST reference goes, so a data set
    // that outlives its context handle (a cached replica across PLS::set_devices, objects finalised in arbitrary order
    // at interpreter exit) stays usable and destroyable.
    std::atomic<int> refs{1};
    bool closed = false;        // plsb_ctx_destroy was called: no new objects on this context
    int device = 0;
    int sm_count = 148;
    size_t smem_optin = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;   // H2D / D2H transfers that overlap kernels on `stream`
    cudaEvent_t xfer_ev[4] = {nullptr, nullptr, nullptr, nullptr};   // upload: copied[2], transposed[2]
    cudaEvent_t col_ev = nullptr;         // fit: "column a of T is complete"
    // scratch for cross-CTA partial sums (grown on demand)
    double *part = nullptr;
    size_t part_bytes = 0;
    // communicator (lazy-loaded NCCL); rank/nranks == 0/1 when absent
    void *comm = nullptr;
    int rank = 0, nranks = 1;
    // peer-memory exchange area for the in-kernel all-reduce of the per-component [g | tt] vector (allreduce.cu)
    struct Xchg {
        bool ready = false;
        bool failed = false;            // an exchange timed out: sharded calls on this context are refused from now on
        char *base = nullptr;           // this rank's area (cudaMalloc: IPC-exportable)
        char *peer[PLSB_MAX_RANKS] = {nullptr};
        bool ipc_opened[PLSB_MAX_RANKS] = {false};
        unsigned long long seq = 0;     // exchanges done so far (identical on every rank)
    } xchg;
    // Pageable host sources (an ordinary Eigen matrix, a numpy array) reach the copy engine through a ring of three
    // pinned, write-combined staging buffers that a small pool of host threads fills (dataset.cu: staged uploads);
    // allocated on first use, kept for the life of the context.
    struct HostStage {
        char *buf[3] = {nullptr, nullptr, nullptr};
        size_t bytes = 0;
        cudaEvent_t done[3] = {nullptr, nullptr, nullptr};   // "the H2D copy that read buf[i] has finished"
        plsb::HostPool *pool = nullptr;
        double pack_s = 0, wait_s = 0, pack_bytes = 0;       // PLSB_TRACE=1: where a staged upload spent its time
        // the way back: two pinned (cached, NOT write-combined: the CPU reads them) buffers for large D2H results
        char *dbuf[2] = {nullptr, nullptr};
        size_t dbytes = 0;
        cudaEvent_t ddone[2] = {nullptr, nullptr};
    } hstage;
    // profiling
    int prof = 0;   // 0 off, 1 an event pair around every launch group, 2 only around whole fits (PLSB_PROF_FIT_TOTAL),
                    // 3 like 1 but the two-lane CV engine keeps both lanes (its kernels are timed IN the overlap)
    struct Pending { int slot; cudaEvent_t a, b; };
    std::vector<Pending> pending;
    std::vector<cudaEvent_t> pool;
    double prof_ms[PLSB_PROF_NSLOTS] = {0};
    int64_t prof_n[PLSB_PROF_NSLOTS] = {0};
    int64_t launches = 0;
};

struct plsb_dataset {
    plsb_ctx *ctx = nullptr;
    int64_t N = 0, K = 0, M = 0;
    int64_t ldk = 0;  // row stride of X (doubles), even, >= K, zero padded
    int64_t ldm = 0;  // row stride of Y (doubles), even, >= M, zero padded
    double *X = nullptr;  // row-major N x ldk
    double *Y = nullptr;  // row-major N x ldm
    bool sharded = false;
    // XX = X'X of THIS data set's rows, accumulated on the tensor cores while the rows were still crossing PCIe
    // (dataset.cu: upload_rows_with_gram).  When present a fit runs from it (fit.cu) instead of re-reading X.
    double *G = nullptr;  // K x ldg
    double *S0 = nullptr; // M x ldk: X'Y of this data set's rows, accumulated the same way (valid with G)
    int64_t ldg = 0;
    bool gram_valid = false;
    bool gram_reduced = false;  // row-sharded: G already summed over the ranks
    int gram_agreed = -1;       // row-sharded: do ALL ranks hold a Gram matrix? (-1: not asked yet)
};

// a fitted model whose matrices live in HBM (plsb_fit_resident): component-major, like the fit driver's buffers
struct plsb_model {
    plsb_ctx *ctx = nullptr;
    int64_t N = 0, K = 0, ldk = 0, M = 0, A = 0;
    int method = 0;
    double *W = nullptr, *P = nullptr, *R = nullptr;   // A x ldk
    double *Q = nullptr;                               // A x M
    double *T = nullptr;                               // A x N (KERNEL_TYPE1 only)
};

namespace plsb {

int live_contexts();             // contexts of this process that have not been torn down
void hstage_release(plsb_ctx *c);  // dataset.cu
void ctx_retain(plsb_ctx *c);
void ctx_release(plsb_ctx *c);   // tears the context down when the last reference goes
int ensure_part(plsb_ctx *c, size_t bytes);
// Stream-ordered allocations from the device's default memory pool.  plsb_ctx_create raises the pool's release
// threshold so freed blocks stay cached: repeated fits / CV calls / uploads do not pay cudaMalloc again
// (multi-GB cudaMalloc + cudaFree cost tens of ms per call).  plsb_ctx_trim returns the cache to the driver.
int dev_alloc(plsb_ctx *c, void **p, size_t bytes);
void dev_free(plsb_ctx *c, void *p);
// free device memory as the wave sizing should see it: driver-free plus what the pool holds unused
int dev_mem_available(plsb_ctx *c, size_t *bytes);
// in-place sum across the communicator's ranks on c->stream (no-op without a communicator)
int allreduce_sum(plsb_ctx *c, double *buf, size_t n);
// out[i] = sum over ranks of (sum_b part[b*ld + i]): the cross-CTA partial-row sum of a pass fused with the
// cross-GPU all-reduce over NVLink peer memory when `across_ranks` (falls back to the sum kernel + NCCL when the
// exchange area is not available).  nb may be 0 (empty shard).
int launch_reduce_partials(plsb_ctx *c, const double *part, int64_t ld, int nb, int64_t len, double *out,
                           bool across_ranks);
int xchg_setup(plsb_ctx *c, int (*allgather)(plsb_ctx *, const void *, void *, size_t));
void xchg_teardown(plsb_ctx *c);
int xchg_check(plsb_ctx *c);  // error if an exchange timed out
// n contiguous doubles from device memory to PAGEABLE host memory at copy-engine speed: chunks go to pinned staging
// buffers on the copy stream (after everything queued on c->stream so far) and a pool of host threads moves chunk i
// out while chunk i+1 is in flight.  as_complex: the destination is n (re, im) pairs, im = 0 — the reference's model
// matrices are complex-typed (pls.h:251-254), so the widening costs no extra pass over host memory.
// A cudaMemcpy into pageable memory runs at 2-3 GB/s here (measured: 256 MB of scores in 0.1 s).
int download_staged(plsb_ctx *c, const double *dev, size_t n, double *host, bool as_complex);
// upload a host column-major matrix into a device row-major one (staged, chunked by columns)
int upload_cm_to_rm(plsb_ctx *c, const double *src, int64_t lds, int64_t rows, int64_t cols, double *dst, int64_t ldd,
                    bool src_on_device);

// RAII-ish profiling scope: records events around a launch group when profiling is on
struct ProfScope {
    plsb_ctx *c;
    cudaStream_t st;
    int idx = -1;
    ProfScope(plsb_ctx *ctx, int slot, cudaStream_t stream = nullptr);   // stream: where the launch group goes (default: the context's)
    ~ProfScope();
};

// ---- layout transforms (pass.cu) ---------------------------------------------------------------
// col-major src (rows x cols, ld = lds) -> row-major dst[(row0+i)*ldd + col0 + j]
int launch_cm_to_rm(plsb_ctx *c, const double *src, int64_t lds, int64_t rows, int64_t cols, double *dst,
                    int64_t ldd, int64_t col0);
// row-major src -> col-major dst
int launch_rm_to_cm(plsb_ctx *c, const double *src, int64_t lds, int64_t rows, int64_t cols, double *dst,
                    int64_t ldd);

// ---- streaming passes over X (pass.cu) --------------------------------------------------------
// out[0..ldk) = X'(X r), out[ldk] = t't, t_out[i] = (X r)_i (optional).  One read of X.
int launch_fused_pass(plsb_ctx *c, const plsb_dataset *ds, const double *r, double *out, double *t_out,
                      bool across_ranks = false);
// S[m*ldk + k] = sum_i X[i,k] Y[i,m]   (M x ldk, zero padded)
int launch_xty_pass(plsb_ctx *c, const plsb_dataset *ds, double *S);
// the DMMA + TMA variant (xty.cu); launch_xty_pass routes to it when supported
bool xty_dmma_supported(const plsb_dataset *ds);
int launch_xty_dmma(plsb_ctx *c, const plsb_dataset *ds, double *S);

// ---- per-component step (step.cu) -------------------------------------------------------------
struct StepArgs {
    int K, ldk, M, A;
    int a;             // component whose pass just finished (-1: initial weight only)
    double *S;         // M x ldk, deflated in place
    const double *g;   // ldk + 1: X'(X r_a) and tt
    double *W, *P, *R; // A x ldk each (column a at [a*ldk])
    double *Q;         // A x M
    long long *trace;  // dev tool (PLSB_STEP_TRACE): globaltimer ns at phase boundaries, or null
    double *wide;      // M > 32 (step_wide.cuh): wide_scratch_doubles(M) doubles of device scratch, else null
};
size_t step_wide_scratch_bytes(int M);   // 0 when M <= 32
int launch_step(plsb_ctx *c, const StepArgs &s);

// y = G r, tt = r'G r for symmetric G (K x K, ld = ldg): KERNEL_TYPE2 per-component product
int launch_symv(plsb_ctx *c, const double *G, int64_t ldg, int K, const double *r, double *out /*ldk+1*/, int ldk);

// ---- FP64 DMMA contraction (gemm.cu) --------------------------------------------------------------
// `lane`: a caller that keeps several streams busy (the two-wave CV engine) passes the stream to launch on and a
// split-K scratch of its own; nullptr = the context's stream and scratch.  No profiling events are recorded for lanes.
struct GemmLane { cudaStream_t stream = nullptr; double *part = nullptr; size_t part_bytes = 0; };
int launch_gemm_atb(plsb_ctx *c, const double *A, int64_t lda, const double *B, int64_t ldb, int64_t Kd, int M, int N,
                    double *C, int64_t ldc, bool syrk, bool b_nmajor, bool transC, int prof_slot, bool accumulate = false,
                    GemmLane *lane = nullptr, bool no_split = false);
// S (component-major A x lds) = X R for row-major X (N x ldk) and component-major R (A x ldk): DMMA (scores.cu)
int launch_scores(plsb_ctx *c, const double *X, int64_t N, int ldk, const double *R, int A, double *S, int64_t lds);
int launch_xtx(plsb_ctx *c, const plsb_dataset *ds, double *G, int64_t ldg);

// ---- batched-fold product of the two-lane CV engine (cv_gemm.cu): persistent DMMA kernel over PRE-TILED operands ----
// Gt: [m-tile][k-tile][32][132] copy of X'X; Rt: [n-tile][k-tile][128][36] the wave's current r vectors, written in
// this layout by the fold kernels (element (f, k) at cv_gemm_rt_index)
constexpr int CVG_BN = 128, CVG_BK = 32, CVG_LDB = 36;
#ifdef __CUDACC__
__host__ __device__ inline size_t cv_gemm_rt_index(int nkt, int f, int k) {
    return (((size_t)(f / CVG_BN) * nkt + (size_t)(k / CVG_BK)) * CVG_BN + (size_t)(f % CVG_BN)) * CVG_LDB + (size_t)(k % CVG_BK);
}
#endif
size_t cv_gemm_gt_doubles(int64_t K);
size_t cv_gemm_rt_doubles(int64_t K, int64_t F);
int launch_cv_gemm_tile_g(plsb_ctx *c, cudaStream_t st, const double *G, int64_t ldg, int64_t K, double *Gt);
int launch_cv_gemm(plsb_ctx *c, cudaStream_t st, const double *Gt, const double *Rt, int64_t K, int64_t ldk, int nf, double *Gr,
                   int stages, int max_ctas, unsigned long long *span = nullptr);

// ---- synthetic data (synth.cu) ----------------------------------------------------------------
int launch_synth(plsb_ctx *c, uint64_t seed, int64_t row0, int64_t N, int64_t K, int64_t M, int64_t L,
                 double *X, int64_t ldk, double *Y, int64_t ldm);

}  // namespace plsb
