// pls_b200/csrc/fit.cu — the fit driver: Model::plsr (reference src/pls.cpp:390-437) on the device.
//
// KERNEL_TYPE1:  S = X'Y (one pass) ; per component: ONE fused pass g = X'(X r), tt = |X r|^2, t = X r,
//                then the single-CTA step (p, q, deflation of S, next w and r).
// KERNEL_TYPE2:  additionally XX = X'X once (FP64 DMMA), per component g = XX r, tt = r'g.
// Row-sharded fits all-reduce S (K*M), XX (K*K) once and [g | tt] (K+1 doubles) per component.
#include "internal.h"

namespace plsb {

struct DevBuf {
    plsb_ctx *c;
    double *p = nullptr;
    explicit DevBuf(plsb_ctx *ctx) : c(ctx) {}
    int alloc(size_t n, bool zero) {
        PLSB_CUDA(cudaMallocAsync(&p, (n ? n : 1) * sizeof(double), c->stream));
        if (zero) PLSB_CUDA(cudaMemsetAsync(p, 0, (n ? n : 1) * sizeof(double), c->stream));
        return PLSB_OK;
    }
    ~DevBuf() { if (p) cudaFreeAsync(p, c->stream); }
};

int launch_xtx(plsb_ctx *c, const plsb_dataset *ds, double *G, int64_t ldg);  // gemm.cu

// device-side fit; outputs stay on the device (A x ldk / A x M / A x N, "component-major")
// T_host (optional, PINNED host memory, N x A column-major): column a of the scores is sent home on the copy
// stream as soon as pass a has produced it, under the following passes, instead of after the fit.
int fit_device(plsb_dataset *ds, int64_t A, int method, double *Wd, double *Pd, double *Rd, double *Qd, double *Td,
               double *T_host = nullptr) {
    plsb_ctx *c = ds->ctx;
    const int64_t K = ds->K, ldk = ds->ldk, M = ds->M;
    const bool shard = ds->sharded && c->nranks > 1;
    ProfScope total(c, PLSB_PROF_FIT_TOTAL);
    DevBuf S(c), g(c), G(c);
    PLSB_TRY(S.alloc((size_t)M * ldk, false));
    PLSB_TRY(g.alloc((size_t)ldk + 2, true));
    const int64_t ldg = round_up(K, 2);
    // A data set uploaded from host memory may carry XX = X'X already (accumulated under the PCIe transfer,
    // dataset.cu).  Then every component is a product with the L2-resident XX (the KERNEL_TYPE2 recurrence, which
    // is the same model: measured against the reference at 1M x 2048 both kernels sit at the same distance from
    // it, profiles/r1c_fullsize_parity.json) and X is read once more only to form the scores T = X R.
    const char *nog = getenv("PLSB_FIT_NO_GRAM");
    bool have_gram = ds->G && ds->gram_valid && !(nog && *nog == '1');
    if (shard) {
        // every rank must take the same branch (the branches differ in their collectives): ask once per data set
        if (ds->gram_agreed < 0) {
            double flag = have_gram ? 1.0 : 0.0;
            PLSB_CUDA(cudaMemcpyAsync(g.p, &flag, sizeof(double), cudaMemcpyHostToDevice, c->stream));
            PLSB_TRY(allreduce_sum(c, g.p, 1));
            PLSB_CUDA(cudaMemcpyAsync(&flag, g.p, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
            PLSB_CUDA(cudaStreamSynchronize(c->stream));
            PLSB_CUDA(cudaMemsetAsync(g.p, 0, sizeof(double), c->stream));
            ds->gram_agreed = flag > (double)c->nranks - 0.5 ? 1 : 0;
        }
        have_gram = have_gram && ds->gram_agreed == 1;
    }
    const bool from_gram = have_gram || method == PLSB_KERNEL_TYPE2;
    if (have_gram && ds->S0)   // X'Y was accumulated under the transfer as well
        PLSB_CUDA(cudaMemcpyAsync(S.p, ds->S0, (size_t)M * ldk * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
    else
        PLSB_TRY(launch_xty_pass(c, ds, S.p));
    if (shard) PLSB_TRY(allreduce_sum(c, S.p, (size_t)M * ldk));
    const double *Gp = nullptr;
    if (have_gram) {
        if (shard && !ds->gram_reduced) {
            PLSB_TRY(allreduce_sum(c, ds->G, (size_t)K * ldg));
            ds->gram_reduced = true;
        }
        Gp = ds->G;
    } else if (method == PLSB_KERNEL_TYPE2) {
        PLSB_TRY(G.alloc((size_t)K * ldg, true));
        PLSB_TRY(launch_xtx(c, ds, G.p, ldg));
        if (shard) PLSB_TRY(allreduce_sum(c, G.p, (size_t)K * ldg));
        Gp = G.p;
    }
    StepArgs st{};
    DevBuf wide(c);
    if (const size_t wb = step_wide_scratch_bytes((int)M)) {
        PLSB_TRY(wide.alloc(wb / sizeof(double), true));
        st.wide = wide.p;
    }
    st.K = (int)K; st.ldk = (int)ldk; st.M = (int)M; st.A = (int)A;
    st.S = S.p; st.g = g.p; st.W = Wd; st.P = Pd; st.R = Rd; st.Q = Qd;
    st.a = -1;
    PLSB_TRY(launch_step(c, st));
    for (int64_t a = 0; a < A; ++a) {
        const double *r = Rd + a * ldk;
        if (!from_gram) {
            // one read of this rank's rows; [g | tt] comes back summed over the CTAs AND over the ranks
            PLSB_TRY(launch_fused_pass(c, ds, r, g.p, Td ? Td + a * ds->N : nullptr, shard));
            if (Td && T_host) {
                PLSB_CUDA(cudaEventRecord(c->col_ev, c->stream));
                PLSB_CUDA(cudaStreamWaitEvent(c->copy_stream, c->col_ev, 0));
                PLSB_CUDA(cudaMemcpyAsync(T_host + a * ds->N, Td + a * ds->N, (size_t)ds->N * sizeof(double),
                                          cudaMemcpyDeviceToHost, c->copy_stream));
            }
        } else {
            PLSB_TRY(launch_symv(c, Gp, ldg, (int)K, r, g.p, (int)ldk));
        }
        st.a = (int)a;
        PLSB_TRY(launch_step(c, st));
    }
    if (from_gram && Td && ds->N > 0) {
        // scores of the training rows, T = X R (pls.cpp:419 for every component at once): DMMA, one read of X per
        // 32 components, in row chunks so that a pinned destination receives chunk i while chunk i+1 is computed
        const int64_t chunk = T_host ? round_up(ceil_div(ds->N, 4), 128) : ds->N;
        for (int64_t r0 = 0; r0 < ds->N; r0 += chunk) {
            const int64_t nr = ds->N - r0 < chunk ? ds->N - r0 : chunk;
            PLSB_TRY(launch_scores(c, ds->X + r0 * ldk, nr, (int)ldk, Rd, (int)A, Td + r0, ds->N));
            if (T_host) {
                PLSB_CUDA(cudaEventRecord(c->col_ev, c->stream));
                PLSB_CUDA(cudaStreamWaitEvent(c->copy_stream, c->col_ev, 0));
                PLSB_CUDA(cudaMemcpy2DAsync(T_host + r0, (size_t)ds->N * 8, Td + r0, (size_t)ds->N * 8, (size_t)nr * 8,
                                            (size_t)A, cudaMemcpyDeviceToHost, c->copy_stream));
            }
        }
    }
    return PLSB_OK;
}

}  // namespace plsb

using namespace plsb;

// the N x A scores to host memory: large blocks bound for pageable memory go through the pinned staging pipeline
static int download_scores(plsb_ctx *c, const double *Td, size_t n, double *T, bool as_complex) {
    cudaPointerAttributes pa{};
    const bool pinned = cudaPointerGetAttributes(&pa, T) == cudaSuccess && pa.type == cudaMemoryTypeHost;
    cudaGetLastError();
    const char *off = getenv("PLSB_NO_STAGE");
    if (as_complex && n * 8 < ((size_t)1 << 20)) {
        // small models (the CLI's files, the one-launch CV cases): no pinned staging buffers, no host thread pool — the
        // reals land in the upper half of the caller's (re, im) buffer and are spread out in place, front to back
        // (pair j is written at 2j, 2j+1 <= n+j, i.e. never over a value that is still to be read)
        if (n == 0) return PLSB_OK;
        PLSB_CUDA(cudaMemcpyAsync(T + n, Td, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        PLSB_CUDA(cudaStreamSynchronize(c->stream));
        for (size_t j = 0; j < n; ++j) {
            const double v = T[n + j];
            T[2 * j] = v;
            T[2 * j + 1] = 0.0;
        }
        return PLSB_OK;
    }
    if (as_complex || (!pinned && n * 8 >= ((size_t)8 << 20) && !(off && *off == '1'))) return download_staged(c, Td, n, T, as_complex);
    PLSB_CUDA(cudaMemcpyAsync(T, Td, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    return PLSB_OK;
}

extern "C" {

int plsb_fit(plsb_dataset_t ds, int64_t A, int method, double *W, double *P, double *R, double *Q, double *T) {
    PLSB_ARG(ds, "dataset is null");
    PLSB_ARG(A >= 1, "A must be >= 1");
    PLSB_ARG(method == PLSB_KERNEL_TYPE1 || method == PLSB_KERNEL_TYPE2, "unknown method %d", method);
    plsb_ctx *c = ds->ctx;
    PLSB_CUDA(cudaSetDevice(c->device));
    const int64_t K = ds->K, ldk = ds->ldk, M = ds->M, N = ds->N;
    DevBuf Wd(c), Pd(c), Rd(c), Qd(c), Td(c);
    PLSB_TRY(Wd.alloc((size_t)A * ldk, true));
    PLSB_TRY(Pd.alloc((size_t)A * ldk, true));
    PLSB_TRY(Rd.alloc((size_t)A * ldk, true));
    PLSB_TRY(Qd.alloc((size_t)A * M, true));
    const bool wantT = T != nullptr && method == PLSB_KERNEL_TYPE1 && N > 0;
    if (wantT) PLSB_TRY(Td.alloc((size_t)A * N, false));
    // a pageable destination would make every cudaMemcpyAsync block the launching thread: stream T only to pinned memory
    bool streamT = false;
    if (wantT) {
        cudaPointerAttributes pa{};
        streamT = cudaPointerGetAttributes(&pa, T) == cudaSuccess && pa.type == cudaMemoryTypeHost;
        cudaGetLastError();
    }
    PLSB_TRY(fit_device(ds, A, method, Wd.p, Pd.p, Rd.p, Qd.p, wantT ? Td.p : nullptr, streamT ? T : nullptr));
    const size_t kb = (size_t)K * sizeof(double), lb = (size_t)ldk * sizeof(double);
    if (W) PLSB_CUDA(cudaMemcpy2DAsync(W, kb, Wd.p, lb, kb, (size_t)A, cudaMemcpyDeviceToHost, c->stream));
    if (P) PLSB_CUDA(cudaMemcpy2DAsync(P, kb, Pd.p, lb, kb, (size_t)A, cudaMemcpyDeviceToHost, c->stream));
    if (R) PLSB_CUDA(cudaMemcpy2DAsync(R, kb, Rd.p, lb, kb, (size_t)A, cudaMemcpyDeviceToHost, c->stream));
    if (Q) PLSB_CUDA(cudaMemcpyAsync(Q, Qd.p, (size_t)A * M * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (wantT && !streamT) PLSB_TRY(download_scores(c, Td.p, (size_t)A * N, T, false));
    if (streamT) PLSB_CUDA(cudaStreamSynchronize(c->copy_stream));
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    if (ds->sharded) PLSB_TRY(xchg_check(c));
    return PLSB_OK;
}

// ---- resident models: the fit's results (W, P, R, Q and the scores T) are formed in HBM and stay there --------------
int plsb_fit_resident(plsb_dataset_t ds, int64_t A, int method, plsb_model_t *out) {
    PLSB_ARG(ds && out, "null argument");
    PLSB_ARG(A >= 1, "A must be >= 1");
    PLSB_ARG(method == PLSB_KERNEL_TYPE1 || method == PLSB_KERNEL_TYPE2, "unknown method %d", method);
    plsb_ctx *c = ds->ctx;
    PLSB_CUDA(cudaSetDevice(c->device));
    plsb_model *m = new plsb_model();
    ctx_retain(c);
    m->ctx = c; m->N = ds->N; m->K = ds->K; m->ldk = ds->ldk; m->M = ds->M; m->A = A; m->method = method;
    const bool wantT = method == PLSB_KERNEL_TYPE1 && ds->N > 0;   // the reference stores T for KERNEL_TYPE1 only (pls.cpp:394, 434)
    int rc = dev_alloc(c, (void **)&m->W, (size_t)A * m->ldk * 8);
    if (rc == PLSB_OK) rc = dev_alloc(c, (void **)&m->P, (size_t)A * m->ldk * 8);
    if (rc == PLSB_OK) rc = dev_alloc(c, (void **)&m->R, (size_t)A * m->ldk * 8);
    if (rc == PLSB_OK) rc = dev_alloc(c, (void **)&m->Q, (size_t)A * m->M * 8);
    if (rc == PLSB_OK && wantT) rc = dev_alloc(c, (void **)&m->T, (size_t)A * m->N * 8);
    if (rc == PLSB_OK) {
        cudaError_t e = cudaMemsetAsync(m->W, 0, (size_t)A * m->ldk * 8, c->stream);
        if (e == cudaSuccess) e = cudaMemsetAsync(m->P, 0, (size_t)A * m->ldk * 8, c->stream);
        if (e == cudaSuccess) e = cudaMemsetAsync(m->R, 0, (size_t)A * m->ldk * 8, c->stream);
        if (e == cudaSuccess) e = cudaMemsetAsync(m->Q, 0, (size_t)A * m->M * 8, c->stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "cudaMemsetAsync", __FILE__, __LINE__);
    }
    if (rc == PLSB_OK) rc = fit_device(ds, A, method, m->W, m->P, m->R, m->Q, m->T);
    if (rc == PLSB_OK) {
        cudaError_t e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "fit", __FILE__, __LINE__);
    }
    if (rc == PLSB_OK && ds->sharded) rc = xchg_check(c);
    if (rc != PLSB_OK) { plsb_model_destroy(m); return rc; }
    *out = m;
    return PLSB_OK;
}

int plsb_model_download(plsb_model_t m, double *W, double *P, double *R, double *Q, double *T) {
    PLSB_ARG(m, "model is null");
    plsb_ctx *c = m->ctx;
    PLSB_CUDA(cudaSetDevice(c->device));
    const size_t kb = (size_t)m->K * 8, lb = (size_t)m->ldk * 8, A = (size_t)m->A;
    if (W) PLSB_CUDA(cudaMemcpy2DAsync(W, kb, m->W, lb, kb, A, cudaMemcpyDeviceToHost, c->stream));
    if (P) PLSB_CUDA(cudaMemcpy2DAsync(P, kb, m->P, lb, kb, A, cudaMemcpyDeviceToHost, c->stream));
    if (R) PLSB_CUDA(cudaMemcpy2DAsync(R, kb, m->R, lb, kb, A, cudaMemcpyDeviceToHost, c->stream));
    if (Q) PLSB_CUDA(cudaMemcpyAsync(Q, m->Q, A * m->M * 8, cudaMemcpyDeviceToHost, c->stream));
    if (T) {
        PLSB_ARG(m->T, "this model holds no scores (KERNEL_TYPE2 stores none, pls.cpp:394)");
        PLSB_TRY(download_scores(c, m->T, A * (size_t)m->N, T, false));
    }
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    return PLSB_OK;
}

int plsb_model_download_scores(plsb_model_t m, double *T, int as_complex) {
    PLSB_ARG(m && T, "null argument");
    PLSB_ARG(m->T, "this model holds no scores (KERNEL_TYPE2 stores none, pls.cpp:394)");
    plsb_ctx *c = m->ctx;
    PLSB_CUDA(cudaSetDevice(c->device));
    PLSB_TRY(download_scores(c, m->T, (size_t)m->A * (size_t)m->N, T, as_complex != 0));
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    return PLSB_OK;
}

int plsb_model_destroy(plsb_model_t m) {
    if (!m) return PLSB_OK;
    cudaSetDevice(m->ctx->device);
    dev_free(m->ctx, m->W); dev_free(m->ctx, m->P); dev_free(m->ctx, m->R); dev_free(m->ctx, m->Q); dev_free(m->ctx, m->T);
    plsb_ctx *c = m->ctx;
    delete m;
    ctx_release(c);
    return PLSB_OK;
}

int plsb_fit_host(plsb_ctx_t c, const double *X, int64_t ldx, int64_t N, int64_t K, const double *Y, int64_t ldy,
                  int64_t M, int64_t A, int method, double *W, double *P, double *R, double *Q, double *T) {
    plsb_dataset_t ds = nullptr;
    PLSB_TRY(plsb_dataset_create(c, X, ldx, N, K, Y, ldy, M, &ds));
    int rc = plsb_fit(ds, A, method, W, P, R, Q, T);
    plsb_dataset_destroy(ds);
    return rc;
}

int plsb_xty(plsb_dataset_t ds, double *XY) {
    PLSB_ARG(ds && XY, "null argument");
    plsb_ctx *c = ds->ctx;
    PLSB_CUDA(cudaSetDevice(c->device));
    DevBuf S(c);
    PLSB_TRY(S.alloc((size_t)ds->M * ds->ldk, false));
    PLSB_TRY(launch_xty_pass(c, ds, S.p));
    if (ds->sharded && c->nranks > 1) PLSB_TRY(allreduce_sum(c, S.p, (size_t)ds->M * ds->ldk));
    PLSB_CUDA(cudaMemcpy2DAsync(XY, (size_t)ds->K * 8, S.p, (size_t)ds->ldk * 8, (size_t)ds->K * 8, (size_t)ds->M,
                                cudaMemcpyDeviceToHost, c->stream));
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    return PLSB_OK;
}

int plsb_xtx(plsb_dataset_t ds, double *XX) {
    PLSB_ARG(ds, "null argument");
    plsb_ctx *c = ds->ctx;
    PLSB_CUDA(cudaSetDevice(c->device));
    const int64_t K = ds->K, ldg = round_up(K, 2);
    DevBuf G(c);
    PLSB_TRY(G.alloc((size_t)K * ldg, false));
    PLSB_TRY(launch_xtx(c, ds, G.p, ldg));
    if (ds->sharded && c->nranks > 1) PLSB_TRY(allreduce_sum(c, G.p, (size_t)K * ldg));
    if (XX) PLSB_CUDA(cudaMemcpy2DAsync(XX, (size_t)K * 8, G.p, (size_t)ldg * 8, (size_t)K * 8, (size_t)K,
                                        cudaMemcpyDeviceToHost, c->stream));
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    return PLSB_OK;
}

// Build XX = X'X and XY = X'Y once and keep them ON the data set (what an upload from host memory accumulates under the
// transfer, dataset.cu): cross-validation calls and fits that follow start from them.  across_ranks (a REPLICATED data
// set on a context with a communicator; collective — every rank calls it): rank g forms the cross products of rows
// [g N/G, (g+1) N/G) only and the partial matrices are summed over the ranks (one all-reduce of K*K + K*M doubles)
// instead of every rank reading all of X.
int plsb_dataset_build_gram(plsb_dataset_t ds, int across_ranks) {
    PLSB_ARG(ds, "dataset is null");
    PLSB_ARG(!ds->sharded, "a row shard keeps partial cross products through its upload; this call is for whole data sets");
    plsb_ctx *c = ds->ctx;
    PLSB_CUDA(cudaSetDevice(c->device));
    if (ds->G && ds->S0 && ds->gram_valid) return PLSB_OK;
    const int64_t K = ds->K, ldk = ds->ldk, M = ds->M;
    const int64_t ldg = round_up(K, 2);
    if (!ds->G) PLSB_TRY(dev_alloc(c, (void **)&ds->G, (size_t)K * ldg * sizeof(double)));
    if (!ds->S0) PLSB_TRY(dev_alloc(c, (void **)&ds->S0, (size_t)M * ldk * sizeof(double)));
    ds->ldg = ldg;
    const bool split = across_ranks && c->nranks > 1;
    plsb_dataset view = *ds;           // this rank's rows (all of them without a communicator)
    if (split) {
        const int64_t r0 = ds->N * c->rank / c->nranks, r1 = ds->N * (c->rank + 1) / c->nranks;
        view.X = ds->X + r0 * ldk; view.Y = ds->Y + r0 * ds->ldm; view.N = r1 - r0;
    }
    PLSB_CUDA(cudaMemsetAsync(ds->G, 0, (size_t)K * ldg * sizeof(double), c->stream));
    PLSB_TRY(launch_xtx(c, &view, ds->G, ldg));
    PLSB_TRY(launch_xty_pass(c, &view, ds->S0));
    if (split) {
        PLSB_TRY(allreduce_sum(c, ds->G, (size_t)K * ldg));
        PLSB_TRY(allreduce_sum(c, ds->S0, (size_t)M * ldk));
    }
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    ds->gram_valid = true;
    ds->gram_reduced = true;
    return PLSB_OK;
}

}  // extern "C"
