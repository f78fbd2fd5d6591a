// pls_b200/csrc/cv_small.cu — cross-validation, small-data regime: the WHOLE data set lives in one SM's shared
// memory and one persistent CTA runs complete A-component fits, fold after fold, in a single launch.
//
// Replaces the per-fold refit loops of Model::cv_LOO (reference src/pls.cpp:478-489) and Model::cv_LSO (533-546)
// when N x K doubles fit in ~200 KB (the reference's own data: nir 60 x 401, toy 10 x 15).  The large-fold engine
// (cv.cu: X'X once, one DMMA GEMM + one fold kernel per component per wave) needs ~3 launches per component and a
// K x K Gram matrix; for 60 x 401 that is ~35 serial launches of latency for microseconds of arithmetic.  Here a
// call is ONE kernel: each CTA pulls X and Y into shared memory once (1-D TMA bulk copies, one mbarrier), then for
// each of its folds runs the reference's KERNEL_TYPE1 recurrence (pls.cpp:400-435) directly on the training rows
//      t = X r (held-out rows masked out of tt and p; their t IS the held-out score),  tt = t't,  p = X't / tt,
//      q = S'r / tt,  S -= tt p q',  next w (dominant eigenvector of S'S), r = w - R (P'w)
// with the single-fit step code (step_core.cuh) and a running residual e_c = e_{c-1} - u q_c' for every prefix.
// No Gram downdate is involved: the arithmetic is the reference's own, on the fold's rows in ascending order.
// P and R (2 x A x K) sit next to X in shared memory when they fit, else in an L2-resident scratch row per CTA.
#include "internal.h"
#include "step_core.cuh"

namespace plsb {

constexpr int CS_NT = 256;

struct CvSmallArgs {
    const double *X, *Y;       // row-major N x ldk / N x ldm
    const int64_t *test_idx;   // device, [local folds][ts]
    double *E;                 // residual block, columns of rows_local entries
    double *pr_scratch;        // [grid][2 * A * ldk] when P, R do not fit in shared memory
    int64_t rows_local;
    int N, K, ldk, ldm, M, A, ts, nf;
    int pr_in_smem;
};

struct CvSmallLayout {
    size_t x, y, s, w, g, t, tm, e, p, r, held, total;
};
static __host__ __device__ inline CvSmallLayout cv_small_layout(int N, int ldk, int ldm, int M, int A, int ts, bool pr) {
    CvSmallLayout l;
    size_t o = 16;                                   // the mbarrier
    l.x = o; o += (size_t)N * ldk * 8;
    l.y = o; o += (size_t)N * ldm * 8;
    l.s = o; o += (size_t)M * ldk * 8;
    l.w = o; o += (size_t)ldk * 8;
    l.g = o; o += (size_t)ldk * 8;
    l.t = o; o += (size_t)((N + 1) & ~1) * 8;
    l.tm = o; o += (size_t)((N + 1) & ~1) * 8;
    l.e = o; o += (size_t)ts * M * 8;
    l.p = o; if (pr) o += (size_t)A * ldk * 8;
    l.r = o; if (pr) o += (size_t)A * ldk * 8;
    l.held = o; o += (size_t)((N + 15) & ~15);
    l.total = o;
    return l;
}

__global__ void __launch_bounds__(CS_NT, 1) cv_small_kernel(const CvSmallArgs v) {
    extern __shared__ __align__(128) char smem[];
    __shared__ StepSmem<CS_NT> sh;
    __shared__ double qd[PLSB_MAX_M];
    const int tid = threadIdx.x;
    const int N = v.N, K = v.K, ldk = v.ldk, ldm = v.ldm, M = v.M, A = v.A, ts = v.ts;
    const CvSmallLayout l = cv_small_layout(N, ldk, ldm, M, A, ts, v.pr_in_smem != 0);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem);
    double *Xs = reinterpret_cast<double *>(smem + l.x), *Ys = reinterpret_cast<double *>(smem + l.y);
    double *S = reinterpret_cast<double *>(smem + l.s), *w = reinterpret_cast<double *>(smem + l.w);
    double *g = reinterpret_cast<double *>(smem + l.g), *t = reinterpret_cast<double *>(smem + l.t);
    double *tmk = reinterpret_cast<double *>(smem + l.tm), *es = reinterpret_cast<double *>(smem + l.e);
    double *Pf = v.pr_in_smem ? reinterpret_cast<double *>(smem + l.p) : v.pr_scratch + (size_t)blockIdx.x * 2 * A * ldk;
    double *Rf = v.pr_in_smem ? reinterpret_cast<double *>(smem + l.r) : Pf + (size_t)A * ldk;
    unsigned char *held = reinterpret_cast<unsigned char *>(smem + l.held);

    // ---- the data set, once per CTA: bulk copies of <= 32 KB each, all counted by one mbarrier ----
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (tid == 0) {
        const uint32_t bx = (uint32_t)((size_t)N * ldk * 8), by = (uint32_t)((size_t)N * ldm * 8);
        mbar_arrive_expect_tx(bar, bx + by);
        for (uint32_t o = 0; o < bx; o += 32768u)
            bulk_g2s(reinterpret_cast<char *>(Xs) + o, reinterpret_cast<const char *>(v.X) + o, bx - o < 32768u ? bx - o : 32768u, bar);
        bulk_g2s(Ys, v.Y, by, bar);
    }
    mbar_wait(bar, 0);

    for (int f = blockIdx.x; f < v.nf; f += gridDim.x) {
        const int64_t *h = v.test_idx + (int64_t)f * ts;
        __syncthreads();                       // the previous fold is done with every shared array
        for (int i = tid; i < N; i += CS_NT) held[i] = 0;
        __syncthreads();
        for (int i = tid; i < ts; i += CS_NT) held[h[i]] = 1;
        __syncthreads();
        // S_f = sum over the training rows of x_i y_i'   (pls.cpp:396 on the fold's rows), e_0 = y_H
        for (int k = tid; k < ldk; k += CS_NT)
            for (int m = 0; m < M; ++m) {
                double s = 0.0;
                if (k < K)
                    for (int i = 0; i < N; ++i)
                        if (!held[i]) s = fma(Xs[(size_t)i * ldk + k], Ys[(size_t)i * ldm + m], s);
                S[(size_t)m * ldk + k] = s;
            }
        for (int e = tid; e < ts * M; e += CS_NT) es[e] = Ys[(size_t)h[e / M] * ldm + (e % M)];
        __syncthreads();
        next_weights<CS_NT>(K, ldk, M, 0, S, Pf, Rf, w, Rf, sh);
        __syncthreads();
        for (int a = 0; a < A; ++a) {
            const double *r = Rf + (size_t)a * ldk;
            // t = X r for every row (pls.cpp:419); held-out rows: their scores u = x_H r
            warp_dots<CS_NT>(N, K, [&](int i) { return (const double *)(Xs + (size_t)i * ldk); }, [&](int) { return r; }, t);
            __syncthreads();
            double acc = 0.0;
            for (int i = tid; i < N; i += CS_NT) {
                const double ti = held[i] ? 0.0 : t[i];
                tmk[i] = ti;
                acc = fma(ti, ti, acc);
            }
            const double tt = block_sum<CS_NT>(acc, sh);     // tt = t't over the training rows (pls.cpp:420)
            __syncthreads();
            for (int k = tid; k < ldk; k += CS_NT) {         // g = X't over the training rows (pls.cpp:421)
                double s = 0.0;
                if (k < K)
                    for (int i = 0; i < N; ++i) s = fma(Xs[(size_t)i * ldk + k], tmk[i], s);
                g[k] = s;
            }
            __syncthreads();
            finish_component<CS_NT>(K, ldk, M, S, g, tt, r, Pf + (size_t)a * ldk, qd, sh);   // p, q, S -= tt p q'
            // residuals of the held-out rows with a+1 components: e_a = e_{a-1} - u q_a'
            for (int e = tid; e < ts * M; e += CS_NT) {
                const int i = e / M, m = e % M;
                const double val = es[e] - t[h[i]] * sh.q[m];
                es[e] = val;
                v.E[((size_t)m * A + a) * v.rows_local + (int64_t)f * ts + i] = val;
            }
            __syncthreads();
            if (a + 1 < A) {
                next_weights<CS_NT>(K, ldk, M, a + 1, S, Pf, Rf, w, Rf + (size_t)(a + 1) * ldk, sh);
                __syncthreads();
            }
        }
    }
}

// Does the small-data kernel apply?  (the whole data set + one fold's state within one SM's shared memory)
bool cv_small_supported(const plsb_ctx *c, const plsb_dataset *ds, int64_t A, int64_t ts, bool *pr_in_smem) {
    if (ds->N > 4096 || ds->ldk > 8192 || ds->M > PLSB_MAX_M || ts >= ds->N) return false;
    const size_t stat = sizeof(StepSmem<CS_NT>) + sizeof(double) * PLSB_MAX_M + 256;
    if (c->smem_optin <= stat) return false;
    const size_t budget = c->smem_optin - stat;
    if (cv_small_layout((int)ds->N, (int)ds->ldk, (int)ds->ldm, (int)ds->M, (int)A, (int)ts, true).total <= budget) {
        *pr_in_smem = true;
        return true;
    }
    *pr_in_smem = false;
    return cv_small_layout((int)ds->N, (int)ds->ldk, (int)ds->ldm, (int)ds->M, (int)A, (int)ts, false).total <= budget;
}

// folds [0, nf) of idx_dev (local numbering) -> Ed (columns of rows_local entries)
int launch_cv_small(plsb_ctx *c, const plsb_dataset *ds, int64_t A, const int64_t *idx_dev, int64_t nf, int64_t ts,
                    int64_t rows_local, double *Ed, bool pr_in_smem) {
    CvSmallArgs v{};
    v.X = ds->X; v.Y = ds->Y; v.test_idx = idx_dev; v.E = Ed; v.rows_local = rows_local;
    v.N = (int)ds->N; v.K = (int)ds->K; v.ldk = (int)ds->ldk; v.ldm = (int)ds->ldm; v.M = (int)ds->M; v.A = (int)A;
    v.ts = (int)ts; v.nf = (int)nf; v.pr_in_smem = pr_in_smem ? 1 : 0;
    const int grid = (int)(nf < c->sm_count ? nf : c->sm_count);
    void *scratch = nullptr;
    if (!pr_in_smem) {
        PLSB_TRY(dev_alloc(c, &scratch, (size_t)grid * 2 * A * ds->ldk * sizeof(double)));
        v.pr_scratch = (double *)scratch;
    }
    const size_t smem = cv_small_layout(v.N, v.ldk, v.ldm, v.M, v.A, v.ts, pr_in_smem).total;
    PLSB_CUDA(cudaFuncSetAttribute(cv_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    {
        ProfScope ps(c, PLSB_PROF_CV_FOLD);
        cv_small_kernel<<<grid, CS_NT, smem, c->stream>>>(v);
    }
    c->launches++;
    PLSB_CUDA(cudaGetLastError());
    dev_free(c, scratch);
    return PLSB_OK;
}

}  // namespace plsb
