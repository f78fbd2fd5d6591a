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

// ------------------------------------------------------------------------------------------------------------------
// One WARP per fold (few responses, K <= 512, N <= 128): the CTA-per-fold kernel above spends its time in CTA barriers —
// eight warps waiting for the one that solves the eigen problem, for M-row dot products that occupy M warps, ... (ncu,
// profiles/r2e: 7 warps stalled on a barrier per issue slot).  Here a fold belongs to ONE warp from start to finish:
// every K-vector of the recurrence (S_f rows, r, w, g = p) lives in the lanes' registers (lane l owns columns l, l + 32,
// ...), reductions are warp shuffles, no CTA barrier is executed after the data set has landed, and the eight warps of
// a CTA run eight folds concurrently out of the same shared-memory copy of X.  Only P_f and R_f (needed again by every
// later component's orthogonalisation) go to an L2-resident scratch row per warp.  Same recurrence, same order of
// the sums over rows (ascending i) as the kernel above.
// ------------------------------------------------------------------------------------------------------------------
constexpr int CW_NW = 8;          // warps (= folds in flight) per CTA
constexpr int CW_NMAX = 128;      // rows

struct CvWarpLayout { size_t x, y, warp0, per_warp, total; };
static __host__ __device__ inline CvWarpLayout cv_warp_layout(int N, int ldk, int ldm, int M, int ts) {
    CvWarpLayout l;
    size_t o = 16;
    l.x = o; o += (size_t)N * ldk * 8;
    l.y = o; o += (size_t)N * ldm * 8;
    // per warp: t[CW_NMAX], es[ts*M], G[16], q[4] (+4 pad), B[128], held[CW_NMAX] bytes
    l.per_warp = (size_t)(CW_NMAX + ((ts * M + 1) & ~1) + 16 + 8 + 128) * 8 + CW_NMAX;
    l.warp0 = o; o += l.per_warp * CW_NW;
    l.total = o;
    return l;
}

template <int MM, int KPL>
__global__ void __launch_bounds__(CW_NW * 32, 1) cv_warp_kernel(const CvSmallArgs v) {
    extern __shared__ __align__(128) char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int N = v.N, K = v.K, ldk = v.ldk, ldm = v.ldm, M = v.M, A = v.A, ts = v.ts;
    const CvWarpLayout l = cv_warp_layout(N, ldk, ldm, M, ts);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem);
    const double *Xs = reinterpret_cast<const double *>(smem + l.x), *Ys = reinterpret_cast<const double *>(smem + l.y);
    char *mine = smem + l.warp0 + (size_t)warp * l.per_warp;
    double *tw = reinterpret_cast<double *>(mine);
    double *es = tw + CW_NMAX;
    double *G = es + ((ts * M + 1) & ~1);
    double *qw = G + 16;
    double *B = qw + 8;
    unsigned char *held = reinterpret_cast<unsigned char *>(B + 128);
    constexpr int LDP = 32 * KPL;
    double *Pf = v.pr_scratch + ((size_t)blockIdx.x * CW_NW + warp) * 2 * A * LDP;
    double *Rf = Pf + (size_t)A * LDP;

    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (tid == 0) {
        const uint32_t bx = (uint32_t)((size_t)N * ldk * 8), by = (uint32_t)((size_t)N * ldm * 8);
        mbar_arrive_expect_tx(bar, bx + by);
        for (uint32_t o = 0; o < bx; o += 32768u)
            bulk_g2s(smem + l.x + o, reinterpret_cast<const char *>(v.X) + o, bx - o < 32768u ? bx - o : 32768u, bar);
        bulk_g2s(smem + l.y, v.Y, by, bar);
    }
    mbar_wait(bar, 0);

    int kc[KPL];            // this lane's columns (clamped to a valid one when dead)
    double lv[KPL];         // 1.0 live / 0.0 dead
#pragma unroll
    for (int j = 0; j < KPL; ++j) {
        const int k = lane + 32 * j;
        kc[j] = k < K ? k : 0;
        lv[j] = k < K ? 1.0 : 0.0;
    }

    for (int f = blockIdx.x + gridDim.x * warp; f < v.nf; f += gridDim.x * CW_NW) {
        const int64_t *h = v.test_idx + (int64_t)f * ts;
        __syncwarp();
        for (int i = lane; i < N; i += 32) held[i] = 0;
        __syncwarp();
        for (int i = lane; i < ts; i += 32) held[h[i]] = 1;
        __syncwarp();
        // S_f = sum over the training rows of x_i y_i'   (pls.cpp:396 on the fold's rows), e_0 = y_H
        double S[MM][KPL];
#pragma unroll
        for (int m = 0; m < MM; ++m)
#pragma unroll
            for (int j = 0; j < KPL; ++j) S[m][j] = 0.0;
        for (int i = 0; i < N; ++i) {
            if (held[i]) continue;                       // warp-uniform
            const double *xr = Xs + (size_t)i * ldk;
            double y[MM];
#pragma unroll
            for (int m = 0; m < MM; ++m) y[m] = m < M ? Ys[(size_t)i * ldm + m] : 0.0;
#pragma unroll
            for (int j = 0; j < KPL; ++j) {
                const double x = xr[kc[j]] * lv[j];
#pragma unroll
                for (int m = 0; m < MM; ++m) S[m][j] = fma(x, y[m], S[m][j]);
            }
        }
        for (int e = lane; e < ts * M; e += 32) es[e] = Ys[(size_t)h[e / M] * ldm + (e % M)];
        __syncwarp();

        double r[KPL];
        for (int an = 0; an <= A; ++an) {
            const int a = an - 1;
            if (a >= 0) {
                // ---- component a with the rotation vector r: t = X r for every row (pls.cpp:419) ----
                for (int i0 = 0; i0 < N; i0 += 8) {
                    double pv[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int i = i0 + u < N ? i0 + u : N - 1;
                        const double *xr = Xs + (size_t)i * ldk;
                        double acc = 0.0;
#pragma unroll
                        for (int j = 0; j < KPL; ++j) acc = fma(xr[kc[j]], r[j], acc);    // r is 0 on dead columns
                        pv[u] = acc;
                    }
                    warp_transpose_reduce<8>(pv);
                    if ((lane & 3) == 0 && i0 + (lane >> 2) < N) tw[i0 + (lane >> 2)] = pv[0];
                }
                __syncwarp();
                double acc = 0.0;
                for (int i = lane; i < N; i += 32) {
                    const double ti = held[i] ? 0.0 : tw[i];
                    acc = fma(ti, ti, acc);
                }
                const double tt = warp_sum(acc);                  // tt = t't over the training rows (pls.cpp:420)
                double g[KPL];
#pragma unroll
                for (int j = 0; j < KPL; ++j) g[j] = 0.0;
                for (int i = 0; i < N; ++i) {                     // g = X't over the training rows (pls.cpp:421)
                    if (held[i]) continue;
                    const double *xr = Xs + (size_t)i * ldk;
                    const double ti = tw[i];
#pragma unroll
                    for (int j = 0; j < KPL; ++j) g[j] = fma(xr[kc[j]] * lv[j], ti, g[j]);
                }
                // p = g / tt, q = S'r / tt, S -= tt p q'   (pls.cpp:427-429)
                double q[MM];
#pragma unroll
                for (int m = 0; m < MM; ++m) {
                    double s = 0.0;
#pragma unroll
                    for (int j = 0; j < KPL; ++j) s = fma(S[m][j], r[j], s);
                    q[m] = warp_sum(s) / tt;
                }
#pragma unroll
                for (int j = 0; j < KPL; ++j) {
                    const double p = g[j] / tt;
                    Pf[(size_t)a * LDP + lane + 32 * j] = p;
                    const double tp = tt * p;
#pragma unroll
                    for (int m = 0; m < MM; ++m) S[m][j] -= tp * q[m];
                }
                if (lane == 0) {
#pragma unroll
                    for (int m = 0; m < MM; ++m) qw[m] = q[m];
                }
                __syncwarp();
                // residuals of the held-out rows with a+1 components: e_a = e_{a-1} - u q_a'
                for (int e = lane; e < ts * M; e += 32) {
                    const int i = e / M, m = e % M;
                    const double val = es[e] - tw[h[i]] * qw[m];
                    es[e] = val;
                    v.E[((size_t)m * A + a) * v.rows_local + (int64_t)f * ts + i] = val;
                }
                __syncwarp();
            }
            if (an >= A) break;
            // ---- weight vector of component an (pls.cpp:403-411) ----
            double w[KPL];
            if (MM == 1) {
#pragma unroll
                for (int j = 0; j < KPL; ++j) w[j] = S[0][j];
            } else {
                int pi = 0;
#pragma unroll
                for (int u = 0; u < MM; ++u)
#pragma unroll
                    for (int w2 = u; w2 < MM; ++w2) {
                        double s = 0.0;
#pragma unroll
                        for (int j = 0; j < KPL; ++j) s = fma(S[u][j], S[w2][j], s);
                        s = warp_sum(s);
                        if (lane == 0 && u < M && w2 < M) { G[u * M + w2] = s; G[w2 * M + u] = s; }
                        ++pi;
                    }
                __syncwarp();
                dominant_eigvec_warp<8>(G, B, B + 64, M, qw);
                __syncwarp();
#pragma unroll
                for (int j = 0; j < KPL; ++j) {
                    double acc = 0.0;
#pragma unroll
                    for (int m = 0; m < MM; ++m)
                        if (m < M) acc = fma(S[m][j], qw[m], acc);
                    w[j] = acc;
                }
                __syncwarp();
            }
            {   // w /= sqrt(w'w)   (pls.cpp:411)
                double acc = 0.0;
#pragma unroll
                for (int j = 0; j < KPL; ++j) acc = fma(w[j], w[j], acc);
                const double nrm = sqrt(warp_sum(acc));
#pragma unroll
                for (int j = 0; j < KPL; ++j) w[j] = w[j] / nrm;
            }
            // ---- r = w - sum_j (p_j'w) r_j   (pls.cpp:412-416; every dot uses the un-modified w) ----
#pragma unroll
            for (int j = 0; j < KPL; ++j) r[j] = w[j];
            for (int j0 = 0; j0 < an; j0 += 4) {
                double c[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int jj = j0 + u < an ? j0 + u : an - 1;
                    const double *pj = Pf + (size_t)jj * LDP + lane;
                    double acc = 0.0;
#pragma unroll
                    for (int j = 0; j < KPL; ++j) acc = fma(pj[32 * j], w[j], acc);
                    c[u] = acc;
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) c[u] = warp_sum(c[u]);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (j0 + u >= an) break;
                    const double *rj = Rf + (size_t)(j0 + u) * LDP + lane;
#pragma unroll
                    for (int j = 0; j < KPL; ++j) r[j] -= c[u] * rj[32 * j];
                }
            }
#pragma unroll
            for (int j = 0; j < KPL; ++j) Rf[(size_t)an * LDP + lane + 32 * j] = r[j];
            __syncwarp();
        }
    }
}

template <int MM, int KPL>
static int launch_cv_warp_t(plsb_ctx *c, const CvSmallArgs &v0, size_t smem, int grid) {
    CvSmallArgs v = v0;
    void *scratch = nullptr;
    PLSB_TRY(dev_alloc(c, &scratch, (size_t)grid * CW_NW * 2 * v.A * 32 * KPL * sizeof(double)));
    v.pr_scratch = (double *)scratch;
    PLSB_CUDA(cudaFuncSetAttribute(cv_warp_kernel<MM, KPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    {
        ProfScope ps(c, PLSB_PROF_CV_FOLD);
        cv_warp_kernel<MM, KPL><<<grid, CW_NW * 32, smem, c->stream>>>(v);
    }
    c->launches++;
    PLSB_CUDA(cudaGetLastError());
    dev_free(c, scratch);
    return PLSB_OK;
}
template <int MM>
static int launch_cv_warp_m(plsb_ctx *c, const CvSmallArgs &v, size_t smem, int grid) {
    if (v.K <= 256) return launch_cv_warp_t<MM, 8>(c, v, smem, grid);
    if (v.K <= 416) return launch_cv_warp_t<MM, 13>(c, v, smem, grid);
    return launch_cv_warp_t<MM, 16>(c, v, smem, grid);
}

bool cv_warp_supported(const plsb_ctx *c, const plsb_dataset *ds, int64_t A, int64_t ts) {
    const char *off = getenv("PLSB_CV_NO_WARP");
    if (off && *off == '1') return false;
    if (ds->N > CW_NMAX || ds->K > 512 || ds->M > 4 || ts >= ds->N) return false;
    if (ds->M == 4 && ds->K > 416) return false;          // S_f alone would be 64 doubles per lane
    return cv_warp_layout((int)ds->N, (int)ds->ldk, (int)ds->ldm, (int)ds->M, (int)ts).total + 1024 <= c->smem_optin;
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
                    int64_t rows_local, double *Ed, bool pr_in_smem, int64_t nf_total) {
    CvSmallArgs v{};
    v.X = ds->X; v.Y = ds->Y; v.test_idx = idx_dev; v.E = Ed; v.rows_local = rows_local;
    v.N = (int)ds->N; v.K = (int)ds->K; v.ldk = (int)ds->ldk; v.ldm = (int)ds->ldm; v.M = (int)ds->M; v.A = (int)A;
    v.ts = (int)ts; v.nf = (int)nf; v.pr_in_smem = pr_in_smem ? 1 : 0;
    const int grid = (int)(nf < c->sm_count ? nf : c->sm_count);
    // more folds than SMs: a warp per fold, eight folds in flight per SM; otherwise a whole CTA per fold finishes sooner
    // (C2 LOO, 60 folds: 0.10 ms against 0.20 ms)
    // The choice goes by the TOTAL fold count of the cross-validation, not by this call's share of it: the two kernels
    // sum in different orders, and fold shards on several GPUs must reproduce the single-GPU residuals bit for bit.
    if (nf_total > c->sm_count && cv_warp_supported(c, ds, A, ts)) {
        const size_t smem = cv_warp_layout(v.N, v.ldk, v.ldm, v.M, v.ts).total;
        switch (v.M) {
            case 1: return launch_cv_warp_m<1>(c, v, smem, grid);
            case 2: return launch_cv_warp_m<2>(c, v, smem, grid);
            case 3: return launch_cv_warp_m<3>(c, v, smem, grid);
            default: return launch_cv_warp_m<4>(c, v, smem, grid);
        }
    }
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
