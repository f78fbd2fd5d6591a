// pls_b200/csrc/cv.cu — batched cross-validation: many independent folds per launch.
//
// Replaces Model::cv_LOO (reference src/pls.cpp:469-491) and Model::cv_LSO (512-549): N resp.
// num_trials refits, each followed by A residual evaluations of the held-out rows.
//
// The reference never re-centres folds (z-scoring happens once, in the CLI), so a fold's cross products
// are low-rank downdates of the full ones:
//      S_f = X'Y - X_H' Y_H ,      (X'X)_f r = (X'X) r - X_H' (X_H r)       H = held-out rows of fold f.
// Per component the whole wave of F folds shares ONE FP64 DMMA product  (X'X) [r_1 .. r_F]  (cv_gemm.cu; gemm.cu for
// the one-lane path), and a one-CTA-per-fold kernel does the K-sized algebra (downdate, tt, p, q, deflation of S_f,
// next w and r): the fused kernel below for 1024 <= K <= 4096, M <= 8, else the phase-by-phase code of the single-fit
// step (step_core.cuh; step_wide.cuh for M > 32).  Two waves are kept in flight so that the tensor-bound product of
// one runs under the HBM-bound fold kernel of the other.  Small data sets take cv_small.cu instead.  The scores of the held-out rows, u = X_H r_a,
// fall out of the downdate, so the residuals for every prefix 1..A are a running update
// e_a = e_{a-1} - u q_a' : no separate prediction pass.
// Components are extracted up to A only (the reference extracts K per fold, pls.cpp:477/531; its first A
// are bit-identical).
#include <chrono>
#include <type_traits>
#include <vector>

#include "internal.h"
#include "step_core.cuh"
#include "step_wide.cuh"

namespace plsb {

constexpr int CV_NT = 256;

struct CvArgs {
    const double *X, *Y, *S0;
    const int64_t *test_idx;   // device, [local folds][ts]
    double *S, *P, *R, *w, *Gr, *U, *E;
    int64_t rows_local;        // columns of E hold rows_local entries
    int K, ldk, ldm, M, A, ts;
    int f0;                    // first local fold of this wave
    int a;                     // component whose GEMM just finished (-1: initialise)
    double *Rt;                // two-lane engine: the next r also goes here, in the tiled layout of cv_gemm.cu (or null)
    int nkt;
    double *wide;              // M > 32: per-fold scratch of wide_scratch_doubles(M) doubles (step_wide.cuh), else null
    int wide_jac_smem;         // M > 32: the Jacobi work area is the kernel's dynamic shared memory
    int prefetch;              // fused kernel: run its HBM streams ahead into L2 with bulk prefetches
    int ring;                  // fused kernel: P_f / R_f rows arrive through a TMA-fed shared-memory ring (dynamic smem)
    int ring_pf;               // ... and the item this many places ahead of each issued one is prefetched into L2
    long long *trace;          // dev tool (PLSB_CV_TRACE=<component>): globaltimer ns at the fused kernel's phase boundaries
    long long *cta_trace;      // dev tool (PLSB_CV_CTA_TRACE=<component>): per CTA (smid, start ns, end ns) of one launch
    unsigned long long *span;  // dev tool (PLSB_CV_TIMELINE=1): [2] = earliest CTA start / latest CTA end of THIS launch (ns)
};
__device__ __forceinline__ void cv_span_mark(unsigned long long *span, int which) {
    if (span && threadIdx.x == 0) {
        unsigned long long t_;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
        if (which == 0) atomicMin(span, t_); else atomicMax(span + 1, t_);
    }
}
#define CV_TRACE(i) do { if (v.trace && blockIdx.x == 7 && threadIdx.x == 0) { long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); v.trace[i] = t_; } } while (0)

// WIDE: M > 32 responses — the M-sized objects live in a per-fold global scratch (step_wide.cuh)
template <bool WIDE>
__global__ void __launch_bounds__(CV_NT) cv_fold_kernel(const CvArgs v) {
    __shared__ typename std::conditional<WIDE, WideSmem<CV_NT>, StepSmem<CV_NT>>::type sh;
    __shared__ double qdump[WIDE ? 1 : PLSB_MAX_M];
    const int tid = threadIdx.x;
    const int f = blockIdx.x;
    const int64_t gf = (int64_t)v.f0 + f;
    const int64_t *h = v.test_idx + gf * v.ts;
    const int K = v.K, ldk = v.ldk, M = v.M;
    double *Sf = v.S + (size_t)f * M * ldk;
    double *Pf = v.P + (size_t)f * v.A * ldk;
    double *Rf = v.R + (size_t)f * v.A * ldk;
    double *wide = WIDE ? v.wide + (size_t)f * wide_scratch_doubles(M) : nullptr;

    if (v.a < 0) {
        // S_f = S0 - X_H' Y_H
        for (int k = tid; k < ldk; k += CV_NT)
            for (int m = 0; m < M; ++m) {
                double s = v.S0[(size_t)m * ldk + k];
                if (k < K)
                    for (int i = 0; i < v.ts; ++i) s -= v.X[h[i] * (int64_t)ldk + k] * v.Y[h[i] * (int64_t)v.ldm + m];
                Sf[(size_t)m * ldk + k] = s;
            }
        __syncthreads();
    } else {
        const double *r = Rf + (size_t)v.a * ldk;
        double *g = v.Gr + (size_t)f * ldk;
        double *U = v.U + (size_t)f * v.ts;
        // downdate: g = (X'X) r - X_H' (X_H r); u = X_H r are the held-out scores of this component
        for (int i0 = 0; i0 < v.ts; i0 += 256) {
            const int ni = v.ts - i0 < 256 ? v.ts - i0 : 256;
            warp_dots<CV_NT>(ni, K, [&](int i) { return v.X + h[i0 + i] * (int64_t)ldk; }, [&](int) { return r; }, sh.c);
            __syncthreads();
            for (int k = tid; k < K; k += CV_NT) {
                double acc = g[k];
                for (int i = 0; i < ni; ++i) acc -= sh.c[i] * v.X[h[i0 + i] * (int64_t)ldk + k];
                g[k] = acc;
            }
            for (int i = tid; i < ni; i += CV_NT) U[i0 + i] = sh.c[i];
            __syncthreads();
        }
        double acc = 0.0;
        for (int k = tid; k < K; k += CV_NT) acc = fma(r[k], g[k], acc);
        const double *qv;
        if constexpr (WIDE) {
            const double tt = wide_block_sum<CV_NT>(acc, sh);
            qv = wide + 2 * (size_t)M * M;
            finish_component_wide<CV_NT>(K, ldk, M, Sf, g, tt, r, Pf + (size_t)v.a * ldk, wide + 2 * (size_t)M * M, wide);
        } else {
            const double tt = block_sum<CV_NT>(acc, sh);
            qv = sh.q;
            finish_component<CV_NT>(K, ldk, M, Sf, g, tt, r, Pf + (size_t)v.a * ldk, qdump, sh);
        }
        // residuals of the held-out rows with a+1 components: e_a = e_{a-1} - u q_a'
        for (int e = tid; e < v.ts * M; e += CV_NT) {
            const int i = e / M, m = e % M;
            const int64_t row = gf * v.ts + i;
            const double prev = (v.a == 0) ? v.Y[h[i] * (int64_t)v.ldm + m]
                                           : v.E[((size_t)m * v.A + (v.a - 1)) * v.rows_local + row];
            v.E[((size_t)m * v.A + v.a) * v.rows_local + row] = prev - U[i] * qv[m];
        }
        __syncthreads();
    }
    const int an = v.a + 1;
    if (an >= v.A) return;
    if constexpr (WIDE) {
        extern __shared__ __align__(16) double cvw_dyn[];
        next_weights_wide<CV_NT>(K, ldk, M, an, Sf, Pf, Rf, v.w + (size_t)f * ldk, Rf + (size_t)an * ldk, wide, sh,
                                 v.wide_jac_smem ? cvw_dyn : nullptr);
    }
    else next_weights<CV_NT>(K, ldk, M, an, Sf, Pf, Rf, v.w + (size_t)f * ldk, Rf + (size_t)an * ldk, sh);
    if (v.Rt) {   // (every thread re-reads the elements it wrote itself: same k -> thread mapping as next_weights)
        const double *rn = Rf + (size_t)an * ldk;
        for (int k = tid; k < K; k += CV_NT) v.Rt[cv_gemm_rt_index(v.nkt, f, k)] = rn[k];
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Fused per-component fold kernel (1024 <= K <= 4096, M <= 8, test_size <= 256): the same algebra as above with every
// thread OWNING its columns {tid, tid + NT, ...} across the phases, so that vectors stay in registers between them
// and the small matrix S_f is swept three times per component instead of ~nine (q = S'r; deflation fused with the
// Gram matrix S'S of the deflated S — its entries are accumulated from the values just computed; w = S q):
// 24 instead of 38 K-vectors of per-fold traffic besides the P/R stream of the orthogonalisation.
// ------------------------------------------------------------------------------------------------------------------
#ifndef PLSB_CV_FNT
#define PLSB_CV_FNT 256
#endif
constexpr int CV_FNT = PLSB_CV_FNT;          // threads of the fused kernel
constexpr int CV_KPT_MAX = 4096 / CV_FNT;    // columns per thread: 16 (K <= 4096) or, as a template parameter, 8 (K <= 2048)

// sums NV per-thread values over the CTA; every thread receives all NV totals (fixed order: deterministic)
template <int NV>
__device__ __forceinline__ void block_sum_vec(double (&v)[NV], double *red /* [CV_FNT/32 + 1][NV] */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int NW = CV_FNT / 32;
#pragma unroll
    for (int i = 0; i < NV; ++i) v[i] = warp_sum(v[i]);
    __syncthreads();   // protect red[] from its previous use
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < NV; ++i) red[warp * NV + i] = v[i];
    }
    __syncthreads();
    if (threadIdx.x < NV) {   // one thread per value adds the warps' sums in warp order
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < NW; ++w) t += red[w * NV + threadIdx.x];
        red[NW * NV + threadIdx.x] = t;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NV; ++i) v[i] = red[NW * NV + i];
}

// out[d] = row_d . x for ndots rows of a K-vector matrix (row d at row_of(d)), with the COLUMN-OWNER layout of the fused
// kernel: thread t holds x at columns {t, t + NT, ...} in registers (xv), forms its share of up to NV dots at once — NV x
// KPT independent loads in flight per thread — and the shares are summed over the CTA in a fixed order.  (One warp per
// dot, four loads in flight per lane, left a lone fold CTA next to the product kernel with ~8 KB in flight per SM.)
// pf_bytes > 0: rows are pf_bytes long and thread 0 keeps CV_PF of them prefetched into L2 ahead of the batch in work.
constexpr int CV_PF = 8;
template <int NV, int KPT, typename FR>
__device__ __forceinline__ void owner_dots(int ndots, int K, FR row_of, const double (&xv)[KPT], double *out, double *red,
                                           uint32_t pf_bytes = 0) {
    const int tid = threadIdx.x;
    for (int d0 = 0; d0 < ndots; d0 += NV) {
        if (pf_bytes && tid == 0) {
#pragma unroll
            for (int u = 0; u < NV; ++u)
                if (d0 + CV_PF + u < ndots) l2_prefetch(row_of(d0 + CV_PF + u), pf_bytes);
        }
        const double *rows[NV];
#pragma unroll
        for (int u = 0; u < NV; ++u) rows[u] = row_of(d0 + u < ndots ? d0 + u : ndots - 1);
        double acc[NV];
#pragma unroll
        for (int u = 0; u < NV; ++u) acc[u] = 0.0;
#pragma unroll
        for (int j = 0; j < KPT; ++j) {
            const int k = tid + CV_FNT * j;
            if (k < K) {
#pragma unroll
                for (int u = 0; u < NV; ++u) acc[u] = fma(rows[u][k], xv[j], acc[u]);
            }
        }
        block_sum_vec<NV>(acc, red);
        if (tid < NV && d0 + tid < ndots) out[d0 + tid] = acc[tid];
    }
    __syncthreads();
}

#ifndef PLSB_CV_MINB
#define PLSB_CV_MINB 4
#endif
// shared state of the fused kernel: what it uses of StepSmem, sized for ITS M <= MM (StepSmem's Gram and eigenvector
// matrices are 32 x 32: 16 KB of the 20 KB — shared memory a co-resident GEMM CTA of the other wave cannot spare)
template <int MM>
struct FusedSmem {
    double q[MM];
    double c[256];
    double gram[MM * MM];
    double B[2 * 64];        // work arrays of the squaring eigen solver (8 x 8 padded, two buffers)
};

// The orthogonalisation stream of the fused kernel (rows of P_f for the dots p_j'w, then rows of R_f for r -= c_j r_j:
// 2 a K-vectors per component, 134 MB per fold at C5 — most of the kernel's traffic) through a ring of shared-memory
// slots filled by 1-D TMA bulk copies: 64 KB in flight per CTA whatever its register budget, where plain loads keep
// ~one row (32 KB) in flight and pay a full HBM round trip per row.  That is what makes ONE fold CTA per SM — all that
// fits next to the product kernel of the other lane — stream at a useful rate.  No producer warp: a slot is refilled
// by the last warp to leave it (same scheme as cv_gemm.cu).  Row a of P_f is written by this very kernel (generic
// proxy), so its dot is taken from global memory by the threads that wrote it and it never goes through the ring.
constexpr int CR_CH = CV_FNT * 8;                 // doubles per slot: 8 columns per thread (16 KB)
constexpr int CR_NSLOT_MAX = 8;                   // slots are a launch parameter (CvArgs::ring = their number, default 4)
static size_t cr_smem(int nslot) { return (size_t)nslot * CR_CH * 8 + nslot * (sizeof(uint64_t) + sizeof(int)) + 16; }

struct FoldRing {
    char *base; uint64_t *full; int *left;
    const double *Pf, *Rf;
    int a, nch, ldk, nP, nitems, ns; // nP: items of the P part (a rows x nch chunks); then (a + 1) rows of R; ns: slots
    int pf;                          // > 0: item it + pf is prefetched into L2 when item it is issued (costs no shared memory)
    __device__ __forceinline__ void prefetch(int it) const {     // one thread
        if (it >= nitems) return;
        const bool isR = it >= nP;
        const int e = isR ? it - nP : it;
        const int row = e / nch, ch = e % nch;
        int cols = ldk - ch * CR_CH; cols = cols > CR_CH ? CR_CH : cols;
        l2_prefetch((isR ? Rf : Pf) + (size_t)row * ldk + (size_t)ch * CR_CH, (uint32_t)cols * 8u);
    }
    __device__ __forceinline__ void issue(int it) const {        // one thread
        const int slot = it % ns;
        const bool isR = it >= nP;
        const int e = isR ? it - nP : it;
        const int row = e / nch, ch = e % nch;
        int cols = ldk - ch * CR_CH; cols = cols > CR_CH ? CR_CH : cols;
        const double *src = (isR ? Rf : Pf) + (size_t)row * ldk + (size_t)ch * CR_CH;
        mbar_arrive_expect_tx(&full[slot], (uint32_t)cols * 8u);
        bulk_g2s(base + (size_t)slot * CR_CH * 8, src, (uint32_t)cols * 8u, &full[slot]);
        if (pf > 0) prefetch(it + pf);
    }
    __device__ __forceinline__ const double *acquire(int it) const {
        const int slot = it % ns;
        mbar_wait(&full[slot], (uint32_t)((it / ns) & 1));
        return reinterpret_cast<const double *>(base + (size_t)slot * CR_CH * 8);
    }
    __device__ __forceinline__ void release(int it) const {      // all threads of the CTA, after their reads of the slot
        __syncwarp();
        if ((threadIdx.x & 31) == 0) {
            const int slot = it % ns;
            __threadfence_block();
            if (atomicSub(&left[slot], 1) == 1) {
                left[slot] = CV_FNT / 32;
                __threadfence_block();
                if (it + ns < nitems) issue(it + ns);
            }
        }
    }
};

// KPT = columns per thread: 16 covers K <= 4096; 8 (K <= 2048) halves the register arrays, so that the M = 8 variant fits
// 64 registers as well and four CTAs share an SM instead of three (a CTA's life is a chain of dependent phases: its
// throughput is the number of CTAs in flight)
template <int MM, int KPT, int MINB = (MM <= 4 ? PLSB_CV_MINB * 256 / CV_FNT : 3)>
__global__ void __launch_bounds__(CV_FNT, MINB) cv_fold_fused_kernel(const CvArgs v) {
    constexpr int CV_KPT = KPT;
    extern __shared__ __align__(128) char cr_dyn[];
    constexpr int NP = MM * (MM + 1) / 2;
    __shared__ FusedSmem<MM> sh;
    __shared__ double red[(CV_FNT / 32 + 1) * (NP > 8 ? NP : 8)];   // block_sum_vec of up to max(NPP, MM, 4) values
    __shared__ double qe[MM];
    const int tid = threadIdx.x;
    const int f = blockIdx.x;
    const int64_t gf = (int64_t)v.f0 + f;
    const int64_t *h = v.test_idx + gf * v.ts;
    const int K = v.K, ldk = v.ldk, M = v.M, a = v.a, ts = v.ts;
    double *Sf = v.S + (size_t)f * M * ldk;
    double *Pf = v.P + (size_t)f * v.A * ldk;
    double *Rf = v.R + (size_t)f * v.A * ldk;
    const double *r = Rf + (size_t)a * ldk;
    const double *g = v.Gr + (size_t)f * ldk;
    double *pa = Pf + (size_t)a * ldk;
    double *U = v.U + (size_t)f * ts;
    cv_span_mark(v.span, 0);
    CV_TRACE(0);
    if (v.cta_trace && tid == 0) {
        unsigned smid; long long t_;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
        v.cta_trace[3 * (size_t)f] = smid; v.cta_trace[3 * (size_t)f + 1] = t_;
    }

    // Everything this CTA is about to read for the first time comes from HBM: run those streams ahead into L2 now (one
    // thread, a handful of bulk-prefetch instructions).  Next to the product kernel of the other lane there is ONE fold
    // CTA per SM, and its chain of dependent phases would otherwise pay a full HBM round trip at the head of each.
    FoldRing ring;
    const bool use_ring = v.ring != 0 && a + 1 < v.A;     // (a CTA that returns before the stream is read must not start it)
    if (use_ring) {
        ring.ns = v.ring; ring.pf = v.ring_pf;
        ring.base = cr_dyn;
        ring.full = reinterpret_cast<uint64_t *>(cr_dyn + (size_t)ring.ns * CR_CH * 8);
        ring.left = reinterpret_cast<int *>(ring.full + ring.ns);
        ring.Pf = Pf; ring.Rf = Rf; ring.a = a; ring.ldk = ldk;
        ring.nch = (ldk + CR_CH - 1) / CR_CH;
        ring.nP = a * ring.nch;
        ring.nitems = (2 * a + 1) * ring.nch;
        if (tid == 0) {
            for (int sl = 0; sl < ring.ns; ++sl) { mbar_init(&ring.full[sl], 1); ring.left[sl] = CV_FNT / 32; }
            mbar_fence_init();
            for (int it = 0; it < ring.ns && it < ring.nitems; ++it) ring.issue(it);   // rows 0.. are final: start now
            for (int it = ring.ns; ring.pf > 0 && it < ring.ns + ring.pf; ++it) ring.prefetch(it);
        }
        __syncthreads();
    }
    const uint32_t row_bytes = (uint32_t)ldk * 8u;
    const bool pf = v.prefetch != 0;
    if (pf && tid == 0) {
        l2_prefetch(r, row_bytes);
        l2_prefetch(g, row_bytes);
        for (int i = 0; i < ts && i < 8; ++i) l2_prefetch(v.X + h[i] * (int64_t)ldk, row_bytes);
        for (int m = 0; m < M; ++m) l2_prefetch(Sf + (size_t)m * ldk, row_bytes);
        for (int j = 0; j < a + 1 && j < CV_PF; ++j) l2_prefetch(Pf + (size_t)j * ldk, row_bytes);
    }
    // ---- u = X_H r (held-out scores), then the downdate g = (X'X) r - X_H' u in registers; tt = r'g ----------------
    {
        double rv[CV_KPT];
#pragma unroll
        for (int j = 0; j < CV_KPT; ++j) {
            const int k = tid + CV_FNT * j;
            rv[j] = k < K ? r[k] : 0.0;
        }
        if (ts == 1) owner_dots<1, KPT>(1, K, [&](int) { return v.X + h[0] * (int64_t)ldk; }, rv, sh.c, red);
        else owner_dots<4, KPT>(ts, K, [&](int i) { return v.X + h[i] * (int64_t)ldk; }, rv, sh.c, red);
    }
    CV_TRACE(1);   // u done
    double gv[CV_KPT];
    double tt1[1] = {0.0};
#pragma unroll
    for (int j = 0; j < CV_KPT; ++j) {
        const int k = tid + CV_FNT * j;
        gv[j] = k < K ? g[k] : 0.0;
    }
    for (int i = 0; i < ts; ++i) {          // all of a thread's columns of one held-out row in flight at once
        const double ci = sh.c[i];
        const double *xr = v.X + h[i] * (int64_t)ldk;
#pragma unroll
        for (int j = 0; j < CV_KPT; ++j) {
            const int k = tid + CV_FNT * j;
            if (k < K) gv[j] -= ci * xr[k];
        }
    }
#pragma unroll
    for (int j = 0; j < CV_KPT; ++j) {
        const int k = tid + CV_FNT * j;
        if (k < K) tt1[0] = fma(r[k], gv[j], tt1[0]);
    }
    for (int i = tid; i < ts; i += CV_FNT) U[i] = sh.c[i];
    block_sum_vec<1>(tt1, red);
    const double tt = tt1[0];
    CV_TRACE(2);   // downdate + tt done

    // ---- p = g / tt (stored), q = S'r / tt ---------------------------------------------------------------------------
    double qm[MM];
#pragma unroll
    for (int m = 0; m < MM; ++m) qm[m] = 0.0;
#pragma unroll
    for (int j = 0; j < CV_KPT; ++j) {
        const int k = tid + CV_FNT * j;
        if (k < K) {
            pa[k] = gv[j] / tt;
            const double rk = r[k];
#pragma unroll
            for (int m = 0; m < MM; ++m)
                if (m < M) qm[m] = fma(Sf[(size_t)m * ldk + k], rk, qm[m]);
        }
    }
    for (int k = K + tid; k < ldk; k += CV_FNT) pa[k] = 0.0;
    block_sum_vec<MM>(qm, red);
#pragma unroll
    for (int m = 0; m < MM; ++m) qm[m] = qm[m] / tt;
    if (tid == 0) {
#pragma unroll
        for (int m = 0; m < MM; ++m) sh.q[m] = qm[m];
    }

    // ---- S -= tt p q' fused with the Gram matrix of the deflated S ---------------------------------------------------
    CV_TRACE(3);   // p, q done
    // (MM = 8: the 36 Gram entries go in two passes of 18 accumulators — the second re-reads the thread's own columns
    //  of the deflated S from L2 — so that the kernel fits 80 registers and three CTAs per SM, one of them next to
    //  the product kernel of the other lane)
    constexpr int NPASS = MM > 4 ? 2 : 1, NPP = (NP + NPASS - 1) / NPASS;
    double gp[NPP];
#pragma unroll
    for (int i = 0; i < NPP; ++i) gp[i] = 0.0;
#pragma unroll
    for (int j = 0; j < CV_KPT; ++j) {
        const int k = tid + CV_FNT * j;
        if (k < K) {
            const double tp = tt * pa[k];
            double sv[MM];
#pragma unroll
            for (int m = 0; m < MM; ++m) {
                sv[m] = 0.0;
                if (m < M) {
                    sv[m] = Sf[(size_t)m * ldk + k] - tp * qm[m];
                    Sf[(size_t)m * ldk + k] = sv[m];
                }
            }
            int pi = 0;
#pragma unroll
            for (int u = 0; u < MM; ++u)
#pragma unroll
                for (int w2 = u; w2 < MM; ++w2) { if (pi < NPP) gp[pi] = fma(sv[u], sv[w2], gp[pi]); ++pi; }
        }
    }
    // residuals of the held-out rows with a+1 components: e_a = e_{a-1} - u q_a'
    __syncthreads();   // sh.q visible
    for (int e = tid; e < ts * M; e += CV_FNT) {
        const int i = e / M, m = e % M;
        const int64_t row = gf * ts + i;
        const double prev = (a == 0) ? v.Y[h[i] * (int64_t)v.ldm + m] : v.E[((size_t)m * v.A + (a - 1)) * v.rows_local + row];
        v.E[((size_t)m * v.A + a) * v.rows_local + row] = prev - sh.c[i] * sh.q[m];
    }
    const int an = a + 1;
    if (an >= v.A) { cv_span_mark(v.span, 1); return; }
    CV_TRACE(4);   // deflation + residuals done

    // ---- next weights: dominant eigenvector of S'S, w = S q / |S q| -------------------------------------------------
    if (MM > 1) {
#pragma unroll
        for (int pass = 0; pass < NPASS; ++pass) {
            if (pass > 0) {
#pragma unroll
                for (int i = 0; i < NPP; ++i) gp[i] = 0.0;
#pragma unroll
                for (int j = 0; j < CV_KPT; ++j) {
                    const int k = tid + CV_FNT * j;
                    if (k < K) {
                        double sv[MM];
#pragma unroll
                        for (int m = 0; m < MM; ++m) sv[m] = m < M ? Sf[(size_t)m * ldk + k] : 0.0;   // written by this thread
                        int pi = 0;
#pragma unroll
                        for (int u = 0; u < MM; ++u)
#pragma unroll
                            for (int w2 = u; w2 < MM; ++w2) {
                                if (pi >= pass * NPP && pi < (pass + 1) * NPP) gp[pi - pass * NPP] = fma(sv[u], sv[w2], gp[pi - pass * NPP]);
                                ++pi;
                            }
                    }
                }
            }
            block_sum_vec<NPP>(gp, red);
            if (tid == 0) {
                int pi = 0;
#pragma unroll
                for (int u = 0; u < MM; ++u)
#pragma unroll
                    for (int w2 = u; w2 < MM; ++w2) {
                        if (pi >= pass * NPP && pi < (pass + 1) * NPP && u < M && w2 < M) {
                            sh.gram[u * M + w2] = gp[pi - pass * NPP]; sh.gram[w2 * M + u] = gp[pi - pass * NPP];
                        }
                        ++pi;
                    }
            }
        }
        __syncthreads();
        if (tid < 32) dominant_eigvec_warp<8>(sh.gram, sh.B, sh.B + 64, M, qe);
        __syncthreads();
    }
    CV_TRACE(5);   // Gram reduced, eigenvector done
    double wv[CV_KPT];
    double n1[1] = {0.0};
#pragma unroll
    for (int j = 0; j < CV_KPT; ++j) {
        const int k = tid + CV_FNT * j;
        double acc = 0.0;
        if (k < K) {
            if (MM == 1) acc = Sf[k];
            else {
#pragma unroll
                for (int m = 0; m < MM; ++m)
                    if (m < M) acc = fma(Sf[(size_t)m * ldk + k], qe[m], acc);
            }
            n1[0] = fma(acc, acc, n1[0]);
        }
        wv[j] = acc;
    }
    block_sum_vec<1>(n1, red);
    const double nrm = sqrt(n1[0]);
    double *w = v.w + (size_t)f * ldk;
#pragma unroll
    for (int j = 0; j < CV_KPT; ++j) {
        const int k = tid + CV_FNT * j;
        if (k < K) { wv[j] = wv[j] / nrm; w[k] = wv[j]; }
    }
    __syncthreads();
    CV_TRACE(6);   // w normalised and stored
    // ---- r = w - sum_j (p_j'w) r_j ------------------------------------------------------------------------------------
    double *rn = Rf + (size_t)an * ldk;
    if (use_ring) {
        int it = 0;
        for (int j0 = 0; j0 < a; j0 += 4) {          // c_j = p_j'w for the rows that were final when the kernel started
            double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (j0 + u < a) {
                    for (int ch = 0; ch < ring.nch; ++ch, ++it) {
                        const double *sl = ring.acquire(it);
                        if (ch == 0) {
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                if (tid + CV_FNT * j < K) acc[u] = fma(sl[tid + CV_FNT * j], wv[j], acc[u]);
                        } else if constexpr (KPT > 8) {
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                if (CR_CH + tid + CV_FNT * j < K) acc[u] = fma(sl[tid + CV_FNT * j], wv[8 + j], acc[u]);
                        }
                        ring.release(it);
                    }
                }
            }
            block_sum_vec<4>(acc, red);
            if (tid < 4 && j0 + tid < a) sh.c[j0 + tid] = acc[tid];
        }
        // row a = this component's p, written above by the same threads that read it back here
        owner_dots<1, KPT>(1, K, [&](int) { return (const double *)pa; }, wv, sh.c + a, red);
        CV_TRACE(7);   // P dots done
        for (int jj = 0; jj < an; ++jj) {
            const double cj = sh.c[jj];
            for (int ch = 0; ch < ring.nch; ++ch, ++it) {
                const double *sl = ring.acquire(it);
                if (ch == 0) {
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (tid + CV_FNT * j < K) wv[j] -= cj * sl[tid + CV_FNT * j];
                } else if constexpr (KPT > 8) {
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (CR_CH + tid + CV_FNT * j < K) wv[8 + j] -= cj * sl[tid + CV_FNT * j];
                }
                ring.release(it);
            }
        }
    } else
    for (int j0 = 0; j0 < an; j0 += 256) {
        const int nj = an - j0 < 256 ? an - j0 : 256;
        if (pf && tid == 0)
            for (int j = 0; j < nj && j < CV_PF; ++j) l2_prefetch(Rf + (size_t)(j0 + j) * ldk, row_bytes);
        owner_dots<4, KPT>(nj, K, [&](int j) { return (const double *)(Pf + (size_t)(j0 + j) * ldk); }, wv, sh.c, red,
                      pf ? row_bytes : 0u);
        for (int jj = 0; jj < nj; ++jj) {   // one row of R_f: all of a thread's columns in flight at once
            if (pf && tid == 0 && jj + CV_PF < nj) l2_prefetch(Rf + (size_t)(j0 + jj + CV_PF) * ldk, row_bytes);
            const double cj = sh.c[jj];
            const double *rr = Rf + (size_t)(j0 + jj) * ldk;
#pragma unroll
            for (int j = 0; j < CV_KPT; ++j) {
                const int k = tid + CV_FNT * j;
                if (k < K) wv[j] -= cj * rr[k];
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < CV_KPT; ++j) {
        const int k = tid + CV_FNT * j;
        if (k < K) {
            rn[k] = wv[j];
            if (v.Rt) v.Rt[cv_gemm_rt_index(v.nkt, f, k)] = wv[j];
        }
    }
    for (int k = K + tid; k < ldk; k += CV_FNT) rn[k] = 0.0;
    CV_TRACE(8);   // R stream done, r stored
    cv_span_mark(v.span, 1);
    if (v.cta_trace && tid == 0) {
        long long t_;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
        v.cta_trace[3 * (size_t)f + 2] = t_;
    }
}

struct CvBuf {
    plsb_ctx *c;
    void *p = nullptr;
    explicit CvBuf(plsb_ctx *ctx) : c(ctx) {}
    ~CvBuf() { dev_free(c, p); }
    int alloc(size_t bytes) { return dev_alloc(c, &p, bytes); }
    double *d() { return (double *)p; }
};

// PRESS partial sums on the device (validation's column sums of squared residuals, reference src/pls.cpp:248-252):
// press[mc] = sum over this call's rows of E[mc][row]^2, two fixed-shape stages (1024-row blocks summed by one CTA
// each, block totals added in block order): deterministic, no atomics.
__global__ void __launch_bounds__(256) press_blocks_kernel(const double *__restrict__ E, int64_t rows, int nblk,
                                                           double *__restrict__ part) {
    __shared__ double red[8];
    const int mc = blockIdx.y, b = blockIdx.x;
    const double *col = E + (size_t)mc * rows;
    const int64_t r0 = (int64_t)b * 1024;
    double acc = 0.0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int64_t i = r0 + threadIdx.x + 256 * j;
        if (i < rows) { const double e = col[i]; acc = fma(e, e, acc); }
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double tsum = 0.0;
#pragma unroll
        for (int wv = 0; wv < 8; ++wv) tsum += red[wv];
        part[(size_t)mc * nblk + b] = tsum;
    }
}
// out is validation's M x A matrix, column-major: out[c*M + m]  (block mc = m*A + c of the residual block)
__global__ void __launch_bounds__(256) press_final_kernel(const double *__restrict__ part, int nblk, int M, int A,
                                                          double *__restrict__ out) {
    __shared__ double red[256];
    const double *p = part + (size_t)blockIdx.x * nblk;
    double acc = 0.0;
    for (int i = threadIdx.x; i < nblk; i += 256) acc += p[i];     // thread t: blocks t, t+256, ... in order
    red[threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double tsum = 0.0;
        for (int i = 0; i < 256; ++i) tsum += red[i];
        const int m = blockIdx.x / A, cc = blockIdx.x % A;
        out[(size_t)cc * M + m] = tsum;
    }
}

static int device_press(plsb_ctx *c, const double *Ed, int64_t rows, int64_t M, int64_t A, double *press_host) {
    const int64_t MA = M * A;
    const int nblk = (int)ceil_div(rows, 1024);
    CvBuf part(c), out(c);
    PLSB_TRY(part.alloc((size_t)MA * nblk * 8));
    PLSB_TRY(out.alloc((size_t)MA * 8));
    press_blocks_kernel<<<dim3((unsigned)nblk, (unsigned)MA), 256, 0, c->stream>>>(Ed, rows, nblk, part.d());
    press_final_kernel<<<(unsigned)MA, 256, 0, c->stream>>>(part.d(), nblk, (int)M, (int)A, out.d());
    c->launches += 2;
    PLSB_CUDA(cudaGetLastError());
    PLSB_CUDA(cudaMemcpyAsync(press_host, out.p, (size_t)MA * 8, cudaMemcpyDeviceToHost, c->stream));
    return PLSB_OK;
}

// small-data regime (cv_small.cu)
bool cv_small_supported(const plsb_ctx *c, const plsb_dataset *ds, int64_t A, int64_t ts, bool *pr_in_smem);
int launch_cv_small(plsb_ctx *c, const plsb_dataset *ds, int64_t A, const int64_t *idx_dev, int64_t nf, int64_t ts,
                    int64_t rows_local, double *Ed, bool pr_in_smem, int64_t nf_total);

}  // namespace plsb

using namespace plsb;

extern "C" int plsb_cv_folds_press(plsb_dataset_t ds, int64_t A, int method, int64_t nfolds_total, int64_t test_size,
                                   const int64_t *test_idx, int64_t fold0, int64_t nfolds_local, double *E, double *press) {
    PLSB_ARG(ds && test_idx && (E || press), "null argument");
    PLSB_ARG(A >= 1, "A must be >= 1");
    PLSB_ARG(method == PLSB_KERNEL_TYPE1 || method == PLSB_KERNEL_TYPE2, "unknown method %d", method);
    PLSB_ARG(test_size >= 1 && test_size < ds->N, "test_size %lld must be in [1, N)", (long long)test_size);
    PLSB_ARG(fold0 >= 0 && nfolds_local >= 0 && fold0 + nfolds_local <= nfolds_total, "bad fold range");
    PLSB_ARG(!ds->sharded, "cross-validation wants the full data set on every GPU (folds are sharded, not rows)");
    const int64_t K = ds->K, ldk = ds->ldk, M = ds->M, ts = test_size;
    if (nfolds_local == 0) {
        if (press) for (int64_t i = 0; i < M * A; ++i) press[i] = 0.0;
        return PLSB_OK;
    }
    plsb_ctx *c = ds->ctx;
    PLSB_CUDA(cudaSetDevice(c->device));
    const char *trace = getenv("PLSB_TRACE");
    const auto t_start = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (trace && *trace == '1') {
            cudaStreamSynchronize(c->stream);
            fprintf(stderr, "[plsb] cv_folds: %s at %.3f ms\n", what,
                    std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_start).count());
        }
    };
    const int64_t *tloc = test_idx + fold0 * ts;
    for (int64_t i = 0; i < nfolds_local * ts; ++i)
        PLSB_ARG(tloc[i] >= 0 && tloc[i] < ds->N, "test index %lld out of range", (long long)tloc[i]);
    const int64_t rows_local = nfolds_local * ts, rows_total = nfolds_total * ts;

    CvBuf G(c), S0(c), idx(c), Ed(c);
    PLSB_TRY(idx.alloc((size_t)rows_local * sizeof(int64_t)));
    PLSB_CUDA(cudaMemcpyAsync(idx.p, tloc, (size_t)rows_local * sizeof(int64_t), cudaMemcpyHostToDevice, c->stream));
    PLSB_TRY(Ed.alloc((size_t)M * A * rows_local * 8));

    lap("indices checked and queued, residual block allocated");
    auto finish = [&]() -> int {
        lap("kernels done");
        if (press) PLSB_TRY(device_press(c, Ed.d(), rows_local, M, A, press));
        if (E) {
            // the whole residual block in one piece, bound for pageable memory: pinned staging + host threads (a plain
            // cudaMemcpy into pageable memory runs at 2-3 GB/s: 4.3 MB of config 3 cost more than its kernel)
            bool staged = false;
            const size_t nE = (size_t)M * A * rows_local;
            if (rows_local == rows_total && nE * 8 >= ((size_t)1 << 20)) {
                cudaPointerAttributes pa{};
                const bool pinned = cudaPointerGetAttributes(&pa, E) == cudaSuccess && pa.type == cudaMemoryTypeHost;
                cudaGetLastError();
                const char *off = getenv("PLSB_NO_STAGE");
                staged = !pinned && !(off && *off == '1');
            }
            if (staged) PLSB_TRY(download_staged(c, Ed.d(), nE, E, false));
            else PLSB_CUDA(cudaMemcpy2DAsync(E + fold0 * ts, (size_t)rows_total * 8, Ed.p, (size_t)rows_local * 8,
                                             (size_t)rows_local * 8, (size_t)(M * A), cudaMemcpyDeviceToHost, c->stream));
        }
        PLSB_CUDA(cudaStreamSynchronize(c->stream));
        lap("results on the host");
        return PLSB_OK;
    };

    // small-data regime: the whole data set in one SM's shared memory, one launch for the call (cv_small.cu)
    {
        bool pr_in_smem = false;
        const char *off = getenv("PLSB_CV_NO_SMALL");
        if (!(off && *off == '1') && cv_small_supported(c, ds, A, ts, &pr_in_smem)) {
            PLSB_TRY(launch_cv_small(c, ds, A, (const int64_t *)idx.p, nfolds_local, ts, rows_local, Ed.d(), pr_in_smem, nfolds_total));
            return finish();
        }
    }

    // the full cross products: taken from the data set when its upload built them under the transfer (dataset.cu)
    const int64_t ldg = round_up(K, 2);
    const double *Gfull = nullptr, *S0full = nullptr;
    if (ds->G && ds->S0 && ds->gram_valid && ds->ldg == ldg) {
        Gfull = ds->G; S0full = ds->S0;
    } else {
        PLSB_TRY(G.alloc((size_t)K * ldg * 8));
        PLSB_CUDA(cudaMemsetAsync(G.p, 0, (size_t)K * ldg * 8, c->stream));
        PLSB_TRY(S0.alloc((size_t)M * ldk * 8));
        PLSB_TRY(launch_xtx(c, ds, G.d(), ldg));
        PLSB_TRY(launch_xty_pass(c, ds, S0.d()));
        Gfull = G.d(); S0full = S0.d();
    }

    // ---- waves and lanes ------------------------------------------------------------------------------------------
    // The batched GEMM is FP64-tensor-bound and leaves HBM idle; the fold kernel is HBM-bound and leaves the tensor
    // pipe idle.  With two LANES (two waves in flight, each with its own S/P/R/w/Gr/U buffers) the GEMM of one lane runs
    // under the fold kernel of the other: all GEMMs go to one high-priority stream, all fold kernels to a low-priority
    // one, chained per lane by events (fold(a) after GEMM(a), GEMM(a+1) after fold(a)):
    //      gemm stream:  G0(0) | G1(0) | G0(1) | G1(1) | ...
    //      fold stream:        | F0(0) | F1(0) | F0(1) | ...
    // The product runs as a persistent kernel over pre-tiled operands (one CTA per SM, 8 warps x 168 registers, two bulk
    // copies per stage: cv_gemm.cu), so the fold CTAs are placed in the quarter of each SM it leaves free.  One lane on the context's stream is kept for
    // per-kernel profiling (profile mode 1), small calls and as the A/B reference (PLSB_CV_LANES=1).
    int lanes = 2;
    if (const char *e = getenv("PLSB_CV_LANES")) { const int l = atoi(e); if (l == 1 || l == 2) lanes = l; }
    if (c->prof == 1 || nfolds_local < 512) lanes = 1;
    // wave size from free memory: per fold S (M), P and R (A each), w, Gr: (M + 2A + 2) * ldk doubles.  One lane: at most
    // 4736 folds = 37 column tiles of the batched GEMM (with K = 4096, 32 row tiles, that is 1184 CTAs = 8 full waves of
    // 148 SMs, with K = 2048 four).  Two lanes: at most 6144 per wave; the wave count is chosen below.
    size_t free_b = 0;
    PLSB_TRY(dev_mem_available(c, &free_b));
    const size_t per_fold = (size_t)(M + 2 * A + 2) * ldk * 8 + (size_t)ts * 8 + (lanes == 2 ? (size_t)ldk * 9 + 64 : 0) +
                            step_wide_scratch_bytes((int)M);
    int64_t F = (int64_t)((double)free_b * 0.6 / (double)per_fold) / lanes;
    if (F > (lanes == 2 ? 6144 : 4736)) F = lanes == 2 ? 6144 : 4736;
    if (const char *e = getenv("PLSB_FOLD_BATCH")) { const int64_t v = atoll(e); if (v > 0 && v < F) F = v; }
    if (F < 1 && lanes == 2) { lanes = 1; F = (int64_t)((double)free_b * 0.6 / (double)per_fold); }
    PLSB_ARG(F >= 1, "not enough device memory for one fold");
    // PLSB_CVG_CTAS: run the product kernel on fewer SMs (one persistent CTA each) and leave the others entirely to the
    // fold kernel of the other lane (three fold CTAs fit a free SM, one fits next to a product CTA).  Measured at C5
    // (profiles/r2l_cv_ab.log): 148 / 136 / 128 / 120 / 112 product CTAs -> 1.81 / 1.76 / 1.75 / 1.82 / 1.80 s per call,
    // i.e. no difference — what the slot model of DESIGN 3.6 predicts: product slots and fold slots trade one for one
    // along the same line.  Default: all SMs.
    int gemm_ctas = c->sm_count;
    if (lanes == 2) {
        if (const char *e = getenv("PLSB_CVG_CTAS")) { const int n = atoi(e); if (n > 0 && n <= c->sm_count) gemm_ctas = n; }
        // Wave size: the one whose product items (row tiles x 128-fold column tiles) fill whole rounds of gemm_ctas CTAs
        // best over the call — equal waves (an even number of them), or whole-tile waves with a ragged last one.  Ties:
        // an even wave count (both lanes busy to the end), then fewer waves.
        const int64_t nmt = ceil_div(K, 128), Fcap = F;
        auto rounds = [&](int64_t folds) { return folds > 0 ? ceil_div(nmt * ceil_div(folds, 128), (int64_t)gemm_ctas) : (int64_t)0; };
        int64_t best_F = 0, best_cost = -1, best_waves = 0;
        auto consider = [&](int64_t Fc) {
            if (Fc < 1 || Fc > Fcap) return;
            const int64_t nfull = nfolds_local / Fc, rem = nfolds_local % Fc, waves = nfull + (rem > 0 ? 1 : 0);
            if (waves < 2) return;
            const int64_t cost = nfull * rounds(Fc) + rounds(rem);
            const bool better = best_cost < 0 || cost < best_cost ||
                                (cost == best_cost && ((waves % 2 == 0) > (best_waves % 2 == 0) ||
                                                       ((waves % 2 == 0) == (best_waves % 2 == 0) && waves < best_waves)));
            if (better) { best_cost = cost; best_F = Fc; best_waves = waves; }
        };
        const int64_t nw0 = 2 * ceil_div(nfolds_local, 2 * Fcap);
        for (int64_t nw = nw0; nw <= (4 * nw0 > nw0 + 12 ? 4 * nw0 : nw0 + 12); nw += 2) {
            const int64_t Fw = ceil_div(nfolds_local, nw);
            consider(Fw);
            if (Fw <= 256) break;
        }
        for (int64_t Fc = 256; Fc <= Fcap; Fc += 128) consider(Fc);
        if (best_F > 0) F = best_F;
    }
    if (F > nfolds_local) F = nfolds_local;

    struct Lane { CvBuf S, P, R, w, Gr, U, Rt, wide; cudaEvent_t evG = nullptr, evF = nullptr; explicit Lane(plsb_ctx *c) : S(c), P(c), R(c), w(c), Gr(c), U(c), Rt(c), wide(c) {} };
    Lane lane0(c), lane1(c);
    Lane *ln[2] = {&lane0, &lane1};
    for (int l = 0; l < lanes; ++l) {
        PLSB_TRY(ln[l]->S.alloc((size_t)F * M * ldk * 8));
        PLSB_TRY(ln[l]->P.alloc((size_t)F * A * ldk * 8));
        PLSB_TRY(ln[l]->R.alloc((size_t)F * A * ldk * 8));
        PLSB_TRY(ln[l]->w.alloc((size_t)F * ldk * 8));
        PLSB_TRY(ln[l]->Gr.alloc((size_t)F * ldk * 8));
        PLSB_TRY(ln[l]->U.alloc((size_t)F * ts * 8));
        if (const size_t wb = step_wide_scratch_bytes((int)M)) PLSB_TRY(ln[l]->wide.alloc((size_t)F * wb));
    }
    CvBuf Gt(c);
    const int nkt = (int)ceil_div(K, CVG_BK);
    if (lanes == 2) {   // operands of the persistent product kernel in its tiled layouts (cv_gemm.cu); the padding stays zero
        PLSB_TRY(Gt.alloc(cv_gemm_gt_doubles(K) * 8));
        PLSB_TRY(launch_cv_gemm_tile_g(c, c->stream, Gfull, ldg, K, Gt.d()));
        for (int l = 0; l < 2; ++l) {
            const size_t nb = cv_gemm_rt_doubles(K, F) * 8;
            PLSB_TRY(ln[l]->Rt.alloc(nb));
            PLSB_CUDA(cudaMemsetAsync(ln[l]->Rt.p, 0, nb, c->stream));
        }
    }

    CvArgs v{};
    v.X = ds->X; v.Y = ds->Y; v.S0 = S0full; v.test_idx = (const int64_t *)idx.p; v.E = Ed.d();
    v.rows_local = rows_local; v.K = (int)K; v.ldk = (int)ldk; v.ldm = (int)ds->ldm; v.M = (int)M; v.A = (int)A;
    v.ts = (int)ts;
    { const char *pe = getenv("PLSB_CV_PREFETCH"); v.prefetch = pe ? atoi(pe) : 0; }   // measured: slower (profiles/r2h)
    const char *v1 = getenv("PLSB_CV_V1");   // A/B switch: the phase-by-phase kernel for every launch
    const bool fused = K >= 1024 && K <= (int64_t)CV_KPT_MAX * CV_FNT && M <= 8 && ts <= 256 && A <= 256 && !(v1 && *v1 == '1');
    // the instance of the fused kernel for this shape: responses 1 / 2 / <= 4 / <= 8, columns per thread 8 (K <= 2048) / 16
    void (*fused_kern)(const CvArgs) = nullptr;
    const bool kpt8 = ldk <= 8 * CV_FNT;
    if (M == 1) fused_kern = kpt8 ? cv_fold_fused_kernel<1, 8> : cv_fold_fused_kernel<1, 16>;
    else if (M == 2) fused_kern = kpt8 ? cv_fold_fused_kernel<2, 8> : cv_fold_fused_kernel<2, 16>;
    else if (M <= 4) fused_kern = kpt8 ? cv_fold_fused_kernel<4, 8> : cv_fold_fused_kernel<4, 16>;
    else {
        const char *mb = getenv("PLSB_CV_M8_MINB");      // A/B: four CTAs per SM at 64 registers (spills) or three at 80
        fused_kern = kpt8 ? ((mb && *mb == '4') ? cv_fold_fused_kernel<8, 8, 4> : cv_fold_fused_kernel<8, 8, 3>) : cv_fold_fused_kernel<8, 16>;
    }
    if (lanes == 2) {
        // An SM runs with ONE shared-memory carve-out at a time: a kernel that prefers a different split of the 256 KB
        // L1/shared array than the resident product kernel (which needs the maximum) is not placed next to it.  Ask for
        // the same carve-out for the fold kernels, so that they can share the SMs.  (PLSB_CV_CARVEOUT=0: A/B)
        const char *ce = getenv("PLSB_CV_CARVEOUT");
        const int carve = (ce && *ce == '0') ? cudaSharedmemCarveoutDefault : cudaSharedmemCarveoutMaxShared;
        PLSB_CUDA(cudaFuncSetAttribute(fused_kern, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
        PLSB_CUDA(cudaFuncSetAttribute(cv_fold_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
        PLSB_CUDA(cudaFuncSetAttribute(cv_fold_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
    }
    {   // the TMA-fed ring for the orthogonalisation stream
        const char *re = getenv("PLSB_CV_RING");
        int ns = 3;                                   // PLSB_CV_RING = 0 (plain loads) or the number of 16 KB slots;
                                                      // three: 53 KB per CTA, so that FOUR 64-register CTAs share a free SM
                                                      // (slot model, DESIGN 3.6: 6.0 -> 5.8 ms per wave-component at C5)
        if (re) { ns = atoi(re); if (ns == 1) ns = 3; if (ns > CR_NSLOT_MAX) ns = CR_NSLOT_MAX; if (ns < 0) ns = 0; }
        v.ring = fused ? ns : 0;
        { const char *pe = getenv("PLSB_CV_RING_PF"); v.ring_pf = pe ? atoi(pe) : 0; }
        if (v.ring) {
            const int fsm_max = (int)cr_smem(CR_NSLOT_MAX);
            PLSB_CUDA(cudaFuncSetAttribute(fused_kern, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm_max));
        }
    }
    const size_t fsm = v.ring ? cr_smem(v.ring) : 0;
    size_t wide_jsm = 0;
    if (M > PLSB_MAX_M) {
        wide_jsm = wide_jacobi_smem_doubles((int)M) * sizeof(double);
        if (wide_jsm + 4096 > c->smem_optin) wide_jsm = 0;
        if (wide_jsm) PLSB_CUDA(cudaFuncSetAttribute(cv_fold_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wide_jsm));
    }
    v.wide_jac_smem = wide_jsm ? 1 : 0;
    CvBuf tracebuf(c);
    int trace_a = -1, traced = 0;
    if (const char *te = getenv("PLSB_CV_TRACE")) {
        trace_a = atoi(te);
        PLSB_TRY(tracebuf.alloc(16 * sizeof(long long)));
        PLSB_CUDA(cudaMemsetAsync(tracebuf.p, 0, 16 * sizeof(long long), c->stream));
    }
    auto dump_trace = [&]() {
        if (!tracebuf.p) return;
        long long hh[16];
        cudaStreamSynchronize(c->stream);
        if (cudaMemcpy(hh, tracebuf.p, sizeof(hh), cudaMemcpyDeviceToHost) != cudaSuccess) return;
        fprintf(stderr, "[plsb] fused fold kernel, component %d, CTA 7, ns since its start:", trace_a);
        for (int i = 1; i <= 8; ++i) fprintf(stderr, " %lld", hh[i] ? hh[i] - hh[0] : -1);
        fprintf(stderr, "  (u | downdate+tt | p,q | deflate | gram+eig | w | P dots | R stream)\n");
    };
    CvBuf spanbuf(c);
    const int span_max = 600;
    int span_n = 0;
    std::vector<char> span_kind;
    if (getenv("PLSB_CV_TIMELINE")) {
        PLSB_TRY(spanbuf.alloc((size_t)2 * span_max * sizeof(unsigned long long)));
        std::vector<unsigned long long> init((size_t)2 * span_max);
        for (int i = 0; i < span_max; ++i) { init[2 * i] = ~0ull; init[2 * i + 1] = 0ull; }
        PLSB_CUDA(cudaMemcpyAsync(spanbuf.p, init.data(), init.size() * sizeof(unsigned long long), cudaMemcpyHostToDevice, c->stream));
        PLSB_CUDA(cudaStreamSynchronize(c->stream));
    }
    auto next_span = [&](char kind) -> unsigned long long * {
        if (!spanbuf.p || span_n >= span_max) return nullptr;
        span_kind.push_back(kind);
        return (unsigned long long *)spanbuf.p + 2 * (span_n++);
    };
    auto dump_timeline = [&]() {
        if (!spanbuf.p || span_n == 0) return;
        std::vector<unsigned long long> hh((size_t)2 * span_n);
        cudaStreamSynchronize(c->stream);
        if (cudaMemcpy(hh.data(), spanbuf.p, hh.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost) != cudaSuccess) return;
        const unsigned long long t0 = hh[0];
        fprintf(stderr, "[plsb] cv timeline (us since the first launch's first CTA): kind start end\n");
        for (int i = 0; i < span_n; ++i)
            fprintf(stderr, "[plsb]   %c %10.1f %10.1f\n", span_kind[i], (double)(hh[2 * i] - t0) / 1e3, (double)(hh[2 * i + 1] - t0) / 1e3);
    };
    CvBuf ctabuf(c);
    int cta_a = -1, cta_traced = 0, cta_nf = 0, cta_which = 1;     // which launch with that component: 0 = lane 0's, 1 = lane 1's, ...
    if (const char *we = getenv("PLSB_CV_CTA_TRACE_WHICH")) cta_which = atoi(we);
    if (const char *te = getenv("PLSB_CV_CTA_TRACE")) {
        cta_a = atoi(te);
        PLSB_TRY(ctabuf.alloc((size_t)3 * 8192 * sizeof(long long)));
        PLSB_CUDA(cudaMemsetAsync(ctabuf.p, 0, (size_t)3 * 8192 * sizeof(long long), c->stream));
    }
    auto dump_cta_trace = [&]() {
        if (!ctabuf.p || cta_nf == 0) return;
        std::vector<long long> hh((size_t)3 * cta_nf);
        cudaStreamSynchronize(c->stream);
        if (cudaMemcpy(hh.data(), ctabuf.p, hh.size() * sizeof(long long), cudaMemcpyDeviceToHost) != cudaSuccess) return;
        const char *path = getenv("PLSB_CV_CTA_TRACE_FILE");
        FILE *fp = fopen(path ? path : "cta_trace.csv", "w");
        if (!fp) return;
        fprintf(fp, "cta,smid,start_ns,end_ns\n");
        for (int i = 0; i < cta_nf; ++i) fprintf(fp, "%d,%lld,%lld,%lld\n", i, hh[3 * i], hh[3 * i + 1], hh[3 * i + 2]);
        fclose(fp);
    };
    auto launch_fold = [&](cudaStream_t st, int a, int nf) -> int {
        v.trace = (a == trace_a && a >= 0 && traced++ == 1) ? (long long *)tracebuf.p : nullptr;   // the second wave's launch
        v.cta_trace = nullptr;
        v.span = (a >= 0 && fused) ? next_span('F') : nullptr;
        if (a == cta_a && a >= 0 && cta_traced++ == cta_which && nf <= 8192) { v.cta_trace = (long long *)ctabuf.p; cta_nf = nf; }
        if (a >= 0 && fused) {
            fused_kern<<<nf, CV_FNT, fsm, st>>>(v);
        } else if (M > PLSB_MAX_M) {
            cv_fold_kernel<true><<<nf, CV_NT, wide_jsm, st>>>(v);
        } else {
            cv_fold_kernel<false><<<nf, CV_NT, 0, st>>>(v);
        }
        c->launches++;
        PLSB_CUDA(cudaGetLastError());
        return PLSB_OK;
    };
    auto bind = [&](Lane &L, int64_t f0) {
        v.S = L.S.d(); v.P = L.P.d(); v.R = L.R.d(); v.w = L.w.d(); v.Gr = L.Gr.d(); v.U = L.U.d(); v.f0 = (int)f0;
        v.Rt = L.Rt.d(); v.nkt = nkt; v.wide = L.wide.d();
    };

    if (lanes == 1) {
        Lane &L = lane0;
        for (int64_t f0 = 0; f0 < nfolds_local; f0 += F) {
            const int nf = (int)(nfolds_local - f0 < F ? nfolds_local - f0 : F);
            bind(L, f0);
            for (int a = -1; a < (int)A; ++a) {
                v.a = a;
                if (a >= 0)  // Gr[f][:] = (X'X) r_f for the whole wave: one DMMA GEMM
                    PLSB_TRY(launch_gemm_atb(c, Gfull, ldg, L.R.d() + (size_t)a * ldk, A * ldk, K, (int)K, nf, L.Gr.d(), ldk,
                                             false, true, true, PLSB_PROF_CV_GEMM, false, nullptr, /*no_split=*/true));
                ProfScope ps(c, PLSB_PROF_CV_FOLD);
                PLSB_TRY(launch_fold(c->stream, a, nf));
            }
        }
        { dump_trace(); dump_cta_trace(); dump_timeline(); return finish(); }
    }

    // ---- two lanes ----------------------------------------------------------------------------------------------
    struct Streams {
        cudaStream_t g = nullptr, f = nullptr;
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        bool own = true;
        ~Streams() {
            if (g && own) { cudaStreamSynchronize(g); cudaStreamDestroy(g); }
            if (f && own) { cudaStreamSynchronize(f); cudaStreamDestroy(f); }
            if (e0) cudaEventDestroy(e0);
            if (e1) cudaEventDestroy(e1);
        }
    } sx;
    struct EvGuard { Lane **ln; ~EvGuard() { for (int l = 0; l < 2; ++l) { if (ln[l]->evG) cudaEventDestroy(ln[l]->evG); if (ln[l]->evF) cudaEventDestroy(ln[l]->evF); } } } evguard{ln};
    {
        int lo = 0, hi = 0;
        PLSB_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));   // lo = least, hi = greatest (numerically lower)
        const char *pe = getenv("PLSB_CV_PRIO");
        const bool prio = !(pe && *pe == '0');
        const char *se = getenv("PLSB_CV_SERIAL");   // measurement aid: the same kernels and buffers, one after the other
        if (se && *se == '1') {
            sx.own = false; sx.g = sx.f = c->stream;
        } else {
            PLSB_CUDA(cudaStreamCreateWithPriority(&sx.g, cudaStreamNonBlocking, prio ? hi : 0));
            PLSB_CUDA(cudaStreamCreateWithPriority(&sx.f, cudaStreamNonBlocking, prio ? lo : 0));
        }
        PLSB_CUDA(cudaEventCreateWithFlags(&sx.e0, cudaEventDisableTiming));
        PLSB_CUDA(cudaEventCreateWithFlags(&sx.e1, cudaEventDisableTiming));
        for (int l = 0; l < 2; ++l) {
            PLSB_CUDA(cudaEventCreateWithFlags(&ln[l]->evG, cudaEventDisableTiming));
            PLSB_CUDA(cudaEventCreateWithFlags(&ln[l]->evF, cudaEventDisableTiming));
        }
    }
    // the product kernel's ring depth: 2 stages (141 KB) next to a fold kernel that brings its own 64 KB ring, else 3
    int gemm_stages = v.ring ? 2 : 3;
    if (const char *e = getenv("PLSB_CVG_STAGES")) { const int n = atoi(e); if (n == 2 || n == 3) gemm_stages = n; }
    // everything queued so far (allocations, index upload, the cross products) happens before the lanes start
    PLSB_CUDA(cudaEventRecord(sx.e0, c->stream));
    PLSB_CUDA(cudaStreamWaitEvent(sx.g, sx.e0, 0));
    PLSB_CUDA(cudaStreamWaitEvent(sx.f, sx.e0, 0));
    for (int64_t f0 = 0; f0 < nfolds_local; f0 += 2 * F) {
        int nfl[2];
        for (int l = 0; l < 2; ++l) {
            const int64_t left = nfolds_local - (f0 + l * F);
            nfl[l] = (int)(left < 0 ? 0 : (left < F ? left : F));
        }
        for (int a = -1; a < (int)A; ++a) {
            v.a = a;
            for (int l = 0; l < 2; ++l) {
                if (nfl[l] == 0) continue;
                Lane &L = *ln[l];
                bind(L, f0 + l * F);
                if (a >= 0) {
                    PLSB_CUDA(cudaStreamWaitEvent(sx.g, L.evF, 0));            // r_a of this lane is ready
                    {
                        ProfScope ps(c, PLSB_PROF_CV_GEMM, sx.g);
                        PLSB_TRY(launch_cv_gemm(c, sx.g, Gt.d(), L.Rt.d(), K, ldk, nfl[l], L.Gr.d(), gemm_stages, gemm_ctas, next_span('G')));
                    }
                    PLSB_CUDA(cudaEventRecord(L.evG, sx.g));
                    PLSB_CUDA(cudaStreamWaitEvent(sx.f, L.evG, 0));
                }
                {
                    ProfScope ps(c, PLSB_PROF_CV_FOLD, sx.f);
                    PLSB_TRY(launch_fold(sx.f, a, nfl[l]));
                }
                PLSB_CUDA(cudaEventRecord(L.evF, sx.f));
            }
        }
    }
    // the context's stream continues (PRESS, D2H of the residual block, frees) once both streams have drained
    PLSB_CUDA(cudaEventRecord(sx.e0, sx.g));
    PLSB_CUDA(cudaEventRecord(sx.e1, sx.f));
    PLSB_CUDA(cudaStreamWaitEvent(c->stream, sx.e0, 0));
    PLSB_CUDA(cudaStreamWaitEvent(c->stream, sx.e1, 0));
    { dump_trace(); dump_cta_trace(); dump_timeline(); return finish(); }
}

extern "C" int plsb_cv_folds(plsb_dataset_t ds, int64_t A, int method, int64_t nfolds_total, int64_t test_size,
                             const int64_t *test_idx, int64_t fold0, int64_t nfolds_local, double *E) {
    PLSB_ARG(E, "null argument");
    return plsb_cv_folds_press(ds, A, method, nfolds_total, test_size, test_idx, fold0, nfolds_local, E, nullptr);
}
