// pls_b200/csrc/step.cu — the serial, K-sized part of one PLS component, done by ONE CTA.
//
// Replaces (reference src/pls.cpp): the loadings / deflation block 427-434 for the component whose
// pass just finished, then the weight step 403-411 (incl. find_dominant_ev / dominant_eigenvector,
// 113-141, on the M x M matrix XY'XY) and the r-orthogonalisation 412-416 for the next component.
// Also here: the KERNEL_TYPE2 product tt = r'XX r, p = XX r (pls.cpp:423-424) as a multi-CTA SYMV.
#include <cooperative_groups.h>

#include "internal.h"
#include "step_cluster.cuh"
#include "step_core.cuh"
#include "step_wide.cuh"

namespace plsb {

constexpr int STEP_NT = 1024;

__global__ void __launch_bounds__(STEP_NT) step_kernel(const StepArgs s) {
    __shared__ StepSmem<STEP_NT> sh;
    const int ldk = s.ldk;
    if (s.a >= 0)
        finish_component<STEP_NT>(s.K, ldk, s.M, s.S, s.g, s.g[ldk], s.R + (size_t)s.a * ldk, s.P + (size_t)s.a * ldk,
                                  s.Q + (size_t)s.a * s.M, sh);
    const int an = s.a + 1;
    if (an >= s.A) return;
    next_weights<STEP_NT>(s.K, ldk, s.M, an, s.S, s.P, s.R, s.W + (size_t)an * ldk, s.R + (size_t)an * ldk, sh);
}

// Wide response blocks (M > 32): the general path of step_wide.cuh, one CTA, M-sized objects in global scratch.
__global__ void __launch_bounds__(STEP_NT) step_wide_kernel(const StepArgs s, const int jac_in_smem) {
    __shared__ WideSmem<STEP_NT> sh;
    extern __shared__ __align__(16) double wide_dyn[];
    const int ldk = s.ldk;
    if (s.a >= 0)
        finish_component_wide<STEP_NT>(s.K, ldk, s.M, s.S, s.g, s.g[ldk], s.R + (size_t)s.a * ldk, s.P + (size_t)s.a * ldk,
                                       s.Q + (size_t)s.a * s.M, s.wide);
    const int an = s.a + 1;
    if (an >= s.A) return;
    next_weights_wide<STEP_NT>(s.K, ldk, s.M, an, s.S, s.P, s.R, s.W + (size_t)an * ldk, s.R + (size_t)an * ldk, s.wide, sh,
                               jac_in_smem ? wide_dyn : nullptr);
}
size_t step_wide_scratch_bytes(int M) { return M > PLSB_MAX_M ? wide_scratch_doubles(M) * sizeof(double) : 0; }

// ------------------------------------------------------------------------------------------------
// The step as a thread-block CLUSTER: 8 CTAs on 8 SMs, each owning a slice of the K axis.
//
// The step is a dependency chain between two streaming passes (and the only thing between them that does
// not shrink when rows are sharded over GPUs), so it is built for latency: one SM reading the ~1 MB of
// P and R for the orthogonalisation, or shuffling 36 Gram entries across 32 warps, costs tens of us
// (measured with clock64: Gram 14 us, eigen solver 59 us, orthogonalisation up to 29 us in a 1-CTA kernel).
// Here every CTA keeps its slice of S in shared memory, dots are reduced warp -> CTA -> cluster through
// distributed shared memory (each CTA reads the other seven's partials after one cluster barrier, in rank
// order: deterministic and identical on every CTA), and the tiny eigen problem is solved redundantly by
// warp 0 of every CTA so no broadcast is needed.
// ------------------------------------------------------------------------------------------------
namespace cg = cooperative_groups;
constexpr int SC_NCTA = 8, SC_NT = 256, SC_KPT = 2;  // K <= 8 * 256 * 2 = 4096
constexpr int SC_CH = 16;  // values reduced per CTA-stage round (staging area: SC_CH x 256 doubles)

struct StepClLay {
    int cpc;      // columns per CTA
    int nvs;      // capacity of one reduction (values)
    size_t bytes;
};
static StepClLay step_cl_layout(int ldk, int M) {
    StepClLay l;
    l.cpc = (int)ceil_div(ldk, SC_NCTA);
    int nv = M * (M + 1) / 2;
    if (nv < 64) nv = 64;
    l.nvs = (int)round_up(nv, 32);
    const int npairs = M * (M + 1) / 2;
    l.bytes = sizeof(double) * ((size_t)M * l.cpc + (size_t)SC_CH * SC_NT + (2 * SC_NCTA + 1) * (size_t)l.nvs + (size_t)M * M + 2048 + M + 8) +
              (size_t)round_up(2 * npairs, 16);
    return l;
}

struct StepClSmem {
    double *S, *red, *gather, *tot, *G, *B, *C, *q;
    uint64_t *bar;          // two mbarriers (one per exchange parity)
    unsigned char *pu, *pv;
    int nvs, phase, rank;
    long long *trace;       // dev tool: timestamps inside exchange number 3 (the P'w dots), or null
};

__device__ __forceinline__ uint32_t cluster_map(const void *smem_ptr, int cta_rank) {
    uint32_t out;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(out) : "r"(smem_u32(smem_ptr)), "r"(cta_rank));
    return out;
}
// asynchronous 8-byte store into another CTA's shared memory; the destination CTA's mbarrier counts the bytes
__device__ __forceinline__ void st_async_f64(uint32_t remote_addr, double v, uint32_t remote_bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(remote_addr),
                 "l"(__double_as_longlong(v)), "r"(remote_bar)
                 : "memory");
}
__device__ __forceinline__ void cluster_barrier() {
    asm volatile("barrier.cluster.arrive.aligned;\n\tbarrier.cluster.wait.aligned;" ::: "memory");
}

__device__ __forceinline__ long long sc_now() {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// CTA-level then cluster-level sum of nv per-thread values; the totals land in sm.tot on every CTA.
// Exchange: every CTA pushes its nv partial sums into a slot of every CTA's `gather` array with st.async; the
// bytes are counted by the destination's mbarrier, so one wait replaces a cluster-wide barrier (whose release
// fence, MEMBAR.ALL.GPU in SASS, would also wait for this kernel's outstanding global stores each time).
// Slots and barriers alternate with the exchange parity: traffic of exchange n+2 can only start after every
// CTA has finished reading exchange n (it has sent n+1 since), so two sets suffice.  Sums run in rank order.
template <typename F>
__device__ __forceinline__ void cl_reduce(StepClSmem &sm, int nv, F val) {
    const int tid = threadIdx.x;
    const int par = sm.phase & 1;
    uint64_t *bar = sm.bar + par;
    double *gather = sm.gather + (size_t)par * SC_NCTA * sm.nvs;
    const bool tr = sm.trace != nullptr && sm.phase == 3 && tid == 0 && blockIdx.x == 0;
    if (tr) sm.trace[11] = sc_now();
    if (tid == 0) mbar_arrive_expect_tx(bar, (uint32_t)(nv * sizeof(double) * SC_NCTA));
    // CTA stage, SC_CH values at a time: every thread parks its partial products in shared memory, then 16 threads
    // per value add the 256 entries in a fixed order.  Deliberately a small LOOP, not an unrolled register
    // butterfly: this kernel runs once per component right after a pass has flushed every cache, so each executed
    // instruction is a cold instruction-cache miss — the unrolled 32-value version (5k instructions) cost 7 us of
    // pure instruction fetch; this one is a few hundred instructions.
    const int g = tid >> 4, l = tid & 15;
    for (int base = 0; base < nv; base += SC_CH) {
        const int n = nv - base < SC_CH ? nv - base : SC_CH;
        double pv[SC_CH];
#pragma unroll
        for (int i = 0; i < SC_CH; ++i) pv[i] = val(base + (i < n ? i : n - 1));   // unconditional: loads stay batched
#pragma unroll
        for (int i = 0; i < SC_CH; ++i) sm.red[i * SC_NT + tid] = pv[i];
        __syncthreads();
        {   // every thread takes part (full-warp shuffles); groups past n re-add the last value and discard it
            const double *row = sm.red + (g < n ? g : n - 1) * SC_NT + l;
            double a0 = 0.0, a1 = 0.0;
#pragma unroll
            for (int u = 0; u < 16; u += 2) {
                a0 += row[16 * u];
                a1 += row[16 * (u + 1)];
            }
            double acc = a0 + a1;
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (l == 0 && g < n) {   // this CTA's total of value base+g: push it to every CTA of the cluster
                double *slot = gather + (size_t)sm.rank * sm.nvs + base + g;
#pragma unroll
                for (int c = 0; c < SC_NCTA; ++c) st_async_f64(cluster_map(slot, c), acc, cluster_map(bar, c));
            }
        }
        __syncthreads();
    }
    if (tr) sm.trace[14] = sc_now();
    mbar_wait(bar, (uint32_t)((sm.phase >> 1) & 1));
    if (tr) sm.trace[15] = sc_now();
    for (int v = tid; v < nv; v += SC_NT) {
        double t = 0.0;
#pragma unroll
        for (int c = 0; c < SC_NCTA; ++c) t += gather[(size_t)c * sm.nvs + v];
        sm.tot[v] = t;
    }
    __syncthreads();
    sm.phase++;
}

#define SC_TRACE(i) do { if (s.trace && threadIdx.x == 0 && blockIdx.x == 0) s.trace[i] = sc_now(); } while (0)
__global__ void __launch_bounds__(SC_NT, 1) step_cluster_kernel(const StepArgs s, const int cpc, const int nvs) {
    extern __shared__ __align__(16) double sc_smem[];
    cg::cluster_group cluster = cg::this_cluster();
    const int K = s.K, ldk = s.ldk, M = s.M;
    const int npairs = M * (M + 1) / 2;
    StepClSmem sm;
    sm.S = sc_smem;                              // M x cpc slice of S
    sm.red = sm.S + (size_t)M * cpc;             // [SC_CH][threads] staging of per-thread partial products
    sm.gather = sm.red + (size_t)SC_CH * SC_NT;  // [2][ctas][nvs] partial sums pushed here by every CTA
    sm.tot = sm.gather + 2 * (size_t)SC_NCTA * nvs;  // [nvs] cluster totals
    sm.G = sm.tot + nvs; sm.B = sm.G + M * M; sm.C = sm.B + 1024; sm.q = sm.C + 1024;  // B, C: up to 32 x 32 padded
    sm.bar = reinterpret_cast<uint64_t *>(sm.q + M);  // 2 of the 8 spare doubles
    sm.pu = reinterpret_cast<unsigned char *>(sm.q + M + 8); sm.pv = sm.pu + npairs;
    sm.nvs = nvs; sm.phase = 0; sm.trace = s.trace;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int rank = (int)cluster.block_rank();
    sm.rank = rank;
    if (tid == 0) {
        mbar_init(&sm.bar[0], 1);
        mbar_init(&sm.bar[1], 1);
        mbar_fence_init();
    }
    const int c0 = rank * cpc;                   // first column of this CTA
    int kl[SC_KPT], kg[SC_KPT];
    int klc[SC_KPT], kgc[SC_KPT];                // the same, clamped to a valid column when the slot is dead
    bool live[SC_KPT], wr[SC_KPT];               // live: a real column; wr: real or padding (gets zero)
    double lv[SC_KPT];                           // 1.0 / 0.0 mask of live
#pragma unroll
    for (int j = 0; j < SC_KPT; ++j) {
        kl[j] = tid + SC_NT * j;
        kg[j] = c0 + kl[j];
        live[j] = kl[j] < cpc && kg[j] < K;
        wr[j] = kl[j] < cpc && kg[j] < ldk;
        klc[j] = live[j] ? kl[j] : 0;
        kgc[j] = live[j] ? kg[j] : 0;
        lv[j] = live[j] ? 1.0 : 0.0;
    }
    SC_TRACE(0);
    // Launched as a programmatic dependent: the set-up below (pair table, S slice, barrier initialisation) overlaps
    // the tail of the partial-row sum that produces g.  S is written only by the previous step, which completed
    // before that kernel could trigger this launch — except for the initial step, whose S comes straight from
    // the kernel before it.
    if (s.a < 0) pdl_wait();
    for (int d = tid; d < npairs; d += SC_NT) {  // (u, v) of Gram entry d, u <= v
        int u = 0, rem = d;
        while (rem >= M - u) { rem -= M - u; ++u; }
        sm.pu[d] = (unsigned char)u;
        sm.pv[d] = (unsigned char)(u + rem);
    }
#pragma unroll
    for (int j = 0; j < SC_KPT; ++j)
        if (wr[j])
            for (int m = 0; m < M; ++m) sm.S[(size_t)m * cpc + kl[j]] = s.S[(size_t)m * ldk + kg[j]];
    __syncthreads();  // pair table visible (each thread touches only its own columns of sm.S)
    cluster_barrier();  // every CTA's mbarriers are initialised before any remote traffic can reach them
    pdl_wait();
    pdl_launch();
    SC_TRACE(1);
    const int an = s.a + 1;
    double x[SC_KPT];  // r_a, later w
    if (s.a >= 0) {
        const double tt = s.g[ldk];
        double p[SC_KPT];
        double *Pa = s.P + (size_t)s.a * ldk;
        const double *Ra = s.R + (size_t)s.a * ldk;
#pragma unroll
        for (int j = 0; j < SC_KPT; ++j) {
            p[j] = live[j] ? s.g[kg[j]] / tt : 0.0;
            x[j] = live[j] ? Ra[kg[j]] : 0.0;
            if (wr[j]) Pa[kg[j]] = p[j];
        }
        // q = S'r / tt   (pls.cpp:428) with S as it was before this component's deflation
        cl_reduce(sm, M, [&](int m) {
            double acc = 0.0;
#pragma unroll
            for (int j = 0; j < SC_KPT; ++j) acc = fma(sm.S[(size_t)m * cpc + klc[j]], x[j], acc);  // x is 0 on dead slots
            return acc;
        });
        SC_TRACE(2);
        if (rank == 0 && tid < M) s.Q[(size_t)s.a * M + tid] = sm.tot[tid] / tt;
        // S -= tt p q'   (pls.cpp:429)
        for (int m = 0; m < M; ++m) {
            const double qm = sm.tot[m] / tt;
#pragma unroll
            for (int j = 0; j < SC_KPT; ++j)
                if (live[j]) sm.S[(size_t)m * cpc + kl[j]] -= (tt * p[j]) * qm;
        }
    }
    // (a CTA may exit as soon as it is done: it has waited for all traffic addressed to it, and nobody reads
    //  its shared memory — data is pushed, never pulled)
    if (an >= s.A) return;
    SC_TRACE(3);
    // the deflated slice goes back to global memory for the next step; nothing below waits on it
#pragma unroll
    for (int j = 0; j < SC_KPT; ++j)
        if (wr[j] && s.a >= 0)
            for (int m = 0; m < M; ++m) s.S[(size_t)m * ldk + kg[j]] = sm.S[(size_t)m * cpc + kl[j]];
    // ---- weight vector of component an (pls.cpp:403-411) ----
    if (M == 1) {
#pragma unroll
        for (int j = 0; j < SC_KPT; ++j) x[j] = live[j] ? sm.S[kl[j]] : 0.0;
    } else {
        cl_reduce(sm, npairs, [&](int d) {
            const double *su = sm.S + (size_t)sm.pu[d] * cpc, *sv = sm.S + (size_t)sm.pv[d] * cpc;
            double acc = 0.0;
#pragma unroll
            for (int j = 0; j < SC_KPT; ++j) acc = fma(su[klc[j]] * lv[j], sv[klc[j]], acc);
            return acc;
        });
        SC_TRACE(4);
        if (warp == 0) {
            for (int d = tid; d < npairs; d += 32) {
                sm.G[sm.pu[d] * M + sm.pv[d]] = sm.tot[d];
                sm.G[sm.pv[d] * M + sm.pu[d]] = sm.tot[d];
            }
            __syncwarp();
            dominant_eigvec_warp(sm.G, sm.B, sm.C, M, sm.q, (s.trace && blockIdx.x == 0) ? s.trace + 10 : nullptr);
        }
        __syncthreads();
        SC_TRACE(5);
#pragma unroll
        for (int j = 0; j < SC_KPT; ++j) {
            double acc = 0.0;
            if (live[j])
                for (int m = 0; m < M; ++m) acc = fma(sm.S[(size_t)m * cpc + kl[j]], sm.q[m], acc);
            x[j] = acc;
        }
    }
    cl_reduce(sm, 1, [&](int) {
        double acc = 0.0;
#pragma unroll
        for (int j = 0; j < SC_KPT; ++j) acc = fma(x[j], x[j], acc);
        return acc;
    });
    SC_TRACE(6);
    {   // w /= sqrt(w'w)   (pls.cpp:411)
        const double nrm = sqrt(sm.tot[0]);
        double *Wn = s.W + (size_t)an * ldk;
#pragma unroll
        for (int j = 0; j < SC_KPT; ++j) {
            x[j] = live[j] ? x[j] / nrm : 0.0;
            if (wr[j]) Wn[kg[j]] = x[j];
        }
    }
    // ---- r = w - sum_j (p_j'w) r_j   (pls.cpp:412-416), subtraction order j = 0..an-1 ----
    double r[SC_KPT];
#pragma unroll
    for (int j = 0; j < SC_KPT; ++j) r[j] = x[j];
    for (int j0 = 0; j0 < an; j0 += nvs) {
        const int nj = an - j0 < nvs ? an - j0 : nvs;
        cl_reduce(sm, nj, [&](int jj) {
            const double *Pj = s.P + (size_t)(j0 + jj) * ldk;
            double acc = 0.0;
#pragma unroll
            for (int j = 0; j < SC_KPT; ++j) acc = fma(Pj[kgc[j]], x[j], acc);  // x is 0 on dead slots
            return acc;
        });
        SC_TRACE(7);
        int jj = 0;
        for (; jj + 8 <= nj; jj += 8) {  // eight independent loads in flight per column
            const double *Rj = s.R + (size_t)(j0 + jj) * ldk;
#pragma unroll
            for (int j = 0; j < SC_KPT; ++j) {
                double rv[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) rv[u] = Rj[kgc[j] + (size_t)u * ldk];
#pragma unroll
                for (int u = 0; u < 8; ++u) r[j] -= sm.tot[jj + u] * rv[u];
            }
        }
        for (; jj < nj; ++jj) {
            const double *Rj = s.R + (size_t)(j0 + jj) * ldk;
#pragma unroll
            for (int j = 0; j < SC_KPT; ++j) r[j] -= sm.tot[jj] * Rj[kgc[j]];
        }
    }
#pragma unroll
    for (int j = 0; j < SC_KPT; ++j) r[j] *= lv[j];
    double *Rn = s.R + (size_t)an * ldk;
#pragma unroll
    for (int j = 0; j < SC_KPT; ++j) if (wr[j]) Rn[kg[j]] = r[j];
    SC_TRACE(9);
}

static int launch_step_cluster(plsb_ctx *c, const StepArgs &s) {
    const StepClLay lay = step_cl_layout(s.ldk, s.M);
    PLSB_CUDA(cudaFuncSetAttribute(step_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lay.bytes));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(SC_NCTA);
    cfg.blockDim = dim3(SC_NT);
    cfg.dynamicSmemBytes = lay.bytes;
    cfg.stream = c->stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = SC_NCTA;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    if (getenv("PLSB_STEP_TRACE")) {  // dev tool (tools/steptrace.py): phase timestamps of this launch on stderr
        static long long *tr = nullptr;
        if (!tr) PLSB_CUDA(cudaMalloc(&tr, 16 * sizeof(long long)));
        PLSB_CUDA(cudaMemsetAsync(tr, 0, 16 * sizeof(long long), c->stream));
        StepArgs st = s;
        st.trace = tr;
        PLSB_CUDA(cudaLaunchKernelEx(&cfg, step_cluster_kernel, st, lay.cpc, lay.nvs));
        long long h[16];
        PLSB_CUDA(cudaMemcpyAsync(h, tr, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
        PLSB_CUDA(cudaStreamSynchronize(c->stream));
        fprintf(stderr, "step a=%d ns:", s.a);
        for (int i = 1; i < 10; ++i) fprintf(stderr, " %lld", h[i] ? h[i] - h[0] : -1);
        fprintf(stderr, " | x3:");
        for (int i = 14; i < 16; ++i) fprintf(stderr, " %lld", h[i] ? h[i] - h[11] : -1);
        fprintf(stderr, " | squarings %lld", h[10]);
        fprintf(stderr, "\n");
        c->launches++;
        return PLSB_OK;
    }
    {
        ProfScope ps(c, PLSB_PROF_STEP);
        PLSB_CUDA(cudaLaunchKernelEx(&cfg, step_cluster_kernel, s, lay.cpc, lay.nvs));
    }
    c->launches++;
    return PLSB_OK;
}

int launch_step(plsb_ctx *c, const StepArgs &s) {
    if (s.M > PLSB_MAX_M) {   // wide response block: the general one-CTA path
        PLSB_ARG(s.wide, "internal: no scratch for the wide-response step");
        size_t jsm = wide_jacobi_smem_doubles(s.M) * sizeof(double);
        if (jsm + 4096 > c->smem_optin) jsm = 0;                 // too wide for one SM: iterate in the global scratch
        if (jsm) PLSB_CUDA(cudaFuncSetAttribute(step_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)jsm));
        {
            ProfScope ps(c, PLSB_PROF_STEP);
            step_wide_kernel<<<1, STEP_NT, jsm, c->stream>>>(s, jsm ? 1 : 0);
        }
        c->launches++;
        PLSB_CUDA(cudaGetLastError());
        return PLSB_OK;
    }
    const bool force_v1 = getenv("PLSB_STEP_V1") != nullptr;  // A/B switch for tools/devbench.py
    if (!force_v1 && s.ldk <= SC_NCTA * SC_NT * SC_KPT && step_cl_layout(s.ldk, s.M).bytes <= c->smem_optin)
        return launch_step_cluster(c, s);
    {
        ProfScope ps(c, PLSB_PROF_STEP);
        step_kernel<<<1, STEP_NT, 0, c->stream>>>(s);
    }
    c->launches++;
    PLSB_CUDA(cudaGetLastError());
    return PLSB_OK;
}

// ------------------------------------------------------------------------------------------------
// KERNEL_TYPE2: out[k] = (G r)_k, out[ldk] = r'G r.  G symmetric K x K (row k contiguous, ld = ldg).
// One warp per row of G; per-CTA partial of r'Gr, summed in CTA order by the last kernel.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) symv_kernel(const double *__restrict__ G, int64_t ldg, int K,
                                                   const double *__restrict__ r, double *__restrict__ out,
                                                   double *__restrict__ part) {
    __shared__ double sm[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int row = blockIdx.x * 8 + warp;
    double contrib = 0.0;
    if (row < K) {
        const double *g = G + (size_t)row * ldg;
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        int k = lane;
        for (; k + 96 < K; k += 128) {
            s0 = fma(g[k], r[k], s0);
            s1 = fma(g[k + 32], r[k + 32], s1);
            s2 = fma(g[k + 64], r[k + 64], s2);
            s3 = fma(g[k + 96], r[k + 96], s3);
        }
        for (; k < K; k += 32) s0 = fma(g[k], r[k], s0);
        const double s = warp_sum((s0 + s1) + (s2 + s3));
        if (lane == 0) out[row] = s;
        contrib = s * r[row];
    }
    if (lane == 0) sm[warp] = contrib;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < 8; ++i) t += sm[i];
        part[blockIdx.x] = t;
    }
}
__global__ void __launch_bounds__(32) symv_finish_kernel(const double *__restrict__ part, int n, double *out_tt, int K,
                                                         int ldk, double *out) {
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < n; ++i) t += part[i];
        *out_tt = t;
    }
    for (int k = K + threadIdx.x; k < ldk; k += 32) out[k] = 0.0;
}

int launch_symv(plsb_ctx *c, const double *G, int64_t ldg, int K, const double *r, double *out, int ldk) {
    const int nb = (int)ceil_div(K, 8);
    PLSB_TRY(ensure_part(c, (size_t)nb * sizeof(double)));
    symv_kernel<<<nb, 256, 0, c->stream>>>(G, ldg, K, r, out, c->part);
    PLSB_CUDA(cudaGetLastError());
    symv_finish_kernel<<<1, 32, 0, c->stream>>>(c->part, nb, out + ldk, K, ldk, out);
    PLSB_CUDA(cudaGetLastError());
    c->launches += 2;
    return PLSB_OK;
}

}  // namespace plsb
