// pls_b200/csrc/cv_gemm.cu — the batched-fold product of the cross-validation engine, Gr[f][:] = (X'X) r_f for a whole
// wave of folds (cv.cu), as a persistent FP64 DMMA kernel built to SHARE each SM with the HBM-bound fold kernel of the
// other lane.
//
// The product is the per-fold, per-component  XX * r  of the reference's KERNEL_TYPE2 recurrence (src/pls.cpp:423-424)
// applied to every fold's current rotation vector at once (the folds of cv_LOO / cv_LSO, pls.cpp:478-489, 533-546).
//
// Why not the general kernel (gemm.cu)?  Its producer warp makes the CTA 9 warps = the whole register file, and its
// row-by-row bulk copies (160 per k-tile) cost the issuing warp ~6000 cycles when a consumer warp has to issue them
// itself (measured: tensor pipe 46 % active, 1.85x slower).  Here both operands are kept PRE-TILED in HBM, padded exactly
// as the fragment loads want them in shared memory:
//      Gt : [m-tile][k-tile][32][132]   X'X, made once per call (symmetric: row k / column m either way)
//      Rt : [n-tile][k-tile][128][36]   the folds' current r vectors, written in this layout by the fold kernels
// so one stage of the ring is TWO bulk copies (33 792 + 36 864 bytes) issued by one lane.  No producer warp and no
// `empty` barriers: the consumer warps count themselves out of a stage with a shared-memory atomic, and the LAST one
// out refills it on the spot with the k-tile three ahead — nobody ever waits for a free stage.  8 warps x 168 registers
// and 212 KB of shared memory leave 22 K registers and 14 KB per SM: one 256-thread fold CTA (64 registers, 3 KB).
// One CTA per SM, looping over the (m-tile, n-tile) items: all CTAs are resident from the start, which is what lets the
// block scheduler place the other stream's CTAs in the left-over resources.  No split-K: the sums do not depend on the
// wave size.
#include "internal.h"

namespace plsb {

constexpr int CG_BM = 128, CG_BN = 128, CG_BK = 32, CG_LDA = 132, CG_LDB = 36, CG_CW = 8;
constexpr int CG_A_BYTES = CG_BK * CG_LDA * 8;      // 33 792
constexpr int CG_B_BYTES = CG_BN * CG_LDB * 8;      // 36 864
constexpr int CG_STAGE = CG_A_BYTES + CG_B_BYTES;

size_t cv_gemm_gt_doubles(int64_t K) { return (size_t)ceil_div(K, CG_BM) * (size_t)ceil_div(K, CG_BK) * (CG_BK * CG_LDA); }
size_t cv_gemm_rt_doubles(int64_t K, int64_t F) { return (size_t)ceil_div(F, CG_BN) * (size_t)ceil_div(K, CG_BK) * (CG_BN * CG_LDB); }

static_assert(CG_BN == CVG_BN && CG_BK == CVG_BK && CG_LDB == CVG_LDB, "internal.h describes the Rt layout");

// Gt <- G (K x K, ld = ldg), zero padded to whole tiles
__global__ void __launch_bounds__(256) cv_gemm_tile_g_kernel(const double *__restrict__ G, int64_t ldg, int K, int nkt,
                                                             double *__restrict__ Gt) {
    const int mt = blockIdx.y, kt = blockIdx.x;
    double *dst = Gt + ((size_t)mt * nkt + kt) * (CG_BK * CG_LDA);
    for (int e = threadIdx.x; e < CG_BK * CG_LDA; e += 256) {
        const int kr = e / CG_LDA, mc = e % CG_LDA;
        const int k = kt * CG_BK + kr, m = mt * CG_BM + mc;
        dst[e] = (mc < CG_BM && k < K && m < K) ? G[(size_t)k * ldg + m] : 0.0;
    }
}

struct CvGemmArgs {
    const double *Gt, *Rt;
    double *Gr;            // [fold][ldk]
    int K, ldk, nf, nkt, nmt, nnt;
    unsigned long long *span;   // dev tool (PLSB_CV_TIMELINE): earliest CTA start / latest CTA end of this launch, or null
};

__device__ __forceinline__ void cg_dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// CG_STAGES: 3 (212 KB of shared memory), or 2 (141 KB) when the fold kernel of the other lane brings its own 64 KB ring
template <int CG_STAGES>
__global__ void __maxnreg__(168) cv_gemm_kernel(const CvGemmArgs g) {
    extern __shared__ __align__(128) char smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)CG_STAGES * CG_STAGE);
    int *left = reinterpret_cast<int *>(full + CG_STAGES);      // consumer warps that have not yet left the stage
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (g.span && tid == 0) {
        unsigned long long t_;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
        atomicMin(g.span, t_);
    }
    const int nitems = g.nmt * g.nnt;
    const int mine = blockIdx.x < nitems ? (nitems - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    const int total = mine * g.nkt;                              // k-tiles this CTA consumes

    // k-tile p of this CTA (all items are nkt deep): item blockIdx.x + (p / nkt) * gridDim.x, n-tile fastest
    auto issue = [&](int p) {   // one lane
        const int s = p % CG_STAGES;
        const int item = blockIdx.x + (p / g.nkt) * gridDim.x, kt = p % g.nkt;
        const int nt = item % g.nnt, mt = item / g.nnt;
        char *sa = smem + (size_t)s * CG_STAGE;
        mbar_arrive_expect_tx(&full[s], (uint32_t)CG_STAGE);
        bulk_g2s(sa, g.Gt + ((size_t)mt * g.nkt + kt) * (CG_BK * CG_LDA), (uint32_t)CG_A_BYTES, &full[s]);
        bulk_g2s(sa + CG_A_BYTES, g.Rt + ((size_t)nt * g.nkt + kt) * (CG_BN * CG_LDB), (uint32_t)CG_B_BYTES, &full[s]);
    };
    if (tid == 0) {
        for (int s = 0; s < CG_STAGES; ++s) { mbar_init(&full[s], 1); left[s] = CG_CW; }
        mbar_fence_init();
        for (int p = 0; p < CG_STAGES && p < total; ++p) issue(p);
    }
    __syncthreads();

    const int wm = (warp >> 2) * 64, wn = (warp & 3) * 32;
    const int fk = lane & 3, fr = lane >> 2;
    int gk = 0;
    for (int it = 0; it < mine; ++it) {
        double acc[8][4][2];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }
        for (int kt = 0; kt < g.nkt; ++kt, ++gk) {
            const int s = gk % CG_STAGES;
            mbar_wait(&full[s], (uint32_t)((gk / CG_STAGES) & 1));
            const double *as = reinterpret_cast<const double *>(smem + (size_t)s * CG_STAGE) + fk * CG_LDA + wm + fr;
            const double *bs = reinterpret_cast<const double *>(smem + (size_t)s * CG_STAGE + CG_A_BYTES) + (wn + fr) * CG_LDB + fk;
#pragma unroll
            for (int kk = 0; kk < CG_BK / 4; ++kk) {
                double a[8], b[4];
#pragma unroll
                for (int i = 0; i < 8; ++i) a[i] = as[kk * 4 * CG_LDA + i * 8];
#pragma unroll
                for (int j = 0; j < 4; ++j) b[j] = bs[j * 8 * CG_LDB + kk * 4];
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) cg_dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
            }
            // this warp is done with stage s (its fragment loads have all been consumed by the DMMAs above); the last
            // warp out refills the stage with the k-tile three ahead
            __syncwarp();
            if (lane == 0) {
                __threadfence_block();
                if (atomicSub(&left[s], 1) == 1) {
                    left[s] = CG_CW;
                    __threadfence_block();
                    if (gk + CG_STAGES < total) issue(gk + CG_STAGES);
                }
            }
        }
        // Gr[f][m] for this item's 128 x 128 tile (rows of the output are folds: "transposed" store)
        const int item = blockIdx.x + it * gridDim.x;
        const int n_base = (item % g.nnt) * CG_BN, m_base = (item / g.nnt) * CG_BM;
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int m = m_base + wm + i * 8 + fr, n = n_base + wn + j * 8 + fk * 2;
                if (m >= g.K) continue;
                if (n < g.nf) g.Gr[(int64_t)n * g.ldk + m] = acc[i][j][0];
                if (n + 1 < g.nf) g.Gr[(int64_t)(n + 1) * g.ldk + m] = acc[i][j][1];
            }
    }
    if (g.span && tid == 0) {
        unsigned long long t_;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
        atomicMax(g.span + 1, t_);
    }
}

int launch_cv_gemm_tile_g(plsb_ctx *c, cudaStream_t st, const double *G, int64_t ldg, int64_t K, double *Gt) {
    const int nkt = (int)ceil_div(K, CG_BK), nmt = (int)ceil_div(K, CG_BM);
    cv_gemm_tile_g_kernel<<<dim3((unsigned)nkt, (unsigned)nmt), 256, 0, st>>>(G, ldg, (int)K, nkt, Gt);
    c->launches++;
    PLSB_CUDA(cudaGetLastError());
    return PLSB_OK;
}

int launch_cv_gemm(plsb_ctx *c, cudaStream_t st, const double *Gt, const double *Rt, int64_t K, int64_t ldk, int nf, double *Gr,
                   int stages, int max_ctas, unsigned long long *span) {
    CvGemmArgs g{};
    g.Gt = Gt; g.Rt = Rt; g.Gr = Gr; g.K = (int)K; g.ldk = (int)ldk; g.nf = nf;
    g.span = span;
    g.nkt = (int)ceil_div(K, CG_BK); g.nmt = (int)ceil_div(K, CG_BM); g.nnt = (int)ceil_div(nf, CG_BN);
    const int ns = stages == 2 ? 2 : 3;
    const size_t smem = (size_t)ns * CG_STAGE + ns * sizeof(uint64_t) + ns * sizeof(int) + 16;
    auto kern = ns == 2 ? cv_gemm_kernel<2> : cv_gemm_kernel<3>;
    PLSB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // An SM runs with one shared-memory carve-out at a time and keeps it while anything is resident.  The 2-stage kernel
    // (141 KB) would be given the 164 KB carve-out, and a fold CTA of the other lane (69 KB) could then not join it on an
    // SM it reached first — seen in the launch timeline (PLSB_CV_TIMELINE) as fold launches that only started once the
    // product's CTAs were leaving.  Ask for the maximum, like the fold kernels do.
    PLSB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    const int nitems = g.nmt * g.nnt;
    int grid = nitems < c->sm_count ? nitems : c->sm_count;
    if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;     // SMs left entirely to the other lane's fold kernel (cv.cu)
    kern<<<grid, CG_CW * 32, smem, st>>>(g);
    c->launches++;
    PLSB_CUDA(cudaGetLastError());
    return PLSB_OK;
}

}  // namespace plsb
