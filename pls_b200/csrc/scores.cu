// pls_b200/csrc/scores.cu — S = X R on the FP64 tensor cores: the scores of PLS::Model for all components at once.
//
// Replaces X_new * R.leftCols(comp) (reference src/pls.cpp:439-442) and, through it, the X-sized part of
// fitted_values / residuals / SSE / cv_NEW_DATA (pls.cpp:449-459, 494-510: the reference re-reads X_new once per
// component count).  Also produces T = X R after a fit driven from the cross-product matrices (fit.cu).
//
// A tall-skinny contraction: 128 rows x up to 32 components per CTA, the K axis streamed in 32-column chunks through
// a cp.async ring (X is read once per 32 components; R's chunk, 8 KB, comes from L2).  8 warps, each 16 rows x
// 8*NT components, mma.sync.m8n8k4.f64 (SASS DMMA).  Both tiles sit in shared memory at a row stride of 36 doubles
// ((36 mod 16) == 4: the 4 k-columns x 8 rows of a fragment fall into 32 distinct banks).
// Arithmetic intensity at 32 components is 8 flop/B, past the FP64 ridge: tensor-bound (2*N*K*A flops), and within
// ~1.7x of the one-read HBM time.
#include "internal.h"

namespace plsb {

constexpr int SC_BM = 128, SC_BK = 32, SC_LD = 36, SC_NT = 256, SC_ST = 2;

__device__ __forceinline__ void sc_cp_async16(void *dst, const void *src, bool pred) {
    const int sz = pred ? 16 : 0;   // src-size 0: the 16 bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void sc_dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// S[(a0 + n)*lds + i] = sum_k X[i*ldk + k] * R[(a0 + n)*ldk + k],  n < na <= 8*NTILES
template <int NTILES>
__global__ void __launch_bounds__(SC_NT, 2) scores_dmma_kernel(const double *__restrict__ X, int64_t N, int ldk,
                                                                const double *__restrict__ R, int a0, int na,
                                                                double *__restrict__ S, int64_t lds) {
    extern __shared__ __align__(128) char smem[];
    double *As = reinterpret_cast<double *>(smem);                 // [ST][BM][LD]
    double *Bs = As + (size_t)SC_ST * SC_BM * SC_LD;               // [ST][32][LD]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t row_base = (int64_t)blockIdx.x * SC_BM;
    const int nk = (ldk + SC_BK - 1) / SC_BK;

    auto load_stage = [&](int st, int kt) {
        const int k0 = kt * SC_BK;
#pragma unroll
        for (int i = 0; i < (SC_BM * SC_BK / 2) / SC_NT; ++i) {
            const int ch = tid + i * SC_NT, row = ch / (SC_BK / 2), cc = (ch % (SC_BK / 2)) * 2;
            const bool p = row_base + row < N && k0 + cc < ldk;
            sc_cp_async16(As + ((size_t)st * SC_BM + row) * SC_LD + cc, p ? X + (row_base + row) * ldk + k0 + cc : X, p);
        }
#pragma unroll
        for (int i = 0; i < (8 * NTILES * SC_BK / 2 + SC_NT - 1) / SC_NT; ++i) {
            const int ch = tid + i * SC_NT, n = ch / (SC_BK / 2), cc = (ch % (SC_BK / 2)) * 2;
            if (n < 8 * NTILES) {
                const bool p = n < na && k0 + cc < ldk;
                sc_cp_async16(Bs + ((size_t)st * 32 + n) * SC_LD + cc, p ? R + (size_t)(a0 + n) * ldk + k0 + cc : R, p);
            }
        }
    };

    double acc[2][NTILES][2];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < NTILES; ++j) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }

    load_stage(0, 0);
    asm volatile("cp.async.commit_group;" ::: "memory");
    const int wm = warp * 16, fk = lane & 3, fr = lane >> 2;
    for (int kt = 0; kt < nk; ++kt) {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();                       // tile kt has landed; everyone is done with tile kt-1
        if (kt + 1 < nk) load_stage((kt + 1) % SC_ST, kt + 1);
        asm volatile("cp.async.commit_group;" ::: "memory");
        const double *as = As + (size_t)(kt % SC_ST) * SC_BM * SC_LD;
        const double *bs = Bs + (size_t)(kt % SC_ST) * 32 * SC_LD;
#pragma unroll
        for (int kk = 0; kk < SC_BK / 4; ++kk) {
            double a[2], b[NTILES];
#pragma unroll
            for (int i = 0; i < 2; ++i) a[i] = as[(wm + i * 8 + fr) * SC_LD + kk * 4 + fk];
#pragma unroll
            for (int j = 0; j < NTILES; ++j) b[j] = bs[(j * 8 + fr) * SC_LD + kk * 4 + fk];
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < NTILES; ++j) sc_dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
    }
    // D fragment: row fr of the 8-row tile, components 2*fk and 2*fk + 1 of the 8-component tile
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int64_t row = row_base + wm + i * 8 + fr;
        if (row >= N) continue;
#pragma unroll
        for (int j = 0; j < NTILES; ++j) {
            const int n = j * 8 + 2 * fk;
            if (n < na) S[(size_t)(a0 + n) * lds + row] = acc[i][j][0];
            if (n + 1 < na) S[(size_t)(a0 + n + 1) * lds + row] = acc[i][j][1];
        }
    }
}

template <int NTILES>
static int run_scores(plsb_ctx *c, const double *X, int64_t N, int ldk, const double *R, int a0, int na, double *S,
                      int64_t lds) {
    const size_t smem = (size_t)SC_ST * (SC_BM + 32) * SC_LD * sizeof(double);
    auto kern = scores_dmma_kernel<NTILES>;
    PLSB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    {
        ProfScope ps(c, PLSB_PROF_SCORES);
        kern<<<(unsigned)ceil_div(N, SC_BM), SC_NT, smem, c->stream>>>(X, N, ldk, R, a0, na, S, lds);
    }
    PLSB_CUDA(cudaGetLastError());
    c->launches += 1;
    return PLSB_OK;
}

// X: row-major N x ldk (ldk even, zero padded), R: component-major A x ldk, S: component-major A x lds (lds >= N)
int launch_scores(plsb_ctx *c, const double *X, int64_t N, int ldk, const double *R, int A, double *S, int64_t lds) {
    if (N == 0 || A == 0) return PLSB_OK;
    for (int a0 = 0; a0 < A; a0 += 32) {
        const int na = A - a0 < 32 ? A - a0 : 32;
        if (na <= 8) PLSB_TRY(run_scores<1>(c, X, N, ldk, R, a0, na, S, lds));
        else if (na <= 16) PLSB_TRY(run_scores<2>(c, X, N, ldk, R, a0, na, S, lds));
        else if (na <= 24) PLSB_TRY(run_scores<3>(c, X, N, ldk, R, a0, na, S, lds));
        else PLSB_TRY(run_scores<4>(c, X, N, ldk, R, a0, na, S, lds));
    }
    return PLSB_OK;
}

}  // namespace plsb
