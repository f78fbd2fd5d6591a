// pls_b200/csrc/step.cu — the serial, K-sized part of one PLS component, done by ONE CTA.
//
// Replaces (reference src/pls.cpp): the loadings / deflation block 427-434 for the component whose
// pass just finished, then the weight step 403-411 (incl. find_dominant_ev / dominant_eigenvector,
// 113-141, on the M x M matrix XY'XY) and the r-orthogonalisation 412-416 for the next component.
// Also here: the KERNEL_TYPE2 product tt = r'XX r, p = XX r (pls.cpp:423-424) as a multi-CTA SYMV.
#include "internal.h"
#include "step_core.cuh"

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

int launch_step(plsb_ctx *c, const StepArgs &s) {
    if (s.M > PLSB_MAX_M) {
        set_error("M=%d responses exceed the in-kernel eigen solver limit (%d)", s.M, PLSB_MAX_M);
        return PLSB_ERR_UNSUPPORTED;
    }
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
