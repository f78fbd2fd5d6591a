// pls_b200/csrc/step.cu — the serial, K-sized part of one PLS component, done by ONE CTA.
//
// Replaces (reference src/pls.cpp): the loadings / deflation block 427-434 for the component whose
// pass just finished, then the weight step 403-411 (incl. find_dominant_ev / dominant_eigenvector,
// 113-141, on the M x M matrix XY'XY) and the r-orthogonalisation 412-416 for the next component.
// Also here: the KERNEL_TYPE2 product tt = r'XX r, p = XX r (pls.cpp:423-424) as a multi-CTA SYMV.
#include "internal.h"
#include "jacobi.cuh"

namespace plsb {

constexpr int STEP_NT = 1024;
constexpr int STEP_NW = STEP_NT / 32;

// dot products: one warp per dot, lanes stride the K axis (coalesced); result in out[d]
template <typename FA, typename FB>
__device__ __forceinline__ void warp_dots(int ndots, int K, FA a_of, FB b_of, double *out) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int d = warp; d < ndots; d += STEP_NW) {
        const double *pa = a_of(d), *pb = b_of(d);
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        int k = lane;
        for (; k + 96 < K; k += 128) {
            s0 = fma(pa[k], pb[k], s0);
            s1 = fma(pa[k + 32], pb[k + 32], s1);
            s2 = fma(pa[k + 64], pb[k + 64], s2);
            s3 = fma(pa[k + 96], pb[k + 96], s3);
        }
        for (; k < K; k += 32) s0 = fma(pa[k], pb[k], s0);
        const double s = warp_sum((s0 + s1) + (s2 + s3));
        if (lane == 0) out[d] = s;
    }
}

__global__ void __launch_bounds__(STEP_NT) step_kernel(const StepArgs s) {
    __shared__ double sm_q[PLSB_MAX_M];                 // q_a (loadings of Y) / eigenvector
    __shared__ double sm_c[256];                        // p_j' w
    __shared__ double sm_gram[PLSB_MAX_M * PLSB_MAX_M];
    __shared__ double sm_V[PLSB_MAX_M * PLSB_MAX_M];
    __shared__ double sm_red[STEP_NW];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int K = s.K, ldk = s.ldk, M = s.M;

    if (s.a >= 0) {
        const double tt = s.g[ldk];
        const double *r = s.R + (size_t)s.a * ldk;
        double *p = s.P + (size_t)s.a * ldk;
        // p = g / tt   (pls.cpp:421/424 then 427)
        for (int k = tid; k < ldk; k += STEP_NT) p[k] = (k < K) ? s.g[k] / tt : 0.0;
        // q = S' r / tt   (pls.cpp:428)
        warp_dots(M, K, [&](int m) { return s.S + (size_t)m * ldk; }, [&](int) { return r; }, sm_q);
        __syncthreads();
        if (tid < M) {
            sm_q[tid] = sm_q[tid] / tt;
            s.Q[(size_t)s.a * M + tid] = sm_q[tid];
        }
        __syncthreads();
        // S -= tt * p q'   (pls.cpp:429)
        for (int k = tid; k < K; k += STEP_NT) {
            const double tp = tt * p[k];
            for (int m = 0; m < M; ++m) s.S[(size_t)m * ldk + k] -= tp * sm_q[m];
        }
        __syncthreads();
    }
    const int an = s.a + 1;
    if (an >= s.A) return;
    double *w = s.W + (size_t)an * ldk;
    double *rn = s.R + (size_t)an * ldk;

    // weight vector (pls.cpp:403-408)
    if (M == 1) {
        for (int k = tid; k < ldk; k += STEP_NT) w[k] = (k < K) ? s.S[k] : 0.0;
    } else {
        const int npairs = M * (M + 1) / 2;
        // Gram = S'S, upper triangle, one warp per entry
        for (int d = warp; d < npairs; d += STEP_NW) {
            int u = 0, rem = d;
            while (rem >= M - u) { rem -= M - u; ++u; }
            const int v = u + rem;
            const double *pa = s.S + (size_t)u * ldk, *pb = s.S + (size_t)v * ldk;
            double s0 = 0.0, s1 = 0.0;
            int k = lane;
            for (; k + 32 < K; k += 64) {
                s0 = fma(pa[k], pb[k], s0);
                s1 = fma(pa[k + 32], pb[k + 32], s1);
            }
            for (; k < K; k += 32) s0 = fma(pa[k], pb[k], s0);
            const double sum = warp_sum(s0 + s1);
            if (lane == 0) {
                sm_gram[u * M + v] = sum;
                sm_gram[v * M + u] = sum;
            }
        }
        __syncthreads();
        if (warp == 0) jacobi_dominant_warp(sm_gram, sm_V, M, sm_q);
        __syncthreads();
        for (int k = tid; k < ldk; k += STEP_NT) {
            double acc = 0.0;
            if (k < K)
                for (int m = 0; m < M; ++m) acc = fma(s.S[(size_t)m * ldk + k], sm_q[m], acc);
            w[k] = acc;
        }
    }
    __syncthreads();
    // w /= sqrt(w'w)   (pls.cpp:411)
    {
        double acc = 0.0;
        for (int k = tid; k < K; k += STEP_NT) acc = fma(w[k], w[k], acc);
        acc = warp_sum(acc);
        if (lane == 0) sm_red[warp] = acc;
        __syncthreads();
        double tot = 0.0;
        for (int i = 0; i < STEP_NW; ++i) tot += sm_red[i];
        const double nrm = sqrt(tot);
        for (int k = tid; k < K; k += STEP_NT) w[k] = w[k] / nrm;
    }
    __syncthreads();
    // r = w - sum_j (p_j'w) r_j   (pls.cpp:412-416; every dot uses the un-modified w)
    for (int j0 = 0; j0 < an; j0 += 256) {
        const int nj = an - j0 < 256 ? an - j0 : 256;
        warp_dots(nj, K, [&](int j) { return s.P + (size_t)(j0 + j) * ldk; }, [&](int) { return (const double *)w; }, sm_c);
        __syncthreads();
        for (int k = tid; k < K; k += STEP_NT) {
            double acc = (j0 == 0) ? w[k] : rn[k];
            for (int j = 0; j < nj; ++j) acc -= sm_c[j] * s.R[(size_t)(j0 + j) * ldk + k];
            rn[k] = acc;
        }
        __syncthreads();
    }
    if (an == 0)
        for (int k = tid; k < K; k += STEP_NT) rn[k] = w[k];
    for (int k = K + tid; k < ldk; k += STEP_NT) rn[k] = 0.0;
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
