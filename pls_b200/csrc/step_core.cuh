// pls_b200/csrc/step_core.cuh — the K-sized per-component algebra shared by the single-fit step
// kernel (one CTA) and the batched cross-validation kernel (one CTA per fold).
//
// Reference src/pls.cpp: 427-429 (p, q, deflation of XY), 403-411 (weight vector incl. the dominant
// eigenvector of XY'XY, 113-141), 412-416 (r = w - sum_j (p_j'w) r_j).
#pragma once
#include "common.cuh"
#include "jacobi.cuh"
#include "step_cluster.cuh"

namespace plsb {

template <int NT>
struct StepSmem {
    double q[PLSB_MAX_M];
    double c[256];
    double gram[PLSB_MAX_M * PLSB_MAX_M];
    double V[PLSB_MAX_M * PLSB_MAX_M];
    double red[NT / 32];
    double bc;  // broadcast slot
};

// Dominant unit eigenvector of the M x M Gram matrix in sh.gram (one warp; role of pls.cpp:406-407, 113-141).  Up to
// 16 responses: the repeated-squaring solver of the cluster step (step_cluster.cuh; ~1-2 us, Jacobi as its own fallback
// for near-degenerate spectra) with sh.V as its two work arrays — the Jacobi sweeps alone were 44 % of the small-data CV
// kernel's time (ncu, profiles/r2e).  Wider blocks keep the Jacobi solver (work arrays of 32 x 32 would not fit).
template <int NT>
__device__ __forceinline__ void dominant_small(StepSmem<NT> &sh, int M) {
    if (M <= 16) dominant_eigvec_warp<16>(sh.gram, sh.V, sh.V + 256, M, sh.q);
    else jacobi_dominant_warp(sh.gram, sh.V, M, sh.q);
}

// one warp per dot product, lanes stride the K axis; out[d] = a_d . b_d
template <int NT, typename FA, typename FB>
__device__ __forceinline__ void warp_dots(int ndots, int K, FA a_of, FB b_of, double *out) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int d = warp; d < ndots; d += NT / 32) {
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

// block-wide sum of one value per thread; every thread gets the result (fixed order: deterministic)
template <int NT>
__device__ __forceinline__ double block_sum(double v, StepSmem<NT> &sh) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();  // protect sh.red from the previous use
    if (lane == 0) sh.red[warp] = v;
    __syncthreads();
    double tot = 0.0;
#pragma unroll
    for (int i = 0; i < NT / 32; ++i) tot += sh.red[i];
    return tot;
}

// Component a is finished: p = g/tt, q = S'r/tt, S -= tt p q'.  g may alias nothing else; q -> Qa (M).
template <int NT>
__device__ __forceinline__ void finish_component(int K, int ldk, int M, double *S, const double *g, double tt,
                                                 const double *r, double *p, double *Qa, StepSmem<NT> &sh) {
    const int tid = threadIdx.x;
    for (int k = tid; k < ldk; k += NT) p[k] = (k < K) ? g[k] / tt : 0.0;
    warp_dots<NT>(M, K, [&](int m) { return (const double *)(S + (size_t)m * ldk); }, [&](int) { return r; }, sh.q);
    __syncthreads();
    if (tid < M) {
        sh.q[tid] = sh.q[tid] / tt;
        Qa[tid] = sh.q[tid];
    }
    __syncthreads();
    for (int k = tid; k < K; k += NT) {
        const double tp = tt * p[k];
        for (int m = 0; m < M; ++m) S[(size_t)m * ldk + k] -= tp * sh.q[m];
    }
    __syncthreads();
}

// Weight and rotation vectors of component `an` from the deflated S and the stored P[0..an), R[0..an)
// (component-major, stride ldk).  w (ldk) and rn (ldk) are outputs; rn gets zero padding.
template <int NT>
__device__ __forceinline__ void next_weights(int K, int ldk, int M, int an, const double *S, const double *P,
                                             const double *R, double *w, double *rn, StepSmem<NT> &sh) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (M == 1) {
        for (int k = tid; k < ldk; k += NT) w[k] = (k < K) ? S[k] : 0.0;
    } else {
        const int npairs = M * (M + 1) / 2;
        for (int d = warp; d < npairs; d += NT / 32) {  // Gram = S'S, upper triangle, one warp per entry
            int u = 0, rem = d;
            while (rem >= M - u) { rem -= M - u; ++u; }
            const int v = u + rem;
            const double *pa = S + (size_t)u * ldk, *pb = S + (size_t)v * ldk;
            double s0 = 0.0, s1 = 0.0;
            int k = lane;
            for (; k + 32 < K; k += 64) {
                s0 = fma(pa[k], pb[k], s0);
                s1 = fma(pa[k + 32], pb[k + 32], s1);
            }
            for (; k < K; k += 32) s0 = fma(pa[k], pb[k], s0);
            const double sum = warp_sum(s0 + s1);
            if (lane == 0) {
                sh.gram[u * M + v] = sum;
                sh.gram[v * M + u] = sum;
            }
        }
        __syncthreads();
        if (warp == 0) dominant_small<NT>(sh, M);
        __syncthreads();
        for (int k = tid; k < ldk; k += NT) {
            double acc = 0.0;
            if (k < K)
                for (int m = 0; m < M; ++m) acc = fma(S[(size_t)m * ldk + k], sh.q[m], acc);
            w[k] = acc;
        }
    }
    __syncthreads();
    {   // w /= sqrt(w'w)   (pls.cpp:411)
        double acc = 0.0;
        for (int k = tid; k < K; k += NT) acc = fma(w[k], w[k], acc);
        const double nrm = sqrt(block_sum<NT>(acc, sh));
        for (int k = tid; k < K; k += NT) w[k] = w[k] / nrm;
    }
    __syncthreads();
    // r = w - sum_j (p_j'w) r_j   (pls.cpp:412-416; every dot uses the un-modified w)
    for (int j0 = 0; j0 < an; j0 += 256) {
        const int nj = an - j0 < 256 ? an - j0 : 256;
        warp_dots<NT>(nj, K, [&](int j) { return P + (size_t)(j0 + j) * ldk; }, [&](int) { return (const double *)w; },
                      sh.c);
        __syncthreads();
        for (int k = tid; k < K; k += NT) {
            double acc = (j0 == 0) ? w[k] : rn[k];
            for (int j = 0; j < nj; ++j) acc -= sh.c[j] * R[(size_t)(j0 + j) * ldk + k];
            rn[k] = acc;
        }
        __syncthreads();
    }
    if (an == 0)
        for (int k = tid; k < K; k += NT) rn[k] = w[k];
    for (int k = K + tid; k < ldk; k += NT) rn[k] = 0.0;
}

}  // namespace plsb
