// pls_b200/csrc/step_cluster.cuh — building blocks of the cluster step kernel (step.cu): a warp-level
// "transposing" reduction of many values at once, and a fast dominant-eigenvector solver for the small
// symmetric PSD matrix XY'XY (reference src/pls.cpp:406-407, 113-141).
#pragma once
#include "common.cuh"
#include "jacobi.cuh"

namespace plsb {

// Reduce N (= 32 or 8) per-lane values over the 32 lanes of a warp with N-1 + (5 - log2 N) shuffles instead
// of 5 N: every halving step trades half of the lane's values with its partner, so the number of live
// values halves while the number of lanes summed doubles.  Afterwards lane l holds the warp total of value
// index l >> (5 - log2 N) in v[0].  Fixed association order: deterministic.
template <int N>
__device__ __forceinline__ void warp_transpose_reduce(double (&v)[N]) {
    const int lane = threadIdx.x & 31;
    int o = 16;
#pragma unroll
    for (int n = N / 2; n >= 1; n >>= 1, o >>= 1) {
        const bool upper = (lane & o) != 0;
#pragma unroll
        for (int i = 0; i < n; ++i) {
            const double send = upper ? v[i] : v[i + n];
            const double keep = upper ? v[i + n] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
#pragma unroll
    for (; o >= 1; o >>= 1) v[0] += __shfl_xor_sync(0xffffffffu, v[0], o);
}

// warp totals of val(0..nv-1) -> dst[0..nv-1] (one writer lane per value)
template <typename F>
__device__ __forceinline__ void warp_reduce_many(int nv, F val, double *dst) {
    const int lane = threadIdx.x & 31;
    int base = 0;
    for (; nv - base > 8; base += 32) {
        double v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = (base + i < nv) ? val(base + i) : 0.0;
        warp_transpose_reduce<32>(v);
        if (base + lane < nv) dst[base + lane] = v[0];
    }
    if (base < nv) {
        double v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = (base + i < nv) ? val(base + i) : 0.0;
        warp_transpose_reduce<8>(v);
        if ((lane & 3) == 0 && base + (lane >> 2) < nv) dst[base + (lane >> 2)] = v[0];
    }
}

// Dominant unit eigenvector of the symmetric PSD m x m matrix G (shared memory, row-major), one warp.
// Repeated squaring: B <- B^2 / trace(B^2) drives B to v1 v1' with the eigenvalue ratio squared every
// step (a few hundred cycles each for m = 8), then two power steps with G itself remove the rounding the
// squarings left.  Near-degenerate spectra (ratio within 1e-10 of 1) fall back to Jacobi.  The result has
// the same direction as find_dominant_ev's pick (largest |eigenvalue|; all are >= 0 here) up to sign.
// G is preserved; B, C are m x m work arrays; q receives the eigenvector.  All 32 lanes must call.
__device__ __forceinline__ void dominant_eigvec_warp(const double *G, double *B, double *C, int m, double *q) {
    const int lane = threadIdx.x & 31;
    const int mm = m * m;
    double tr = 0.0;
    for (int i = lane; i < m; i += 32) tr += G[i * m + i];
    tr = warp_sum(tr);
    if (!(tr > 0.0)) {  // S is exactly zero (components past the rank): any unit vector
        if (lane < m) q[lane] = lane == 0 ? 1.0 : 0.0;
        __syncwarp();
        return;
    }
    for (int e = lane; e < mm; e += 32) B[e] = G[e] / tr;
    __syncwarp();
    int extra = 0;
    bool converged = false;
    for (int it = 0; it < 48; ++it) {
        double t = 0.0;
        for (int e = lane; e < mm; e += 32) {
            const int i = e / m, j = e % m;
            double acc = 0.0;
            for (int k = 0; k < m; ++k) acc = fma(B[i * m + k], B[k * m + j], acc);
            C[e] = acc;
            if (i == j) t += acc;
        }
        t = warp_sum(t);  // trace(B^2) = sum of squared normalised eigenvalues: -> 1 as B -> rank one
        __syncwarp();
        for (int e = lane; e < mm; e += 32) B[e] = C[e] / t;
        __syncwarp();
        if (1.0 - t <= 1e-9 && ++extra > 2) { converged = true; break; }
    }
    if (!converged) {
        for (int e = lane; e < mm; e += 32) B[e] = G[e];
        __syncwarp();
        jacobi_dominant_warp(B, C, m, q);
        return;
    }
    int col = 0;
    double best = -1.0;
    for (int i = 0; i < m; ++i) {
        const double d = B[i * m + i];
        if (d > best) { best = d; col = i; }
    }
    double x = lane < m ? B[lane * m + col] : 0.0;
#pragma unroll 1
    for (int rep = 0; rep < 3; ++rep) {  // normalise, then v <- G v (twice)
        const double n2 = warp_sum(x * x);
        if (lane < m) q[lane] = x / sqrt(n2);
        __syncwarp();
        if (rep == 2) break;
        double acc = 0.0;
        if (lane < m)
            for (int k = 0; k < m; ++k) acc = fma(G[lane * m + k], q[k], acc);
        __syncwarp();
        x = acc;
    }
}

}  // namespace plsb
