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

// warp totals of val(0..nv-1) -> dst[0..nv-1] (one writer lane per value).  val must be callable for any
// index in [0, nv): out-of-range slots of a chunk are evaluated at a clamped index and then zeroed, so that
// the loads inside val stay unconditional and the compiler can put a whole chunk of them in flight at once
// (a guarded load per value serialises into one memory round trip per value on this critical path).
template <typename F>
__device__ __forceinline__ void warp_reduce_many(int nv, F val, double *dst) {
    const int lane = threadIdx.x & 31;
    int base = 0;
    for (; nv - base > 8; base += 32) {
        double v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = val(base + i < nv ? base + i : nv - 1);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = base + i < nv ? v[i] : 0.0;
        warp_transpose_reduce<32>(v);
        if (base + lane < nv) dst[base + lane] = v[0];
    }
    if (base < nv) {
        double v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = val(base + i < nv ? base + i : nv - 1);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = base + i < nv ? v[i] : 0.0;
        warp_transpose_reduce<8>(v);
        if ((lane & 3) == 0 && base + (lane >> 2) < nv) dst[base + (lane >> 2)] = v[0];
    }
}

// 2^(-ilogb(x)) for a positive normal double: exact power-of-two scale that brings x into [1, 2)
__device__ __forceinline__ double pow2_scale(double x) {
    const long long e = (__double_as_longlong(x) >> 52) & 0x7ff;
    return __longlong_as_double((2046LL - e) << 52);
}

// Repeated squaring of the MP x MP (zero padded) symmetric PSD matrix in buf[0] until it is numerically rank
// one.  The true iterate is sc * buf[cur] with sc a power of two (applied lazily and exactly inside the next
// product), so an iteration has no division and no shuffle on its dependency chain: rows from shared memory ->
// FMA chains -> store -> every lane re-reads the diagonal for the trace.  Lane owns entries (i, j) with
// j = lane % MP fixed (row j in registers, rows i broadcast); (B^2)_ij = row_i . row_j by symmetry.
// Converged when trace(B^2) / trace(B)^2 (= sum of squared normalised eigenvalues) is within 1e-9 of 1, plus
// two more squarings (the eigenvector's contamination is the square root of that gap).  Returns the index of
// the buffer holding the result, or -1 if 48 squarings were not enough (near-degenerate spectrum).
template <int MP>
__device__ __forceinline__ int squaring_loop(double *buf0, double *buf1, double tr0, long long *iters) {
    constexpr int T = MP * MP / 32, IS = 32 / MP;
    const int lane = threadIdx.x & 31;
    const int j = lane % MP, i0 = lane / MP;
    double *buf[2] = {buf0, buf1};
    int cur = 0, extra = 0;
    double sc = pow2_scale(tr0), trB = sc * tr0;
    for (int it = 0; it < 48; ++it) {
        const double *B = buf[cur];
        double *C = buf[cur ^ 1];
        const double sc2 = sc * sc;
        double bj[MP];
#pragma unroll
        for (int k = 0; k < MP; ++k) bj[k] = B[j * MP + k];
#pragma unroll
        for (int t = 0; t < T; ++t) {
            const int i = i0 + t * IS;
            double a0 = 0.0, a1 = 0.0;
#pragma unroll
            for (int k = 0; k < MP; k += 2) {
                a0 = fma(B[i * MP + k], bj[k], a0);
                a1 = fma(B[i * MP + k + 1], bj[k + 1], a1);
            }
            C[i * MP + j] = (a0 + a1) * sc2;
        }
        __syncwarp();
        double d0 = 0.0, d1 = 0.0;
#pragma unroll
        for (int k = 0; k < MP; k += 2) {
            d0 += C[k * MP + k];
            d1 += C[(k + 1) * MP + k + 1];
        }
        const double trC = d0 + d1;      // trace of the true B^2
        const double t2 = trB * trB;
        cur ^= 1;
        if (iters && lane == 0) *iters = it + 1;
        if (!(trC > 0.0)) return -1;
        if (t2 - trC <= 1e-9 * t2 && ++extra > 2) return cur;
        sc = pow2_scale(trC);
        trB = sc * trC;
        // (the buffer read in this iteration is overwritten in the next one: the __syncwarp above orders it)
    }
    return -1;
}

// Dominant unit eigenvector of the symmetric PSD m x m matrix G (shared memory, row-major), one warp.
// Repeated squaring (above), then two power steps with G itself remove the rounding the squarings left.
// Near-degenerate spectra (ratio within ~1e-10 of 1) fall back to Jacobi.  The result has the direction of
// find_dominant_ev's pick (largest |eigenvalue|; all are >= 0 here) up to sign.
// G is preserved; B0, B1 (>= MP*MP doubles each, MP = m rounded up to 8/16/32) are work arrays; q receives the
// eigenvector.  All 32 lanes must call.
// MAXMP (8, 16 or 32): the caller guarantees m <= MAXMP; only the squaring loops up to that size are instantiated
// (the 32-wide one keeps a 32-double row in registers: not something to inline into a 64-register kernel).
template <int MAXMP = 32>
__device__ __forceinline__ void dominant_eigvec_warp(const double *G, double *B0, double *B1, int m, double *q,
                                                     long long *iters = nullptr) {
    const int lane = threadIdx.x & 31;
    const int mp = (m <= 8 || MAXMP == 8) ? 8 : ((m <= 16 || MAXMP == 16) ? 16 : 32);
    double tr = 0.0;
    for (int i = 0; i < m; ++i) tr += G[i * m + i];
    if (!(tr > 0.0) || !(tr < 1e300)) {  // S is exactly zero (components past the rank) or not finite
        if (lane < m) q[lane] = lane == 0 ? 1.0 : 0.0;
        __syncwarp();
        return;
    }
    for (int e = lane; e < mp * mp; e += 32) {
        const int i = e / mp, j = e % mp;
        B0[e] = (i < m && j < m) ? G[i * m + j] : 0.0;
    }
    __syncwarp();
    int res;
    if (mp == 8) res = squaring_loop<8>(B0, B1, tr, iters);
    else if constexpr (MAXMP >= 16) {
        if (mp == 16) res = squaring_loop<16>(B0, B1, tr, iters);
        else if constexpr (MAXMP >= 32) res = squaring_loop<32>(B0, B1, tr, iters);
        else res = -1;
    } else res = -1;
    __syncwarp();
    if (res < 0) {
        for (int e = lane; e < m * m; e += 32) B0[e] = G[e];
        __syncwarp();
        jacobi_dominant_warp(B0, B1, m, q);
        return;
    }
    const double *B = res == 0 ? B0 : B1;
    int col = 0;
    double best = -1.0;
    for (int i = 0; i < m; ++i) {
        const double d = B[i * mp + i];
        if (d > best) { best = d; col = i; }
    }
    double x = lane < m ? B[lane * mp + col] : 0.0;
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
