// pls_b200/csrc/step_wide.cuh — the per-component algebra for WIDE response blocks (M > PLSB_MAX_M = 32).
//
// The reference finds the weight vector through Eigen::EigenSolver on the M x M matrix XY'XY (src/pls.cpp:406-407,
// 113-141), which has no size limit.  The fast kernels (step_cluster.cuh, step_core.cuh, cv.cu) keep that matrix, its
// eigenvectors and q in shared memory with one lane per row: M <= 32.  This file is the general path: the same
// phases (p, q, deflation of S: pls.cpp:427-429; Gram, dominant eigenvector, w: 403-411; r: 412-416) with the M-sized
// objects in a GLOBAL scratch area (L2-resident) and a parallel-ordering Jacobi run by the whole CTA — pivot pairs of
// one tournament round are disjoint, so all threads update columns (A J, V J), then rows (J' A), of every pair at once.
// Built for correctness on rare shapes, not for latency: ~3 CTA barriers per round, M - 1 rounds per sweep.
#pragma once
#include "common.cuh"

namespace plsb {

// doubles of scratch per problem (one fit, or one fold): gram M*M, V M*M, q M, rotations 2*ceil(M/2), pair table
// (as doubles) 2*ceil(M/2), + 8
__host__ __device__ inline size_t wide_scratch_doubles(int M) {
    const size_t np = (size_t)(M + 1) / 2;
    return 2 * (size_t)M * M + (size_t)M + 4 * np + 8;
}

template <int NT>
struct WideSmem {
    double c[256];
    double red[NT / 32];
    double bc[2];
};

template <int NT>
__device__ __forceinline__ double wide_block_sum(double v, WideSmem<NT> &sh) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) sh.red[warp] = v;
    __syncthreads();
    double tot = 0.0;
#pragma unroll
    for (int i = 0; i < NT / 32; ++i) tot += sh.red[i];
    return tot;
}

// A (m x m, row-major, symmetric, global) is destroyed; V receives the eigenvectors (columns); q the unit eigenvector
// of the eigenvalue with the largest magnitude (first index on ties, pls.cpp:119-126).  All NT threads of the CTA.
template <int NT>
__device__ void jacobi_dominant_cta(double *A, double *V, int m, double *q, double *rot, double *pairs, WideSmem<NT> &sh) {
    const int tid = threadIdx.x;
    for (int i = tid; i < m * m; i += NT) V[i] = ((i / m) == (i % m)) ? 1.0 : 0.0;
    __syncthreads();
    const int mp = m + (m & 1), np = mp / 2, ring = mp - 1;
    double prev_off = 0.0;
    for (int sweep = 0; sweep < 64 && m > 1; ++sweep) {
        double off = 0.0, diag = 0.0;
        for (int i = tid; i < m * m; i += NT) {
            const int rr = i / m, cc = i % m;
            const double v = A[i];
            if (rr == cc) diag = fma(v, v, diag);
            else if (rr < cc) off = fma(v, v, off);
        }
        off = wide_block_sum<NT>(off, sh);
        diag = wide_block_sum<NT>(diag, sh);
        if (off <= 1e-300 || off <= 4e-32 * diag || (sweep > 1 && off >= 0.0625 * prev_off && off <= 1e-24 * diag)) break;
        prev_off = off;
        for (int t = 0; t < ring; ++t) {
            // pair i of this round (circle method: player mp-1 stays put); rotation from the current A
            for (int i = tid; i < np; i += NT) {
                const int x = i == 0 ? mp - 1 : (t + i) % ring;
                const int y = i == 0 ? t % ring : (t + ring - i) % ring;
                int p = x < y ? x : y, r = x < y ? y : x;
                double c = 1.0, s = 0.0;
                if (r < m) {
                    const double apr = A[p * m + r];
                    if (apr != 0.0) {
                        const double theta = (A[r * m + r] - A[p * m + p]) / (2.0 * apr);
                        const double tn = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(fma(theta, theta, 1.0)));
                        c = 1.0 / sqrt(fma(tn, tn, 1.0));
                        s = tn * c;
                    }
                } else {
                    r = p;   // dummy opponent: identity
                }
                pairs[2 * i] = (double)p; pairs[2 * i + 1] = (double)r;
                rot[2 * i] = c; rot[2 * i + 1] = s;
            }
            __syncthreads();
            // A <- A J, V <- V J : (pair i, row k)
            for (int e = tid; e < np * m; e += NT) {
                const int i = e / m, k = e % m;
                const int p = (int)pairs[2 * i], r = (int)pairs[2 * i + 1];
                if (p == r) continue;
                const double c = rot[2 * i], s = rot[2 * i + 1];
                const double akp = A[k * m + p], akr = A[k * m + r];
                A[k * m + p] = c * akp - s * akr;
                A[k * m + r] = s * akp + c * akr;
                const double vkp = V[k * m + p], vkr = V[k * m + r];
                V[k * m + p] = c * vkp - s * vkr;
                V[k * m + r] = s * vkp + c * vkr;
            }
            __syncthreads();
            // A <- J' A : (pair i, column k)
            for (int e = tid; e < np * m; e += NT) {
                const int i = e / m, k = e % m;
                const int p = (int)pairs[2 * i], r = (int)pairs[2 * i + 1];
                if (p == r) continue;
                const double c = rot[2 * i], s = rot[2 * i + 1];
                const double apk = A[p * m + k], ark = A[r * m + k];
                A[p * m + k] = c * apk - s * ark;
                A[r * m + k] = s * apk + c * ark;
            }
            __syncthreads();
        }
    }
    __syncthreads();
    if (tid == 0) {
        int idx = 0;
        double best = 0.0;
        for (int i = 0; i < m; ++i) {
            const double v = fabs(A[i * m + i]);
            if (v > best) { best = v; idx = i; }
        }
        double nrm = 0.0;
        for (int k = 0; k < m; ++k) nrm = fma(V[k * m + idx], V[k * m + idx], nrm);
        sh.bc[0] = (double)idx;
        sh.bc[1] = sqrt(nrm);
    }
    __syncthreads();
    const int idx = (int)sh.bc[0];
    const double nrm = sh.bc[1];
    for (int k = tid; k < m; k += NT) q[k] = V[k * m + idx] / nrm;
    __syncthreads();
}

// 2^(-ilogb(x)) for a positive normal double (the same exact power-of-two scaling as step_cluster.cuh's solver)
__device__ __forceinline__ double wide_pow2_scale(double x) {
    const long long e = (__double_as_longlong(x) >> 52) & 0x7ff;
    return __longlong_as_double((2046LL - e) << 52);
}

// Dominant unit eigenvector of the symmetric PSD m x m matrix G by REPEATED SQUARING, run by the whole CTA — the
// algorithm of step_cluster.cuh's warp solver (B <- B^2 with lazy exact power-of-two scaling until trace(B^2)/trace(B)^2
// is within 1e-9 of 1, three more squarings, then two power steps with G itself to remove the rounding the squarings
// left) for matrices too wide for one warp.  work: SHARED memory, wide_squaring_smem_doubles(m) doubles (a copy of G and
// two iterates at an odd row stride: conflict-free).  A squaring is m^3 FMAs over NT threads: ~2 us at m = 64, where a
// Jacobi sweep costs 63 rounds of three CTA barriers each and ten to sixty sweeps are needed.  Returns false when 48
// squarings were not enough (near-degenerate spectrum) or the matrix is zero: the caller falls back to Jacobi.
__host__ __device__ inline size_t wide_squaring_smem_doubles(int m) { return 3 * (size_t)m * (size_t)(m | 1); }

template <int NT>
__device__ bool dominant_eigvec_cta(const double *G, double *work, int m, double *q, WideSmem<NT> &sh) {
    const int tid = threadIdx.x;
    const int ld = m | 1;
    double *Gs = work, *buf[2] = {work + (size_t)m * ld, work + 2 * (size_t)m * ld};
    for (int e = tid; e < m * m; e += NT) {
        const double v = G[e];
        Gs[(e / m) * ld + e % m] = v;
        buf[0][(e / m) * ld + e % m] = v;
    }
    __syncthreads();
    double tr = 0.0;
    for (int i = 0; i < m; ++i) tr += Gs[i * ld + i];
    if (!(tr > 0.0) || !(tr < 1e300)) return false;
    int cur = 0, extra = 0, res = -1;
    double sc = wide_pow2_scale(tr), trB = sc * tr;
    for (int it = 0; it < 48; ++it) {
        const double *B = buf[cur];
        double *C = buf[cur ^ 1];
        const double sc2 = sc * sc;
        for (int e = tid; e < m * m; e += NT) {
            const int i = e / m, j = e % m;               // (B^2)_ij = row_i . row_j by symmetry
            const double *bi = B + (size_t)i * ld, *bj = B + (size_t)j * ld;
            double a0 = 0.0, a1 = 0.0;
            int k = 0;
            for (; k + 1 < m; k += 2) {
                a0 = fma(bi[k], bj[k], a0);
                a1 = fma(bi[k + 1], bj[k + 1], a1);
            }
            if (k < m) a0 = fma(bi[k], bj[k], a0);
            C[(size_t)i * ld + j] = (a0 + a1) * sc2;
        }
        __syncthreads();
        double trC = 0.0;
        for (int k = 0; k < m; ++k) trC += C[(size_t)k * ld + k];     // every thread: the same sum in the same order
        const double t2 = trB * trB;
        cur ^= 1;
        if (!(trC > 0.0)) break;
        if (t2 - trC <= 1e-9 * t2 && ++extra > 2) { res = cur; break; }
        sc = wide_pow2_scale(trC);
        trB = sc * trC;
        __syncthreads();       // the buffer read in this iteration is overwritten in the next one
    }
    __syncthreads();
    if (res < 0) return false;
    const double *B = buf[res];
    int col = 0;
    double best = -1.0;
    for (int i = 0; i < m; ++i) {
        const double d = B[(size_t)i * ld + i];
        if (d > best) { best = d; col = i; }
    }
    double *x = buf[res ^ 1];          // free now: x at [0, m), G x at [m, 2m)
    for (int k = tid; k < m; k += NT) x[k] = B[(size_t)k * ld + col];
    __syncthreads();
    for (int rep = 0; rep < 3; ++rep) {  // normalise, then v <- G v (twice)
        double n2 = 0.0;
        for (int k = 0; k < m; ++k) n2 = fma(x[k], x[k], n2);
        const double nrm = sqrt(n2);
        __syncthreads();
        for (int k = tid; k < m; k += NT) x[k] = x[k] / nrm;
        __syncthreads();
        if (rep == 2) break;
        for (int i = tid; i < m; i += NT) {
            double acc = 0.0;
            for (int k = 0; k < m; ++k) acc = fma(Gs[(size_t)i * ld + k], x[k], acc);
            x[m + i] = acc;
        }
        __syncthreads();
        for (int k = tid; k < m; k += NT) x[k] = x[m + k];
        __syncthreads();
    }
    for (int k = tid; k < m; k += NT) q[k] = x[k];
    __syncthreads();
    return true;
}

// one warp per dot product (the same order of operations as step_core.cuh's warp_dots)
template <int NT, typename FA, typename FB>
__device__ __forceinline__ void wide_warp_dots(int ndots, int K, FA a_of, FB b_of, double *out) {
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

// Component a is finished: p = g/tt, q = S'r/tt (-> Qa and scratch q), S -= tt p q'.
template <int NT>
__device__ void finish_component_wide(int K, int ldk, int M, double *S, const double *g, double tt, const double *r,
                                      double *p, double *Qa, double *wide) {
    const int tid = threadIdx.x;
    double *q = wide + 2 * (size_t)M * M;
    for (int k = tid; k < ldk; k += NT) p[k] = (k < K) ? g[k] / tt : 0.0;
    wide_warp_dots<NT>(M, K, [&](int m) { return (const double *)(S + (size_t)m * ldk); }, [&](int) { return r; }, q);
    __syncthreads();
    for (int m = tid; m < M; m += NT) {
        const double v = q[m] / tt;
        q[m] = v;
        Qa[m] = v;
    }
    __syncthreads();
    for (int k = tid; k < K; k += NT) {
        const double tp = tt * p[k];
        for (int m = 0; m < M; ++m) S[(size_t)m * ldk + k] -= tp * q[m];
    }
    __syncthreads();
}

// Weight and rotation vectors of component `an` (M > 1 assumed: wide).
// jac: optional SHARED-memory work area of wide_jacobi_smem_doubles(M) doubles (the kernel's dynamic shared memory) for
// the Jacobi iteration — every tournament round is a few dependent loads and stores per thread, and out of global
// memory a round costs ~4 us of L2 round trips (measured: 2.0 ms per component at M = 40, 8.2 ms at M = 64); null:
// the global scratch is used (M too wide for one SM's shared memory).
__host__ __device__ inline size_t wide_jacobi_smem_doubles(int M) {
    const size_t jac = 2 * (size_t)M * M + 4 * (size_t)((M + 1) / 2), sq = wide_squaring_smem_doubles(M);
    return jac > sq ? jac : sq;
}

template <int NT>
__device__ void next_weights_wide(int K, int ldk, int M, int an, const double *S, const double *P, const double *R,
                                  double *w, double *rn, double *wide, WideSmem<NT> &sh, double *jac = nullptr) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double *gram = wide, *V = wide + (size_t)M * M, *q = V + (size_t)M * M, *rot = q + M;
    double *pairs = rot + 2 * ((M + 1) / 2);
    if (jac) { gram = jac; V = jac + (size_t)M * M; rot = V + (size_t)M * M; pairs = rot + 2 * ((M + 1) / 2); }
    const int npairs = M * (M + 1) / 2;
    for (int d = warp; d < npairs; d += NT / 32) {   // Gram = S'S, upper triangle, one warp per entry
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
        if (lane == 0) {              // the Gram matrix itself always goes to the global scratch (wide[0 .. M*M))
            wide[u * M + v] = sum;
            wide[v * M + u] = sum;
        }
    }
    __syncthreads();
    // shared-memory work area: repeated squaring first (the Gram matrix still sits in the global scratch), Jacobi as
    // the fallback for near-degenerate spectra — and the only solver when M is too wide for one SM's shared memory
    bool solved = false;
    if (jac) solved = dominant_eigvec_cta<NT>(wide, jac, M, q, sh);
    if (!solved) {
        if (jac) {
            for (int e = tid; e < M * M; e += NT) gram[e] = wide[e];
            __syncthreads();
        }
        jacobi_dominant_cta<NT>(gram, V, M, q, rot, pairs, sh);
    }
    for (int k = tid; k < ldk; k += NT) {
        double acc = 0.0;
        if (k < K)
            for (int m = 0; m < M; ++m) acc = fma(S[(size_t)m * ldk + k], q[m], acc);
        w[k] = acc;
    }
    __syncthreads();
    {   // w /= sqrt(w'w)   (pls.cpp:411)
        double acc = 0.0;
        for (int k = tid; k < K; k += NT) acc = fma(w[k], w[k], acc);
        const double nrm = sqrt(wide_block_sum<NT>(acc, sh));
        for (int k = tid; k < K; k += NT) w[k] = w[k] / nrm;
    }
    __syncthreads();
    // r = w - sum_j (p_j'w) r_j   (pls.cpp:412-416; every dot uses the un-modified w)
    for (int j0 = 0; j0 < an; j0 += 256) {
        const int nj = an - j0 < 256 ? an - j0 : 256;
        wide_warp_dots<NT>(nj, K, [&](int j) { return P + (size_t)(j0 + j) * ldk; }, [&](int) { return (const double *)w; }, sh.c);
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
