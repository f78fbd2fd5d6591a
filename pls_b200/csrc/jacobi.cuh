// pls_b200/csrc/jacobi.cuh — dominant eigenvector of a small symmetric PSD matrix, one warp.
//
// Role of find_dominant_ev / dominant_eigenvector (reference src/pls.cpp:113-141) applied to the
// M x M matrix XY'XY (pls.cpp:406-407).  The reference calls Eigen::EigenSolver (Hessenberg + Francis
// QR); the matrix is symmetric, so Jacobi rotations yield the same unit eigenvector up to sign.  The sign
// cancels in coefficients / fitted values and is normalised away when scores / loadings are compared.
//
// Parallel (round-robin tournament) ordering: a sweep is m-1 rounds of floor(m/2) DISJOINT pivot pairs.
// Disjoint rotations commute, so one round applies them all at once — lanes compute one rotation each,
// then every lane updates its own row (A J, V J) and its own column (J' A).  This sits on the serial
// critical path of every component (the step kernel), so what counts is the length of the dependency
// chain: m-1 rounds per sweep instead of m(m-1)/2 rotations.
#pragma once
#include "common.cuh"

#define PLSB_MAX_M 32

namespace plsb {

// A (m x m, row-major, symmetric) is destroyed; V receives the eigenvectors (columns); q the unit
// eigenvector of the eigenvalue with the largest magnitude (first index on ties, pls.cpp:119-126).
// Must be called by all 32 lanes of one warp; A, V, q in shared memory; m <= PLSB_MAX_M.
__device__ __forceinline__ void jacobi_dominant_warp(double *A, double *V, int m, double *q) {
    const int lane = threadIdx.x & 31;
    const unsigned full = 0xffffffffu;
    for (int i = lane; i < m * m; i += 32) V[i] = ((i / m) == (i % m)) ? 1.0 : 0.0;
    __syncwarp();
    const int mp = m + (m & 1);       // players in the tournament (one dummy when m is odd)
    const int np = mp / 2;            // pivot pairs per round
    const int ring = mp - 1;
    double prev_off = 0.0;
    for (int sweep = 0; sweep < 64 && m > 1; ++sweep) {
        double off = 0.0, diag = 0.0;
        for (int i = lane; i < m * m; i += 32) {
            const int rr = i / m, cc = i % m;
            const double v = A[i];
            if (rr == cc) diag = fma(v, v, diag);
            else if (rr < cc) off = fma(v, v, off);
        }
        off = warp_sum(off);
        diag = warp_sum(diag);
        // converged: off-diagonal norm below 2e-16 of the diagonal's (squares compared), or rounding has
        // stopped the quadratic convergence (no 4x drop over a whole sweep) — without the second test the
        // loop never meets a threshold under one ulp and burns all 64 sweeps on the critical path
        if (off <= 1e-300 || off <= 4e-32 * diag || (sweep > 1 && off >= 0.0625 * prev_off && off <= 1e-24 * diag)) break;
        prev_off = off;
        for (int t = 0; t < ring; ++t) {
            // lane i < np owns the pair (x, y) of this round (circle method: player mp-1 stays put)
            int p = 0, r = 0;
            double c = 1.0, s = 0.0;
            if (lane < np) {
                const int x = lane == 0 ? mp - 1 : (t + lane) % ring;
                const int y = lane == 0 ? t % ring : (t + ring - lane) % ring;
                p = x < y ? x : y;
                r = x < y ? y : x;
                if (r < m) {
                    const double apr = A[p * m + r];
                    if (apr != 0.0) {
                        const double theta = (A[r * m + r] - A[p * m + p]) / (2.0 * apr);
                        const double tn = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(fma(theta, theta, 1.0)));
                        c = 1.0 / sqrt(fma(tn, tn, 1.0));
                        s = tn * c;
                    }
                } else {
                    r = p;  // dummy opponent: identity
                }
            }
            __syncwarp();
            // A <- A J, V <- V J   (lane = row; the np column pairs are disjoint)
            for (int i = 0; i < np; ++i) {
                const int pi = __shfl_sync(full, p, i), ri = __shfl_sync(full, r, i);
                const double ci = __shfl_sync(full, c, i), si = __shfl_sync(full, s, i);
                if (lane < m && pi != ri) {
                    const double akp = A[lane * m + pi], akr = A[lane * m + ri];
                    A[lane * m + pi] = ci * akp - si * akr;
                    A[lane * m + ri] = si * akp + ci * akr;
                    const double vkp = V[lane * m + pi], vkr = V[lane * m + ri];
                    V[lane * m + pi] = ci * vkp - si * vkr;
                    V[lane * m + ri] = si * vkp + ci * vkr;
                }
            }
            __syncwarp();
            // A <- J' A   (lane = column)
            for (int i = 0; i < np; ++i) {
                const int pi = __shfl_sync(full, p, i), ri = __shfl_sync(full, r, i);
                const double ci = __shfl_sync(full, c, i), si = __shfl_sync(full, s, i);
                if (lane < m && pi != ri) {
                    const double apk = A[pi * m + lane], ark = A[ri * m + lane];
                    A[pi * m + lane] = ci * apk - si * ark;
                    A[ri * m + lane] = si * apk + ci * ark;
                }
            }
            __syncwarp();
        }
    }
    __syncwarp();
    int idx = 0;
    double best = 0.0;
    for (int i = 0; i < m; ++i) {
        const double v = fabs(A[i * m + i]);
        if (v > best) { best = v; idx = i; }
    }
    double nrm = 0.0;
    for (int k = 0; k < m; ++k) nrm = fma(V[k * m + idx], V[k * m + idx], nrm);
    nrm = sqrt(nrm);
    if (lane < m) q[lane] = V[lane * m + idx] / nrm;
    __syncwarp();
}

}  // namespace plsb
