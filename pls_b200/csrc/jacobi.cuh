// pls_b200/csrc/jacobi.cuh — dominant eigenvector of a small symmetric PSD matrix, one warp.
//
// Role of find_dominant_ev / dominant_eigenvector (reference src/pls.cpp:113-141) applied to the
// M x M matrix XY'XY (pls.cpp:406-407).  The reference calls Eigen::EigenSolver (Hessenberg + Francis
// QR); the matrix is symmetric, so cyclic Jacobi yields the same unit eigenvector up to sign.  The sign
// cancels in coefficients / fitted values and is normalised away when scores / loadings are compared.
#pragma once
#include "common.cuh"

#define PLSB_MAX_M 32

namespace plsb {

// A (m x m, row-major, symmetric) is destroyed; V receives the eigenvectors (columns); q the unit
// eigenvector of the eigenvalue with the largest magnitude (first index on ties, pls.cpp:119-126).
// Must be called by all 32 lanes of one warp; A, V, q in shared memory.
__device__ __forceinline__ void jacobi_dominant_warp(double *A, double *V, int m, double *q) {
    const int lane = threadIdx.x & 31;
    for (int i = lane; i < m * m; i += 32) V[i] = ((i / m) == (i % m)) ? 1.0 : 0.0;
    __syncwarp();
    for (int sweep = 0; sweep < 64; ++sweep) {
        double off = 0.0, diag = 0.0;
        for (int i = lane; i < m * m; i += 32) {
            const int rr = i / m, cc = i % m;
            const double v = A[i];
            if (rr == cc) diag = fma(v, v, diag);
            else if (rr < cc) off = fma(v, v, off);
        }
        off = warp_sum(off);
        diag = warp_sum(diag);
        if (off <= 1e-60 || off <= 1e-34 * diag) break;
        for (int p = 0; p < m - 1; ++p)
            for (int r = p + 1; r < m; ++r) {
                const double apr = A[p * m + r];
                if (apr == 0.0) continue;  // uniform across the warp
                const double theta = (A[r * m + r] - A[p * m + p]) / (2.0 * apr);
                const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(fma(theta, theta, 1.0)));
                const double c = 1.0 / sqrt(fma(t, t, 1.0)), s = t * c;
                __syncwarp();
                if (lane < m) {  // A <- A J, V <- V J  (columns p, r; lane = row)
                    const int k = lane;
                    const double akp = A[k * m + p], akr = A[k * m + r];
                    A[k * m + p] = c * akp - s * akr;
                    A[k * m + r] = s * akp + c * akr;
                    const double vkp = V[k * m + p], vkr = V[k * m + r];
                    V[k * m + p] = c * vkp - s * vkr;
                    V[k * m + r] = s * vkp + c * vkr;
                }
                __syncwarp();
                if (lane < m) {  // A <- J' A  (rows p, r; lane = column)
                    const int k = lane;
                    const double apk = A[p * m + k], ark = A[r * m + k];
                    A[p * m + k] = c * apk - s * ark;
                    A[r * m + k] = s * apk + c * ark;
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
