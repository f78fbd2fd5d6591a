/* oracle/pls_oracle_ext.c — TEST INFRASTRUCTURE ONLY (golden-vector generation; never loaded by the product).
 *
 * The fit of pls_oracle.c:orc_fit (reference src/pls.cpp:390-437, KERNEL_TYPE1) carried out in x87 EXTENDED
 * precision (long double, 64-bit significand): every vector of the recurrence (S, w, r, t, p, q) is kept in long
 * double and every sum is accumulated in long double; only X and Y are the given FP64 bits.  It is ~11 bits more
 * precise than any FP64 evaluation of the same algorithm, so the distance of an FP64 result from it measures that
 * evaluation's rounding error — the yardstick for "how far apart may two correct FP64 fits be" at sizes where the
 * late components are ill-conditioned (tests/golden/make_golden_fullsize.py records these distances for the
 * reference, the C oracle and — via the committed fixtures — the CUDA path).
 * OpenMP over rows / columns; sums inside one t_i or p_j run in index order whatever the thread count.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

typedef long double ld;
#define XE(i, j) X[(size_t)(j) * (size_t)N + (size_t)(i)]

static void jacobi_dominant_ld(const ld *G, int m, ld *q) {
    ld *A = (ld *)malloc(sizeof(ld) * m * m), *V = (ld *)malloc(sizeof(ld) * m * m);
    for (int i = 0; i < m * m; ++i) { A[i] = G[i]; V[i] = 0.0L; }
    for (int i = 0; i < m; ++i) V[i * m + i] = 1.0L;
    for (int sweep = 0; sweep < 100; ++sweep) {
        ld off = 0.0L, diag = 0.0L;
        for (int p = 0; p < m; ++p) {
            diag += A[p * m + p] * A[p * m + p];
            for (int r = p + 1; r < m; ++r) off += A[p * m + r] * A[p * m + r];
        }
        if (off <= 1e-80L || off <= 1e-42L * diag) break;
        for (int p = 0; p < m - 1; ++p)
            for (int r = p + 1; r < m; ++r) {
                ld apr = A[p * m + r];
                if (apr == 0.0L) continue;
                ld theta = (A[r * m + r] - A[p * m + p]) / (2.0L * apr);
                ld t = (theta >= 0 ? 1.0L : -1.0L) / (fabsl(theta) + sqrtl(theta * theta + 1.0L));
                ld c = 1.0L / sqrtl(t * t + 1.0L), s = t * c;
                for (int k = 0; k < m; ++k) { ld akp = A[k * m + p], akr = A[k * m + r]; A[k * m + p] = c * akp - s * akr; A[k * m + r] = s * akp + c * akr; }
                for (int k = 0; k < m; ++k) { ld apk = A[p * m + k], ark = A[r * m + k]; A[p * m + k] = c * apk - s * ark; A[r * m + k] = s * apk + c * ark; }
                for (int k = 0; k < m; ++k) { ld vkp = V[k * m + p], vkr = V[k * m + r]; V[k * m + p] = c * vkp - s * vkr; V[k * m + r] = s * vkp + c * vkr; }
            }
    }
    int idx = 0; ld best = 0.0L;
    for (int i = 0; i < m; ++i) if (fabsl(A[i * m + i]) > best) { best = fabsl(A[i * m + i]); idx = i; }
    ld nrm = 0.0L;
    for (int k = 0; k < m; ++k) nrm += V[k * m + idx] * V[k * m + idx];
    nrm = sqrtl(nrm);
    for (int k = 0; k < m; ++k) q[k] = V[k * m + idx] / nrm;
    free(A); free(V);
}

/* outputs rounded to FP64 at the very end: W,P,R (K x A), Q (M x A), T (N x A or NULL), column-major */
int orc_fit_ext(const double *X, int64_t N, int64_t K, const double *Y, int64_t M, int64_t A,
                double *W, double *P, double *R, double *Q, double *T) {
    ld *S = (ld *)calloc((size_t)K * M, sizeof(ld));
    ld *w = (ld *)malloc(sizeof(ld) * K), *r = (ld *)malloc(sizeof(ld) * K), *p = (ld *)malloc(sizeof(ld) * K);
    ld *t = (ld *)malloc(sizeof(ld) * (N > 0 ? N : 1));
    ld *q = (ld *)malloc(sizeof(ld) * M), *gram = (ld *)malloc(sizeof(ld) * M * M), *qe = (ld *)malloc(sizeof(ld) * M);
    ld *Pl = (ld *)calloc((size_t)K * A, sizeof(ld)), *Rl = (ld *)calloc((size_t)K * A, sizeof(ld));
#pragma omp parallel for collapse(2) schedule(static)
    for (int64_t m = 0; m < M; ++m)
        for (int64_t j = 0; j < K; ++j) {
            ld s = 0.0L;
            for (int64_t i = 0; i < N; ++i) s += (ld)XE(i, j) * (ld)Y[m * N + i];
            S[m * K + j] = s;
        }
    for (int64_t a = 0; a < A; ++a) {
        if (M == 1) {
            for (int64_t j = 0; j < K; ++j) w[j] = S[j];
        } else {
            for (int64_t u = 0; u < M; ++u)
                for (int64_t v = 0; v < M; ++v) {
                    ld s = 0.0L;
                    for (int64_t j = 0; j < K; ++j) s += S[u * K + j] * S[v * K + j];
                    gram[u * M + v] = s;
                }
            jacobi_dominant_ld(gram, (int)M, qe);
            for (int64_t j = 0; j < K; ++j) {
                ld s = 0.0L;
                for (int64_t m = 0; m < M; ++m) s += S[m * K + j] * qe[m];
                w[j] = s;
            }
        }
        ld nw = 0.0L;
        for (int64_t j = 0; j < K; ++j) nw += w[j] * w[j];
        nw = sqrtl(nw);
        for (int64_t j = 0; j < K; ++j) { w[j] /= nw; r[j] = w[j]; }
        for (int64_t b = 0; b < a; ++b) {
            ld c = 0.0L;
            for (int64_t j = 0; j < K; ++j) c += Pl[b * K + j] * w[j];
            for (int64_t j = 0; j < K; ++j) r[j] -= c * Rl[b * K + j];
        }
        /* t = X r, row blocks so that the long-double t block stays in cache while the columns stream */
#pragma omp parallel for schedule(static)
        for (int64_t i0 = 0; i0 < N; i0 += 2048) {
            const int64_t i1 = i0 + 2048 < N ? i0 + 2048 : N;
            for (int64_t i = i0; i < i1; ++i) t[i] = 0.0L;
            for (int64_t j = 0; j < K; ++j) {
                const ld rj = r[j];
                for (int64_t i = i0; i < i1; ++i) t[i] += (ld)XE(i, j) * rj;
            }
        }
        ld tt = 0.0L;
        for (int64_t i = 0; i < N; ++i) tt += t[i] * t[i];
#pragma omp parallel for schedule(static)
        for (int64_t j = 0; j < K; ++j) {
            ld s = 0.0L;
            for (int64_t i = 0; i < N; ++i) s += (ld)XE(i, j) * t[i];
            p[j] = s / tt;
        }
        for (int64_t m = 0; m < M; ++m) {
            ld s = 0.0L;
            for (int64_t j = 0; j < K; ++j) s += r[j] * S[m * K + j];
            q[m] = s / tt;
        }
        for (int64_t m = 0; m < M; ++m)
            for (int64_t j = 0; j < K; ++j) S[m * K + j] -= (p[j] * q[m]) * tt;
        for (int64_t j = 0; j < K; ++j) {
            Pl[a * K + j] = p[j]; Rl[a * K + j] = r[j];
            W[a * K + j] = (double)w[j]; P[a * K + j] = (double)p[j]; R[a * K + j] = (double)r[j];
        }
        for (int64_t m = 0; m < M; ++m) Q[a * M + m] = (double)q[m];
        if (T) for (int64_t i = 0; i < N; ++i) T[a * N + i] = (double)t[i];
    }
    free(S); free(w); free(r); free(p); free(t); free(q); free(gram); free(qe); free(Pl); free(Rl);
    return 0;
}
