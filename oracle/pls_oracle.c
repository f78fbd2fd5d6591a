/* oracle/pls_oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C CPU restatement of the reference's PLS fit / predict / cross-validation path
 * (tjhladish/PLS @ d976786, src/pls.cpp; every function cites the lines it follows).  It exists
 * so that the parity tests and bench.py's cpu_baseline leg have a checker that travels to the GPU
 * box without Eigen.  Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline leg) may
 * load it; the product library (libplsb.so) never links or calls it and has no CPU fallback.
 *
 * PARITY PINNING: this restatement is pinned against the reference itself — oracle/_ref/libplsref.so,
 * the unmodified reference compiled by oracle/Makefile — in tests/test_oracle_vs_reference.py (run
 * wherever /root/reference or a prebuilt _ref/ exists) and against the committed golden vectors in
 * tests/golden/ that were generated from that library by tests/golden/make_golden.py.
 *
 * Deliberate deviations from the reference's *code* (never from its results beyond rounding):
 *  - the dominant eigenvector of the symmetric M x M matrix XY'XY is found with cyclic Jacobi
 *    instead of Eigen::EigenSolver (Hessenberg + Francis QR, pls.cpp:406-407); same unit vector up
 *    to sign, and sign cancels in every product the model exposes except the column signs of
 *    W,P,R,Q,T (compare those after sign normalisation, as the north star allows);
 *  - wilcoxon ranks with a stable sort; the reference uses unstable std::sort (pls.h:58-69), which
 *    can only differ on exact non-zero ties of ||e1|-|e2||;
 *  - CV folds extract A components, not K (pls.cpp:477, 531 extract K; first A are bit-identical).
 * All matrices are column-major with leading dimension == rows (Eigen's layout).
 */
#include "pls_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define XE(i, j) X[(size_t)(j) * (size_t)N + (size_t)(i)]

/* ---- dominant eigenvector of a symmetric PSD m x m matrix (role of pls.cpp:113-141) ---------- */
static void jacobi_dominant(const double *G, int m, double *q) {
    double a[16 * 16], v[16 * 16];
    double *A = a, *V = v;
    int heap = 0;
    if (m > 16) { A = (double *)malloc(sizeof(double) * m * m); V = (double *)malloc(sizeof(double) * m * m); heap = 1; }
    for (int i = 0; i < m * m; ++i) { A[i] = G[i]; V[i] = 0.0; }
    for (int i = 0; i < m; ++i) V[i * m + i] = 1.0;
    for (int sweep = 0; sweep < 64; ++sweep) {
        double off = 0.0, diag = 0.0;
        for (int p = 0; p < m; ++p) {
            diag += A[p * m + p] * A[p * m + p];
            for (int r = p + 1; r < m; ++r) off += A[p * m + r] * A[p * m + r];
        }
        if (off <= 1e-60 || off <= 1e-34 * diag) break;
        for (int p = 0; p < m - 1; ++p)
            for (int r = p + 1; r < m; ++r) {
                double apr = A[p * m + r];
                if (apr == 0.0) continue;
                double theta = (A[r * m + r] - A[p * m + p]) / (2.0 * apr);
                double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < m; ++k) { /* A <- A J */
                    double akp = A[k * m + p], akr = A[k * m + r];
                    A[k * m + p] = c * akp - s * akr;
                    A[k * m + r] = s * akp + c * akr;
                }
                for (int k = 0; k < m; ++k) { /* A <- J' A */
                    double apk = A[p * m + k], ark = A[r * m + k];
                    A[p * m + k] = c * apk - s * ark;
                    A[r * m + k] = s * apk + c * ark;
                }
                for (int k = 0; k < m; ++k) {
                    double vkp = V[k * m + p], vkr = V[k * m + r];
                    V[k * m + p] = c * vkp - s * vkr;
                    V[k * m + r] = s * vkp + c * vkr;
                }
            }
    }
    /* largest |eigenvalue|, first index on ties (pls.cpp:119-126) */
    int idx = 0; double best = 0.0;
    for (int i = 0; i < m; ++i) if (fabs(A[i * m + i]) > best) { best = fabs(A[i * m + i]); idx = i; }
    double nrm = 0.0;
    for (int k = 0; k < m; ++k) nrm += V[k * m + idx] * V[k * m + idx];
    nrm = sqrt(nrm);
    for (int k = 0; k < m; ++k) q[k] = V[k * m + idx] / nrm; /* EigenSolver.h:360 normalises columns */
    if (heap) { free(A); free(V); }
}

/* ---- the fit: Dayal & MacGregor modified kernel algorithms #1/#2 (pls.cpp:390-437) ----------- */
int orc_fit(const double *X, int64_t N, int64_t K, const double *Y, int64_t M, int64_t A, int method,
            double *W, double *P, double *R, double *Q, double *T) {
    double *S = (double *)calloc((size_t)K * M, sizeof(double));   /* XY, K x M col-major */
    double *G = NULL;                                                /* XX, K x K */
    double *w = (double *)malloc(sizeof(double) * K), *r = (double *)malloc(sizeof(double) * K);
    double *p = (double *)malloc(sizeof(double) * K), *t = (double *)malloc(sizeof(double) * (N > 0 ? N : 1));
    double *q = (double *)malloc(sizeof(double) * M), *gram = (double *)malloc(sizeof(double) * M * M);
    double *qe = (double *)malloc(sizeof(double) * M);
    /* XY = X' * Y  (pls.cpp:396) */
    for (int64_t m = 0; m < M; ++m)
        for (int64_t j = 0; j < K; ++j) {
            double s = 0.0;
            for (int64_t i = 0; i < N; ++i) s += XE(i, j) * Y[m * N + i];
            S[m * K + j] = s;
        }
    if (method == 1) { /* XX = X' * X  (pls.cpp:398) */
        G = (double *)malloc(sizeof(double) * (size_t)K * K);
        for (int64_t a = 0; a < K; ++a)
            for (int64_t b = a; b < K; ++b) {
                double s = 0.0;
                for (int64_t i = 0; i < N; ++i) s += XE(i, a) * XE(i, b);
                G[a * K + b] = s; G[b * K + a] = s;
            }
    }
    for (int64_t a = 0; a < A; ++a) {
        if (M == 1) {                      /* pls.cpp:403-404 */
            for (int64_t j = 0; j < K; ++j) w[j] = S[j];
        } else {                           /* pls.cpp:406-408 */
            for (int64_t u = 0; u < M; ++u)
                for (int64_t v = 0; v < M; ++v) {
                    double s = 0.0;
                    for (int64_t j = 0; j < K; ++j) s += S[u * K + j] * S[v * K + j];
                    gram[u * M + v] = s;
                }
            jacobi_dominant(gram, (int)M, qe);
            for (int64_t j = 0; j < K; ++j) {
                double s = 0.0;
                for (int64_t m = 0; m < M; ++m) s += S[m * K + j] * qe[m];
                w[j] = s;
            }
        }
        double nw = 0.0;                   /* pls.cpp:411 */
        for (int64_t j = 0; j < K; ++j) nw += w[j] * w[j];
        nw = sqrt(nw);
        for (int64_t j = 0; j < K; ++j) { w[j] /= nw; r[j] = w[j]; }
        for (int64_t b = 0; b < a; ++b) {  /* pls.cpp:414-416: dot uses the original w */
            double c = 0.0;
            for (int64_t j = 0; j < K; ++j) c += P[b * K + j] * w[j];
            for (int64_t j = 0; j < K; ++j) r[j] -= c * R[b * K + j];
        }
        double tt = 0.0;
        if (method == 0) {                 /* pls.cpp:419-421 */
            for (int64_t i = 0; i < N; ++i) t[i] = 0.0;
            for (int64_t j = 0; j < K; ++j) { double rj = r[j]; for (int64_t i = 0; i < N; ++i) t[i] += XE(i, j) * rj; }
            for (int64_t i = 0; i < N; ++i) tt += t[i] * t[i];
            for (int64_t j = 0; j < K; ++j) { double s = 0.0; for (int64_t i = 0; i < N; ++i) s += XE(i, j) * t[i]; p[j] = s; }
        } else {                           /* pls.cpp:423-424 */
            for (int64_t j = 0; j < K; ++j) { double s = 0.0; for (int64_t k = 0; k < K; ++k) s += r[k] * G[j * K + k]; p[j] = s; }
            for (int64_t j = 0; j < K; ++j) tt += p[j] * r[j];
        }
        for (int64_t j = 0; j < K; ++j) p[j] /= tt;                 /* pls.cpp:427 */
        for (int64_t m = 0; m < M; ++m) {                           /* pls.cpp:428 */
            double s = 0.0;
            for (int64_t j = 0; j < K; ++j) s += r[j] * S[m * K + j];
            q[m] = s / tt;
        }
        for (int64_t m = 0; m < M; ++m)                             /* pls.cpp:429 */
            for (int64_t j = 0; j < K; ++j) S[m * K + j] -= (p[j] * q[m]) * tt;
        for (int64_t j = 0; j < K; ++j) { W[a * K + j] = w[j]; P[a * K + j] = p[j]; R[a * K + j] = r[j]; }
        for (int64_t m = 0; m < M; ++m) Q[a * M + m] = q[m];
        if (method == 0 && T) for (int64_t i = 0; i < N; ++i) T[a * N + i] = t[i];   /* pls.cpp:434 */
    }
    free(S); free(G); free(w); free(r); free(p); free(t); free(q); free(gram); free(qe);
    return 0;
}

/* ---- prediction helpers (pls.cpp:439-467) ------------------------------------------------------ */
void orc_coefficients(const double *R, const double *Q, int64_t K, int64_t M, int64_t comp, double *B) {
    for (int64_t m = 0; m < M; ++m)      /* B = R_c Q_c'  (pls.cpp:446) */
        for (int64_t j = 0; j < K; ++j) {
            double s = 0.0;
            for (int64_t a = 0; a < comp; ++a) s += R[a * K + j] * Q[a * M + m];
            B[m * K + j] = s;
        }
}
void orc_scores(const double *R, int64_t K, int64_t comp, const double *Xn, int64_t Nn, double *S) {
    for (int64_t a = 0; a < comp; ++a)   /* X_new R_c  (pls.cpp:441) */
        for (int64_t i = 0; i < Nn; ++i) {
            double s = 0.0;
            for (int64_t j = 0; j < K; ++j) s += Xn[j * Nn + i] * R[a * K + j];
            S[a * Nn + i] = s;
        }
}
void orc_fitted_values(const double *R, const double *Q, int64_t K, int64_t M, int64_t comp,
                       const double *Xn, int64_t Nn, double *F) {
    double *B = (double *)malloc(sizeof(double) * K * M);
    orc_coefficients(R, Q, K, M, comp, B);
    for (int64_t m = 0; m < M; ++m)      /* X_new B  (pls.cpp:450) */
        for (int64_t i = 0; i < Nn; ++i) {
            double s = 0.0;
            for (int64_t j = 0; j < K; ++j) s += Xn[j * Nn + i] * B[m * K + j];
            F[m * Nn + i] = s;
        }
    free(B);
}
void orc_residuals(const double *R, const double *Q, int64_t K, int64_t M, int64_t comp,
                   const double *Xn, int64_t Nn, const double *Yn, double *E) {
    orc_fitted_values(R, Q, K, M, comp, Xn, Nn, E);
    for (int64_t i = 0; i < Nn * M; ++i) E[i] = Yn[i] - E[i];      /* pls.cpp:454 */
}
void orc_sse(const double *R, const double *Q, int64_t K, int64_t M, int64_t comp, const double *Xn,
             int64_t Nn, const double *Yn, double *out) {
    double *E = (double *)malloc(sizeof(double) * Nn * M);
    orc_residuals(R, Q, K, M, comp, Xn, Nn, Yn, E);
    for (int64_t m = 0; m < M; ++m) {    /* pls.cpp:458 */
        double s = 0.0;
        for (int64_t i = 0; i < Nn; ++i) s += E[m * Nn + i] * E[m * Nn + i];
        out[m] = s;
    }
    free(E);
}
void orc_sst(const double *X, int64_t N, int64_t K, double *out) {  /* pls.cpp:69-77 */
    for (int64_t j = 0; j < K; ++j) {
        if (N < 2) { out[j] = 0.0; continue; }
        double mean = 0.0;
        for (int64_t i = 0; i < N; ++i) mean += XE(i, j);
        mean /= (double)N;
        double s = 0.0;
        for (int64_t i = 0; i < N; ++i) { double d = XE(i, j) - mean; s += d * d; }
        out[j] = s;
    }
}
void orc_explained_variance(const double *R, const double *Q, int64_t K, int64_t M, int64_t comp,
                            const double *Xn, int64_t Nn, const double *Yn, double *out) {
    double *sst = (double *)malloc(sizeof(double) * M);
    orc_sse(R, Q, K, M, comp, Xn, Nn, Yn, out);
    orc_sst(Yn, Nn, M, sst);
    for (int64_t m = 0; m < M; ++m) out[m] = 1.0 - out[m] / sst[m];  /* pls.cpp:465 */
    free(sst);
}
void orc_colwise_stdev(const double *X, int64_t N, int64_t K, double *out) {  /* pls.cpp:79-87 */
    orc_sst(X, N, K, out);
    for (int64_t j = 0; j < K; ++j) out[j] = sqrt(out[j] / ((double)N - 1.0));
}
void orc_colwise_z_scores(const double *X, int64_t N, int64_t K, double *Z) {  /* pls.cpp:93-111 */
    double *sd = (double *)malloc(sizeof(double) * K);
    orc_colwise_stdev(X, N, K, sd);
    for (int64_t j = 0; j < K; ++j) {
        double mean = 0.0;
        for (int64_t i = 0; i < N; ++i) mean += XE(i, j);
        mean /= (double)N;
        /* divides by the ORIGINAL stdev: a constant column yields NaN, as in the reference (B-4) */
        for (int64_t i = 0; i < N; ++i) Z[j * N + i] = (XE(i, j) - mean) / sd[j];
    }
    free(sd);
}
void orc_predict_residuals(const double *R, const double *Q, int64_t K, int64_t M, int64_t A,
                           const double *Xn, int64_t Nn, const double *Yn, double *E) {
    double *res = (double *)malloc(sizeof(double) * Nn * M);       /* pls.cpp:494-510 */
    for (int64_t c = 1; c <= A; ++c) {
        orc_residuals(R, Q, K, M, c, Xn, Nn, Yn, res);
        for (int64_t m = 0; m < M; ++m)
            for (int64_t i = 0; i < Nn; ++i) E[m * Nn * A + (c - 1) * Nn + i] = res[m * Nn + i];
    }
    free(res);
}

/* ---- cross-validation with explicit fold lists (pls.cpp:469-491 LOO, 512-549 LSO) ------------ */
int orc_cv_folds(const double *X, int64_t N, int64_t K, const double *Y, int64_t M, int64_t A, int method,
                 int64_t nfolds, int64_t test_size, const int64_t *test_idx, int64_t train_size,
                 const int64_t *train_idx, double *E) {
    const int64_t rows = nfolds * test_size;
    double *Xv = (double *)malloc(sizeof(double) * train_size * K), *Yv = (double *)malloc(sizeof(double) * train_size * M);
    double *Xp = (double *)malloc(sizeof(double) * test_size * K), *Yp = (double *)malloc(sizeof(double) * test_size * M);
    double *W = (double *)malloc(sizeof(double) * K * A), *P = (double *)malloc(sizeof(double) * K * A);
    double *R = (double *)malloc(sizeof(double) * K * A), *Q = (double *)malloc(sizeof(double) * M * A);
    double *res = (double *)malloc(sizeof(double) * test_size * M);
    int64_t *tr = (int64_t *)malloc(sizeof(int64_t) * train_size);
    char *is_test = (char *)malloc(N);
    for (int64_t f = 0; f < nfolds; ++f) {
        const int64_t *te = test_idx + f * test_size;
        if (train_idx) memcpy(tr, train_idx + f * train_size, sizeof(int64_t) * train_size);
        else {
            memset(is_test, 0, N);
            for (int64_t i = 0; i < test_size; ++i) is_test[te[i]] = 1;
            int64_t o = 0;
            for (int64_t i = 0; i < N; ++i) if (!is_test[i]) tr[o++] = i;
        }
        for (int64_t j = 0; j < K; ++j) {
            for (int64_t i = 0; i < train_size; ++i) Xv[j * train_size + i] = XE(tr[i], j);
            for (int64_t i = 0; i < test_size; ++i) Xp[j * test_size + i] = XE(te[i], j);
        }
        for (int64_t m = 0; m < M; ++m) {
            for (int64_t i = 0; i < train_size; ++i) Yv[m * train_size + i] = Y[m * N + tr[i]];
            for (int64_t i = 0; i < test_size; ++i) Yp[m * test_size + i] = Y[m * N + te[i]];
        }
        orc_fit(Xv, train_size, K, Yv, M, A, method, W, P, R, Q, NULL);
        for (int64_t c = 1; c <= A; ++c) {
            orc_residuals(R, Q, K, M, c, Xp, test_size, Yp, res);
            for (int64_t m = 0; m < M; ++m)
                for (int64_t i = 0; i < test_size; ++i)
                    E[m * rows * A + (c - 1) * rows + f * test_size + i] = res[m * test_size + i];
        }
    }
    free(Xv); free(Yv); free(Xp); free(Yp); free(W); free(P); free(R); free(Q); free(res); free(tr); free(is_test);
    return 0;
}

/* ---- validation statistics (pls.cpp:152-289) ---------------------------------------------------- */
void orc_validation(const double *E, int64_t M, int64_t rows, int64_t A, int out_type, double *out) {
    for (int64_t m = 0; m < M; ++m)      /* pls.cpp:244-258; out_type 0 = RESS, 1 = MSE */
        for (int64_t c = 0; c < A; ++c) {
            double s = 0.0;
            for (int64_t i = 0; i < rows; ++i) { double e = E[m * rows * A + c * rows + i]; s += e * e; }
            out[c * M + m] = out_type == 0 ? s : s / (double)rows;
        }
}
double orc_normalcdf(double z) {          /* pls.cpp:152-160 */
    const double c1 = 0.196854, c2 = 0.115194, c3 = 0.000344, c4 = 0.019527;
    double zs = fabs(z);
    double p = 0.5 / pow(1 + c1 * zs + c2 * zs * zs + c3 * zs * zs * zs + c4 * zs * zs * zs * zs, 4);
    return z < 0 ? p : 1.0 - p;
}
typedef struct { double v; int64_t i; } orc_vi;
static int orc_vi_cmp(const void *a, const void *b) {
    const orc_vi *x = (const orc_vi *)a, *y = (const orc_vi *)b;
    if (x->v < y->v) return -1;
    if (x->v > y->v) return 1;
    return x->i < y->i ? -1 : (x->i > y->i ? 1 : 0);
}
double orc_wilcoxon(const double *e1, const double *e2, int64_t n) {   /* pls.cpp:190-211 */
    orc_vi *s = (orc_vi *)malloc(sizeof(orc_vi) * (n > 0 ? n : 1));
    int *sg = (int *)malloc(sizeof(int) * (n > 0 ? n : 1));
    for (int64_t i = 0; i < n; ++i) {
        double del = fabs(e1[i]) - fabs(e2[i]);
        sg[i] = (0 < del) - (del < 0);
        s[i].v = fabs(del); s[i].i = i;
    }
    qsort(s, n, sizeof(orc_vi), orc_vi_cmp);
    double d = 0.0;
    for (int64_t i = 0; i < n; ++i) d += (double)(i + 1) * sg[s[i].i];
    double t = (double)(n * (n + 1)) / 2.0;
    double v = (t - d) / 2.0, ev = t / 2.0;
    double sv = sqrt((double)(n * (n + 1) * (2 * n + 1)) / 24.0);
    double z = (v - ev) / sv;
    free(s); free(sg);
    return 1.0 - orc_normalcdf(z);
}
void orc_optimal_num_components(const double *E, int64_t M, int64_t rows, int64_t A, double alpha,
                                int64_t *out) {                          /* pls.cpp:265-289 */
    double *press = (double *)malloc(sizeof(double) * M * A);
    orc_validation(E, M, rows, A, 0, press);
    for (int64_t m = 0; m < M; ++m) {
        int64_t best = 0;
        for (int64_t c = 1; c < A; ++c) if (press[c * M + m] < press[best * M + m]) best = c;  /* first minimum */
        int64_t pick = best;
        const double *err1 = E + m * rows * A + best * rows;
        for (int64_t alt = 0; alt < best; ++alt) {
            const double *err2 = E + m * rows * A + alt * rows;
            if (orc_wilcoxon(err1, err2, rows) > alpha) { pick = alt; break; }
        }
        out[m] = pick + 1;
    }
    free(press);
}

/* ---- deterministic synthetic data; integer hash => identical bits on CPU and GPU */
static uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
double orc_synth_u(uint64_t seed, uint64_t idx) {  /* Irwin-Hall(3) on 20-bit integers: mean 0, var 0.25 */
    uint64_t h = splitmix64(splitmix64(seed) + idx);
    uint64_t s = (h & 0xFFFFF) + ((h >> 20) & 0xFFFFF) + ((h >> 40) & 0xFFFFF);
    return (double)s / 1048576.0 - 1.5;
}
/* X = (Z C) Bx + 0.5 Ex,  Y = (Z C) By + 0.1 Ey,  C = diag(4/(4+l)); sums in index order, no FMA
 * (compile with -ffp-contract=off).  Z uses global row indices so any row block can be generated. */
void orc_synth(uint64_t seed, int64_t row0, int64_t rows, int64_t K, int64_t M, int64_t L,
               double *X, int64_t ldx, double *Y, int64_t ldy) {
    double *z = (double *)malloc(sizeof(double) * L);
    double *Bx = (double *)malloc(sizeof(double) * L * K), *By = (double *)malloc(sizeof(double) * L * M);
    for (int64_t l = 0; l < L; ++l) {
        for (int64_t j = 0; j < K; ++j) Bx[l * K + j] = orc_synth_u(seed + 1, (uint64_t)(l * K + j));
        for (int64_t m = 0; m < M; ++m) By[l * M + m] = orc_synth_u(seed + 2, (uint64_t)(l * M + m));
    }
    for (int64_t ii = 0; ii < rows; ++ii) {
        uint64_t i = (uint64_t)(row0 + ii);
        for (int64_t l = 0; l < L; ++l) z[l] = orc_synth_u(seed, i * (uint64_t)L + (uint64_t)l) * (4.0 / (4.0 + (double)l));
        for (int64_t j = 0; j < K; ++j) {
            double acc = 0.0;
            for (int64_t l = 0; l < L; ++l) acc = acc + z[l] * Bx[l * K + j];
            X[j * ldx + ii] = acc + 0.5 * orc_synth_u(seed + 3, i * (uint64_t)K + (uint64_t)j);
        }
        for (int64_t m = 0; m < M; ++m) {
            double acc = 0.0;
            for (int64_t l = 0; l < L; ++l) acc = acc + z[l] * By[l * M + m];
            Y[m * ldy + ii] = acc + 0.1 * orc_synth_u(seed + 4, i * (uint64_t)M + (uint64_t)m);
        }
    }
    free(z); free(Bx); free(By);
}
