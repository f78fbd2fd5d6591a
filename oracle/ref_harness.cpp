// oracle/ref_harness.cpp — TEST INFRASTRUCTURE ONLY.
//
// A plain-C ABI wrapped around the UNMODIFIED reference implementation
// (/root/reference/src/pls.cpp + include/PLS/pls.h, Eigen from lib/eigen).
// It is compiled by oracle/Makefile from the sources where they lie under
// /root/reference into oracle/_ref/libplsref.so (git-ignored, travels to the
// GPU box).  No reference source is copied into this repository; this file
// only *calls* the reference's public API (plus its private members through
// the access trick below, needed because W,P,R,Q,T have no public getter and
// loadingsX/loadingsY are declared but never defined, pls.h:207-211).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load the resulting library.
//
// All matrices cross this ABI as column-major FP64 (Eigen's native layout,
// leading dimension == rows).  Complex-typed reference members are returned
// as their real parts; the maximum |imag| seen is reported by
// plsref_model_max_imag so tests can assert it is exactly zero.

#include <Eigen/Dense>
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <numeric>
#include <random>
#include <string>
#include <vector>

#define class struct
#define private public
#include <PLS/pls.h>
#undef private
#undef class

using PLS::Model;
using PLS::Residual;

namespace {

typedef Eigen::Map<const Mat2D> CMap;
typedef Eigen::Map<Mat2D> MMap;

inline PLS::METHOD to_method(int m) { return m == 1 ? PLS::KERNEL_TYPE2 : PLS::KERNEL_TYPE1; }

void put_real(const Mat2Dc &src, double *dst) {
    if (!dst) return;
    for (Eigen::Index j = 0; j < src.cols(); ++j)
        for (Eigen::Index i = 0; i < src.rows(); ++i) dst[j * src.rows() + i] = src(i, j).real();
}

double max_imag(const Mat2Dc &m) {
    double v = 0;
    for (Eigen::Index j = 0; j < m.cols(); ++j)
        for (Eigen::Index i = 0; i < m.rows(); ++i) v = std::max(v, std::abs(m(i, j).imag()));
    return v;
}

// Ev (vector of M matrices rows x A) -> M stacked column-major blocks
void put_residual(const Residual &res, double *E) {
    const auto ev = res.errors();
    size_t off = 0;
    for (const auto &m : ev) {
        std::memcpy(E + off, m.data(), sizeof(double) * m.size());
        off += m.size();
    }
}

Residual make_residual(const double *E, int64_t M, int64_t rows, int64_t A, const char *label) {
    std::vector<Mat2D> ev;
    for (int64_t y = 0; y < M; ++y) ev.push_back(CMap(E + y * rows * A, rows, A));
    return Residual(ev, label);
}

} // namespace

extern "C" {

// ---- model lifecycle (pls.cpp:340-359) ------------------------------------------------------
void *plsref_model_new(const double *X, int64_t N, int64_t K, const double *Y, int64_t M,
                       int method, int64_t A) {
    Mat2D Xm = CMap(X, N, K), Ym = CMap(Y, N, M);
    return new Model(Xm, Ym, to_method(method), static_cast<size_t>(A));
}
void plsref_model_free(void *h) { delete static_cast<Model *>(h); }

// refit on new data of the same K, M (pls.cpp:390-437)
void plsref_model_plsr(void *h, const double *X, int64_t N, int64_t K, const double *Y, int64_t M,
                       int method) {
    Mat2D Xm = CMap(X, N, K), Ym = CMap(Y, N, M);
    static_cast<Model *>(h)->plsr(Xm, Ym, to_method(method));
}

// W,P,R: K x A; Q: M x A; T: N x A (type 1 only; pass NULL to skip). Real parts.
void plsref_model_state(void *h, double *W, double *P, double *R, double *Q, double *T) {
    Model *m = static_cast<Model *>(h);
    put_real(m->W, W); put_real(m->P, P); put_real(m->R, R); put_real(m->Q, Q);
    if (T && m->T.size()) put_real(m->T, T);
}
double plsref_model_max_imag(void *h) {
    Model *m = static_cast<Model *>(h);
    return std::max({max_imag(m->W), max_imag(m->P), max_imag(m->R), max_imag(m->Q), max_imag(m->T)});
}

// ---- prediction helpers (pls.cpp:439-467) ---------------------------------------------------
void plsref_scores(void *h, const double *Xn, int64_t Nn, int64_t K, int64_t comp, double *S) {
    put_real(static_cast<Model *>(h)->scores(CMap(Xn, Nn, K), comp), S);
}
void plsref_coefficients(void *h, int64_t comp, double *B) {
    put_real(static_cast<Model *>(h)->coefficients(comp), B);
}
void plsref_fitted_values(void *h, const double *Xn, int64_t Nn, int64_t K, int64_t M, int64_t comp, double *F) {
    MMap(F, Nn, M) = static_cast<Model *>(h)->fitted_values(CMap(Xn, Nn, K), comp);
}
void plsref_residuals(void *h, const double *Xn, int64_t Nn, int64_t K, const double *Yn, int64_t M,
                      int64_t comp, double *E) {
    MMap(E, Nn, M) = static_cast<Model *>(h)->residuals(CMap(Xn, Nn, K), CMap(Yn, Nn, M), comp);
}
void plsref_sse(void *h, const double *Xn, int64_t Nn, int64_t K, const double *Yn, int64_t M,
                int64_t comp, double *out) {
    Eigen::Map<Row>(out, M) = static_cast<Model *>(h)->SSE(CMap(Xn, Nn, K), CMap(Yn, Nn, M), comp);
}
void plsref_explained_variance(void *h, const double *Xn, int64_t Nn, int64_t K, const double *Yn,
                               int64_t M, int64_t comp, double *out) {
    Eigen::Map<Row>(out, M) =
        static_cast<Model *>(h)->explained_variance(CMap(Xn, Nn, K), CMap(Yn, Nn, M), comp);
}

// ---- cross-validation (pls.cpp:469-549) -----------------------------------------------------
// The reference's own drivers, unmodified: inner models extract K components (pls.cpp:477, 531).
// E = M stacked (rows x A) column-major blocks; rows = N (LOO), num_trials*test_size (LSO).
void plsref_cv_loo(void *h, double *E) { put_residual(static_cast<Model *>(h)->cv_LOO(), E); }
void plsref_cv_lso(void *h, double frac, int64_t trials, uint32_t seed, double *E) {
    std::mt19937 rng(seed);
    put_residual(static_cast<Model *>(h)->cv_LSO(frac, trials, rng), E);
}
void plsref_cv_new_data(void *h, const double *Xn, int64_t Nn, int64_t K, const double *Yn, int64_t M,
                        double *E) {
    put_residual(static_cast<Model *>(h)->cv_NEW_DATA(CMap(Xn, Nn, K), CMap(Yn, Nn, M)), E);
}

// The fold index lists cv_LSO would draw (rand_nchoosek pls.cpp:218-227, driver pls.cpp:522-534):
// train[trials*train_size], test[trials*test_size].
void plsref_lso_folds(int64_t N, double frac, int64_t trials, uint32_t seed, int64_t *train, int64_t *test) {
    const size_t test_size = static_cast<size_t>(frac * N + 0.5);
    const size_t train_size = N - test_size;
    std::vector<Eigen::Index> sample(train_size), complement(test_size), full(N);
    std::iota(full.begin(), full.end(), 0);
    std::mt19937 rng(seed);
    for (int64_t rep = 0; rep < trials; ++rep) {
        PLS::rand_nchoosek(rng, full, sample, complement);
        for (size_t i = 0; i < train_size; ++i) train[rep * train_size + i] = sample[i];
        for (size_t i = 0; i < test_size; ++i) test[rep * test_size + i] = complement[i];
    }
}

// What the caller's generator yields next after cv_LSO(frac, trials, rng) returns: cv_LSO takes `rng` by reference
// (pls.cpp:512) and advances it by one rand_nchoosek per trial (pls.cpp:534).  The Model is the reference's own; the
// residuals are discarded.
uint32_t plsref_cv_lso_rng_after(void *h, double frac, int64_t trials, uint32_t seed) {
    std::mt19937 rng(seed);
    static_cast<Model *>(h)->cv_LSO(frac, trials, rng);
    return (uint32_t)rng();
}

// Per-fold refit with the reference's own plsr, but with A (not K) inner components — the
// residuals for the first A components are bit-identical. Explicit fold lists:
// test_idx[nfolds*test_size]; train_idx[nfolds*train_size] (NULL => ascending complement, LOO order).
// rows of E: fold-major, test rows in list order (the Ev layout of pls.cpp:475, 520).
// nthreads > 1 spreads folds over std::threads? -- no: the reference is single threaded; callers
// wanting all cores call this from several host threads over disjoint fold ranges.
void plsref_cv_folds(const double *X, int64_t N, int64_t K, const double *Y, int64_t M, int method,
                     int64_t A, int64_t nfolds, int64_t test_size, const int64_t *test_idx,
                     int64_t train_size, const int64_t *train_idx, double *E, int64_t e_rows,
                     int64_t e_row0) {
    CMap Xm(X, N, K), Ym(Y, N, M);
    Mat2D Xv(train_size, K), Yv(train_size, M), Xp(test_size, K), Yp(test_size, M);
    std::vector<char> is_test(N);
    for (int64_t f = 0; f < nfolds; ++f) {
        const int64_t *te = test_idx + f * test_size;
        if (train_idx) {
            const int64_t *tr = train_idx + f * train_size;
            for (int64_t i = 0; i < train_size; ++i) { Xv.row(i) = Xm.row(tr[i]); Yv.row(i) = Ym.row(tr[i]); }
        } else {
            std::fill(is_test.begin(), is_test.end(), 0);
            for (int64_t i = 0; i < test_size; ++i) is_test[te[i]] = 1;
            int64_t o = 0;
            for (int64_t i = 0; i < N; ++i) if (!is_test[i]) { Xv.row(o) = Xm.row(i); Yv.row(o) = Ym.row(i); ++o; }
        }
        for (int64_t i = 0; i < test_size; ++i) { Xp.row(i) = Xm.row(te[i]); Yp.row(i) = Ym.row(te[i]); }
        Model m(Xv, Yv, to_method(method), static_cast<size_t>(A));
        for (int64_t c = 1; c <= A; ++c) {
            Mat2D res = m.residuals(Xp, Yp, c);
            for (int64_t y = 0; y < M; ++y)
                for (int64_t i = 0; i < test_size; ++i)
                    E[y * e_rows * A + (c - 1) * e_rows + e_row0 + f * test_size + i] = res(i, y);
        }
    }
}

// ---- validation statistics (pls.cpp:152-289) ------------------------------------------------
void plsref_validation(const double *E, int64_t M, int64_t rows, int64_t A, int out_type, double *out) {
    Mat2D v = PLS::validation(make_residual(E, M, rows, A, "X"), out_type == 0 ? PLS::RESS : PLS::MSE);
    std::memcpy(out, v.data(), sizeof(double) * v.size()); // M x A column-major
}
void plsref_optimal_num_components(const double *E, int64_t M, int64_t rows, int64_t A, double alpha,
                                   int64_t *out) {
    Colsz r = PLS::optimal_num_components(make_residual(E, M, rows, A, "X"), alpha);
    for (int64_t i = 0; i < M; ++i) out[i] = static_cast<int64_t>(r[i]);
}
double plsref_wilcoxon(const double *e1, const double *e2, int64_t n) {
    return PLS::wilcoxon(Eigen::Map<const Col>(e1, n), Eigen::Map<const Col>(e2, n));
}
double plsref_normalcdf(double z) { return PLS::normalcdf(z); }

// ---- column statistics (pls.cpp:69-111) -----------------------------------------------------
void plsref_colwise_z_scores(const double *X, int64_t N, int64_t K, double *Z) {
    MMap(Z, N, K) = PLS::colwise_z_scores(CMap(X, N, K));
}
void plsref_colwise_stdev(const double *X, int64_t N, int64_t K, double *out) {
    Eigen::Map<Row>(out, K) = PLS::colwise_stdev(CMap(X, N, K));
}
void plsref_sst(const double *X, int64_t N, int64_t K, double *out) {
    Eigen::Map<Row>(out, K) = PLS::SST(CMap(X, N, K));
}

// ---- csv (pls.cpp:37-67) ---------------------------------------------------------------------
// returns rows*cols; fills dims; data copied column-major into out if out != NULL and cap suffices
int64_t plsref_read_matrix_file(const char *path, int64_t *rows, int64_t *cols, double *out, int64_t cap) {
    Mat2D m = PLS::read_matrix_file(path);
    *rows = m.rows(); *cols = m.cols();
    if (out && cap >= m.size()) std::memcpy(out, m.data(), sizeof(double) * m.size());
    return m.size();
}

// ---- timing helper for the CPU baseline (bench.py only) -------------------------------------
// Times `reps` full constructions Model(X,Y,method,A) (copy of X,Y + plsr, pls.cpp:340-353).
double plsref_time_fit(const double *X, int64_t N, int64_t K, const double *Y, int64_t M, int method,
                       int64_t A, int reps) {
    Mat2D Xm = CMap(X, N, K), Ym = CMap(Y, N, M);
    auto t0 = std::chrono::steady_clock::now();
    double sink = 0;
    for (int r = 0; r < reps; ++r) {
        Model m(Xm, Ym, to_method(method), static_cast<size_t>(A));
        sink += m.Q(0, 0).real();
    }
    auto t1 = std::chrono::steady_clock::now();
    if (sink == 12345.678) std::cerr << "";
    return std::chrono::duration<double>(t1 - t0).count();
}

} // extern "C"
