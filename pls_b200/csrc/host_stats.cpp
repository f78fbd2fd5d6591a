// pls_b200/csrc/host_stats.cpp — the host-side, non-arithmetic-heavy part of the validation path.
//
// These stay on the CPU on purpose: optimal_num_components is an exact integer decided by
// ranks produced with libstdc++'s (unstable) std::sort, and the leave-some-out folds are defined by
// libstdc++'s std::shuffle driven by the caller's std::mt19937 — both must be the very same library
// routines the reference calls to reproduce its results exactly.  Inputs are the residual blocks the
// GPU produced.  Reference: src/pls.cpp:152-160 (normalcdf), 190-211 (wilcoxon), 218-227
// (rand_nchoosek), 235-261 (validation), 265-289 (optimal_num_components); include/PLS/pls.h:58-69.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <numeric>
#include <random>
#include <thread>
#include <vector>

#include "../../include/plsb.h"

// fn(m) for every response m; on a few threads when the block is large (each response's sums and ranks are computed by
// one thread in the reference's order, so results do not depend on the thread count)
template <typename Fn>
static void over_responses(int64_t M, int64_t work_per_response, Fn fn) {
    const int64_t nthr = work_per_response >= (1 << 18) ? std::min<int64_t>(M, std::max(1u, std::thread::hardware_concurrency())) : 1;
    if (nthr <= 1) { for (int64_t m = 0; m < M; ++m) fn(m); return; }
    std::vector<std::thread> th;
    for (int64_t t = 0; t < nthr; ++t)
        th.emplace_back([=] { for (int64_t m = t; m < M; m += nthr) fn(m); });
    for (auto &x : th) x.join();
}

extern "C" {

double plsb_normalcdf(double z) {
    // Abramowitz & Stegun 26.2.19 as the reference evaluates it (pls.cpp:152-160)
    const double c1 = 0.196854, c2 = 0.115194, c3 = 0.000344, c4 = 0.019527;
    const double a = std::fabs(z);
    const double poly = 1 + c1 * a + c2 * a * a + c3 * a * a * a + c4 * a * a * a * a;
    const double tail = 0.5 / std::pow(poly, 4);
    return z < 0 ? tail : 1.0 - tail;
}

double plsb_wilcoxon(const double *e1, const double *e2, int64_t n) {
    // signed-rank test on | |e1| - |e2| | (pls.cpp:190-211); ranks from an index sort with the same
    // comparator and the same std::sort as pls.h:58-69
    std::vector<double> mag((size_t)n);
    std::vector<int> sgn((size_t)n);
    for (int64_t i = 0; i < n; ++i) {
        const double d = std::fabs(e1[i]) - std::fabs(e2[i]);
        sgn[(size_t)i] = (0 < d) - (d < 0);
        mag[(size_t)i] = std::fabs(d);
    }
    std::vector<size_t> order((size_t)n);
    std::iota(order.begin(), order.end(), 0);
    std::sort(order.begin(), order.end(),
              [&mag](const auto &lhs, const auto &rhs) { return *(mag.begin() + lhs) < *(mag.begin() + rhs); });
    double d = 0;
    for (int64_t i = 0; i < n; ++i) d += static_cast<double>(i + 1) * sgn[order[(size_t)i]];
    const size_t un = (size_t)n;
    const double t = static_cast<double>(un * (un + 1)) / 2.0;
    const double v = (t - d) / 2.0;
    const double ev = t / 2.0;
    const double sv = std::sqrt(static_cast<double>(un * (un + 1) * (2 * un + 1)) / 24.0);
    return 1.0 - plsb_normalcdf((v - ev) / sv);
}

int plsb_validation(const double *E, int64_t M, int64_t rows, int64_t A, int out_type, double *out) {
    // out: M x A column-major; out_type 0 = RESS, 1 = MSE (pls.cpp:235-261)
    if (!E || !out || M < 0 || rows < 0 || A < 0) return PLSB_ERR_ARG;
    over_responses(M, rows * A, [=](int64_t m) {
        for (int64_t c = 0; c < A; ++c) {
            const double *e = E + ((size_t)m * A + c) * rows;
            double s = 0.0;
            for (int64_t i = 0; i < rows; ++i) s += e[i] * e[i];
            out[c * M + m] = out_type == 0 ? s : s / (double)rows;
        }
    });
    return PLSB_OK;
}

int plsb_optimal_num_components(const double *E, int64_t M, int64_t rows, int64_t A, double alpha, int64_t *out) {
    if (!E || !out || A < 1) return PLSB_ERR_ARG;
    std::vector<double> press((size_t)(M * A));
    plsb_validation(E, M, rows, A, 0, press.data());
    const double *pr = press.data();
    over_responses(M, rows * A, [=](int64_t m) {
        int64_t ref = 0;  // first minimum, like Eigen's minCoeff
        for (int64_t c = 1; c < A; ++c)
            if (pr[(size_t)(c * M + m)] < pr[(size_t)(ref * M + m)]) ref = c;
        int64_t pick = ref;
        const double *e_ref = E + ((size_t)m * A + ref) * rows;
        for (int64_t alt = 0; alt < ref; ++alt) {
            const double *e_alt = E + ((size_t)m * A + alt) * rows;
            if (plsb_wilcoxon(e_ref, e_alt, rows) > alpha) { pick = alt; break; }
        }
        out[m] = pick + 1;
    });
    return PLSB_OK;
}

int plsb_lso_folds(int64_t N, double test_fraction, int64_t num_trials, uint32_t seed, int64_t *test_idx,
                   int64_t *train_idx) {
    // the fold lists cv_LSO draws (pls.cpp:515-538): one permutation, reshuffled per trial by
    // std::shuffle; first train_size entries train, the rest are held out
    if (N < 1 || num_trials < 0 || !test_idx) return PLSB_ERR_ARG;
    const int64_t test_size = (int64_t)(test_fraction * (double)N + 0.5);
    const int64_t train_size = N - test_size;
    if (test_size < 0 || train_size < 0) return PLSB_ERR_ARG;
    std::mt19937 rng(seed);
    std::vector<int64_t> full((size_t)N);
    std::iota(full.begin(), full.end(), 0);
    for (int64_t t = 0; t < num_trials; ++t) {
        std::shuffle(full.begin(), full.end(), rng);
        if (train_idx) std::copy(full.begin(), full.begin() + train_size, train_idx + t * train_size);
        std::copy(full.begin() + train_size, full.end(), test_idx + t * test_size);
    }
    return PLSB_OK;
}

}  // extern "C"
