// tests/cpp/api_bench.cpp — wall-clock timing of the C++ drop-in API (include/PLS/pls.h) as a user of the
// reference calls it: an ordinary (pageable) Eigen::MatrixXd goes into PLS::Model's constructor
// (reference src/pls.cpp:340-353: copy of X, Y + plsr), results are read through the public accessors.
//
//   api_bench fit  N K M A reps [seed L]   Model(X, Y, KERNEL_TYPE1, A) on synthetic data, seconds per construction
//   api_bench loo  N K M A reps [seed L]   the same model's cv_LOO() + validation(MSE) + optimal_num_components()
//
// The synthetic matrices are generated on the device by the library's own generator (bit-identical to the test
// oracle's) and downloaded into Eigen matrices before the clock starts.  Prints one JSON object on stdout.
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include <PLS/pls.h>

#include "../../include/plsb.h"

using namespace PLS;
using clk = std::chrono::steady_clock;

static double since(clk::time_point t0) { return std::chrono::duration<double>(clk::now() - t0).count(); }

static void need(int rc, const char *what) {
    if (rc != PLSB_OK) { std::cerr << what << ": " << plsb_last_error() << std::endl; exit(1); }
}

int main(int argc, char **argv) {
    if (argc < 7) { std::cerr << "usage: api_bench fit|loo N K M A reps [seed L]" << std::endl; return 2; }
    const std::string mode = argv[1];
    const int64_t N = atoll(argv[2]), K = atoll(argv[3]), M = atoll(argv[4]);
    const size_t A = (size_t)atoll(argv[5]);
    const int reps = atoi(argv[6]);
    const uint64_t seed = argc > 7 ? strtoull(argv[7], nullptr, 10) : 20260004ull;
    const int64_t L = argc > 8 ? atoll(argv[8]) : 16;

    Mat2D X(N, K), Y(N, M);
    {   // inputs: device generator -> pageable Eigen matrices (outside every timed region)
        const int dev = PLS::devices()[0];
        plsb_ctx_t c = nullptr;
        need(plsb_ctx_create(dev, &c), "plsb_ctx_create");
        plsb_dataset_t ds = nullptr;
        need(plsb_dataset_create_synth(c, seed, 0, N, K, M, L, &ds), "plsb_dataset_create_synth");
        need(plsb_dataset_download(ds, X.data(), N, Y.data(), N), "plsb_dataset_download");
        plsb_dataset_destroy(ds);
        plsb_ctx_trim(c);
        plsb_ctx_destroy(c);
    }
    std::vector<double> fit_s, cv_s;
    double checksum = 0.0;
    long long opt0 = -1;
    {
        Model warm(X, Y, KERNEL_TYPE1, A);       // first call: contexts, pinned staging, memory pool
        checksum = warm.coefficients().real().cwiseAbs().sum();
        if (mode == "loo") (void)warm.cv_LOO();
    }
    for (int r = 0; r < reps; ++r) {
        const auto t0 = clk::now();
        Model m(X, Y, KERNEL_TYPE1, A);
        fit_s.push_back(since(t0));
        if (mode == "loo") {
            const auto t1 = clk::now();
            const Residual res = m.cv_LOO();
            const Mat2D mse = validation(res, MSE);
            const Colsz opt = optimal_num_components(res);
            cv_s.push_back(since(t1));
            opt0 = (long long)opt(0);
            checksum += mse.sum() * 0.0;
        }
    }
    auto mean = [](const std::vector<double> &v) { double s = 0; for (double x : v) s += x; return v.empty() ? 0.0 : s / v.size(); };
    auto minv = [](const std::vector<double> &v) { double s = 1e300; for (double x : v) s = x < s ? x : s; return v.empty() ? 0.0 : s; };
    std::cout.precision(9);
    std::cout << "{\"mode\": \"" << mode << "\", \"N\": " << N << ", \"K\": " << K << ", \"M\": " << M << ", \"A\": " << A
              << ", \"reps\": " << reps << ", \"gpus\": " << PLS::devices().size()
              << ", \"model_ctor_s_mean\": " << mean(fit_s) << ", \"model_ctor_s_min\": " << minv(fit_s);
    if (mode == "loo")
        std::cout << ", \"cv_loo_validation_optimal_s_mean\": " << mean(cv_s) << ", \"cv_loo_validation_optimal_s_min\": " << minv(cv_s)
                  << ", \"optimal_num_components_0\": " << opt0;
    std::cout << ", \"coef_abs_sum\": " << checksum << "}" << std::endl;
    return 0;
}
