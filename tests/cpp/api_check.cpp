// tests/cpp/api_check.cpp — drives the C++ drop-in API (include/PLS/pls.h) the way a user of the
// reference would, and dumps every result as a PLSBMAT1 file; tests/test_gpu_cpp_api.py compares the
// dumps with the golden vectors produced by the unmodified reference (tests/golden/make_golden.py).
//
//   api_check X.bin Y.bin A method outdir lso_frac lso_trials [zscore]
#include <cstdlib>
#include <iostream>
#include <random>
#include <string>

#include <PLS/pls.h>

using namespace PLS;

static void dump(const std::string &dir, const std::string &name, const Mat2D &m) { write_matrix_binary(dir + "/" + name + ".bin", m); }
static void dump(const std::string &dir, const std::string &name, const Mat2Dc &m) {
    if (m.imag().cwiseAbs().size() && m.imag().cwiseAbs().maxCoeff() != 0.0) {
        std::cerr << name << ": non-zero imaginary part" << std::endl;
        exit(3);
    }
    dump(dir, name, Mat2D(m.real()));
}
static void dump_residual(const std::string &dir, const std::string &name, const Residual &r) {
    const std::vector<Mat2D> e = r.errors();
    for (size_t m = 0; m < e.size(); ++m) dump(dir, name + "_" + std::to_string(m), e[m]);
    dump(dir, name + "_ress", validation(r, RESS));
    dump(dir, name + "_mse", validation(r, MSE));
    const Colsz opt = optimal_num_components(r);
    dump(dir, name + "_opt", Mat2D(opt.cast<double>()));
}

int main(int argc, char **argv) {
    if (argc < 8) { std::cerr << "usage: api_check X.bin Y.bin A method outdir lso_frac lso_trials [zscore]" << std::endl; return 2; }
    Mat2D X = read_matrix_binary(argv[1]), Y = read_matrix_binary(argv[2]);
    const size_t A = atoi(argv[3]);
    const METHOD method = atoi(argv[4]) == 1 ? KERNEL_TYPE2 : KERNEL_TYPE1;
    const std::string dir = argv[5];
    const double frac = atof(argv[6]);
    const size_t trials = atoi(argv[7]);
    if (argc > 8 && std::string(argv[8]) == "zscore") {
        dump(dir, "x_stdev", Mat2D(colwise_stdev(X)));
        dump(dir, "x_sst", Mat2D(SST(X)));
        X = colwise_z_scores(X);
        Y = colwise_z_scores(Y);
        dump(dir, "Xz", X);
        dump(dir, "Yz", Y);
    }
    Model m(X, Y, method, A);
    dump(dir, "P", m.loadingsX());
    dump(dir, "Q", m.loadingsY());
    dump(dir, "R", m.scores(Mat2D::Identity(X.cols(), X.cols())));  // scores(I) = R
    dump(dir, "scores_A", m.scores(X));
    dump(dir, "coef_A", m.coefficients());
    dump(dir, "coef_1", m.coefficients(1));
    dump(dir, "fitted_A", m.fitted_values(X));
    dump(dir, "resid_A", m.residuals(X, Y));
    Mat2D sse(A, Y.cols()), ev(A, Y.cols());
    for (size_t c = 1; c <= A; ++c) {
        sse.row(c - 1) = m.SSE(X, Y, c);
        ev.row(c - 1) = m.explained_variance(X, Y, c);
    }
    dump(dir, "sse", sse);
    dump(dir, "ev", ev);
    if (getenv("PLSB_TEST_RECONFIGURE")) {
        // reconfigure the engine between the fit and the validation: the model's device copy of its training data was
        // made on contexts the engine now lets go of, and must follow the model to the new ones
        PLS::set_devices(PLS::devices());
        Model other(X.topRows(X.rows() / 2), Y.topRows(X.rows() / 2), method, A);   // new contexts in use
        (void)other;
    }
    dump_residual(dir, "loo", m.cv_LOO());
    if (trials > 0) {
        std::mt19937 rng;
        dump_residual(dir, "lso", m.cv_LSO(frac, trials, rng));
        dump(dir, "rng_after", Mat2D(Mat2D::Constant(1, 1, (double)rng())));
    }
    const Eigen::Index nn = std::min<Eigen::Index>(7, X.rows());
    const Residual nd = m.cv_NEW_DATA(X.topRows(nn), Y.topRows(nn));
    if (nd.method() != "NEW DATA") return 4;
    dump_residual(dir, "newdata", nd);
    // north_star spellings
    PLS_Model m2 = plsr(X, Y, A, method);
    dump(dir, "alias_coef_A", m2.coefficients());
    dump(dir, "alias_loo_ress", loo_validation(m2, RESS));
    // refit path: plsr on other data of the same shape, then back
    Mat2D X2 = X.colwise().reverse(), Y2 = Y.colwise().reverse();
    m2.plsr(X2, Y2, method);
    dump(dir, "refit_coef_A", m2.coefficients());
    std::cerr << "api_check ok" << std::endl;
    return 0;
}
