// pls_b200/host/model.cpp — PLS::Model on the B200 engine: every method marshals Eigen column-major
// host buffers into include/plsb.h calls and wraps the real results back into the reference's
// (complex-typed) return types.  No arithmetic of the fit / predict / CV path happens here.
//
// Reference being replaced: src/pls.cpp:323-359 (constructors), 390-437 (plsr), 439-467 (scores,
// coefficients, fitted_values, residuals, SSE, explained_variance), 469-549 (cv_LOO, cv_NEW_DATA,
// cv_LSO), 551-580 (print_explained_variance, print_state).
#include <cassert>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iomanip>
#include <thread>

#include <PLS/pls.h>

#include "engine.h"

namespace PLS {

using detail::DeviceMirror;
using detail::Engine;
using detail::ok;

namespace {
// real K x A (column-major) -> the reference's complex matrix type with zero imaginary parts
Mat2Dc as_complex(const Mat2D &m) { return m.cast<std::complex<float_type>>(); }
// Eigen::Map-free access: Mat2D is column-major with ld = rows
inline int64_t ld(const Mat2D &m) { return m.rows() > 0 ? (int64_t)m.rows() : 1; }
}  // namespace

// ---- constructors (pls.cpp:323-359) ---------------------------------------------------------------
// every constructor leaves K x A zero matrices P, W, R and an M x A zero Q behind before any fit runs
void Model::init_state(size_t num_predictors, size_t num_responses) {
    assert(A <= num_predictors);
    for (Mat2Dc *m : {&P, &W, &R}) *m = Mat2Dc::Zero(num_predictors, A);
    Q = Mat2Dc::Zero(num_responses, A);
}

Model::Model(const size_t &num_predictors, const size_t &num_responses, const METHOD &algorithm,
             const size_t &max_components)
    : A(max_components), method(algorithm) {
    init_state(num_predictors, num_responses);
}

Model::Model(const size_t &num_predictors, const size_t &num_responses, const METHOD &algorithm)
    : Model(num_predictors, num_responses, algorithm, num_predictors) {}

// The fitting constructor: the reference copies X, Y into `_X,_Y` and fits (pls.cpp:340-353).  Here the copy is the
// upload: the data set the fit creates in HBM is kept as the model's own copy of its training data (cv_LOO / cv_LSO
// run from it), and no second copy is made in host memory.
Model::Model(const Mat2D &X, const Mat2D &Y, const METHOD &algorithm, const size_t &max_components)
    : A(max_components), method(algorithm) {
    assert(X.rows() != 0);
    assert(X.rows() == Y.rows());
    init_state(X.cols(), Y.cols());
    mirror = std::make_shared<DeviceMirror>();
    mirror->N = X.rows(); mirror->K = X.cols(); mirror->M = Y.cols();
    fit_on(*mirror, X, Y, algorithm);
}

Model::Model(const Mat2D &X, const Mat2D &Y, const METHOD &algorithm) : Model(X, Y, algorithm, X.cols()) {}

// ---- fit (pls.cpp:390-437) ------------------------------------------------------------------------
// A public plsr() refits P, W, R, Q, T on the data it is given and leaves the model's own training data alone, as
// in the reference (`_X,_Y` are const there): the device copy of foreign data lives for this call only.
void Model::plsr(const Mat2D &X, const Mat2D &Y, const METHOD &algorithm) {
    DeviceMirror scratch;
    scratch.N = X.rows(); scratch.K = X.cols(); scratch.M = Y.cols();
    fit_on(scratch, X, Y, algorithm);
}

void Model::fit_on(DeviceMirror &data, const Mat2D &X, const Mat2D &Y, const METHOD &algorithm) {
    method = algorithm;
    Engine &e = Engine::get();
    std::lock_guard<std::recursive_mutex> lk(e.mu);
    const int64_t N = X.rows(), K = X.cols(), M = Y.cols();
    const int meth = algorithm == KERNEL_TYPE2 ? PLSB_KERNEL_TYPE2 : PLSB_KERNEL_TYPE1;
    const bool wantT = algorithm == KERNEL_TYPE1;
    Mat2D Wr(K, A), Pr(K, A), Rr(K, A), Qr(M, A), Tr;
    const size_t ndev = e.ndev();
    bool direct_T = false;
    const char *trace = getenv("PLSB_TRACE");
    const auto t_start = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (trace && *trace == '1')
            fprintf(stderr, "[pls] fit_on: %s at %.3f s\n", what, std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count());
    };
    if (ndev > 1 && N >= e.shard_min_rows) {
        // row-sharded fit: every GPU streams its rows, K-length partials are all-reduced.  The shards
        // stay: together they are the model's copy of the training data (cv_* gathers them device to device).
        e.ensure_comm();
        data.load_shards(X.data(), ld(X), Y.data(), ld(Y));
        if (wantT) Tr.resize(N, A);
        std::vector<Mat2D> Tpart(ndev);
        e.parallel([&](size_t i) {
            const int64_t rows = data.shard_row0[i + 1] - data.shard_row0[i];
            if (wantT) Tpart[i].resize(rows, A);
            const bool lead = i == 0;  // the K-sized results are replicated: read them from rank 0 only
            return plsb_fit(data.shard[i], (int64_t)A, meth, lead ? Wr.data() : nullptr, lead ? Pr.data() : nullptr,
                            lead ? Rr.data() : nullptr, lead ? Qr.data() : nullptr, wantT ? Tpart[i].data() : nullptr);
        }, "plsb_fit(sharded)");
        if (wantT)
            for (size_t i = 0; i < ndev; ++i) Tr.middleRows(data.shard_row0[i], Tpart[i].rows()) = Tpart[i];
    } else {
        // one GPU: the fit's results are formed in HBM (resident model); the K-sized matrices come home as they are, the
        // N x A scores go straight from the transfer's staging buffers into the complex-typed member T
        data.load_full(X.data(), ld(X), Y.data(), ld(Y));
        lap("data set in HBM");
        plsb_model_t mh = nullptr;
        ok(plsb_fit_resident(data.full[0], (int64_t)A, meth, &mh), "plsb_fit_resident");
        lap("fit done");
        int rc = plsb_model_download(mh, Wr.data(), Pr.data(), Rr.data(), Qr.data(), nullptr);
        if (rc == PLSB_OK && wantT) {
            T.resize(N, A);
            rc = plsb_model_download_scores(mh, reinterpret_cast<double *>(T.data()), 1);
            direct_T = true;
        }
        plsb_model_destroy(mh);
        ok(rc, "plsb_model_download");
    }
    lap("fit done, results on the host");
    W = as_complex(Wr);
    P = as_complex(Pr);
    R = as_complex(Rr);
    Q = as_complex(Qr);
    if (wantT && !direct_T) {  // KERNEL_TYPE2 leaves T as it was (pls.cpp:394, 434)
        T.resize(N, A);
        Engine::widen(Tr.data(), T.data(), (size_t)N * A);
    }
    lap("complex-typed members filled");
}

// ---- prediction helpers (pls.cpp:439-467) ---------------------------------------------------------
// Rows of X_new are independent, so with several GPUs a large prediction is row-sharded like the fit (SURVEY
// 8(f)-2): device i takes rows [b[i], b[i+1]) — its own share of the PCIe transfer and of the pass over X_new —
// and the model (R, Q: K*A + M*A doubles) is simply handed to every device.  No communication.
namespace {
std::vector<int64_t> row_shards(Engine &e, int64_t rows) {
    size_t g = e.ndev();
    if (g < 2 || rows < e.shard_min_rows || rows < (int64_t)g) g = 1;
    std::vector<int64_t> b(g + 1);
    for (size_t i = 0; i <= g; ++i) b[i] = (int64_t)(i * (size_t)rows / g);
    return b;
}
template <typename Fn>
void over_shards(Engine &e, const std::vector<int64_t> &b, const char *where, Fn fn) {   // fn(i) returns a plsb status
    const size_t g = b.size() - 1;
    if (g == 1) { ok(fn(0), where); return; }
    e.parallel([&](size_t i) { return i < g ? fn(i) : (int)PLSB_OK; }, where);
}
}  // namespace

const Mat2Dc Model::scores(const Mat2D &X_new, const size_t comp) const {
    assert(A >= comp);
    Engine &e = Engine::get();
    std::lock_guard<std::recursive_mutex> lk(e.mu);
    const Mat2D Rr = R.real();
    Mat2D S(X_new.rows(), comp);
    if (comp > 0) {
        const std::vector<int64_t> b = row_shards(e, X_new.rows());
        std::vector<Mat2D> part(b.size() - 1);
        over_shards(e, b, "plsb_scores", [&](size_t i) {
            part[i].resize(b[i + 1] - b[i], comp);
            return plsb_scores(e.ctx[i], Rr.data(), Rr.rows(), (int64_t)comp, X_new.data() + b[i], ld(X_new), b[i + 1] - b[i],
                               part[i].data());
        });
        for (size_t i = 0; i + 1 < b.size(); ++i) S.middleRows(b[i], b[i + 1] - b[i]) = part[i];
    }
    return as_complex(S);
}

const Mat2Dc Model::loadingsX(const size_t comp) const {
    assert(A >= comp);
    return P.leftCols(comp);
}

const Mat2Dc Model::loadingsY(const size_t comp) const {
    assert(A >= comp);
    return Q.leftCols(comp);
}

const Mat2Dc Model::coefficients(const size_t comp) const {
    assert(A >= comp);
    Engine &e = Engine::get();
    std::lock_guard<std::recursive_mutex> lk(e.mu);
    const Mat2D Rr = R.real(), Qr = Q.real();
    Mat2D B(Rr.rows(), Qr.rows());
    ok(plsb_coefficients(e.primary(), Rr.data(), Qr.data(), Rr.rows(), Qr.rows(), (int64_t)comp, B.data()),
       "plsb_coefficients");
    return as_complex(B);
}

const Mat2D Model::fitted_values(const Mat2D &X_new, const size_t comp) const {
    Engine &e = Engine::get();
    std::lock_guard<std::recursive_mutex> lk(e.mu);
    const Mat2D Rr = R.real(), Qr = Q.real();
    Mat2D F = Mat2D::Zero(X_new.rows(), Qr.rows());
    if (comp > 0) {
        const std::vector<int64_t> b = row_shards(e, X_new.rows());
        std::vector<Mat2D> part(b.size() - 1);
        over_shards(e, b, "plsb_fitted", [&](size_t i) {
            part[i].resize(b[i + 1] - b[i], Qr.rows());
            return plsb_fitted(e.ctx[i], Rr.data(), Qr.data(), Rr.rows(), Qr.rows(), (int64_t)comp, X_new.data() + b[i],
                               ld(X_new), b[i + 1] - b[i], nullptr, 1, part[i].data());
        });
        for (size_t i = 0; i + 1 < b.size(); ++i) F.middleRows(b[i], b[i + 1] - b[i]) = part[i];
    }
    return F;
}

const Mat2D Model::residuals(const Mat2D &X_new, const Mat2D &Y_new, const size_t comp) const {
    if (comp == 0) return Y_new;
    Engine &e = Engine::get();
    std::lock_guard<std::recursive_mutex> lk(e.mu);
    const Mat2D Rr = R.real(), Qr = Q.real();
    Mat2D E(X_new.rows(), Qr.rows());
    const std::vector<int64_t> b = row_shards(e, X_new.rows());
    std::vector<Mat2D> part(b.size() - 1);
    over_shards(e, b, "plsb_fitted(residuals)", [&](size_t i) {
        part[i].resize(b[i + 1] - b[i], Qr.rows());
        return plsb_fitted(e.ctx[i], Rr.data(), Qr.data(), Rr.rows(), Qr.rows(), (int64_t)comp, X_new.data() + b[i], ld(X_new),
                           b[i + 1] - b[i], Y_new.data() + b[i], ld(Y_new), part[i].data());
    });
    for (size_t i = 0; i + 1 < b.size(); ++i) E.middleRows(b[i], b[i + 1] - b[i]) = part[i];
    return E;
}

const Row Model::SSE(const Mat2D &X_new, const Mat2D &Y_new, const size_t comp) const {
    if (comp == 0) return Y_new.colwise().squaredNorm();
    Engine &e = Engine::get();
    std::lock_guard<std::recursive_mutex> lk(e.mu);
    const Mat2D Rr = R.real(), Qr = Q.real();
    const std::vector<int64_t> b = row_shards(e, X_new.rows());
    std::vector<Row> part(b.size() - 1, Row::Zero(Qr.rows()));
    over_shards(e, b, "plsb_sse", [&](size_t i) {
        return plsb_sse(e.ctx[i], Rr.data(), Qr.data(), Rr.rows(), Qr.rows(), (int64_t)comp, X_new.data() + b[i], ld(X_new),
                        b[i + 1] - b[i], Y_new.data() + b[i], ld(Y_new), part[i].data());
    });
    Row sse = part[0];
    for (size_t i = 1; i < part.size(); ++i) sse += part[i];   // shard order: deterministic
    return sse;
}

const Row Model::explained_variance(const Mat2D &X_new, const Mat2D &Y_new, const size_t comp) const {
    return (1.0 - (SSE(X_new, Y_new, comp).array() / SST(Y_new).array())).matrix();
}

// ---- cross-validation (pls.cpp:469-549) -----------------------------------------------------------
// All folds go to the GPU(s) as one batched job; folds are split across the engine's devices with no
// communication and every device writes its own rows of the residual block.
Residual Model::cv_on_device(const std::vector<int64_t> &test_idx, size_t nfolds, size_t test_size,
                             const char *label) const {
    Engine &e = Engine::get();
    std::lock_guard<std::recursive_mutex> lk(e.mu);
    if (!mirror) detail::die(label, PLSB_ERR_ARG, "the model holds no training data");
    DeviceMirror &mir = *mirror;
    const int64_t M = mir.M, rows = (int64_t)(nfolds * test_size);
    const int meth = method == KERNEL_TYPE2 ? PLSB_KERNEL_TYPE2 : PLSB_KERNEL_TYPE1;
    std::vector<double> E((size_t)(M * (int64_t)A * rows));
    const size_t ndev = std::min(e.ndev(), std::max<size_t>(nfolds, 1));
    mir.prepare();
    // replicas first, on every GPU at once, device to device from the copy the fit left behind — and only then the
    // folds: a replica made while its source GPU is already running folds would wait for that GPU's stream
    if (ndev > 1)
        e.parallel([&](size_t i) -> int { return i < ndev ? mir.ensure_full(i) : PLSB_OK; }, "plsb_dataset_concat");
    auto work = [&](size_t i) -> int {
        if (i >= ndev) return PLSB_OK;
        const int64_t f0 = (int64_t)(i * nfolds / ndev), f1 = (int64_t)((i + 1) * nfolds / ndev);
        int rc = mir.ensure_full(i);
        if (rc == PLSB_OK)
            rc = plsb_cv_folds(mir.full[i], (int64_t)A, meth, (int64_t)nfolds, (int64_t)test_size, test_idx.data(), f0, f1 - f0,
                               E.data());
        return rc;
    };
    if (ndev == 1) ok(work(0), "plsb_cv_folds");
    else e.parallel(work, "plsb_cv_folds");
    std::vector<Mat2D> Ev((size_t)M);
    for (int64_t m = 0; m < M; ++m)
        Ev[(size_t)m] = Eigen::Map<const Mat2D>(E.data() + (size_t)(m * (int64_t)A * rows), rows, (Eigen::Index)A);
    return Residual(Ev, label);
}

Residual Model::cv_LOO() const {
    const size_t N = mirror ? (size_t)mirror->N : 0;
    std::vector<int64_t> idx(N);
    std::iota(idx.begin(), idx.end(), int64_t(0));
    return cv_on_device(idx, N, 1, "LOO");
}

Residual Model::cv_LSO(const float_type test_fraction, const size_t num_trials, std::mt19937 &rng) const {
    const size_t N = mirror ? (size_t)mirror->N : 0;
    const size_t test_size = static_cast<size_t>(test_fraction * N + 0.5);
    const size_t train_size = N - test_size;
    assert((test_size != 0) and (train_size != 0));
    // fold lists come from the caller's generator through rand_nchoosek, one persistent permutation
    // reshuffled per trial (pls.cpp:522-534): the folds and the final state of `rng` match the reference
    std::vector<Eigen::Index> sample(train_size), complement(test_size), full(N);
    std::iota(full.begin(), full.end(), 0);
    std::vector<int64_t> idx(num_trials * test_size);
    for (size_t rep = 0; rep < num_trials; ++rep) {
        rand_nchoosek(rng, full, sample, complement);
        std::copy(complement.begin(), complement.end(), idx.begin() + rep * test_size);
    }
    return cv_on_device(idx, num_trials, test_size, "LSO");
}

Residual Model::cv_NEW_DATA(const Mat2D &X_new, const Mat2D &Y_new) const {
    assert((X_new.cols() == R.rows()) and (Y_new.cols() == Q.rows()));
    Engine &e = Engine::get();
    std::lock_guard<std::recursive_mutex> lk(e.mu);
    const Mat2D Rr = R.real(), Qr = Q.real();
    const int64_t M = Y_new.cols(), rows = X_new.rows();
    std::vector<double> E((size_t)(M * (int64_t)A * rows));
    const std::vector<int64_t> b = row_shards(e, rows);
    if (b.size() == 2) {
        ok(plsb_predict_residuals(e.primary(), Rr.data(), Qr.data(), Rr.rows(), M, (int64_t)A, X_new.data(), ld(X_new), rows,
                                  Y_new.data(), ld(Y_new), E.data()),
           "plsb_predict_residuals");
    } else {
        over_shards(e, b, "plsb_predict_residuals", [&](size_t i) {
            const int64_t ri = b[i + 1] - b[i];
            std::vector<double> Ei((size_t)(M * (int64_t)A * ri));
            const int rc = plsb_predict_residuals(e.ctx[i], Rr.data(), Qr.data(), Rr.rows(), M, (int64_t)A, X_new.data() + b[i],
                                                  ld(X_new), ri, Y_new.data() + b[i], ld(Y_new), Ei.data());
            if (rc != PLSB_OK) return rc;
            for (int64_t mc = 0; mc < M * (int64_t)A; ++mc)   // column (m, c) of the block: rows [b[i], b[i+1])
                std::copy(Ei.begin() + mc * ri, Ei.begin() + (mc + 1) * ri, E.begin() + mc * rows + b[i]);
            return (int)PLSB_OK;
        });
    }
    std::vector<Mat2D> Ev((size_t)M);
    for (int64_t m = 0; m < M; ++m)
        Ev[(size_t)m] = Eigen::Map<const Mat2D>(E.data() + (size_t)(m * (int64_t)A * rows), rows, (Eigen::Index)A);
    return Residual(Ev, "NEW DATA");
}

// ---- text output (pls.cpp:551-580): Eigen's operator<< at the stream's precision, like the reference
void Model::print_explained_variance(const Mat2D &X, const Mat2D &Y, std::ostream &os) const {
    // one pass over X gives the SSE of every component count (the reference makes 2A passes)
    Engine &e = Engine::get();
    Mat2D sse(Y.cols(), A);
    {
        std::lock_guard<std::recursive_mutex> lk(e.mu);
        const Mat2D Rr = R.real(), Qr = Q.real();
        sse.setZero(Qr.rows(), A);
        const std::vector<int64_t> b = row_shards(e, X.rows());
        std::vector<Mat2D> part(b.size() - 1);
        over_shards(e, b, "plsb_sse_prefixes", [&](size_t i) {
            part[i].resize(Qr.rows(), A);
            return plsb_sse_prefixes(e.ctx[i], Rr.data(), Qr.data(), Rr.rows(), Qr.rows(), (int64_t)A, X.data() + b[i], ld(X),
                                     b[i + 1] - b[i], Y.data() + b[i], ld(Y), part[i].data());
        });
        for (const Mat2D &p : part) sse += p;
    }
    const Row sst = SST(Y);
    const size_t wd = ceil(std::log10(A));
    for (size_t ncomp = 1; ncomp <= A; ncomp++) {
        const Row s = sse.col(ncomp - 1).transpose();
        const Row ev = (1.0 - (s.array() / sst.array())).matrix();
        os << std::setw(wd) << ncomp << " components explained variance: " << ev;
        os << "  - SSE: " << s << std::endl;
    }
}

void Model::print_state(std::ostream &os) const {
    os << "P:" << std::endl << P << std::endl;
    os << "W:" << std::endl << W << std::endl;
    os << "R:" << std::endl << R << std::endl;
    os << "Q:" << std::endl << Q << std::endl;
    os << "T:" << std::endl << T << std::endl;
    os << "coefficients:" << std::endl << coefficients() << std::endl;
}

}  // namespace PLS
