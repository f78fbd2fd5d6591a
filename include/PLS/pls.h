// include/PLS/pls.h — source-compatible stand-in for tjhladish/PLS's public header (v2.0.0,
// reference include/PLS/pls.h:1-270), backed by the B200 engine libplsb.so instead of Eigen arithmetic.
//
// Every name, argument order, default argument and return type a caller of the reference can spell is
// kept; Eigen matrices remain the HOST containers of the API and nothing more: all
// fitting, prediction and cross-validation arithmetic is done by sm_100a CUDA kernels reached through
// the C-ABI in include/plsb.h (pls_b200/host/model.cpp is the marshalling layer).  There is no CPU
// fallback: without a B200 the first call that needs the engine prints the error and exits(1), the
// reference's only failure idiom (pls.cpp:56-57).
//
// Differences a caller can observe (documented in DESIGN.md):
//   * loadingsX / loadingsY are defined here (the reference only declares them, pls.h:207-211).
//   * cross-validation extracts A components per fold, not K (identical results).
//   * the spellings used by BASELINE.json's north_star (PLS_Model, loo_validation, ...) are provided
//     as thin aliases at the bottom of this file.
#ifndef PLS_H
#define PLS_H

#include <Eigen/Core>
#include <algorithm>
#include <iostream>
#include <memory>
#include <numeric>
#include <random>
#include <string>
#include <vector>

// ---- global typedefs (reference pls.h:22-33; FP64 only, the MPREAL switch pls.h:11-21 is dead) -----
typedef double float_type;
typedef Eigen::MatrixXd Mat2D;
typedef Eigen::VectorXd Col;
typedef Eigen::RowVectorXd Row;
typedef Eigen::MatrixXcd Mat2Dc;
typedef Eigen::VectorXcd Colc;
typedef Eigen::VectorXi Coli;
typedef Eigen::Matrix<size_t, Eigen::Dynamic, 1> Colsz;
typedef Eigen::RowVectorXi Rowi;
typedef Eigen::Matrix<size_t, 1, Eigen::Dynamic> Rowsz;

namespace PLS {

struct Model;
namespace detail { struct DeviceMirror; }

// ---- cross-validation residuals (reference pls.h:41-53) ------------------------------------------
// errors()[y](obs, c-1) = residual of response y for held-out observation obs with c components.
typedef std::vector<Mat2D> ResidualData;

class Residual {
  public:
    const std::vector<Mat2D> errors() const { return _residual; }
    const std::string method() const { return _method_label; }

  private:
    friend struct Model;
    Residual(const std::vector<Mat2D> &residual, const std::string &method)
        : _residual(residual), _method_label(method) {}
    const std::vector<Mat2D> _residual;
    const std::string _method_label;
};

// ---- small host helpers (reference pls.h:58-85) ----------------------------------------------------
// Index permutation that sorts v ascending.  std::sort (unstable) on purpose: wilcoxon's ranks, and with
// them optimal_num_components, depend on how libstdc++ orders ties.
template <typename T>
std::vector<size_t> ordered(const T &v) {
    std::vector<size_t> idx(v.size());
    std::iota(idx.begin(), idx.end(), size_t(0));
    std::sort(idx.begin(), idx.end(), [&v](const auto &a, const auto &b) { return *(v.begin() + a) < *(v.begin() + b); });
    return idx;
}

template <typename EIGENTYPE>
std::vector<float_type> to_cvector(const EIGENTYPE &data) {
    return std::vector<float_type>(data.begin(), data.end());
}

template <typename EIGENTYPE>
inline EIGENTYPE to_evector(const std::vector<float_type> &data) {
    EIGENTYPE out(data.size());
    std::copy(data.begin(), data.end(), out.data());
    return out;
}

// ---- text input (reference pls.h:89-93, pls.cpp:23-67) -------------------------------------------
std::vector<std::string> split(const std::string &s, const char separator = ',');
// headerless separator-delimited numbers, one matrix row per line; exit(1) on ragged rows
Mat2D read_matrix_file(const std::string &filename, const char separator = ',');
// raw little-endian FP64 column-major file with a 24-byte header {u64 magic "PLSBMAT1", rows, cols}:
// the input format for matrices too large for CSV parsing.  Not in the reference.
Mat2D read_matrix_binary(const std::string &filename);
void write_matrix_binary(const std::string &filename, const Mat2D &mat);

// ---- column statistics (reference pls.h:97-111, pls.cpp:69-111) ----------------------------------
Row SST(const Mat2D &mat, const Row &means);
Row SST(const Mat2D &mat);
Row colwise_stdev(const Mat2D &mat, const Row &means);
Row colwise_stdev(const Mat2D &mat);
Row z_scores(const Row &obs, const Row &mean, const Row &stdev);
Mat2D colwise_z_scores(const Mat2D &mat, const Row &mean, const Row &stdev);
Mat2D colwise_z_scores(const Mat2D &mat);

// ---- rank statistics (reference pls.h:114-126, pls.cpp:152-227) ----------------------------------
float_type normalcdf(const float_type z);
float_type wilcoxon(const Col &err_1, const Col &err_2);
void rand_nchoosek(std::mt19937 &rng, std::vector<Eigen::Index> &full, std::vector<Eigen::Index> &sample,
                   std::vector<Eigen::Index> &complement);

// ---- model selection (reference pls.h:131-159, pls.cpp:235-305) ----------------------------------
typedef enum { KERNEL_TYPE1, KERNEL_TYPE2 } METHOD;
typedef enum { RESS, MSE } VALIDATION_OUTPUT;

Mat2D validation(const Residual &residual, const PLS::VALIDATION_OUTPUT out_type);
Colsz optimal_num_components(const Residual &residual, const float_type ALPHA = 0.1);
void print_validation(const Residual &residual, const VALIDATION_OUTPUT out_type, std::ostream &os = std::cerr);

// ---- the regression object (reference pls.h:184-266, pls.cpp:323-580) ------------------------------
// Dayal & MacGregor (1997) modified kernel PLS #1 / #2.  X is N x K, Y is N x M, A components;
// W, P, R are K x A, Q is M x A, T is N x A (KERNEL_TYPE1 only).  Complex-typed results carry zero
// imaginary parts, exactly as the reference's do for real input.
struct Model {
    // fit immediately with at most max_components components (pls.cpp:340-353)
    Model(const Mat2D &X, const Mat2D &Y, const METHOD &algorithm, const size_t &max_components);
    // fit immediately with A = K components (pls.cpp:356-359)
    Model(const Mat2D &X, const Mat2D &Y, const METHOD &algorithm = KERNEL_TYPE1);

    // refit on new data of the same K and M; A stays what the constructor set (pls.cpp:390-437)
    void plsr(const Mat2D &X, const Mat2D &Y, const METHOD &algorithm);

    // X_new * R[:, :comp]  (pls.cpp:439-442)
    const Mat2Dc scores(const Mat2D &X_new, const size_t comp) const;
    const Mat2Dc scores(const Mat2D &X_new) const { return scores(X_new, A); }

    // P[:, :comp] and Q[:, :comp] (the reference's legend pls.h:169-170; undefined there)
    const Mat2Dc loadingsX(const size_t comp) const;
    const Mat2Dc loadingsX() const { return loadingsX(A); }
    const Mat2Dc loadingsY(const size_t comp) const;
    const Mat2Dc loadingsY() const { return loadingsY(A); }

    // B = R[:, :comp] * Q[:, :comp]'  (pls.cpp:444-447)
    const Mat2Dc coefficients(const size_t comp) const;
    const Mat2Dc coefficients() const { return coefficients(A); }

    // X * B  (pls.cpp:449-451)
    const Mat2D fitted_values(const Mat2D &X, const size_t comp) const;
    const Mat2D fitted_values(const Mat2D &X) const { return fitted_values(X, A); }

    // Y - X * B  (pls.cpp:453-455)
    const Mat2D residuals(const Mat2D &X, const Mat2D &Y, const size_t comp) const;
    const Mat2D residuals(const Mat2D &X, const Mat2D &Y) const { return residuals(X, Y, A); }

    // column sums of squared residuals (pls.cpp:457-459)
    const Row SSE(const Mat2D &X, const Mat2D &Y, const size_t comp) const;
    const Row SSE(const Mat2D &X, const Mat2D &Y) const { return SSE(X, Y, A); }

    // 1 - SSE / SST(Y)  (pls.cpp:461-467)
    const Row explained_variance(const Mat2D &X, const Mat2D &Y, const size_t comp) const;
    const Row explained_variance(const Mat2D &X, const Mat2D &Y) const { return explained_variance(X, Y, A); }

    // leave-one-out, new-data and leave-some-out validation (pls.cpp:469-549).  LOO and LSO run as a
    // batched many-folds-per-launch job on the GPU(s); the LSO fold lists are drawn on the host from
    // `rng` with the reference's rand_nchoosek so the folds (and the state `rng` is left in) match.
    Residual cv_LOO() const;
    Residual cv_NEW_DATA(const Mat2D &X, const Mat2D &Y) const;
    Residual cv_LSO(const float_type test_fraction, const size_t num_trials, std::mt19937 &rng) const;

    void print_explained_variance(const Mat2D &X, const Mat2D &Y, std::ostream &os = std::cerr) const;
    void print_state(std::ostream &os = std::cerr) const;

  private:
    size_t A;            // number of components
    Mat2Dc P, W, R, Q, T;
    PLS::METHOD method;
    // The training data kept for cv_LOO / cv_LSO — the reference's `const Mat2D _X, _Y` (pls.h:251, copied at
    // pls.cpp:344) — live in HBM only: the copy the constructor's fit uploads IS the model's copy (no second copy
    // in host memory; later changes to the caller's matrices do not reach it, as in the reference).
    std::shared_ptr<detail::DeviceMirror> mirror;
    void init_state(size_t num_predictors, size_t num_responses);

    Model(const size_t &num_predictors, const size_t &num_responses, const METHOD &algorithm = KERNEL_TYPE1);
    Model(const size_t &num_predictors, const size_t &num_responses, const METHOD &algorithm,
          const size_t &max_components);
    void fit_on(detail::DeviceMirror &data, const Mat2D &X, const Mat2D &Y, const METHOD &algorithm);
    Residual cv_on_device(const std::vector<int64_t> &test_idx, size_t nfolds, size_t test_size,
                          const char *label) const;
};

// ---- engine controls (not in the reference; environment variables give the same control to
// unmodified callers: PLSB_DEVICES="0,1,.."  PLSB_SHARD_MIN_ROWS=<rows>) --------------------------------
// GPUs used by this process: cross-validation folds are split over them with no communication and a
// fit on at least `shard_min_rows` rows is row-sharded over them (NCCL all-reduce of K-length partials).
void set_devices(const std::vector<int> &devices);
std::vector<int> devices();

}  // namespace PLS

// ---- north_star spellings (BASELINE.json) → this header's symbols ------------------
typedef PLS::Model PLS_Model;
namespace PLS {
inline Model plsr(const Mat2D &X, const Mat2D &Y, const size_t ncomp, const METHOD method = KERNEL_TYPE1) {
    return Model(X, Y, method, ncomp);
}
inline Mat2D loo_validation(const Model &m, const VALIDATION_OUTPUT out_type) { return validation(m.cv_LOO(), out_type); }
inline Mat2D lso_validation(const Model &m, const float_type test_fraction, const size_t num_trials, std::mt19937 &rng,
                            const VALIDATION_OUTPUT out_type) {
    return validation(m.cv_LSO(test_fraction, num_trials, rng), out_type);
}
inline Mat2D new_data_validation(const Model &m, const Mat2D &X, const Mat2D &Y, const VALIDATION_OUTPUT out_type) {
    return validation(m.cv_NEW_DATA(X, Y), out_type);
}
}  // namespace PLS

#endif  // PLS_H
