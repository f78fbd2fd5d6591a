// pls_b200/host/stats.cpp — the host-side free functions of namespace PLS: text / binary input, column
// statistics, rank statistics and the model-selection summaries.
//
// Reference: src/pls.cpp:23-67 (split, read_matrix_file), 69-111 (SST, colwise_stdev, z_scores,
// colwise_z_scores), 152-227 (normalcdf, wilcoxon, rand_nchoosek), 235-305 (validation,
// optimal_num_components, print_validation).  The rank statistics and summaries are thin wrappers over
// the C entry points of libplsb.so (host_stats.cpp) so that C++, ctypes and any other binding share one
// implementation; they stay on the CPU because their results are decided by libstdc++'s std::sort /
// std::shuffle.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>

#include <PLS/pls.h>

#include "engine.h"

namespace PLS {

// ---- input ------------------------------------------------------------------------------------------
std::vector<std::string> split(const std::string &s, const char separator) {
    // every separator ends a field; the text after the last one (possibly empty) is the final field
    std::vector<std::string> fields;
    size_t start = 0;
    for (size_t pos = s.find(separator); pos != std::string::npos; pos = s.find(separator, start)) {
        fields.emplace_back(s, start, pos - start);
        start = pos + 1;
    }
    fields.emplace_back(s, start);
    return fields;
}

Mat2D read_matrix_file(const std::string &filename, const char separator) {
    std::ifstream in(filename);
    std::vector<float_type> values;  // row-major as read
    size_t ncols = 0, nrows = 0;
    std::string line;
    while (in.is_open() && std::getline(in, line)) {
        const std::vector<std::string> fields = split(line, separator);
        if (nrows != 0 && fields.size() != ncols) {
            std::cerr << "Error: row " << nrows << " has " << fields.size() << " columns, but previous row(s) have "
                      << ncols << " columns." << std::endl;
            exit(1);
        }
        ncols = fields.size();
        for (const std::string &f : fields) values.push_back(std::stod(f));  // throws on junk, like the reference
        ++nrows;
    }
    if (nrows == 0) {  // the reference dereferences an empty vector here; fail loudly instead
        std::cerr << "Error: could not read any rows from " << filename << std::endl;
        exit(1);
    }
    Mat2D X(nrows, ncols);
    for (size_t r = 0; r < nrows; ++r)
        for (size_t c = 0; c < ncols; ++c) X(r, c) = values[r * ncols + c];
    return X;
}

static const char kMagic[8] = {'P', 'L', 'S', 'B', 'M', 'A', 'T', '1'};

Mat2D read_matrix_binary(const std::string &filename) {
    std::ifstream in(filename, std::ios::binary);
    char magic[8];
    uint64_t dims[2] = {0, 0};
    if (!in.read(magic, 8) || std::memcmp(magic, kMagic, 8) != 0 || !in.read(reinterpret_cast<char *>(dims), 16)) {
        std::cerr << "Error: " << filename << " is not a PLSBMAT1 matrix file" << std::endl;
        exit(1);
    }
    Mat2D X(dims[0], dims[1]);
    if (!in.read(reinterpret_cast<char *>(X.data()), (std::streamsize)(dims[0] * dims[1] * sizeof(double)))) {
        std::cerr << "Error: " << filename << " is truncated (" << dims[0] << " x " << dims[1] << " expected)" << std::endl;
        exit(1);
    }
    return X;
}

void write_matrix_binary(const std::string &filename, const Mat2D &mat) {
    std::ofstream out(filename, std::ios::binary);
    const uint64_t dims[2] = {(uint64_t)mat.rows(), (uint64_t)mat.cols()};
    out.write(kMagic, 8);
    out.write(reinterpret_cast<const char *>(dims), 16);
    out.write(reinterpret_cast<const char *>(mat.data()), (std::streamsize)(dims[0] * dims[1] * sizeof(double)));
}

// ---- column statistics: on the device (pls_b200/csrc/stats.cu), two-pass formulas like the reference --------
namespace {
inline int64_t ldof(const Mat2D &m) { return m.rows() > 0 ? (int64_t)m.rows() : 1; }
plsb_ctx_t engine_ctx() { return detail::Engine::get().primary(); }
}  // namespace

Row SST(const Mat2D &mat, const Row &means) {
    if (mat.rows() < 2) return Row::Zero(mat.cols());
    Row out(mat.cols());
    std::lock_guard<std::recursive_mutex> lk(detail::Engine::get().mu);
    detail::ok(plsb_colstats(engine_ctx(), mat.data(), ldof(mat), mat.rows(), mat.cols(), means.data(), nullptr, out.data(),
                             nullptr), "plsb_colstats");
    return out;
}

Row SST(const Mat2D &mat) {
    if (mat.rows() < 2) return Row::Zero(mat.cols());
    Row out(mat.cols());
    std::lock_guard<std::recursive_mutex> lk(detail::Engine::get().mu);
    detail::ok(plsb_colstats(engine_ctx(), mat.data(), ldof(mat), mat.rows(), mat.cols(), nullptr, nullptr, out.data(), nullptr),
               "plsb_colstats");
    return out;
}

Row colwise_stdev(const Mat2D &mat, const Row &means) {
    Row sd(mat.cols());
    std::lock_guard<std::recursive_mutex> lk(detail::Engine::get().mu);
    detail::ok(plsb_colstats(engine_ctx(), mat.data(), ldof(mat), mat.rows(), mat.cols(), means.data(), nullptr, nullptr,
                             sd.data()), "plsb_colstats");
    return sd;
}

Row colwise_stdev(const Mat2D &mat) {
    Row sd(mat.cols());
    std::lock_guard<std::recursive_mutex> lk(detail::Engine::get().mu);
    detail::ok(plsb_colstats(engine_ctx(), mat.data(), ldof(mat), mat.rows(), mat.cols(), nullptr, nullptr, nullptr, sd.data()),
               "plsb_colstats");
    return sd;
}

Row z_scores(const Row &obs, const Row &mean, const Row &stdev) {
    Row z(obs.size());  // one observation: K subtractions and divisions, not worth a launch
    for (Eigen::Index c = 0; c < obs.size(); ++c) z[c] = (obs[c] - mean[c]) / stdev[c];
    return z;
}

Mat2D colwise_z_scores(const Mat2D &mat, const Row &mean, const Row &stdev) {
    // Parity note: the reference prepares a zero -> one patched copy of stdev but then
    // divides by the ORIGINAL stdev, so a constant column becomes NaN.  Replicated, not fixed.
    Mat2D z(mat.rows(), mat.cols());
    std::lock_guard<std::recursive_mutex> lk(detail::Engine::get().mu);
    detail::ok(plsb_colwise_z_scores(engine_ctx(), mat.data(), ldof(mat), mat.rows(), mat.cols(), mean.data(), stdev.data(),
                                     z.data(), ldof(z)), "plsb_colwise_z_scores");
    return z;
}

Mat2D colwise_z_scores(const Mat2D &mat) {
    Mat2D z(mat.rows(), mat.cols());
    std::lock_guard<std::recursive_mutex> lk(detail::Engine::get().mu);
    detail::ok(plsb_colwise_z_scores(engine_ctx(), mat.data(), ldof(mat), mat.rows(), mat.cols(), nullptr, nullptr, z.data(),
                                     ldof(z)), "plsb_colwise_z_scores");
    return z;
}

// ---- rank statistics ----------------------------------------------------------------------------------
float_type normalcdf(const float_type z) { return plsb_normalcdf(z); }

float_type wilcoxon(const Col &err_1, const Col &err_2) {
    return plsb_wilcoxon(err_1.data(), err_2.data(), (int64_t)err_1.size());
}

void rand_nchoosek(std::mt19937 &rng, std::vector<Eigen::Index> &full, std::vector<Eigen::Index> &sample,
                   std::vector<Eigen::Index> &complement) {
    // libstdc++'s std::shuffle on the persistent permutation: the first |sample| entries train, the rest test
    std::shuffle(full.begin(), full.end(), rng);
    const auto cut = full.begin() + static_cast<std::ptrdiff_t>(sample.size());
    std::copy(full.begin(), cut, sample.begin());
    std::copy(cut, full.end(), complement.begin());
}

// ---- summaries ------------------------------------------------------------------------------------------
namespace {
// Residual::errors() -> the C-ABI's stacked block layout: M column-major (rows x A) matrices
std::vector<double> stack(const std::vector<Mat2D> &errors) {
    const size_t per = (size_t)(errors[0].rows() * errors[0].cols());
    std::vector<double> buf(errors.size() * per);
    for (size_t m = 0; m < errors.size(); ++m) std::copy(errors[m].data(), errors[m].data() + per, buf.begin() + m * per);
    return buf;
}
}  // namespace

Mat2D validation(const Residual &residual, const VALIDATION_OUTPUT out_type) {
    const std::vector<Mat2D> errors = residual.errors();
    if (errors.empty()) return Mat2D::Zero(0, 0);
    const int64_t M = (int64_t)errors.size(), rows = errors[0].rows(), A = errors[0].cols();
    Mat2D out(M, A);
    const std::vector<double> buf = stack(errors);
    detail::ok(plsb_validation(buf.data(), M, rows, A, out_type == RESS ? 0 : 1, out.data()), "plsb_validation");
    if (out_type != RESS && out_type != MSE) return out.cwiseSqrt();  // unreachable through the enum (App. B-6)
    return out;
}

Colsz optimal_num_components(const Residual &residual, const float_type ALPHA) {
    const std::vector<Mat2D> errors = residual.errors();
    Colsz best(errors.size());
    if (errors.empty()) return best;
    const int64_t M = (int64_t)errors.size(), rows = errors[0].rows(), A = errors[0].cols();
    std::vector<int64_t> pick((size_t)M);
    const std::vector<double> buf = stack(errors);
    detail::ok(plsb_optimal_num_components(buf.data(), M, rows, A, ALPHA, pick.data()), "plsb_optimal_num_components");
    for (int64_t m = 0; m < M; ++m) best[m] = (size_t)pick[(size_t)m];
    return best;
}

void print_validation(const Residual &residual, const VALIDATION_OUTPUT out_type, std::ostream &os) {
    os << residual.method() << " Validation:" << std::endl;
    Mat2D em = validation(residual, out_type);
    if (out_type == MSE) {
        os << "RMSE ";
        em = em.cwiseSqrt();
    } else if (out_type == RESS) {
        os << "PRESS ";
    } else {
        os << "UNKNOWN ";
    }
    os << " Matrix (rows = Y variable; cols = # of components):" << std::endl << em << std::endl;
    os << "Optimal number of components (by Y variable):\t" << optimal_num_components(residual) << std::endl;
}

}  // namespace PLS
