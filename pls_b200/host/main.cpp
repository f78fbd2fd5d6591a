// pls_b200/host/main.cpp — the `PLS X.csv Y.csv ncomp` command line tool on the B200 engine.
//
// Same argument handling, processing order, output stream (stderr) and exit codes as the reference's
// src/main.cpp:10-44: z-score both matrices, fit KERNEL_TYPE1 with ncomp components, print the model
// state and explained variance, then leave-one-out and leave-some-out (30 % held out, 10 N trials,
// default-seeded mt19937) validation reports.  One extension: a file whose name ends in ".bin" is read
// with read_matrix_binary (column-major FP64), the input format for matrices too large for CSV.
#include <cstdlib>
#include <iostream>
#include <random>
#include <string>

#include <PLS/pls.h>

using namespace PLS;

static Mat2D load(const std::string &name) {
    const bool binary = name.size() > 4 && name.compare(name.size() - 4, 4, ".bin") == 0;
    return binary ? read_matrix_binary(name) : read_matrix_file(name);
}

int main(int argc, char *argv[]) {
    if (argc != 4) {
        std::cerr << "Usage: ./pls X_data.csv Y_data.csv num_components" << std::endl;
        std::cerr << "NB: X and Y csvs must be comma delimited, square numerical data, with no headers." << std::endl;
        exit(100);
    }
    const Mat2D X = colwise_z_scores(load(argv[1]));
    const Mat2D Y = colwise_z_scores(load(argv[2]));
    const size_t ncomp = atoi(argv[3]);

    Model plsm(X, Y, METHOD::KERNEL_TYPE1, ncomp);
    plsm.print_state();
    plsm.print_explained_variance(X, Y);

    print_validation(plsm.cv_LOO(), MSE);

    std::mt19937 rng;
    print_validation(plsm.cv_LSO(0.3, 10 * X.rows(), rng), MSE);
    return 0;
}
