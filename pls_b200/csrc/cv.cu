// pls_b200/csrc/cv.cu — batched cross-validation (placeholder until the Gram-downdate engine lands)
#include "internal.h"
using namespace plsb;
extern "C" int plsb_cv_folds(plsb_dataset_t ds, int64_t A, int method, int64_t nfolds_total, int64_t test_size,
                             const int64_t *test_idx, int64_t fold0, int64_t nfolds_local, double *E) {
    (void)ds; (void)A; (void)method; (void)nfolds_total; (void)test_size; (void)test_idx; (void)fold0;
    (void)nfolds_local; (void)E;
    set_error("plsb_cv_folds: not implemented yet");
    return PLSB_ERR_UNSUPPORTED;
}
