/* oracle/pls_oracle.h — TEST INFRASTRUCTURE ONLY (see pls_oracle.c header). */
#ifndef PLS_ORACLE_H
#define PLS_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* method: 0 = KERNEL_TYPE1, 1 = KERNEL_TYPE2 (pls.h:131). All matrices column-major, ld == rows. */
int orc_fit(const double *X, int64_t N, int64_t K, const double *Y, int64_t M, int64_t A, int method,
            double *W, double *P, double *R, double *Q, double *T);
void orc_coefficients(const double *R, const double *Q, int64_t K, int64_t M, int64_t comp, double *B);
void orc_scores(const double *R, int64_t K, int64_t comp, const double *Xn, int64_t Nn, double *S);
void orc_fitted_values(const double *R, const double *Q, int64_t K, int64_t M, int64_t comp,
                       const double *Xn, int64_t Nn, double *F);
void orc_residuals(const double *R, const double *Q, int64_t K, int64_t M, int64_t comp,
                   const double *Xn, int64_t Nn, const double *Yn, double *E);
void orc_sse(const double *R, const double *Q, int64_t K, int64_t M, int64_t comp, const double *Xn,
             int64_t Nn, const double *Yn, double *out);
void orc_sst(const double *X, int64_t N, int64_t K, double *out);
void orc_explained_variance(const double *R, const double *Q, int64_t K, int64_t M, int64_t comp,
                            const double *Xn, int64_t Nn, const double *Yn, double *out);
void orc_colwise_stdev(const double *X, int64_t N, int64_t K, double *out);
void orc_colwise_z_scores(const double *X, int64_t N, int64_t K, double *Z);
/* residuals for comp = 1..A of a stored model on new data: E = M stacked (Nn x A) blocks */
void orc_predict_residuals(const double *R, const double *Q, int64_t K, int64_t M, int64_t A,
                           const double *Xn, int64_t Nn, const double *Yn, double *E);
/* explicit-fold cross-validation; train_idx NULL => ascending complement of the test list */
int orc_cv_folds(const double *X, int64_t N, int64_t K, const double *Y, int64_t M, int64_t A, int method,
                 int64_t nfolds, int64_t test_size, const int64_t *test_idx, int64_t train_size,
                 const int64_t *train_idx, double *E);
void orc_validation(const double *E, int64_t M, int64_t rows, int64_t A, int out_type, double *out);
double orc_normalcdf(double z);
double orc_wilcoxon(const double *e1, const double *e2, int64_t n);
void orc_optimal_num_components(const double *E, int64_t M, int64_t rows, int64_t A, double alpha,
                                int64_t *out);
/* deterministic synthetic latent-factor data; rows [row0,row0+rows) of an N-row set */
void orc_synth(uint64_t seed, int64_t row0, int64_t rows, int64_t K, int64_t M, int64_t L,
               double *X, int64_t ldx, double *Y, int64_t ldy);
double orc_synth_u(uint64_t seed, uint64_t idx);

#ifdef __cplusplus
}
#endif
#endif
