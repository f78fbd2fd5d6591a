// pls_b200/csrc/predict.cu — prediction helpers of PLS::Model on the device.
//
// Replaces (reference src/pls.cpp): scores 439-442 (X_new R_c), coefficients 444-447 + fitted_values
// 449-451 + residuals 453-455, and cv_NEW_DATA 494-510.  The reference rebuilds R_c Q_c' and re-reads
// X_new for every component count; here ONE pass over X_new yields the scores S = X_new R for all A
// components and the residuals for every prefix c follow from a running sum of s_a q_a'.
#include <vector>

#include "internal.h"

namespace plsb {

constexpr int PR_RW = 4;  // rows per warp
constexpr int PR_AC = 4;  // components per register block

// S[a*Nn + i] = sum_k Xn[i*ldk + k] * R[a*ldk + k]
__global__ void __launch_bounds__(256) scores_kernel(const double *__restrict__ Xn, int64_t Nn, int ldk, int K,
                                                     const double *__restrict__ R, int A, double *__restrict__ S) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t row0 = ((int64_t)blockIdx.x * 8 + warp) * PR_RW;
    if (row0 >= Nn) return;
    const double *xr[PR_RW];
#pragma unroll
    for (int i = 0; i < PR_RW; ++i) xr[i] = Xn + (row0 + i < Nn ? row0 + i : row0) * (int64_t)ldk;
    for (int a0 = 0; a0 < A; a0 += PR_AC) {
        double acc[PR_RW][PR_AC];
#pragma unroll
        for (int i = 0; i < PR_RW; ++i)
#pragma unroll
            for (int j = 0; j < PR_AC; ++j) acc[i][j] = 0.0;
        const double *rr[PR_AC];
#pragma unroll
        for (int j = 0; j < PR_AC; ++j) rr[j] = R + (size_t)(a0 + j < A ? a0 + j : a0) * ldk;
        for (int k = lane; k < K; k += 32) {
            double x[PR_RW], r[PR_AC];
#pragma unroll
            for (int i = 0; i < PR_RW; ++i) x[i] = xr[i][k];
#pragma unroll
            for (int j = 0; j < PR_AC; ++j) r[j] = rr[j][k];
#pragma unroll
            for (int i = 0; i < PR_RW; ++i)
#pragma unroll
                for (int j = 0; j < PR_AC; ++j) acc[i][j] = fma(x[i], r[j], acc[i][j]);
        }
#pragma unroll
        for (int i = 0; i < PR_RW; ++i)
#pragma unroll
            for (int j = 0; j < PR_AC; ++j) {
                const double v = warp_sum(acc[i][j]);
                if (lane == 0 && row0 + i < Nn && a0 + j < A) S[(size_t)(a0 + j) * Nn + row0 + i] = v;
            }
    }
}

// running prefix over components: acc_c = sum_{a<=c} S[a][i] Q[a][m]
// mode 0: E[(m*A + c)*Nn + i] = Y[i][m] - acc_c for every c     (cv_NEW_DATA layout)
// mode 1: F[m*Nn + i] = acc_{comp-1}            (fitted values)
// mode 2: F[m*Nn + i] = Y[i][m] - acc_{comp-1}  (residuals)
__global__ void __launch_bounds__(256) prefix_residual_kernel(const double *__restrict__ S, const double *__restrict__ Q,
                                                              const double *__restrict__ Y, int ldm, int64_t Nn, int M,
                                                              int A, int comp, int mode, double *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int m = blockIdx.y;
    if (i >= Nn) return;
    const double y = (mode != 1) ? Y[i * ldm + m] : 0.0;
    double acc = 0.0;
    for (int c = 0; c < comp; ++c) {
        acc = fma(S[(size_t)c * Nn + i], Q[(size_t)c * M + m], acc);
        if (mode == 0) out[((size_t)m * A + c) * Nn + i] = y - acc;
    }
    if (mode == 1) out[(size_t)m * Nn + i] = acc;
    if (mode == 2) out[(size_t)m * Nn + i] = y - acc;
}

// B[m*K + k] = sum_{a<comp} R[a*K + k] * Q[a*M + m]   (K x M column-major; pls.cpp:444-447)
__global__ void __launch_bounds__(256) coefficients_kernel(const double *__restrict__ R, const double *__restrict__ Q,
                                                           int K, int M, int comp, double *__restrict__ B) {
    const int k = blockIdx.x * 256 + threadIdx.x, m = blockIdx.y;
    if (k >= K) return;
    double acc = 0.0;
    for (int a = 0; a < comp; ++a) acc = fma(R[(size_t)a * K + k], Q[(size_t)a * M + m], acc);
    B[(size_t)m * K + k] = acc;
}

// out[col] = sum_i E[col*Nn + i]^2, one CTA per column, fixed summation tree (pls.cpp:457-459, 251)
__global__ void __launch_bounds__(256) colsumsq_kernel(const double *__restrict__ E, int64_t Nn, double *__restrict__ out) {
    __shared__ double red[8];
    const double *e = E + (size_t)blockIdx.x * Nn;
    double s0 = 0.0, s1 = 0.0;
    int64_t i = threadIdx.x;
    for (; i + 256 < Nn; i += 512) {
        s0 = fma(e[i], e[i], s0);
        s1 = fma(e[i + 256], e[i + 256], s1);
    }
    for (; i < Nn; i += 256) s0 = fma(e[i], e[i], s0);
    const double w = warp_sum(s0 + s1);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = w;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int j = 0; j < 8; ++j) t += red[j];
        out[blockIdx.x] = t;
    }
}

struct Tmp {
    plsb_ctx *c;
    double *p = nullptr;
    explicit Tmp(plsb_ctx *ctx) : c(ctx) {}
    ~Tmp() { dev_free(c, p); }
    int alloc(size_t n) { return dev_alloc(c, (void **)&p, (n ? n : 1) * sizeof(double)); }
};

// upload Xn (row-major, padded) and the first `comp` columns of R; compute S (comp x Nn on device)
static int scores_device(plsb_ctx *c, const double *R, int64_t K, int64_t comp, const double *Xn, int64_t ldx,
                         int64_t Nn, Tmp &Xd, Tmp &Rd, Tmp &Sd) {
    const int64_t ldk = round_up(K, 2);
    PLSB_TRY(Xd.alloc((size_t)Nn * ldk));
    PLSB_TRY(Rd.alloc((size_t)comp * ldk));
    PLSB_TRY(Sd.alloc((size_t)comp * Nn));
    if (ldk != K) {
        PLSB_CUDA(cudaMemsetAsync(Xd.p, 0, (size_t)(Nn ? Nn : 1) * ldk * 8, c->stream));
        PLSB_CUDA(cudaMemsetAsync(Rd.p, 0, (size_t)comp * ldk * 8, c->stream));
    }
    PLSB_TRY(upload_cm_to_rm(c, Xn, ldx, Nn, K, Xd.p, ldk, false));
    PLSB_CUDA(cudaMemcpy2DAsync(Rd.p, (size_t)ldk * 8, R, (size_t)K * 8, (size_t)K * 8, (size_t)comp,
                                cudaMemcpyHostToDevice, c->stream));
    // DMMA path (scores.cu); PLSB_SCORES_FMA=1 keeps the FMA kernel for A/B runs
    const char *fma = getenv("PLSB_SCORES_FMA");
    if (Nn > 0 && !(fma && *fma == '1')) {
        PLSB_TRY(launch_scores(c, Xd.p, Nn, (int)ldk, Rd.p, (int)comp, Sd.p, Nn));
    } else if (Nn > 0) {
        scores_kernel<<<(unsigned)ceil_div(Nn, 8 * PR_RW), 256, 0, c->stream>>>(Xd.p, Nn, (int)ldk, (int)K, Rd.p,
                                                                               (int)comp, Sd.p);
        c->launches++;
        PLSB_CUDA(cudaGetLastError());
    }
    return PLSB_OK;
}

static int predict_common(plsb_ctx *c, const double *R, const double *Q, int64_t K, int64_t M, int64_t A, int64_t comp,
                          const double *Xn, int64_t ldx, int64_t Nn, const double *Yn, int64_t ldy, int mode,
                          double *out, bool want_sse = false) {
    PLSB_ARG(c && R && Q && out && (Nn == 0 || Xn), "null argument");
    PLSB_ARG(comp >= 1 && comp <= A, "component count %lld out of range", (long long)comp);
    PLSB_ARG(mode == 1 || Nn == 0 || Yn, "Y is null");
    PLSB_CUDA(cudaSetDevice(c->device));
    if (Nn == 0) {
        if (want_sse) for (int64_t i = 0; i < M * (mode == 0 ? A : 1); ++i) out[i] = 0.0;
        return PLSB_OK;
    }
    Tmp Xd(c), Rd(c), Sd(c), Qd(c), Yd(c), Od(c);
    PLSB_TRY(scores_device(c, R, K, comp, Xn, ldx, Nn, Xd, Rd, Sd));
    const int64_t ldm = round_up(M, 2);
    PLSB_TRY(Qd.alloc((size_t)comp * M));
    PLSB_CUDA(cudaMemcpyAsync(Qd.p, Q, (size_t)comp * M * 8, cudaMemcpyHostToDevice, c->stream));
    if (mode != 1) {
        PLSB_TRY(Yd.alloc((size_t)Nn * ldm));
        PLSB_TRY(upload_cm_to_rm(c, Yn, ldy, Nn, M, Yd.p, ldm, false));
    }
    const size_t on = (mode == 0) ? (size_t)M * A * Nn : (size_t)M * Nn;
    PLSB_TRY(Od.alloc(on));
    dim3 grid((unsigned)ceil_div(Nn, 256), (unsigned)M);
    prefix_residual_kernel<<<grid, 256, 0, c->stream>>>(Sd.p, Qd.p, Yd.p, (int)ldm, Nn, (int)M, (int)A, (int)comp, mode,
                                                        Od.p);
    c->launches++;
    PLSB_CUDA(cudaGetLastError());
    if (want_sse) {  // column sums of squares of the residual block: (m, c) -> out[c*M + m]
        const int64_t nc = (mode == 0) ? A : 1;  // component counts per response in the block
        Tmp Sd2(c);
        PLSB_TRY(Sd2.alloc((size_t)M * nc));
        colsumsq_kernel<<<(unsigned)(M * nc), 256, 0, c->stream>>>(Od.p, Nn, Sd2.p);
        c->launches++;
        PLSB_CUDA(cudaGetLastError());
        std::vector<double> tmp((size_t)M * nc);
        PLSB_CUDA(cudaMemcpyAsync(tmp.data(), Sd2.p, (size_t)M * nc * 8, cudaMemcpyDeviceToHost, c->stream));
        PLSB_CUDA(cudaStreamSynchronize(c->stream));
        for (int64_t m = 0; m < M; ++m)
            for (int64_t a = 0; a < nc; ++a) out[a * M + m] = tmp[(size_t)(m * nc + a)];
        return PLSB_OK;
    }
    PLSB_CUDA(cudaMemcpyAsync(out, Od.p, on * 8, cudaMemcpyDeviceToHost, c->stream));
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    return PLSB_OK;
}

}  // namespace plsb

using namespace plsb;

extern "C" {

int plsb_scores(plsb_ctx_t c, const double *R, int64_t K, int64_t comp, const double *Xn, int64_t ldx, int64_t Nn,
                double *S) {
    PLSB_ARG(c && R && S && (Nn == 0 || Xn), "null argument");
    PLSB_ARG(comp >= 1, "component count must be >= 1");
    PLSB_CUDA(cudaSetDevice(c->device));
    if (Nn == 0) return PLSB_OK;
    Tmp Xd(c), Rd(c), Sd(c);
    PLSB_TRY(scores_device(c, R, K, comp, Xn, ldx, Nn, Xd, Rd, Sd));
    PLSB_CUDA(cudaMemcpyAsync(S, Sd.p, (size_t)comp * Nn * 8, cudaMemcpyDeviceToHost, c->stream));
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    return PLSB_OK;
}

int plsb_fitted(plsb_ctx_t c, const double *R, const double *Q, int64_t K, int64_t M, int64_t comp, const double *Xn,
                int64_t ldx, int64_t Nn, const double *Yn, int64_t ldy, double *F) {
    return predict_common(c, R, Q, K, M, comp, comp, Xn, ldx, Nn, Yn, ldy, Yn ? 2 : 1, F);
}

int plsb_predict_residuals(plsb_ctx_t c, const double *R, const double *Q, int64_t K, int64_t M, int64_t A,
                           const double *Xn, int64_t ldx, int64_t Nn, const double *Yn, int64_t ldy, double *E) {
    return predict_common(c, R, Q, K, M, A, A, Xn, ldx, Nn, Yn, ldy, 0, E);
}

int plsb_sse_prefixes(plsb_ctx_t c, const double *R, const double *Q, int64_t K, int64_t M, int64_t A, const double *Xn,
                      int64_t ldx, int64_t Nn, const double *Yn, int64_t ldy, double *sse) {
    return predict_common(c, R, Q, K, M, A, A, Xn, ldx, Nn, Yn, ldy, 0, sse, true);
}

int plsb_sse(plsb_ctx_t c, const double *R, const double *Q, int64_t K, int64_t M, int64_t comp, const double *Xn,
             int64_t ldx, int64_t Nn, const double *Yn, int64_t ldy, double *sse) {
    PLSB_ARG(Nn == 0 || Yn, "Y is null");
    return predict_common(c, R, Q, K, M, comp, comp, Xn, ldx, Nn, Yn, ldy, 2, sse, true);
}

int plsb_coefficients(plsb_ctx_t c, const double *R, const double *Q, int64_t K, int64_t M, int64_t comp, double *B) {
    PLSB_ARG(c && R && Q && B, "null argument");
    PLSB_ARG(comp >= 0 && K >= 1 && M >= 1, "bad shape");
    PLSB_CUDA(cudaSetDevice(c->device));
    Tmp Rd(c), Qd(c), Bd(c);
    PLSB_TRY(Rd.alloc((size_t)comp * K));
    PLSB_TRY(Qd.alloc((size_t)comp * M));
    PLSB_TRY(Bd.alloc((size_t)K * M));
    PLSB_CUDA(cudaMemcpyAsync(Rd.p, R, (size_t)comp * K * 8, cudaMemcpyHostToDevice, c->stream));
    PLSB_CUDA(cudaMemcpyAsync(Qd.p, Q, (size_t)comp * M * 8, cudaMemcpyHostToDevice, c->stream));
    dim3 grid((unsigned)ceil_div(K, 256), (unsigned)M);
    coefficients_kernel<<<grid, 256, 0, c->stream>>>(Rd.p, Qd.p, (int)K, (int)M, (int)comp, Bd.p);
    c->launches++;
    PLSB_CUDA(cudaGetLastError());
    PLSB_CUDA(cudaMemcpyAsync(B, Bd.p, (size_t)K * M * 8, cudaMemcpyDeviceToHost, c->stream));
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    return PLSB_OK;
}

}  // extern "C"
