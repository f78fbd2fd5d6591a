// pls_b200/csrc/stats.cu — column statistics and z-scoring on the device.
//
// Replaces (reference src/pls.cpp): SST 69-77, colwise_stdev 79-87, colwise_z_scores 93-111 — the
// pre-processing the CLI applies to X and Y before the fit (main.cpp:24-25).  Same two-pass formulas as the
// reference (column mean, then centred sum of squares; N-1 in the variance), not Welford, so results agree
// to rounding.  A column with zero variance becomes NaN exactly as in the reference.
// All kernels stream the row-major device matrix once with coalesced 128-bit loads; per-CTA partial rows are
// summed in CTA order (launch_reduce_partials): no atomics, reproducible.
#include "internal.h"

namespace plsb {

// part[cta][k] = sum over the CTA's rows of (X[i][k] - mu[k])^2, or of X[i][k] when mu == nullptr
__global__ void __launch_bounds__(256) colsum_kernel(const double *__restrict__ X, int64_t N, int ldk,
                                                     const double *__restrict__ mu, double *__restrict__ part, int64_t ldp) {
    const int64_t rows_per = (N + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per, r1 = r0 + rows_per < N ? r0 + rows_per : N;
    double *prow = part + (size_t)blockIdx.x * ldp;
    for (int c0 = 2 * (int)threadIdx.x; c0 < ldk; c0 += 512) {
        const double m0 = mu ? mu[c0] : 0.0, m1 = mu ? mu[c0 + 1] : 0.0;
        double a0 = 0.0, a1 = 0.0, b0 = 0.0, b1 = 0.0;
        int64_t i = r0;
        for (; i + 1 < r1; i += 2) {
            const double2 x = *reinterpret_cast<const double2 *>(X + i * ldk + c0);
            const double2 y = *reinterpret_cast<const double2 *>(X + (i + 1) * ldk + c0);
            if (mu) {
                a0 = fma(x.x - m0, x.x - m0, a0); a1 = fma(x.y - m1, x.y - m1, a1);
                b0 = fma(y.x - m0, y.x - m0, b0); b1 = fma(y.y - m1, y.y - m1, b1);
            } else {
                a0 += x.x; a1 += x.y; b0 += y.x; b1 += y.y;
            }
        }
        if (i < r1) {
            const double2 x = *reinterpret_cast<const double2 *>(X + i * ldk + c0);
            if (mu) { a0 = fma(x.x - m0, x.x - m0, a0); a1 = fma(x.y - m1, x.y - m1, a1); }
            else { a0 += x.x; a1 += x.y; }
        }
        prow[c0] = a0 + b0;
        prow[c0 + 1] = a1 + b1;
    }
}

__global__ void __launch_bounds__(256) scale_vec_kernel(double *v, int n, double f) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i < n) v[i] *= f;
}

// sd[k] = sqrt(sst[k] / (N - 1))   (pls.cpp:79-83)
__global__ void __launch_bounds__(256) stdev_kernel(const double *sst, int n, double dof, double *sd) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i < n) sd[i] = sqrt(sst[i] / dof);
}

// X[i][k] = (X[i][k] - mu[k]) / sd[k] for k < K (padding columns stay zero)   (pls.cpp:93-104)
__global__ void __launch_bounds__(256) zscore_kernel(double *__restrict__ X, int64_t N, int ldk, int K,
                                                     const double *__restrict__ mu, const double *__restrict__ sd) {
    const int64_t total = N * (int64_t)(ldk / 2);
    for (int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x; e < total; e += (int64_t)gridDim.x * 256) {
        const int64_t i = e / (ldk / 2);
        const int c0 = 2 * (int)(e % (ldk / 2));
        double2 *p = reinterpret_cast<double2 *>(X + i * ldk + c0);
        double2 x = *p;
        x.x = (x.x - mu[c0]) / sd[c0];
        x.y = c0 + 1 < K ? (x.y - mu[c0 + 1]) / sd[c0 + 1] : 0.0;
        *p = x;
    }
}

// mean (ldk), sst (ldk) of a device row-major matrix; either output may be null.  means_in: use these instead of
// computing them (the reference's SST(mat, means) overload).
int device_colstats(plsb_ctx *c, const double *X, int64_t N, int ldk, int K, const double *means_in, double *mean,
                    double *sst) {
    const int64_t grid = (int64_t)c->sm_count * 4;
    PLSB_TRY(ensure_part(c, (size_t)grid * ldk * sizeof(double)));
    double *mu = mean;
    void *tmp = nullptr;
    if (!mu) { PLSB_TRY(dev_alloc(c, &tmp, (size_t)ldk * sizeof(double))); mu = (double *)tmp; }
    if (means_in) {
        PLSB_CUDA(cudaMemcpyAsync(mu, means_in, (size_t)ldk * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
    } else {
        colsum_kernel<<<(unsigned)grid, 256, 0, c->stream>>>(X, N, ldk, nullptr, c->part, ldk);
        PLSB_CUDA(cudaGetLastError());
        c->launches++;
        PLSB_TRY(launch_reduce_partials(c, c->part, ldk, (int)grid, ldk, mu, false));
        scale_vec_kernel<<<(unsigned)ceil_div(ldk, 256), 256, 0, c->stream>>>(mu, ldk, N > 0 ? 1.0 / (double)N : 0.0);
        c->launches++;
    }
    if (sst) {
        if (N < 2) {  // the reference returns zeros for fewer than two rows (pls.cpp:70-71)
            PLSB_CUDA(cudaMemsetAsync(sst, 0, (size_t)ldk * sizeof(double), c->stream));
        } else {
            colsum_kernel<<<(unsigned)grid, 256, 0, c->stream>>>(X, N, ldk, mu, c->part, ldk);
            PLSB_CUDA(cudaGetLastError());
            c->launches++;
            PLSB_TRY(launch_reduce_partials(c, c->part, ldk, (int)grid, ldk, sst, false));
        }
    }
    dev_free(c, tmp);
    (void)K;
    return PLSB_OK;
}

int device_zscore(plsb_ctx *c, double *X, int64_t N, int ldk, int K, double *mean, double *sd) {
    PLSB_TRY(device_colstats(c, X, N, ldk, K, nullptr, mean, sd));  // sd holds SST for now
    stdev_kernel<<<(unsigned)ceil_div(ldk, 256), 256, 0, c->stream>>>(sd, ldk, (double)N - 1.0, sd);
    if (N > 0) zscore_kernel<<<(unsigned)(c->sm_count * 8), 256, 0, c->stream>>>(X, N, ldk, K, mean, sd);
    c->launches += 2;
    PLSB_CUDA(cudaGetLastError());
    return PLSB_OK;
}

}  // namespace plsb

using namespace plsb;

namespace {
struct Buf {
    plsb_ctx *c;
    double *p = nullptr;
    explicit Buf(plsb_ctx *ctx) : c(ctx) {}
    ~Buf() { dev_free(c, p); }
    int alloc(size_t n, bool zero) {
        PLSB_TRY(dev_alloc(c, (void **)&p, (n ? n : 1) * sizeof(double)));
        if (zero) PLSB_CUDA(cudaMemsetAsync(p, 0, (n ? n : 1) * sizeof(double), c->stream));
        return PLSB_OK;
    }
};
}  // namespace

extern "C" {

int plsb_colstats(plsb_ctx_t c, const double *X, int64_t ldx, int64_t N, int64_t K, const double *means_in,
                  double *mean, double *sst, double *stdev) {
    PLSB_ARG(c && (N == 0 || X) && K >= 1 && ldx >= N, "bad argument");
    PLSB_CUDA(cudaSetDevice(c->device));
    const int64_t ldk = round_up(K, 2);
    Buf Xd(c), mu(c), ss(c), mi(c);
    PLSB_TRY(Xd.alloc((size_t)N * ldk, ldk != K));
    PLSB_TRY(mu.alloc((size_t)ldk, true));
    PLSB_TRY(ss.alloc((size_t)ldk, true));
    PLSB_TRY(upload_cm_to_rm(c, X, ldx, N, K, Xd.p, ldk, false));
    if (means_in) {
        PLSB_TRY(mi.alloc((size_t)ldk, true));
        PLSB_CUDA(cudaMemcpyAsync(mi.p, means_in, (size_t)K * 8, cudaMemcpyHostToDevice, c->stream));
    }
    PLSB_TRY(device_colstats(c, Xd.p, N, (int)ldk, (int)K, means_in ? mi.p : nullptr, mu.p, ss.p));
    if (mean) PLSB_CUDA(cudaMemcpyAsync(mean, mu.p, (size_t)K * 8, cudaMemcpyDeviceToHost, c->stream));
    if (sst) PLSB_CUDA(cudaMemcpyAsync(sst, ss.p, (size_t)K * 8, cudaMemcpyDeviceToHost, c->stream));
    if (stdev) {
        stdev_kernel<<<(unsigned)ceil_div(ldk, 256), 256, 0, c->stream>>>(ss.p, (int)ldk, (double)N - 1.0, ss.p);
        c->launches++;
        PLSB_CUDA(cudaMemcpyAsync(stdev, ss.p, (size_t)K * 8, cudaMemcpyDeviceToHost, c->stream));
    }
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    return PLSB_OK;
}

int plsb_colwise_z_scores(plsb_ctx_t c, const double *X, int64_t ldx, int64_t N, int64_t K, const double *mean_in,
                          const double *stdev_in, double *Z, int64_t ldz) {
    PLSB_ARG(c && (N == 0 || (X && Z)) && K >= 1 && ldx >= N && ldz >= N, "bad argument");
    PLSB_ARG((mean_in == nullptr) == (stdev_in == nullptr), "give both mean and stdev, or neither");
    PLSB_CUDA(cudaSetDevice(c->device));
    const int64_t ldk = round_up(K, 2);
    Buf Xd(c), mu(c), sd(c);
    PLSB_TRY(Xd.alloc((size_t)N * ldk, ldk != K));
    PLSB_TRY(mu.alloc((size_t)ldk, true));
    PLSB_TRY(sd.alloc((size_t)ldk, true));
    PLSB_TRY(upload_cm_to_rm(c, X, ldx, N, K, Xd.p, ldk, false));
    if (mean_in) {
        PLSB_CUDA(cudaMemcpyAsync(mu.p, mean_in, (size_t)K * 8, cudaMemcpyHostToDevice, c->stream));
        PLSB_CUDA(cudaMemcpyAsync(sd.p, stdev_in, (size_t)K * 8, cudaMemcpyHostToDevice, c->stream));
        if (N > 0) zscore_kernel<<<(unsigned)(c->sm_count * 8), 256, 0, c->stream>>>(Xd.p, N, (int)ldk, (int)K, mu.p, sd.p);
        c->launches++;
        PLSB_CUDA(cudaGetLastError());
    } else {
        PLSB_TRY(device_zscore(c, Xd.p, N, (int)ldk, (int)K, mu.p, sd.p));
    }
    // back to the caller's column-major layout, in column blocks
    if (N > 0) {
        int64_t kc = (int64_t)(256u << 20) / (N * 8);
        if (kc < 1) kc = 1;
        if (kc > K) kc = K;
        Buf stage(c);
        PLSB_TRY(stage.alloc((size_t)N * kc, false));
        for (int64_t c0 = 0; c0 < K; c0 += kc) {
            const int64_t nc = K - c0 < kc ? K - c0 : kc;
            PLSB_TRY(launch_rm_to_cm(c, Xd.p + c0, ldk, N, nc, stage.p, N));
            PLSB_CUDA(cudaMemcpy2DAsync(Z + c0 * ldz, (size_t)ldz * 8, stage.p, (size_t)N * 8, (size_t)N * 8, (size_t)nc,
                                        cudaMemcpyDeviceToHost, c->stream));
            PLSB_CUDA(cudaStreamSynchronize(c->stream));
        }
    }
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    return PLSB_OK;
}

int plsb_dataset_zscore(plsb_dataset_t ds, double *x_mean, double *x_stdev, double *y_mean, double *y_stdev) {
    PLSB_ARG(ds, "dataset is null");
    plsb_ctx *c = ds->ctx;
    PLSB_ARG(!ds->sharded, "z-scoring a row shard needs the global column statistics: standardise before sharding");
    PLSB_CUDA(cudaSetDevice(c->device));
    Buf mx(c), sx(c), my(c), sy(c);
    PLSB_TRY(mx.alloc((size_t)ds->ldk, true)); PLSB_TRY(sx.alloc((size_t)ds->ldk, true));
    PLSB_TRY(my.alloc((size_t)ds->ldm, true)); PLSB_TRY(sy.alloc((size_t)ds->ldm, true));
    ds->gram_valid = false; ds->gram_agreed = -1;   // X changes in place: a Gram matrix accumulated during the upload no longer applies
    PLSB_TRY(device_zscore(c, ds->X, ds->N, (int)ds->ldk, (int)ds->K, mx.p, sx.p));
    PLSB_TRY(device_zscore(c, ds->Y, ds->N, (int)ds->ldm, (int)ds->M, my.p, sy.p));
    if (x_mean) PLSB_CUDA(cudaMemcpyAsync(x_mean, mx.p, (size_t)ds->K * 8, cudaMemcpyDeviceToHost, c->stream));
    if (x_stdev) PLSB_CUDA(cudaMemcpyAsync(x_stdev, sx.p, (size_t)ds->K * 8, cudaMemcpyDeviceToHost, c->stream));
    if (y_mean) PLSB_CUDA(cudaMemcpyAsync(y_mean, my.p, (size_t)ds->M * 8, cudaMemcpyDeviceToHost, c->stream));
    if (y_stdev) PLSB_CUDA(cudaMemcpyAsync(y_stdev, sy.p, (size_t)ds->M * 8, cudaMemcpyDeviceToHost, c->stream));
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    return PLSB_OK;
}

}  // extern "C"
