// pls_b200/csrc/dataset.cu — device-resident data sets: the role of Model's `_X,_Y` copies
// (reference src/pls.cpp:344, include/PLS/pls.h:251), kept in HBM in the row-major layout the
// streaming passes want.
#include <string.h>

#include <chrono>
#include <thread>

#include "host_pool.h"
#include "internal.h"

namespace plsb {

// ---- pageable host sources ------------------------------------------------------------------------------------------
// cudaMemcpy from pageable memory is staged by the driver through its own bounce buffers on ONE thread (6-12 GB/s: a
// fifth of PCIe 5 x16).  A caller of the drop-in API hands over an ordinary Eigen matrix (pls.cpp:340-353), so the
// staging is done here instead: a pool of host threads packs block i+1 into a pinned write-combined buffer while the
// copy engine sends block i, three buffers deep.  Pinned (or registered) sources skip it and are copied in place.
static bool host_is_pageable(const void *p) {
    cudaPointerAttributes a{};
    const cudaError_t e = cudaPointerGetAttributes(&a, p);
    cudaGetLastError();
    return e != cudaSuccess || a.type == cudaMemoryTypeUnregistered;
}

static void hstage_pool(plsb_ctx *c) {
    plsb_ctx::HostStage &h = c->hstage;
    if (!h.pool) {
        int n = 0;
        if (const char *e = getenv("PLSB_PACK_THREADS")) n = atoi(e);
        if (n <= 0) {
            const int hw = (int)std::thread::hardware_concurrency(), live = live_contexts() > 0 ? live_contexts() : 1;
            n = hw / live;
            if (n > 16) n = 16;
            if (n < 2) n = 2;
        }
        h.pool = new HostPool(n);
    }
}

// device -> MAPPED pinned host memory by a kernel (posted PCIe writes, coalesced): for results of a few MB the copy
// engine's fixed set-up cost on this platform (~0.5 ms per transfer, measured: 4.3 MB in 0.75 ms) dominates the DMA
__global__ void __launch_bounds__(256) store_to_host_kernel(const double *__restrict__ src, double *__restrict__ dst, size_t n) {
    const size_t n2 = n / 2;
    const double2 *s2 = reinterpret_cast<const double2 *>(src);
    double2 *d2 = reinterpret_cast<double2 *>(dst);
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n2; i += (size_t)gridDim.x * 256) d2[i] = s2[i];
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) dst[n - 1] = src[n - 1];
}

int download_staged(plsb_ctx *c, const double *dev, size_t n, double *host, bool as_complex) {
    if (n == 0) return PLSB_OK;
    plsb_ctx::HostStage &h = c->hstage;
    hstage_pool(c);
    const size_t chunk = (size_t)32 << 20;
    if (!h.dbuf[0]) {
        for (int i = 0; i < 2; ++i) {
            PLSB_CUDA(cudaHostAlloc((void **)&h.dbuf[i], chunk, cudaHostAllocDefault));
            PLSB_CUDA(cudaEventCreateWithFlags(&h.ddone[i], cudaEventDisableTiming));
        }
        h.dbytes = chunk;
    }
    const size_t per = chunk / 8, nch = (n + per - 1) / per;
    if (nch == 1) {   // one chunk: nothing to overlap — copy on the context's own stream (no cross-stream hand-over)
        const char *ke = getenv("PLSB_D2H_KERNEL");
        if (n * 8 <= ((size_t)16 << 20) && ((uintptr_t)dev & 15) == 0 && !(ke && *ke == '0')) {
            store_to_host_kernel<<<(unsigned)(c->sm_count * 2), 256, 0, c->stream>>>(dev, (double *)h.dbuf[0], n);
            c->launches++;
            PLSB_CUDA(cudaGetLastError());
        } else {
            PLSB_CUDA(cudaMemcpyAsync(h.dbuf[0], dev, n * 8, cudaMemcpyDeviceToHost, c->stream));
        }
        PLSB_CUDA(cudaStreamSynchronize(c->stream));
        const double *src = (const double *)h.dbuf[0];
        h.pool->run([&](int t, int nt) {
            const size_t a = n / (size_t)nt * (size_t)t, b = (t + 1 == nt) ? n : n / (size_t)nt * (size_t)(t + 1);
            if (!as_complex) memcpy(host + a, src + a, (b - a) * 8);
            else {
                double *d = host + 2 * a;
                for (size_t j = a; j < b; ++j) { *d++ = src[j]; *d++ = 0.0; }
            }
        });
        return PLSB_OK;
    }
    PLSB_CUDA(cudaEventRecord(c->col_ev, c->stream));                 // the source is complete
    PLSB_CUDA(cudaStreamWaitEvent(c->copy_stream, c->col_ev, 0));
    auto move_out = [&](size_t i) {
        const size_t off = i * per, cnt = n - off < per ? n - off : per;
        const double *src = (const double *)h.dbuf[i & 1];
        h.pool->run([&](int t, int nt) {
            const size_t a = cnt / (size_t)nt * (size_t)t, b = (t + 1 == nt) ? cnt : cnt / (size_t)nt * (size_t)(t + 1);
            if (!as_complex) memcpy(host + off + a, src + a, (b - a) * 8);
            else {
                double *d = host + 2 * (off + a);
                for (size_t j = a; j < b; ++j) { *d++ = src[j]; *d++ = 0.0; }
            }
        });
    };
    const char *trace = getenv("PLSB_TRACE");
    const bool tr = trace && *trace == '1';
    const auto t0 = std::chrono::steady_clock::now();
    double wait_s = 0, move_s = 0;
    auto now_s = [&] { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); };
    for (size_t i = 0; i < nch; ++i) {
        const size_t off = i * per, cnt = n - off < per ? n - off : per;
        PLSB_CUDA(cudaMemcpyAsync(h.dbuf[i & 1], dev + off, cnt * 8, cudaMemcpyDeviceToHost, c->copy_stream));
        PLSB_CUDA(cudaEventRecord(h.ddone[i & 1], c->copy_stream));
        if (i > 0) {
            PLSB_CUDA(cudaEventSynchronize(h.ddone[(i - 1) & 1]));
            move_out(i - 1);
        }
    }
    const double ta = now_s();
    PLSB_CUDA(cudaEventSynchronize(h.ddone[(nch - 1) & 1]));
    const double tb = now_s();
    move_out(nch - 1);
    wait_s = tb - ta; move_s = now_s() - tb;
    if (tr) fprintf(stderr, "[plsb] staged download of %.1f MB: %.3f ms in all; last chunk: %.3f ms waiting for the copy, %.3f ms moving it out\n",
                    n * 8e-6, now_s() * 1e3, wait_s * 1e3, move_s * 1e3);
    return PLSB_OK;
}

static int hstage_ensure(plsb_ctx *c, size_t bytes) {
    plsb_ctx::HostStage &h = c->hstage;
    hstage_pool(c);
    if (bytes <= h.bytes) return PLSB_OK;
    for (int i = 0; i < 3; ++i) {
        if (h.done[i]) cudaEventSynchronize(h.done[i]);
        if (h.buf[i]) cudaFreeHost(h.buf[i]);
        h.buf[i] = nullptr;
    }
    h.bytes = 0;
    bytes = (size_t)round_up((int64_t)bytes, 1 << 20);
    const char *wc = getenv("PLSB_STAGE_WC");
    const unsigned flags = (wc && *wc == '0') ? cudaHostAllocDefault : cudaHostAllocWriteCombined;
    for (int i = 0; i < 3; ++i) {
        PLSB_CUDA(cudaHostAlloc((void **)&h.buf[i], bytes, flags));
        if (!h.done[i]) PLSB_CUDA(cudaEventCreateWithFlags(&h.done[i], cudaEventDisableTiming));
    }
    h.bytes = bytes;
    return PLSB_OK;
}

void hstage_release(plsb_ctx *c) {
    plsb_ctx::HostStage &h = c->hstage;
    for (int i = 0; i < 3; ++i) {
        if (h.done[i]) cudaEventDestroy(h.done[i]);
        if (h.buf[i]) cudaFreeHost(h.buf[i]);
    }
    for (int i = 0; i < 2; ++i) {
        if (h.ddone[i]) cudaEventDestroy(h.ddone[i]);
        if (h.dbuf[i]) cudaFreeHost(h.dbuf[i]);
    }
    delete h.pool;
    h = plsb_ctx::HostStage();
}

// dst (packed, segment j at j * seg) <- nseg segments of `seg` bytes, segment j at src + j * stride; the bytes are
// dealt to the pool's threads in contiguous, cache-line-aligned ranges of the destination
static void pack_segments(HostPool &pool, char *dst, const char *src, size_t seg, size_t stride, size_t nseg) {
    const size_t total = seg * nseg;
    pool.run([&](int t, int n) {
        size_t a = total / (size_t)n * (size_t)t, b = (t + 1 == n) ? total : total / (size_t)n * (size_t)(t + 1);
        a &= ~(size_t)63;
        if (t + 1 < n) b &= ~(size_t)63;
        while (a < b) {
            const size_t j = a / seg, off = a % seg;
            const size_t len = seg - off < b - a ? seg - off : b - a;
            memcpy(dst + a, src + j * stride + off, len);
            a += len;
        }
    });
}

static bool stage_wanted(const void *src, size_t total_bytes) {
    const char *e = getenv("PLSB_NO_STAGE");
    if (e && *e == '1') return false;
    return total_bytes >= ((size_t)8 << 20) && host_is_pageable(src);
}

// One block of a host column-major matrix (nseg column segments of nr rows, packed column-major with ld = nr) to
// the device buffer `buf` on the copy stream: straight from a pinned source, or packed by the host threads first.
static cudaError_t send_block(plsb_ctx *c, bool staged, int64_t i, double *buf, const double *src, int64_t lds, int64_t nr,
                              int64_t nseg) {
    if (!staged)
        return cudaMemcpy2DAsync(buf, (size_t)nr * 8, src, (size_t)lds * 8, (size_t)nr * 8, (size_t)nseg,
                                 cudaMemcpyHostToDevice, c->copy_stream);
    plsb_ctx::HostStage &h = c->hstage;
    const int hb = (int)(i % 3);
    const auto t0 = std::chrono::steady_clock::now();
    cudaError_t e = cudaEventSynchronize(h.done[hb]);   // the copy that read this buffer three blocks ago
    if (e != cudaSuccess) return e;
    const auto t1 = std::chrono::steady_clock::now();
    pack_segments(*h.pool, h.buf[hb], (const char *)src, (size_t)nr * 8, (size_t)lds * 8, (size_t)nseg);
    h.wait_s += std::chrono::duration<double>(t1 - t0).count();
    h.pack_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count();
    h.pack_bytes += (double)nr * 8.0 * (double)nseg;
    e = cudaMemcpyAsync(buf, h.buf[hb], (size_t)nr * 8 * (size_t)nseg, cudaMemcpyHostToDevice, c->copy_stream);
    if (e == cudaSuccess) e = cudaEventRecord(h.done[hb], c->copy_stream);
    return e;
}

int upload_cm_to_rm(plsb_ctx *c, const double *src, int64_t lds, int64_t rows, int64_t cols, double *dst, int64_t ldd,
                    bool src_on_device) {
    if (rows == 0 || cols == 0) return PLSB_OK;
    if (src_on_device) return launch_cm_to_rm(c, src, lds, rows, cols, dst, ldd, 0);
    // Column blocks of <= 128 MB go through TWO staging buffers: the copy engine fills one (copy_stream) while the
    // transpose kernel empties the other (stream), so PCIe never waits for a kernel.  The transpose runs at HBM
    // speed; the copy engine is the only thing that matters here.
    int64_t kc = (int64_t)(128u << 20) / (rows * 8);
    if (kc < 1) kc = 1;
    if (kc > cols) kc = cols;
    const bool staged = stage_wanted(src, (size_t)rows * cols * 8);
    if (staged) PLSB_TRY(hstage_ensure(c, (size_t)rows * kc * 8));
    double *stage = nullptr;
    PLSB_TRY(dev_alloc(c, (void **)&stage, (size_t)2 * rows * kc * sizeof(double)));
    cudaEvent_t *copied = c->xfer_ev, *moved = c->xfer_ev + 2;
    int rc = PLSB_OK;
    cudaError_t e = cudaEventRecord(moved[0], c->stream);     // orders the first copies after the allocation
    if (e == cudaSuccess) e = cudaEventRecord(moved[1], c->stream);
    int64_t i = 0;
    for (int64_t c0 = 0; c0 < cols && rc == PLSB_OK && e == cudaSuccess; c0 += kc, ++i) {
        const int64_t nc = cols - c0 < kc ? cols - c0 : kc;
        const int b = (int)(i & 1);
        double *buf = stage + (size_t)b * rows * kc;
        e = cudaStreamWaitEvent(c->copy_stream, moved[b], 0);  // block i-2 has left this buffer
        if (e == cudaSuccess) e = send_block(c, staged, i, buf, src + c0 * lds, lds, rows, nc);
        if (e == cudaSuccess) e = cudaEventRecord(copied[b], c->copy_stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(c->stream, copied[b], 0);
        if (e != cudaSuccess) break;
        rc = launch_cm_to_rm(c, buf, rows, rows, nc, dst, ldd, c0);
        if (rc == PLSB_OK) e = cudaEventRecord(moved[b], c->stream);
    }
    if (e != cudaSuccess && rc == PLSB_OK) rc = cuda_fail(e, "upload (H2D)", __FILE__, __LINE__);
    e = cudaStreamSynchronize(c->copy_stream);
    if (rc == PLSB_OK && e != cudaSuccess) rc = cuda_fail(e, "upload sync", __FILE__, __LINE__);
    e = cudaStreamSynchronize(c->stream);
    if (rc == PLSB_OK && e != cudaSuccess) rc = cuda_fail(e, "upload sync", __FILE__, __LINE__);
    dev_free(c, stage);
    return rc;
}

// Host column-major X -> device row-major X in ROW blocks (strided 2-D H2D copies on the copy stream, two staging
// buffers), and while block i+1 crosses PCIe the tensor cores add block i's contribution to XX = X'X (SYRK, DMMA).
// The upload is PCIe-bound (~55 GB/s) and X'X needs N*K*(K+128) flops at ~36 TFLOP/s: for K <= ~3000 the whole
// cross-product build hides under the transfer, and a fit that follows needs no further pass over X except the
// one that forms the scores (fit.cu).  SURVEY 8(f)-3: chunked H2D overlapped with the cross-product pass.
static bool gram_on_upload(int64_t N, int64_t K) {
    const char *e = getenv("PLSB_UPLOAD_GRAM");
    if (e && *e == '0') return false;
    if (e && *e == '1') return N > 0;
    return K >= 256 && K <= 3072 && N >= 16 * K && N >= 65536;
}

static int upload_rows_with_gram(plsb_ctx *c, const double *src, int64_t lds, plsb_dataset *ds) {
    const int64_t N = ds->N, K = ds->K, ldk = ds->ldk;
    ds->ldg = round_up(K, 2);
    PLSB_TRY(dev_alloc(c, (void **)&ds->G, (size_t)K * ds->ldg * sizeof(double)));
    PLSB_CUDA(cudaMemsetAsync(ds->G, 0, (size_t)K * ds->ldg * sizeof(double), c->stream));
    PLSB_TRY(dev_alloc(c, (void **)&ds->S0, (size_t)ds->M * ldk * sizeof(double)));
    PLSB_CUDA(cudaMemsetAsync(ds->S0, 0, (size_t)ds->M * ldk * sizeof(double), c->stream));
    int64_t nb = (int64_t)(128u << 20) / (K * 8) / 1024 * 1024;
    if (nb < 1024) nb = 1024;
    if (nb > N) nb = N;
    const bool staged = stage_wanted(src, (size_t)N * K * 8);
    if (staged) PLSB_TRY(hstage_ensure(c, (size_t)nb * K * 8));
    double *stage = nullptr;
    PLSB_TRY(dev_alloc(c, (void **)&stage, (size_t)2 * nb * K * sizeof(double)));
    cudaEvent_t *copied = c->xfer_ev, *moved = c->xfer_ev + 2;
    int rc = PLSB_OK;
    cudaError_t e = cudaEventRecord(moved[0], c->stream);
    if (e == cudaSuccess) e = cudaEventRecord(moved[1], c->stream);
    int64_t i = 0;
    for (int64_t r0 = 0; r0 < N && rc == PLSB_OK && e == cudaSuccess; r0 += nb, ++i) {
        const int64_t nr = N - r0 < nb ? N - r0 : nb;
        const int b = (int)(i & 1);
        double *buf = stage + (size_t)b * nb * K;
        e = cudaStreamWaitEvent(c->copy_stream, moved[b], 0);
        if (e == cudaSuccess)   // K column segments of nr rows each, packed column-major (ld = nr)
            e = send_block(c, staged, i, buf, src + r0, lds, nr, K);
        if (e == cudaSuccess) e = cudaEventRecord(copied[b], c->copy_stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(c->stream, copied[b], 0);
        if (e != cudaSuccess) break;
        rc = launch_cm_to_rm(c, buf, nr, nr, K, ds->X + r0 * ldk, ldk, 0);
        if (rc == PLSB_OK) e = cudaEventRecord(moved[b], c->stream);
        if (rc == PLSB_OK && e == cudaSuccess)
            rc = launch_gemm_atb(c, ds->X + r0 * ldk, ldk, ds->X + r0 * ldk, ldk, nr, (int)K, (int)K, ds->G, ds->ldg, true,
                                 false, false, PLSB_PROF_XTX, true);
        for (int64_t m0 = 0; m0 < ds->M && rc == PLSB_OK; m0 += 128) {   // S0 += X_b' Y_b (Y is already on the device)
            const int mc = (int)(ds->M - m0 < 128 ? ds->M - m0 : 128);
            rc = launch_gemm_atb(c, ds->X + r0 * ldk, ldk, ds->Y + r0 * ds->ldm + m0, ds->ldm, nr, (int)K, mc,
                                 ds->S0 + m0 * ldk, ldk, false, false, true, PLSB_PROF_PASS_XTY, true);
        }
    }
    if (e != cudaSuccess && rc == PLSB_OK) rc = cuda_fail(e, "upload (H2D, row blocks)", __FILE__, __LINE__);
    e = cudaStreamSynchronize(c->copy_stream);
    if (rc == PLSB_OK && e != cudaSuccess) rc = cuda_fail(e, "upload sync", __FILE__, __LINE__);
    e = cudaStreamSynchronize(c->stream);
    if (rc == PLSB_OK && e != cudaSuccess) rc = cuda_fail(e, "upload sync", __FILE__, __LINE__);
    dev_free(c, stage);
    ds->gram_valid = rc == PLSB_OK;
    ds->gram_reduced = false;
    return rc;
}

static int dataset_alloc(plsb_ctx *c, int64_t N, int64_t K, int64_t M, plsb_dataset **out) {
    PLSB_ARG(c && out, "null argument");
    PLSB_ARG(!c->closed, "the context has been destroyed");
    PLSB_ARG(N >= 0 && K >= 1 && M >= 1, "bad shape N=%lld K=%lld M=%lld", (long long)N, (long long)K, (long long)M);
    PLSB_CUDA(cudaSetDevice(c->device));
    plsb_dataset *ds = new plsb_dataset();
    ctx_retain(c);
    ds->ctx = c; ds->N = N; ds->K = K; ds->M = M;
    ds->ldk = round_up(K, 2);
    ds->ldm = round_up(M, 2);
    const size_t nx = (size_t)(N > 0 ? N : 1) * ds->ldk, ny = (size_t)(N > 0 ? N : 1) * ds->ldm;
    cudaError_t e = cudaMallocAsync(&ds->X, nx * sizeof(double), c->stream);
    if (e == cudaSuccess) e = cudaMallocAsync(&ds->Y, ny * sizeof(double), c->stream);
    if (e == cudaSuccess && ds->ldk != K) e = cudaMemsetAsync(ds->X, 0, nx * sizeof(double), c->stream);
    if (e == cudaSuccess && ds->ldm != M) e = cudaMemsetAsync(ds->Y, 0, ny * sizeof(double), c->stream);
    if (e != cudaSuccess) {
        dev_free(c, ds->X);
        dev_free(c, ds->Y);
        delete ds;
        const int rc = cuda_fail(e, "dataset alloc", __FILE__, __LINE__);
        ctx_release(c);
        return rc;
    }
    *out = ds;
    return PLSB_OK;
}

static int dataset_fill(plsb_ctx *c, const double *X, int64_t ldx, int64_t N, int64_t K, const double *Y, int64_t ldy,
                        int64_t M, bool on_device, plsb_dataset_t *out) {
    PLSB_ARG(N == 0 || (X && Y), "X / Y is null");
    PLSB_ARG(ldx >= N && ldy >= N, "leading dimension smaller than N");
    plsb_dataset *ds = nullptr;
    PLSB_TRY(dataset_alloc(c, N, K, M, &ds));
    int rc;
    {
        ProfScope ps(c, PLSB_PROF_UPLOAD);
        rc = upload_cm_to_rm(c, Y, ldy, N, M, ds->Y, ds->ldm, on_device);
        if (rc == PLSB_OK) {
            if (!on_device && gram_on_upload(N, K)) rc = upload_rows_with_gram(c, X, ldx, ds);
            else rc = upload_cm_to_rm(c, X, ldx, N, K, ds->X, ds->ldk, on_device);
        }
    }
    if (rc != PLSB_OK) { plsb_dataset_destroy(ds); return rc; }
    if (const char *tr = getenv("PLSB_TRACE")) {
        plsb_ctx::HostStage &h = c->hstage;
        if (*tr == '1' && h.pack_bytes > 0)
            fprintf(stderr, "[plsb] staged upload: %.2f GB packed by %d host threads in %.3f s (%.1f GB/s), %.3f s waiting for "
                            "the copy engine\n", h.pack_bytes / 1e9, h.pool ? h.pool->size() : 0, h.pack_s,
                    h.pack_bytes / 1e9 / (h.pack_s > 0 ? h.pack_s : 1), h.wait_s);
        h.pack_s = h.wait_s = h.pack_bytes = 0.0;
    }
    *out = ds;
    return PLSB_OK;
}

}  // namespace plsb

using namespace plsb;

extern "C" {

int plsb_dataset_create(plsb_ctx_t c, const double *X, int64_t ldx, int64_t N, int64_t K, const double *Y,
                        int64_t ldy, int64_t M, plsb_dataset_t *ds) {
    return dataset_fill(c, X, ldx, N, K, Y, ldy, M, false, ds);
}
int plsb_dataset_create_device(plsb_ctx_t c, const double *dX, int64_t ldx, int64_t N, int64_t K, const double *dY,
                               int64_t ldy, int64_t M, plsb_dataset_t *ds) {
    return dataset_fill(c, dX, ldx, N, K, dY, ldy, M, true, ds);
}
int plsb_dataset_create_synth(plsb_ctx_t c, uint64_t seed, int64_t row0, int64_t N, int64_t K, int64_t M, int64_t L,
                              plsb_dataset_t *out) {
    plsb_dataset *ds = nullptr;
    PLSB_TRY(dataset_alloc(c, N, K, M, &ds));
    int rc = launch_synth(c, seed, row0, N, K, M, L, ds->X, ds->ldk, ds->Y, ds->ldm);
    if (rc == PLSB_OK) {
        cudaError_t e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "synth", __FILE__, __LINE__);
    }
    if (rc != PLSB_OK) { plsb_dataset_destroy(ds); return rc; }
    *out = ds;
    return PLSB_OK;
}
int plsb_dataset_destroy(plsb_dataset_t ds) {
    if (!ds) return PLSB_OK;
    cudaSetDevice(ds->ctx->device);
    cudaStreamSynchronize(ds->ctx->stream);
    dev_free(ds->ctx, ds->X);
    dev_free(ds->ctx, ds->Y);
    dev_free(ds->ctx, ds->G);
    dev_free(ds->ctx, ds->S0);
    plsb_ctx *c = ds->ctx;
    delete ds;
    ctx_release(c);
    return PLSB_OK;
}
int plsb_dataset_concat(plsb_ctx_t c, int64_t nparts, const plsb_dataset_t *parts, plsb_dataset_t *out) {
    PLSB_ARG(c && parts && out && nparts >= 1, "null argument");
    int64_t N = 0;
    for (int64_t i = 0; i < nparts; ++i) {
        PLSB_ARG(parts[i], "part %lld is null", (long long)i);
        PLSB_ARG(parts[i]->K == parts[0]->K && parts[i]->M == parts[0]->M, "parts disagree on K or M");
        N += parts[i]->N;
    }
    plsb_dataset *ds = nullptr;
    PLSB_TRY(dataset_alloc(c, N, parts[0]->K, parts[0]->M, &ds));
    // rows are contiguous in the row-major layout: one copy per part and matrix (device to device; peer copies over
    // NVLink when the part lives on another GPU).  Padding columns are copied with the rows.
    int rc = PLSB_OK;
    int64_t r0 = 0;
    for (int64_t i = 0; i < nparts && rc == PLSB_OK; ++i) {
        const plsb_dataset *p = parts[i];
        cudaError_t e = cudaSetDevice(p->ctx->device);
        if (e == cudaSuccess) e = cudaStreamSynchronize(p->ctx->stream);     // the part's own work has finished
        if (e == cudaSuccess) e = cudaSetDevice(c->device);
        if (e == cudaSuccess && p->N > 0)
            e = cudaMemcpyPeerAsync(ds->X + r0 * ds->ldk, c->device, p->X, p->ctx->device, (size_t)p->N * ds->ldk * 8, c->stream);
        if (e == cudaSuccess && p->N > 0)
            e = cudaMemcpyPeerAsync(ds->Y + r0 * ds->ldm, c->device, p->Y, p->ctx->device, (size_t)p->N * ds->ldm * 8, c->stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "dataset concat", __FILE__, __LINE__);
        r0 += p->N;
    }
    if (rc == PLSB_OK) {
        cudaError_t e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "dataset concat", __FILE__, __LINE__);
    }
    if (rc != PLSB_OK) { plsb_dataset_destroy(ds); return rc; }
    *out = ds;
    return PLSB_OK;
}
int plsb_dataset_shape(plsb_dataset_t ds, int64_t *N, int64_t *K, int64_t *M) {
    PLSB_ARG(ds, "dataset is null");
    if (N) *N = ds->N;
    if (K) *K = ds->K;
    if (M) *M = ds->M;
    return PLSB_OK;
}
int plsb_dataset_set_sharded(plsb_dataset_t ds, int sharded) {
    PLSB_ARG(ds, "dataset is null");
    ds->sharded = sharded != 0;
    ds->gram_agreed = -1;
    return PLSB_OK;
}

static int download_rm_to_cm(plsb_ctx *c, const double *src, int64_t lds, int64_t rows, int64_t cols, double *dst,
                             int64_t ldd) {
    if (rows == 0 || cols == 0) return PLSB_OK;
    int64_t kc = (int64_t)(256u << 20) / (rows * 8);
    if (kc < 1) kc = 1;
    if (kc > cols) kc = cols;
    double *stage = nullptr;
    PLSB_TRY(dev_alloc(c, (void **)&stage, (size_t)rows * kc * sizeof(double)));
    int rc = PLSB_OK;
    for (int64_t c0 = 0; c0 < cols && rc == PLSB_OK; c0 += kc) {
        const int64_t nc = cols - c0 < kc ? cols - c0 : kc;
        rc = launch_rm_to_cm(c, src + c0, lds, rows, nc, stage, rows);
        if (rc != PLSB_OK) break;
        cudaError_t e = cudaMemcpy2DAsync(dst + c0 * ldd, (size_t)ldd * 8, stage, (size_t)rows * 8, (size_t)rows * 8,
                                          (size_t)nc, cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "cudaMemcpy2DAsync(D2H)", __FILE__, __LINE__);
    }
    dev_free(c, stage);
    return rc;
}

int plsb_dataset_download(plsb_dataset_t ds, double *X, int64_t ldx, double *Y, int64_t ldy) {
    PLSB_ARG(ds, "dataset is null");
    plsb_ctx *c = ds->ctx;
    PLSB_CUDA(cudaSetDevice(c->device));
    if (X) { PLSB_ARG(ldx >= ds->N, "ldx < N"); PLSB_TRY(download_rm_to_cm(c, ds->X, ds->ldk, ds->N, ds->K, X, ldx)); }
    if (Y) { PLSB_ARG(ldy >= ds->N, "ldy < N"); PLSB_TRY(download_rm_to_cm(c, ds->Y, ds->ldm, ds->N, ds->M, Y, ldy)); }
    return PLSB_OK;
}

}  // extern "C"
