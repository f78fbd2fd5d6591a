// pls_b200/csrc/api.cu — C-ABI entry points of libplsb.so (include/plsb.h): context, data sets,
// the fit driver (reference src/pls.cpp:390-437) and the measurement hooks.
#include <dlfcn.h>
#include <stdarg.h>
#include <string.h>

#include <mutex>

#include "internal.h"

// ---- minimal NCCL surface, bound lazily with dlopen so the library loads without NCCL ------------
typedef struct { char internal[128]; } plsb_nccl_uid;
namespace {
struct NcclApi {
    void *h = nullptr;
    int (*GetUniqueId)(plsb_nccl_uid *) = nullptr;
    int (*CommInitRank)(void **, int, plsb_nccl_uid, int) = nullptr;
    int (*CommDestroy)(void *) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
} g_nccl;
const int kNcclDouble = 8, kNcclSum = 0, kNcclInt8 = 0;  // ncclFloat64, ncclSum (nccl.h enum values, stable since 2.0)

int nccl_load() {
    if (g_nccl.h) return PLSB_OK;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *n : names) {
        g_nccl.h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl.h) break;
    }
    if (!g_nccl.h) { plsb::set_error("NCCL not found: %s", dlerror()); return PLSB_ERR_NCCL; }
    g_nccl.GetUniqueId = (int (*)(plsb_nccl_uid *))dlsym(g_nccl.h, "ncclGetUniqueId");
    g_nccl.CommInitRank = (int (*)(void **, int, plsb_nccl_uid, int))dlsym(g_nccl.h, "ncclCommInitRank");
    g_nccl.CommDestroy = (int (*)(void *))dlsym(g_nccl.h, "ncclCommDestroy");
    g_nccl.AllReduce = (int (*)(const void *, void *, size_t, int, int, void *, cudaStream_t))dlsym(g_nccl.h, "ncclAllReduce");
    g_nccl.AllGather = (int (*)(const void *, void *, size_t, int, void *, cudaStream_t))dlsym(g_nccl.h, "ncclAllGather");
    g_nccl.GetErrorString = (const char *(*)(int))dlsym(g_nccl.h, "ncclGetErrorString");
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce) {
        plsb::set_error("NCCL symbols missing");
        return PLSB_ERR_NCCL;
    }
    return PLSB_OK;
}
int nccl_fail(int rc, const char *what) {
    plsb::set_error("NCCL error %d (%s) in %s", rc, g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?", what);
    return PLSB_ERR_NCCL;
}
thread_local char g_err[1024] = "";

// host-to-host all-gather of `bytes` per rank through the communicator (set-up traffic only)
int host_allgather(plsb_ctx *c, const void *mine, void *all, size_t bytes) {
    if (!g_nccl.AllGather) { plsb::set_error("ncclAllGather missing"); return PLSB_ERR_NCCL; }
    char *d = nullptr;
    PLSB_CUDA(cudaMalloc(&d, bytes * (size_t)(c->nranks + 1)));
    int rc = PLSB_OK;
    cudaError_t e = cudaMemcpyAsync(d, mine, bytes, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) {
        int n = g_nccl.AllGather(d, d + bytes, bytes, kNcclInt8, c->comm, c->stream);
        if (n) rc = nccl_fail(n, "ncclAllGather");
    }
    if (rc == PLSB_OK && e == cudaSuccess) e = cudaMemcpyAsync(all, d + bytes, bytes * (size_t)c->nranks, cudaMemcpyDeviceToHost, c->stream);
    if (rc == PLSB_OK && e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d);
    if (rc == PLSB_OK && e != cudaSuccess) rc = plsb::cuda_fail(e, "host_allgather", __FILE__, __LINE__);
    return rc;
}
}  // namespace

namespace plsb {

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
int cuda_fail(cudaError_t e, const char *what, const char *file, int line) {
    set_error("CUDA error %d (%s) at %s:%d in %s", (int)e, cudaGetErrorString(e), file, line, what);
    return PLSB_ERR_CUDA;
}

int dev_alloc(plsb_ctx *c, void **p, size_t bytes) {
    PLSB_CUDA(cudaMallocAsync(p, bytes ? bytes : 8, c->stream));
    return PLSB_OK;
}
void dev_free(plsb_ctx *c, void *p) {
    if (p) cudaFreeAsync(p, c->stream);
}
int dev_mem_available(plsb_ctx *c, size_t *bytes) {
    size_t free_b = 0, total_b = 0;
    PLSB_CUDA(cudaMemGetInfo(&free_b, &total_b));
    cudaMemPool_t pool;
    uint64_t reserved = 0, used = 0;
    if (cudaDeviceGetDefaultMemPool(&pool, c->device) == cudaSuccess &&
        cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &reserved) == cudaSuccess &&
        cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &used) == cudaSuccess && reserved > used)
        free_b += (size_t)(reserved - used);
    *bytes = free_b;
    return PLSB_OK;
}

bool pdl_enabled() {
    static const bool on = [] { const char *e = getenv("PLSB_NO_PDL"); return !(e && *e == '1'); }();
    return on;
}

int ensure_part(plsb_ctx *c, size_t bytes) {
    if (bytes <= c->part_bytes) return PLSB_OK;
    if (c->part) {
        dev_free(c, c->part);  // stream-ordered: kernels already queued on the stream keep their scratch
        c->part = nullptr;
        c->part_bytes = 0;
    }
    bytes = (size_t)round_up((int64_t)bytes, 1 << 20);
    PLSB_TRY(dev_alloc(c, (void **)&c->part, bytes));
    c->part_bytes = bytes;
    return PLSB_OK;
}

ProfScope::ProfScope(plsb_ctx *ctx, int slot, cudaStream_t stream) : c(ctx), st(stream ? stream : ctx->stream) {
    if (slot < 0 || !c->prof || (c->prof == 2 && slot != PLSB_PROF_FIT_TOTAL)) return;
    plsb_ctx::Pending p;
    p.slot = slot;
    for (cudaEvent_t *e : {&p.a, &p.b}) {
        if (!c->pool.empty()) { *e = c->pool.back(); c->pool.pop_back(); }
        else if (cudaEventCreate(e) != cudaSuccess) return;
    }
    cudaEventRecord(p.a, st);
    c->pending.push_back(p);
    idx = (int)c->pending.size() - 1;
}
ProfScope::~ProfScope() {
    if (idx >= 0) cudaEventRecord(c->pending[idx].b, st);
}

static int prof_drain(plsb_ctx *c) {
    if (c->pending.empty()) return PLSB_OK;
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    for (auto &p : c->pending) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
            c->prof_ms[p.slot] += ms;
            c->prof_n[p.slot] += 1;
        }
        c->pool.push_back(p.a);
        c->pool.push_back(p.b);
    }
    c->pending.clear();
    return PLSB_OK;
}

static std::atomic<int> g_live_ctx{0};
int live_contexts() { return g_live_ctx.load(std::memory_order_relaxed); }

void ctx_retain(plsb_ctx *c) { c->refs.fetch_add(1, std::memory_order_relaxed); }

void ctx_release(plsb_ctx *c) {
    if (c->refs.fetch_sub(1, std::memory_order_acq_rel) != 1) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    xchg_teardown(c);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    for (auto &p : c->pending) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    for (auto e : c->pool) cudaEventDestroy(e);
    if (c->part) cudaFreeAsync(c->part, c->stream);
    cudaStreamSynchronize(c->stream);
    cudaStreamDestroy(c->stream);
    if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
    for (int i = 0; i < 4; ++i) if (c->xfer_ev[i]) cudaEventDestroy(c->xfer_ev[i]);
    if (c->col_ev) cudaEventDestroy(c->col_ev);
    hstage_release(c);
    g_live_ctx.fetch_sub(1, std::memory_order_relaxed);
    delete c;
}

}  // namespace plsb

using namespace plsb;

extern "C" {

const char *plsb_last_error(void) { return g_err; }
const char *plsb_version(void) { return "pls_b200 0.1 (sm_100a)"; }

int plsb_device_count(int *count) {
    PLSB_ARG(count, "count is null");
    PLSB_CUDA(cudaGetDeviceCount(count));
    return PLSB_OK;
}

int plsb_ctx_create(int device, plsb_ctx_t *out) {
    PLSB_ARG(out, "ctx out pointer is null");
    int n = 0;
    PLSB_CUDA(cudaGetDeviceCount(&n));
    PLSB_ARG(device >= 0 && device < n, "device %d out of range (%d visible)", device, n);
    PLSB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    PLSB_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        set_error("device %d is sm_%d%d; libplsb.so carries sm_100a code only (no fallback)", device, prop.major, prop.minor);
        return PLSB_ERR_UNSUPPORTED;
    }
    plsb_ctx *c = new plsb_ctx();
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    c->smem_optin = prop.sharedMemPerBlockOptin;
    cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
    for (int i = 0; i < 4 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&c->xfer_ev[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->col_ev, cudaEventDisableTiming);
    if (e != cudaSuccess) { delete c; return cuda_fail(e, "cudaStreamCreate", __FILE__, __LINE__); }
    {   // keep freed blocks in the pool (see dev_alloc)
        cudaMemPool_t pool;
        uint64_t keep = UINT64_MAX;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess)
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    g_live_ctx.fetch_add(1, std::memory_order_relaxed);
    *out = c;
    return PLSB_OK;
}

int plsb_ctx_destroy(plsb_ctx_t c) {
    if (!c || c->closed) return PLSB_OK;
    c->closed = true;
    ctx_release(c);
    return PLSB_OK;
}

int plsb_ctx_trim(plsb_ctx_t c) {
    PLSB_ARG(c, "ctx is null");
    PLSB_CUDA(cudaSetDevice(c->device));
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    PLSB_CUDA(cudaStreamSynchronize(c->copy_stream));
    hstage_release(c);   // pinned staging buffers + host thread pool of large transfers: rebuilt on the next one
    cudaMemPool_t pool;
    PLSB_CUDA(cudaDeviceGetDefaultMemPool(&pool, c->device));
    PLSB_CUDA(cudaMemPoolTrimTo(pool, 0));
    return PLSB_OK;
}

int plsb_ctx_sync(plsb_ctx_t c) {
    PLSB_ARG(c, "ctx is null");
    PLSB_CUDA(cudaSetDevice(c->device));
    PLSB_CUDA(cudaStreamSynchronize(c->stream));
    return PLSB_OK;
}

int plsb_comm_unique_id(void *id128) {
    PLSB_ARG(id128, "id buffer is null");
    PLSB_TRY(nccl_load());
    plsb_nccl_uid id;
    int rc = g_nccl.GetUniqueId(&id);
    if (rc) return nccl_fail(rc, "ncclGetUniqueId");
    memcpy(id128, &id, sizeof(id));
    return PLSB_OK;
}

int plsb_comm_init(plsb_ctx_t c, int rank, int nranks, const void *id128) {
    PLSB_ARG(c && id128, "ctx / id is null");
    PLSB_ARG(nranks >= 1 && rank >= 0 && rank < nranks, "bad rank %d of %d", rank, nranks);
    PLSB_TRY(nccl_load());
    PLSB_CUDA(cudaSetDevice(c->device));
    plsb_nccl_uid id;
    memcpy(&id, id128, sizeof(id));
    int rc = g_nccl.CommInitRank(&c->comm, nranks, id, rank);
    if (rc) return nccl_fail(rc, "ncclCommInitRank");
    c->rank = rank;
    c->nranks = nranks;
    // peer-memory exchange area for the in-kernel all-reduce (falls back to NCCL per call if it cannot be set up)
    PLSB_TRY(xchg_setup(c, host_allgather));
    return PLSB_OK;
}

int plsb_comm_info(plsb_ctx_t c, int *rank, int *nranks, int *peer_exchange) {
    PLSB_ARG(c, "ctx is null");
    if (rank) *rank = c->rank;
    if (nranks) *nranks = c->nranks;
    if (peer_exchange) *peer_exchange = (c->nranks > 1 && c->xchg.ready && !c->xchg.failed) ? 1 : 0;
    return PLSB_OK;
}

int plsb_profile_enable(plsb_ctx_t c, int on) {
    PLSB_ARG(c, "ctx is null");
    PLSB_TRY(prof_drain(c));
    c->prof = (on == 2 || on == 3) ? on : (on != 0 ? 1 : 0);
    return PLSB_OK;
}
int plsb_profile_reset(plsb_ctx_t c) {
    PLSB_ARG(c, "ctx is null");
    PLSB_TRY(prof_drain(c));
    for (int i = 0; i < PLSB_PROF_NSLOTS; ++i) { c->prof_ms[i] = 0; c->prof_n[i] = 0; }
    return PLSB_OK;
}
int plsb_profile_get(plsb_ctx_t c, int slot, double *ms_total, int64_t *launches) {
    PLSB_ARG(c && slot >= 0 && slot < PLSB_PROF_NSLOTS, "bad profile slot");
    PLSB_TRY(prof_drain(c));
    if (ms_total) *ms_total = c->prof_ms[slot];
    if (launches) *launches = c->prof_n[slot];
    return PLSB_OK;
}
int plsb_launch_count(plsb_ctx_t c, int64_t *count) {
    PLSB_ARG(c && count, "null argument");
    *count = c->launches;
    return PLSB_OK;
}

}  // extern "C"

// allreduce helper used by the fit driver (sum of FP64 partials across the row shards)
namespace plsb {
int allreduce_sum(plsb_ctx *c, double *buf, size_t n) {
    if (c->nranks <= 1 || !c->comm) return PLSB_OK;
    int rc = g_nccl.AllReduce(buf, buf, n, kNcclDouble, kNcclSum, c->comm, c->stream);
    if (rc) return nccl_fail(rc, "ncclAllReduce");
    return PLSB_OK;
}
}  // namespace plsb

extern "C" int plsb_comm_allreduce_test(plsb_ctx_t c, double *host_buf, int64_t n) {
    PLSB_ARG(c && host_buf && n > 0, "bad argument");
    PLSB_CUDA(cudaSetDevice(c->device));
    double *d = nullptr;
    PLSB_TRY(plsb::dev_alloc(c, (void **)&d, sizeof(double) * n));
    PLSB_CUDA(cudaMemcpyAsync(d, host_buf, sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
    int rc = plsb::allreduce_sum(c, d, (size_t)n);
    if (rc == PLSB_OK) {
        cudaError_t e = cudaMemcpyAsync(host_buf, d, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "allreduce_test copy", __FILE__, __LINE__);
    }
    plsb::dev_free(c, d);
    return rc;
}
