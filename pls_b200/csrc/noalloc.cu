// pls_b200/csrc/noalloc.cu — plsb_colstats_static: column mean / SST / stdev without a single allocation.
//
// No counterpart in the reference (its SST / colwise_stdev, src/pls.cpp:69-87, run on Eigen temporaries).
//
// Who this is for.  plsb_colstats (stats.cu) takes a context, stages the matrix into a device buffer it allocates
// and keeps its per-CTA partial sums in a scratch buffer the context grows on demand.  Some programs that want
// the same numbers are not in a position to allocate device memory at the time they call us:
//   - a plug-in inside a host application that reserved (nearly) the whole GPU for its own arena at start-up
//     and hands out slices of it: a cudaMalloc from a guest library fails there, or succeeds and starves the host;
//   - a step of a latency-sensitive loop: cudaMalloc / cudaFree may synchronise the whole device and stall
//     every other stream the application is running;
//   - code under a "no driver allocations after initialisation" rule (memory budgets audited up front).
// Such callers own device buffers already (the matrix lives in one).  They pass those in, and the only memory
// this entry point needs beyond them — the partial column sums of the two-pass formulas — lives in a
// file-scope __device__ array below.  Nothing here calls cudaMalloc, cudaFree, a memory pool or any library
// routine that allocates; it takes no plsb_ctx_t (creating one allocates streams, events and pools).
//
// What that costs.  The array has a fixed size, PLSB_STATIC_SCRATCH_BYTES = (64 + 1) * 8192 * 8 = 4.06 MiB, and
// every GPU on which this library is loaded carries it from module load until the process exits, used or not.
// It bounds the width: K <= PLSB_STATIC_MAX_K = 8192 columns (N is unbounded); wider matrices are refused
// with PLSB_ERR_ARG, not truncated.
//
// Sharing.  There is ONE array per GPU, and every call on that GPU uses it, whatever stream it is given.  The
// contents left by the previous call are stale but harmless: each call writes every partial sum it later reads
// (row chunks with no rows write 0.0), so nothing has to be cleared in between and nothing is accumulated with
// atomics.  Two calls whose kernels can run at the same time on one GPU (different streams, different host
// threads) race on the array and return garbage: the caller serialises them (one stream, or an event).
//
// Set-up.  In host code the array's name is only a handle for the runtime, not a device address, so the first
// call on each device asks the runtime for the real address (cudaGetSymbolAddress, with that device current)
// and caches it; later calls on that device reuse it.  Make that first call per device during initialisation,
// before any stream capture is begun.
#include <mutex>

#include "internal.h"

#define PLSB_STATIC_MAX_K 8192
#define PLSB_STATIC_CHUNKS 64
#define PLSB_STATIC_SCRATCH_DOUBLES ((PLSB_STATIC_CHUNKS + 1) * PLSB_STATIC_MAX_K)

// [chunk][k] partial sums (row stride K of the call), then K doubles of column means when the caller does not
// want them back (offset PLSB_STATIC_CHUNKS * PLSB_STATIC_MAX_K)
static __device__ __align__(256) double g_static_scratch[PLSB_STATIC_SCRATCH_DOUBLES];

namespace {

// part[chunk * K + k] = sum over the chunk's rows of X[i + k * ldx]            (mu == nullptr)
//                     = sum over the chunk's rows of (X[i + k * ldx] - mu[k])^2 (otherwise)
// X is COLUMN-MAJOR on the device (a column is contiguous: the 256 threads of a CTA read 2 KB runs of it).  One
// work item = (chunk, column); items are dealt to CTAs grid-stride.  Inside an item every thread keeps one
// accumulator, warps combine by shuffles, thread 0 adds the eight warp sums in warp order: a fixed order, so
// two calls on the same data give the same bits.
__global__ void __launch_bounds__(256) colsum_cm_static_kernel(const double *__restrict__ X, int64_t ldx, int64_t N, int K,
                                                               const double *__restrict__ mu, double *__restrict__ part) {
    __shared__ double wsum[8];
    const int64_t rows_per = (N + PLSB_STATIC_CHUNKS - 1) / PLSB_STATIC_CHUNKS;
    const int64_t items = (int64_t)PLSB_STATIC_CHUNKS * K;
    for (int64_t it = blockIdx.x; it < items; it += gridDim.x) {
        const int chunk = (int)(it / K), k = (int)(it % K);
        const int64_t r0 = (int64_t)chunk * rows_per;
        const int64_t r1 = r0 + rows_per < N ? r0 + rows_per : N;
        const double *col = X + (size_t)k * ldx;
        const double m = mu ? mu[k] : 0.0;
        double acc = 0.0;
        for (int64_t i = r0 + threadIdx.x; i < r1; i += 256) {
            const double d = col[i] - m;
            acc = mu ? fma(d, d, acc) : acc + d;
        }
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double s = 0.0;
            for (int w = 0; w < 8; ++w) s += wsum[w];
            part[(size_t)chunk * K + k] = s;
        }
        __syncthreads();
    }
}

// out[k] = scale * (part[0][k] + part[1][k] + ... in chunk order)
__global__ void __launch_bounds__(256) fold_chunks_static_kernel(const double *__restrict__ part, int K, double scale,
                                                                 double *__restrict__ out) {
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k >= K) return;
    double s = 0.0;
    for (int c = 0; c < PLSB_STATIC_CHUNKS; ++c) s += part[(size_t)c * K + k];
    out[k] = s * scale;
}

// sst[k] = sum of the chunk partials (or 0 for fewer than two rows, pls.cpp:70-71); stdev[k] = sqrt(sst / (N-1))
__global__ void __launch_bounds__(256) finish_static_kernel(const double *__restrict__ part, int K, int64_t N,
                                                            double *__restrict__ sst, double *__restrict__ stdev) {
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k >= K) return;
    double s = 0.0;
    if (N >= 2)
        for (int c = 0; c < PLSB_STATIC_CHUNKS; ++c) s += part[(size_t)c * K + k];
    if (sst) sst[k] = s;
    if (stdev) stdev[k] = sqrt(s / ((double)N - 1.0));
}

// device address of g_static_scratch on each GPU, resolved once per device by the first call there
constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
double *g_scratch_addr[kMaxDevices] = {};
int g_sm_count[kMaxDevices] = {};

// `device` is current when this runs
int static_scratch_for(int device, double **addr, int *sms) {
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_scratch_addr[device]) {
        double *p = nullptr;
        PLSB_CUDA(cudaGetSymbolAddress((void **)&p, g_static_scratch));
        int n = 0;
        PLSB_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device));
        g_sm_count[device] = n > 0 ? n : 148;
        g_scratch_addr[device] = p;
    }
    *addr = g_scratch_addr[device];
    *sms = g_sm_count[device];
    return PLSB_OK;
}

struct DeviceGuard {   // put the caller's current device back on every exit path
    int prev = -1;
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace

extern "C" int plsb_colstats_static(int device, void *stream, const double *dX, int64_t ldx, int64_t N, int64_t K,
                                    double *d_mean, double *d_sst, double *d_stdev) {
    PLSB_ARG(device >= 0 && device < kMaxDevices, "plsb_colstats_static: device %d out of range", device);
    PLSB_ARG(N >= 0 && K >= 1 && (N == 0 || dX) && ldx >= N, "plsb_colstats_static: bad argument");
    PLSB_ARG(K <= PLSB_STATIC_MAX_K,
             "plsb_colstats_static: K = %lld exceeds the %d columns the built-in scratch array holds "
             "(use plsb_colstats, which allocates)", (long long)K, PLSB_STATIC_MAX_K);
    if (!d_mean && !d_sst && !d_stdev) return PLSB_OK;
    DeviceGuard guard;
    PLSB_CUDA(cudaGetDevice(&guard.prev));
    PLSB_CUDA(cudaSetDevice(device));
    double *scratch = nullptr;
    int sms = 0;
    PLSB_TRY(static_scratch_for(device, &scratch, &sms));
    cudaStream_t st = (cudaStream_t)stream;
    double *mu = d_mean ? d_mean : scratch + (size_t)PLSB_STATIC_CHUNKS * PLSB_STATIC_MAX_K;
    const int64_t items = (int64_t)PLSB_STATIC_CHUNKS * K;
    const unsigned grid = (unsigned)(items < (int64_t)sms * 8 ? items : (int64_t)sms * 8);
    const unsigned kgrid = (unsigned)((K + 255) / 256);

    colsum_cm_static_kernel<<<grid, 256, 0, st>>>(dX, ldx, N, (int)K, nullptr, scratch);
    fold_chunks_static_kernel<<<kgrid, 256, 0, st>>>(scratch, (int)K, N > 0 ? 1.0 / (double)N : 0.0, mu);
    if (d_sst || d_stdev) {
        if (N >= 2) colsum_cm_static_kernel<<<grid, 256, 0, st>>>(dX, ldx, N, (int)K, mu, scratch);
        finish_static_kernel<<<kgrid, 256, 0, st>>>(scratch, (int)K, N, d_sst, d_stdev);
    }
    PLSB_CUDA(cudaGetLastError());
    return PLSB_OK;
}
