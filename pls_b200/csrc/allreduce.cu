// pls_b200/csrc/allreduce.cu — the cross-CTA partial-row sum that ends every streaming pass, fused with the
// cross-GPU all-reduce of the row-sharded fit: ONE kernel sums this GPU's per-CTA partial rows of
// [g | tt] (reference src/pls.cpp:420-421: tt = t't, p = X't, here per row shard) and then adds the other GPUs'
// vectors read straight from their memory over NVLink (peer mappings; CUDA IPC between processes).
//
// Why not NCCL for this step: the message is K+1 doubles (16 KB at C4) once per component, on the serial path
// between two passes.  A separate sum kernel + ncclAllReduce costs two launches and NCCL's latency (~30-40 us);
// here the exchange is a flag write, a flag wait and (ranks-1) x 16 KB of peer loads inside the kernel that is
// running anyway.  Large one-off reductions (X'Y, X'X) stay on NCCL.
//
// Protocol (per exchange number seq = 1, 2, ...; every rank runs the same sequence) — a PUSH exchange, so that
// every wait polls local memory and every remote access is a fire-and-forget store:
//   1. every CTA sums its 32 columns over the partial rows and stores them into slot[seq & 1][my rank] of EVERY
//      rank's exchange area (remote stores over NVLink; own area included);
//   2. the last CTA to finish (device-scope counter) fences system-wide and writes flag[my rank] = seq into every
//      rank's area;
//   3. every CTA waits until its LOCAL flags of all ranks are >= seq (one polling thread per rank), then adds the
//      LOCAL slots in RANK order: all ranks get bit-identical sums, so the replicated K-sized step stays
//      replicated exactly.
// Two slot sets suffice: a rank can push exchange seq+2 only after every peer has pushed seq+1 to it, which a peer
// does after it has finished reading exchange seq.  Flags only grow, so there is no reset race.  A peer that never
// shows up trips a 10 s timeout and raises an error flag instead of hanging the GPU.
#include <unistd.h>

#include "internal.h"

namespace plsb {

constexpr size_t XCHG_CTR_OFF = 0;                                    // local arrival counter (u32)
constexpr size_t XCHG_ERR_OFF = 64;                                   // local error flag (int)
constexpr size_t XCHG_FLAG_OFF = 128;                                 // u64 flag per source rank
constexpr size_t XCHG_SLOT_OFF = XCHG_FLAG_OFF + PLSB_MAX_RANKS * 8;  // [2][ranks][PLSB_XCHG_LEN] doubles
constexpr size_t XCHG_BYTES = XCHG_SLOT_OFF + 2 * (size_t)PLSB_MAX_RANKS * PLSB_XCHG_LEN * sizeof(double);

struct XchgArgs {
    int rank, nranks;
    unsigned long long seq;
    char *area[PLSB_MAX_RANKS];   // every rank's exchange area (area[rank] is local)
    unsigned *counter;   // device-scope arrival counter (local)
    int *err;            // local error flag (timeout)
    unsigned long long timeout_ns;
};

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_sys(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ double ld_relaxed_sys(const double *p) {
    double v;
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// 32 columns x 8 row groups per CTA: row group y sums partial rows y, y+8, .. with eight independent loads in
// flight, then the eight group sums are added in group order.  (A one-thread-per-column loop over ~300 rows is a
// chain of L2 round trips: 54 us between two 2.4 ms passes; this shape takes a few us.)
template <bool ACROSS>
__global__ void __launch_bounds__(256) reduce_partials_kernel(const double *__restrict__ part, int64_t ld, int nb,
                                                              int64_t len, double *__restrict__ out, const XchgArgs x) {
    __shared__ double sm[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * 32 + tx;
    pdl_wait();      // the partial rows come from the kernel before this one
    pdl_launch();
    double acc = 0.0;
    if (i < len) {
        const double *p = part + i;
        int b = ty;
        for (; b + 56 < nb; b += 64) {
            const double v0 = p[(int64_t)b * ld], v1 = p[(int64_t)(b + 8) * ld], v2 = p[(int64_t)(b + 16) * ld],
                         v3 = p[(int64_t)(b + 24) * ld], v4 = p[(int64_t)(b + 32) * ld], v5 = p[(int64_t)(b + 40) * ld],
                         v6 = p[(int64_t)(b + 48) * ld], v7 = p[(int64_t)(b + 56) * ld];
            acc += ((v0 + v1) + (v2 + v3)) + ((v4 + v5) + (v6 + v7));
        }
        for (; b < nb; b += 8) acc += p[(int64_t)b * ld];
    }
    sm[ty][tx] = acc;
    __syncthreads();
    double mine = 0.0;
    if (ty == 0 && i < len) {
#pragma unroll
        for (int y = 0; y < 8; ++y) mine += sm[y][tx];
        if (!ACROSS) out[i] = mine;
    }
    if (!ACROSS) return;

    const size_t set = (size_t)(x.seq & 1) * PLSB_MAX_RANKS * PLSB_XCHG_LEN;
    if (ty == 0 && i < len) {
#pragma unroll
        for (int r = 0; r < PLSB_MAX_RANKS; ++r)
            if (r < x.nranks)
                reinterpret_cast<double *>(x.area[r] + XCHG_SLOT_OFF)[set + (size_t)x.rank * PLSB_XCHG_LEN + i] = mine;
    }
    __threadfence_system();
    __syncthreads();
    char *local = x.area[x.rank];
    if (threadIdx.x == 0) {
        const unsigned prev = atomicAdd(x.counter, 1u);
        if (prev == gridDim.x - 1) {   // this rank's vector has been pushed everywhere: announce it
            *x.counter = 0;
            __threadfence_system();
            for (int r = 0; r < x.nranks; ++r)
                st_relaxed_sys(reinterpret_cast<unsigned long long *>(x.area[r] + XCHG_FLAG_OFF) + x.rank, x.seq);
        }
    }
    if ((int)threadIdx.x < x.nranks) {   // one polling thread per source rank, all on local memory
        const unsigned long long *flag = reinterpret_cast<const unsigned long long *>(local + XCHG_FLAG_OFF) + threadIdx.x;
        const unsigned long long t0 = global_ns();
        // a peer that never shows up costs ONE timeout: once the flag is up, this and every later exchange of the
        // call fall straight through (the results are garbage from there on; xchg_check turns the flag into the
        // call's error and puts the context's communicator in a failed state)
        while (ld_acquire_sys(flag) < x.seq) {
            if (*reinterpret_cast<volatile int *>(x.err)) break;
            if (global_ns() - t0 > x.timeout_ns) { *reinterpret_cast<volatile int *>(x.err) = 1; break; }
        }
    }
    __syncthreads();
    if (ty == 0 && i < len) {
        const double *slots = reinterpret_cast<const double *>(local + XCHG_SLOT_OFF) + set + i;
        double v[PLSB_MAX_RANKS];
#pragma unroll
        for (int r = 0; r < PLSB_MAX_RANKS; ++r) v[r] = r < x.nranks ? ld_relaxed_sys(slots + (size_t)r * PLSB_XCHG_LEN) : 0.0;
        double tot = 0.0;
#pragma unroll
        for (int r = 0; r < PLSB_MAX_RANKS; ++r) tot += v[r];   // rank order; absent ranks add an exact zero
        out[i] = tot;
    }
}

int launch_reduce_partials(plsb_ctx *c, const double *part, int64_t ld, int nb, int64_t len, double *out,
                           bool across_ranks) {
    const unsigned grid = (unsigned)ceil_div(len, 32);
    XchgArgs x{};
    if (across_ranks && c->xchg.failed) {
        set_error("the communicator of this context is in a failed state (an earlier peer all-reduce timed out): "
                  "destroy the contexts and create them again");
        return PLSB_ERR_NCCL;
    }
    const bool fused = across_ranks && c->nranks > 1 && c->xchg.ready && len <= PLSB_XCHG_LEN && grid <= (unsigned)c->sm_count;
    if (fused) {
        x.rank = c->rank; x.nranks = c->nranks; x.seq = ++c->xchg.seq;
        for (int r = 0; r < c->nranks; ++r) x.area[r] = c->xchg.peer[r];
        x.counter = reinterpret_cast<unsigned *>(c->xchg.base + XCHG_CTR_OFF);
        x.err = reinterpret_cast<int *>(c->xchg.base + XCHG_ERR_OFF);
        static const unsigned long long tmo = [] {
            const char *e = getenv("PLSB_XCHG_TIMEOUT_MS");
            const long long ms = e ? atoll(e) : 0;
            return (unsigned long long)(ms > 0 ? ms : 10000) * 1000000ull;
        }();
        x.timeout_ns = tmo;
    }
    {
        ProfScope ps(c, PLSB_PROF_REDUCE);
        if (fused) PLSB_CUDA(launch_dep(reduce_partials_kernel<true>, dim3(grid), dim3(256), 0, c->stream, part, ld, nb, len, out, x));
        else PLSB_CUDA(launch_dep(reduce_partials_kernel<false>, dim3(grid), dim3(256), 0, c->stream, part, ld, nb, len, out, x));
    }
    c->launches++;
    PLSB_CUDA(cudaGetLastError());
    if (across_ranks && !fused) PLSB_TRY(allreduce_sum(c, out, (size_t)len));
    return PLSB_OK;
}

// ---- set-up: allocate this rank's area, exchange handles through the communicator, map the peers' areas -----
struct XchgHello {
    cudaIpcMemHandle_t handle;
    long long pid;
    char *ptr;
    int device;
    int ok;
};

int xchg_setup(plsb_ctx *c, int (*allgather)(plsb_ctx *, const void *, void *, size_t)) {
    if (c->nranks <= 1 || c->nranks > PLSB_MAX_RANKS || getenv("PLSB_NO_P2P")) return PLSB_OK;
    XchgHello me{};
    me.pid = (long long)getpid();
    me.device = c->device;
    me.ok = 0;
    if (cudaMalloc(&c->xchg.base, XCHG_BYTES) == cudaSuccess && cudaMemset(c->xchg.base, 0, XCHG_BYTES) == cudaSuccess &&
        cudaIpcGetMemHandle(&me.handle, c->xchg.base) == cudaSuccess)
        me.ok = 1;
    cudaGetLastError();
    me.ptr = c->xchg.base;
    XchgHello all[PLSB_MAX_RANKS];
    PLSB_TRY(allgather(c, &me, all, sizeof(XchgHello)));
    bool good = true;
    for (int r = 0; r < c->nranks; ++r) good = good && all[r].ok;
    for (int r = 0; r < c->nranks && good; ++r) {
        if (r == c->rank) { c->xchg.peer[r] = c->xchg.base; continue; }
        if (all[r].pid == me.pid) {   // another context of this process: plain peer access
            int can = 0;
            cudaDeviceCanAccessPeer(&can, c->device, all[r].device);
            if (all[r].device == c->device) { good = false; break; }  // two ranks on one GPU: kernels could starve each other
            {
                if (!can) { good = false; break; }
                cudaError_t e = cudaDeviceEnablePeerAccess(all[r].device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { good = false; break; }
                cudaGetLastError();
            }
            c->xchg.peer[r] = all[r].ptr;
        } else {
            void *p = nullptr;
            if (cudaIpcOpenMemHandle(&p, all[r].handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { good = false; break; }
            c->xchg.peer[r] = (char *)p;
            c->xchg.ipc_opened[r] = true;
        }
    }
    cudaGetLastError();
    // every rank must take the same decision: a second gather of the local verdicts
    int verdict = good ? 1 : 0, verdicts[PLSB_MAX_RANKS * 16];
    int padded[16] = {verdict};
    PLSB_TRY(allgather(c, padded, verdicts, sizeof(padded)));
    for (int r = 0; r < c->nranks; ++r) good = good && verdicts[r * 16];
    c->xchg.ready = good;
    c->xchg.seq = 0;
    return PLSB_OK;
}

void xchg_teardown(plsb_ctx *c) {
    for (int r = 0; r < PLSB_MAX_RANKS; ++r)
        if (c->xchg.ipc_opened[r] && c->xchg.peer[r]) cudaIpcCloseMemHandle(c->xchg.peer[r]);
    if (c->xchg.base) cudaFree(c->xchg.base);
    c->xchg = plsb_ctx::Xchg();
}

// raise the exchange's timeout flag, if any, as an error of the call that just synchronised
int xchg_check(plsb_ctx *c) {
    if (!c->xchg.ready) return PLSB_OK;
    int err = 0;
    PLSB_CUDA(cudaMemcpy(&err, c->xchg.base + XCHG_ERR_OFF, sizeof(int), cudaMemcpyDeviceToHost));
    if (err) {
        // The ranks no longer agree on the exchange number and their results are garbage: no further sharded
        // call may run on this communicator.  The flag itself is cleared so that the report is made once.
        c->xchg.failed = true;
        cudaMemset(c->xchg.base + XCHG_ERR_OFF, 0, sizeof(int));
        set_error("peer all-reduce timed out: a rank of the communicator did not reach the exchange");
        return PLSB_ERR_NCCL;
    }
    return PLSB_OK;
}

}  // namespace plsb
