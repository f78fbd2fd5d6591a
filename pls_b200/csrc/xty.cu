// pls_b200/csrc/xty.cu — S = X'Y as a streaming FP64 tensor-core (DMMA) kernel with TMA-staged row tiles.
//
// Replaces XY = X.transpose() * Y (reference src/pls.cpp:396).  The contraction is 2*N*K*M flops over one read of
// X (8*N*(K+M) bytes): HBM-bound, but with plain FMAs the FP64 pipe saturates first (8 DFMA per loaded element at
// M = 8; measured 4.0 TB/s).  Here every 4 rows x 8 columns of X meet 4 rows x 8 responses of Y in ONE
// mma.sync.m8n8k4.f64 (SASS DMMA): one LDS.64 + one DMMA per warp per 256 bytes of X.
//
// Structure: one persistent CTA per SM, 16 consumer warps + 1 producer warp.  The producer keeps a ring of
// shared-memory stages filled with 1-D TMA bulk copies (one per row, so rows land at a PADDED stride that makes
// the A-fragment loads bank-conflict free; full/empty mbarriers, no CTA-wide barrier in the loop).  Consumer warp
// w owns column groups {w, w+16, ...} (8 columns each) and keeps their 8 x 8 accumulator tiles in registers for
// the whole pass.  Per-CTA partial rows are summed in CTA order by reduce_partials (deterministic, no atomics).
#include "internal.h"

namespace plsb {

constexpr int XT_CW = 16;                 // consumer warps
constexpr int XT_NT = (XT_CW + 1) * 32;   // + the producer warp

struct XtyArgs {
    const double *X, *Y;
    double *part;
    int64_t N, tiles;
    int ldk, ldm, ldp, stages;
    int rows;          // rows per tile (multiple of 4)
    int xs;            // padded shared-memory row stride of the X tile (doubles)
    int stage_bytes, ytile_off;
    int m0, mc;
};

__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void dmma_884(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

template <int CG>
__global__ void __launch_bounds__(XT_NT, 1) xty_dmma_kernel(const XtyArgs a) {
    extern __shared__ __align__(128) char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)a.stages * a.stage_bytes);
    uint64_t *empty = full + 8;
    if (tid == 0) {
        for (int s = 0; s < a.stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], XT_CW);
        }
        mbar_fence_init();
    }
    __syncthreads();

    if (warp == XT_CW) {
        // ---- producer: one bulk copy per row (padded destination stride) + one for the Y rows of the tile ----
        int j = 0;
        for (int64_t tile = blockIdx.x; tile < a.tiles; tile += gridDim.x, ++j) {
            const int s = j % a.stages;
            if (j >= a.stages) mbar_wait(&empty[s], (uint32_t)(((j / a.stages) - 1) & 1));
            const int64_t row0 = tile * a.rows;
            int rows = (int)(a.N - row0 < a.rows ? a.N - row0 : a.rows);
            char *st = smem + (size_t)s * a.stage_bytes;
            if (lane == 0) mbar_arrive_expect_tx(&full[s], (uint32_t)rows * (uint32_t)(a.ldk + a.ldm) * 8u);
            __syncwarp();
            for (int r = lane; r < rows; r += 32)
                bulk_g2s(st + (size_t)r * a.xs * 8, a.X + (row0 + r) * a.ldk, (uint32_t)a.ldk * 8u, &full[s]);
            if (lane == 0) bulk_g2s(st + a.ytile_off, a.Y + row0 * a.ldm, (uint32_t)rows * (uint32_t)a.ldm * 8u, &full[s]);
        }
        return;
    }

    // ---- consumers ------------------------------------------------------------------------------------------
    const int gq = lane >> 2, kq = lane & 3;  // fragment coordinates: A(m = gq, k = kq), B(k = kq, n = gq)
    double acc[CG][2];
    int col[CG];
#pragma unroll
    for (int i = 0; i < CG; ++i) {
        acc[i][0] = 0.0;
        acc[i][1] = 0.0;
        col[i] = (warp + XT_CW * i) * 8 + gq;
    }
    const bool resp_ok = gq < a.mc;
    int j = 0;
    for (int64_t tile = blockIdx.x; tile < a.tiles; tile += gridDim.x, ++j) {
        const int s = j % a.stages;
        mbar_wait(&full[s], (uint32_t)((j / a.stages) & 1));
        const char *st = smem + (size_t)s * a.stage_bytes;
        const double *xs = reinterpret_cast<const double *>(st);
        const double *ys = reinterpret_cast<const double *>(st + a.ytile_off) + a.m0 + gq;
        const int64_t left = a.N - tile * a.rows;  // valid rows in this tile (may exceed a.rows)
        for (int r4 = 0; r4 < a.rows; r4 += 4) {
            if (r4 >= left) break;
            const int row = r4 + kq;
            const bool live = row < left;   // rows past N hold stale shared memory: feed zeros
            const double b = (live && resp_ok) ? ys[(size_t)row * a.ldm] : 0.0;
            const double *xr = xs + (size_t)row * a.xs;
            double av[CG];
#pragma unroll
            for (int i = 0; i < CG; ++i) av[i] = (live && col[i] < a.ldk) ? xr[col[i]] : 0.0;
#pragma unroll
            for (int i = 0; i < CG; ++i) dmma_884(acc[i][0], acc[i][1], av[i], b);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }
    // D fragment: row (column of X) = gq, responses 2*kq, 2*kq+1.  Partial row layout: part[cta][m*ldk + col]
    double *prow = a.part + (size_t)blockIdx.x * a.ldp;
#pragma unroll
    for (int i = 0; i < CG; ++i) {
        if (col[i] < a.ldk) {
            if (2 * kq < a.mc) prow[(size_t)(2 * kq) * a.ldk + col[i]] = acc[i][0];
            if (2 * kq + 1 < a.mc) prow[(size_t)(2 * kq + 1) * a.ldk + col[i]] = acc[i][1];
        }
    }
}

template <int CG>
static int run_xty_dmma_cg(plsb_ctx *c, const plsb_dataset *ds, double *S) {
    XtyArgs a{};
    a.X = ds->X; a.Y = ds->Y; a.N = ds->N; a.ldk = (int)ds->ldk; a.ldm = (int)ds->ldm;
    a.ldp = 8 * a.ldk;
    a.xs = a.ldk + (20 - a.ldk % 16) % 16;            // (xs mod 16) == 4: the 4 k-rows of a fragment hit disjoint banks
    const size_t row_bytes = (size_t)a.xs * 8 + (size_t)a.ldm * 8;
    int rows = (int)(65536 / row_bytes) / 4 * 4;      // ~64 KB stages
    if (rows < 4) rows = 4;
    if (rows > 128) rows = 128;
    a.rows = rows;
    a.ytile_off = (int)round_up((int64_t)rows * a.xs * 8, 128);
    a.stage_bytes = (int)round_up(a.ytile_off + (int64_t)rows * a.ldm * 8, 128);
    int stages = (int)((c->smem_optin - 256) / (size_t)a.stage_bytes);
    if (stages > 4) stages = 4;
    if (stages < 1) { set_error("xty: K=%lld does not fit shared memory", (long long)ds->K); return PLSB_ERR_UNSUPPORTED; }
    a.stages = stages;
    a.tiles = ceil_div(ds->N, rows);
    const size_t smem = (size_t)stages * a.stage_bytes + 128;
    int64_t grid = c->sm_count;
    if (grid > a.tiles) grid = a.tiles;
    PLSB_TRY(ensure_part(c, (size_t)grid * a.ldp * sizeof(double)));
    a.part = c->part;
    auto kern = xty_dmma_kernel<CG>;
    PLSB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int m0 = 0; m0 < ds->M; m0 += 8) {
        a.m0 = m0;
        a.mc = (int)(ds->M - m0 < 8 ? ds->M - m0 : 8);
        {
            ProfScope ps(c, PLSB_PROF_PASS_XTY);
            kern<<<(unsigned)grid, XT_NT, smem, c->stream>>>(a);
        }
        PLSB_CUDA(cudaGetLastError());
        c->launches += 1;
        PLSB_TRY(launch_reduce_partials(c, c->part, a.ldp, (int)grid, (int64_t)a.mc * a.ldk, S + (size_t)m0 * a.ldk, false));
    }
    return PLSB_OK;
}

// ldk in [256, 2048]: rows of >= 2 KB (one bulk copy each) and <= 16 accumulator tiles per warp
bool xty_dmma_supported(const plsb_dataset *ds) { return ds->ldk >= 256 && ds->ldk <= 2048 && ds->N > 0; }

int launch_xty_dmma(plsb_ctx *c, const plsb_dataset *ds, double *S) {
    const int64_t ncg = ceil_div(ds->ldk, 8);
    if (ncg <= 4 * XT_CW) return run_xty_dmma_cg<4>(c, ds, S);
    if (ncg <= 8 * XT_CW) return run_xty_dmma_cg<8>(c, ds, S);
    return run_xty_dmma_cg<16>(c, ds, S);
}

}  // namespace plsb
