// pls_b200/csrc/gemm.cu — FP64 tensor-core (DMMA) contraction C = A' B over a long reduction axis.
//
// Replaces XX = X' * X (reference src/pls.cpp:398; Eigen GEBP there) and serves the batched-fold
// product XX * R_batch of the cross-validation engine.  Operands are "k-major": A is Kd x M and B is
// Kd x N, both row-major (the reduction index is the slow one), which is exactly how X (rows = samples)
// and the fold-minor vector batches sit in HBM.  FP64 has no tcgen05 kind on sm_100a: the tensor path
// is warp-level mma.sync m8n8k4 (SASS DMMA.8x8x4; the wider PTX shapes m16n8k4/k8/k16 decompose into the
// same instruction on sm_100a).
//
// dmma_tma_kernel (default): a warp-specialised pipeline.  One PRODUCER warp keeps a ring of shared-memory
// stages filled with 1-D TMA bulk copies (cp.async.bulk + mbarrier complete_tx, SASS UBLKCP) — one copy per
// k-row of a k-major tile, one per column of an n-major B tile, each landing at a PADDED stride so that the
// fragment loads are bank-conflict free — and the CONSUMER warps do nothing but LDS + DMMA between a `full`
// wait and an `empty` arrive: no CTA-wide barrier and no address arithmetic in the k-loop.  (The first
// version, dmma_atb_kernel below, staged with cp.async from all warps behind a __syncthreads() per k-tile: the
// ~25 integer instructions per 16-byte copy drained the tensor pipe after every barrier — 81 % pipe-active, ncu.)
//
// Split-K over the reduction axis fills the 148 SMs; partial tiles go to scratch and are summed in
// split order by a second kernel (no atomics: deterministic).  SYRK mode computes only tiles on or
// above the diagonal and mirrors them.
#include "internal.h"

namespace plsb {

#ifndef PLSB_GM_BK       // tuning knobs (tools/gemm_variants.sh builds A/B libraries with other values)
#define PLSB_GM_BK 32
#endif
#ifndef PLSB_GM_STAGES
#define PLSB_GM_STAGES 3
#endif
#ifndef PLSB_GM_NT
#define PLSB_GM_NT 256
#endif
constexpr int GM_BM = 128, GM_BN = 128, GM_BK = PLSB_GM_BK, GM_LD = 132, GM_NT = PLSB_GM_NT, GM_STAGES = PLSB_GM_STAGES;
constexpr int GM_WM = GM_BM / (GM_NT / 32 / 4);   // rows per warp tile: 4 warps across the 128 columns, the rest down
constexpr int GM_MI = GM_WM / 8;                  // 8-row fragments per warp tile
constexpr int GM_LDBN = GM_BK + 4;                       // n-major B tile: [BN][BK + 4]
constexpr int GM_BS_KMAJOR = GM_BK * GM_LD, GM_BS_NMAJOR = GM_BN * GM_LDBN;

struct GemmArgs {
    const double *A; int64_t lda;   // Kd x M
    const double *B; int64_t ldb;   // Kd x N
    int64_t Kd; int M, N;
    int tiles_m, tiles_n, ntiles, splitk;
    int64_t k_per_split;            // multiple of GM_BK
    double *part;                   // [splitk][ntiles][BM*BN]
    int syrk;
    int transC;                     // write C[n*ldc + m] instead of C[m*ldc + n]
    int accum;                      // C += product (through the partial-tile reduction) instead of C = product
    double *C; int64_t ldc;         // direct output when splitk == 1
};

__device__ __forceinline__ void cp_async16(void *dst, const void *src, bool pred) {
    const int sz = pred ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// map a linear SYRK tile index to (tm, tn) with tn >= tm
__device__ __forceinline__ void syrk_tile(int t, int T, int &tm, int &tn) {
    int row = 0;
    while (t >= T - row) { t -= T - row; ++row; }
    tm = row; tn = row + t;
}

// B_NMAJOR: B is given as N x Kd (each column of the product contiguous along the reduction axis), the
// layout of a batch of per-fold vectors.
template <bool B_NMAJOR>
__global__ void __launch_bounds__(GM_NT, 1) dmma_atb_kernel(const GemmArgs g) {
    extern __shared__ __align__(128) char smem[];
    double *As = reinterpret_cast<double *>(smem);                       // [STAGES][BK][LD]
    double *Bs = As + (size_t)GM_STAGES * GM_BK * GM_LD;
    constexpr int BS_STAGE = B_NMAJOR ? GM_BS_NMAJOR : GM_BS_KMAJOR;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x % g.ntiles, split = blockIdx.x / g.ntiles;
    int tm, tn;
    if (g.syrk) syrk_tile(tile, g.tiles_m, tm, tn);
    else { tm = tile / g.tiles_n; tn = tile % g.tiles_n; }
    const int m_base = tm * GM_BM, n_base = tn * GM_BN;
    const int64_t k_begin = (int64_t)split * g.k_per_split;
    int64_t k_end = k_begin + g.k_per_split;
    if (k_end > g.Kd) k_end = g.Kd;
    const int nk = k_begin < k_end ? (int)((k_end - k_begin + GM_BK - 1) / GM_BK) : 0;

    auto load_stage = [&](int st, int kt) {
        const int64_t k0 = k_begin + (int64_t)kt * GM_BK;
#pragma unroll
        for (int i = 0; i < (GM_BK * GM_BM / 2) / GM_NT; ++i) {
            const int ch = tid + i * GM_NT;        // 16-byte chunk id in the BK x 128 tile
            const int kr = ch / (GM_BM / 2), cc = (ch % (GM_BM / 2)) * 2;
            const int64_t k = k0 + kr;
            const bool kin = k < k_end;
            const bool pa = kin && (m_base + cc) < g.M;
            cp_async16(As + ((size_t)st * GM_BK + kr) * GM_LD + cc, pa ? g.A + k * g.lda + m_base + cc : g.A, pa);
            if (!B_NMAJOR) {
                const bool pb = kin && (n_base + cc) < g.N;
                cp_async16(Bs + (size_t)st * BS_STAGE + kr * GM_LD + cc, pb ? g.B + k * g.ldb + n_base + cc : g.B, pb);
            } else {
                const int n = ch / (GM_BK / 2), kc = (ch % (GM_BK / 2)) * 2;
                const bool pb = (k0 + kc) < k_end && (n_base + n) < g.N;
                cp_async16(Bs + (size_t)st * BS_STAGE + n * GM_LDBN + kc,
                           pb ? g.B + (int64_t)(n_base + n) * g.ldb + k0 + kc : g.B, pb);
            }
        }
    };

    double acc[GM_MI][4][2];
#pragma unroll
    for (int i = 0; i < GM_MI; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }

    for (int s = 0; s < GM_STAGES - 1; ++s) {
        if (s < nk) load_stage(s, s);
        cp_async_commit();
    }
    const int wm = (warp >> 2) * GM_WM, wn = (warp & 3) * 32;
    const int fk = lane & 3, fr = lane >> 2;
    for (int kt = 0; kt < nk; ++kt) {
        cp_async_wait<GM_STAGES - 2>();
        __syncthreads();
        {   // prefetch the tile STAGES-1 ahead into the slot freed in the previous iteration
            const int nx = kt + GM_STAGES - 1;
            if (nx < nk) load_stage(nx % GM_STAGES, nx);
            cp_async_commit();
        }
        const double *as = As + (size_t)(kt % GM_STAGES) * GM_BK * GM_LD;
        const double *bs = Bs + (size_t)(kt % GM_STAGES) * BS_STAGE;
#pragma unroll
        for (int kk = 0; kk < GM_BK / 4; ++kk) {
            double a[GM_MI], b[4];
#pragma unroll
            for (int i = 0; i < GM_MI; ++i) a[i] = as[(kk * 4 + fk) * GM_LD + wm + i * 8 + fr];
#pragma unroll
            for (int j = 0; j < 4; ++j)
                b[j] = B_NMAJOR ? bs[(wn + j * 8 + fr) * GM_LDBN + kk * 4 + fk] : bs[(kk * 4 + fk) * GM_LD + wn + j * 8 + fr];
#pragma unroll
            for (int i = 0; i < GM_MI; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
    }
    cp_async_wait<0>();
    if (g.splitk == 1 && !g.syrk && !g.accum) {  // no partials needed: write the product directly
#pragma unroll
        for (int i = 0; i < GM_MI; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int m = m_base + wm + i * 8 + fr, n = n_base + wn + j * 8 + fk * 2;
                if (m >= g.M) continue;
                if (!g.transC) {
                    if (n < g.N) g.C[(int64_t)m * g.ldc + n] = acc[i][j][0];
                    if (n + 1 < g.N) g.C[(int64_t)m * g.ldc + n + 1] = acc[i][j][1];
                } else {
                    if (n < g.N) g.C[(int64_t)n * g.ldc + m] = acc[i][j][0];
                    if (n + 1 < g.N) g.C[(int64_t)(n + 1) * g.ldc + m] = acc[i][j][1];
                }
            }
        return;
    }
    double *out = g.part + ((size_t)split * g.ntiles + tile) * (GM_BM * GM_BN);
#pragma unroll
    for (int i = 0; i < GM_MI; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int row = wm + i * 8 + fr, col = wn + j * 8 + fk * 2;
            *reinterpret_cast<double2 *>(out + row * GM_BN + col) = make_double2(acc[i][j][0], acc[i][j][1]);
        }
}

// ------------------------------------------------------------------------------------------------------------------
// Warp-specialised TMA pipeline (see the header).  CW consumer warps in a (CW/4) x 4 grid over the 128 x 128 tile
// (CW = 8: 64 x 32 per warp), + 1 producer warp.  Stage = A tile [BK][GT_LD] + B tile ([BK][GT_LD] k-major or
// [BN][GT_LDBN] n-major), 3 stages of depth 32.
// ------------------------------------------------------------------------------------------------------------------
constexpr int GT_BK = 32, GT_LD = 132, GT_LDBN = GT_BK + 4, GT_STAGES = 3;
constexpr int GT_A_BYTES = GT_BK * GT_LD * 8;
__host__ __device__ constexpr int gt_b_bytes(bool nm) { return nm ? GM_BN * GT_LDBN * 8 : GT_BK * GT_LD * 8; }
__host__ __device__ constexpr int gt_stage_bytes(bool nm) { return GT_A_BYTES + gt_b_bytes(nm); }

__device__ __forceinline__ void gt_mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// SELF = false: CW consumer warps + one producer warp (warp CW).  SELF = true: no producer warp — the CW consumer warps
// take turns (warp gk mod CW issues the copies of k-tile gk + 2 at the top of iteration gk).  A CTA of 8 warps x 168
// registers leaves a quarter of the SM's register file and 16 KB of its shared memory free, so that a memory-bound kernel
// of another stream (the per-fold kernel of the other CV wave, cv.cu) can be co-resident and use the HBM bandwidth this
// tensor-bound kernel leaves idle; 9 warps are allocated as 12 and take the whole register file.
//
// The kernel is PERSISTENT over work items (output tile x k-split): CTA b takes items b, b + gridDim.x, ...  The ring of
// stages and its mbarrier phases run on across items (one global k-tile counter), so the copies of the next item's first
// tiles are in flight while the warps write the finished tile out.  Launched with one CTA per work item it is the plain
// kernel; launched with one CTA per SM (`persist`) ALL its CTAs are resident from the start, which is what lets the block
// scheduler place the CTAs of a concurrent kernel of another stream in the resources left over — it dispatches the
// blocks of a second kernel only once every block of the first has been dispatched.
struct GtItem {
    int m_base, n_base, a_cols, b_cols, tile, split, nk;
    int64_t k_begin, k_end;
};
__device__ __forceinline__ GtItem gt_item(const GemmArgs &g, int w) {
    GtItem it;
    it.tile = w % g.ntiles; it.split = w / g.ntiles;
    int tm, tn;
    if (g.syrk) syrk_tile(it.tile, g.tiles_m, tm, tn);
    else { tm = it.tile / g.tiles_n; tn = it.tile % g.tiles_n; }
    it.m_base = tm * GM_BM; it.n_base = tn * GM_BN;
    it.k_begin = (int64_t)it.split * g.k_per_split;
    it.k_end = it.k_begin + g.k_per_split;
    if (it.k_end > g.Kd) it.k_end = g.Kd;
    it.nk = it.k_begin < it.k_end ? (int)((it.k_end - it.k_begin + GT_BK - 1) / GT_BK) : 0;
    int a_cols = g.M - it.m_base; it.a_cols = a_cols > GM_BM ? GM_BM : (a_cols + 1) & ~1;
    int b_cols = g.N - it.n_base; it.b_cols = b_cols > GM_BN ? GM_BN : b_cols;   // n-major: rows of the B tile
    return it;
}

// length of work item w's slice of the reduction axis (what the loops need; the rest of GtItem is recomputed where used,
// to keep the cursor of the pipeline in a few registers)
__device__ __forceinline__ int gt_klen(const GemmArgs &g, int w) {
    const int64_t k_begin = (int64_t)(w / g.ntiles) * g.k_per_split;
    int64_t k_end = k_begin + g.k_per_split;
    if (k_end > g.Kd) k_end = g.Kd;
    return k_end > k_begin ? (int)(k_end - k_begin) : 0;
}

template <bool B_NMAJOR, int CW, bool SELF>
__global__ void __launch_bounds__((CW + (SELF ? 0 : 1)) * 32) __maxnreg__(CW == 8 ? 168 : 96) dmma_tma_kernel(const GemmArgs g) {
    extern __shared__ __align__(128) char smem[];
    constexpr int STAGE = gt_stage_bytes(B_NMAJOR);
    constexpr int WM = GM_BM / (CW / 4), MI = WM / 8;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)GT_STAGES * STAGE);
    uint64_t *empty = full + GT_STAGES;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nwork = g.ntiles * g.splitk;
    if (tid == 0) {
        for (int s = 0; s < GT_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], CW);
        }
        mbar_fence_init();
    }
    __syncthreads();

    // ---- producer cursor: the next k-tile to copy, as (work item, k-tile, global k-tile count).  Every warp that may
    // issue copies advances its own copy of the cursor identically; `issue` says whether THIS warp issues the copies:
    // bulk copies of the valid part of each tile row (rows / columns past the matrix edge are never copied: the
    // consumers mask the k tail, and out-of-range output rows / columns are discarded) ----
    int pw = blockIdx.x, pkt = 0, pgk = 0;
    int pnk = pw < nwork ? (gt_klen(g, pw) + GT_BK - 1) / GT_BK : 0;
    auto produce_next = [&](bool issue) -> bool {
        while (pw < nwork && pkt >= pnk) {
            pw += gridDim.x; pkt = 0;
            if (pw < nwork) pnk = (gt_klen(g, pw) + GT_BK - 1) / GT_BK;
        }
        if (pw >= nwork) return false;
        if (issue) {
            const GtItem pit = gt_item(g, pw);
            const int s = pgk % GT_STAGES;
            if (pgk >= GT_STAGES) mbar_wait(&empty[s], (uint32_t)(((pgk / GT_STAGES) - 1) & 1));
            const int64_t k0 = pit.k_begin + (int64_t)pkt * GT_BK;
            const int kv = (int)(pit.k_end - k0 < GT_BK ? pit.k_end - k0 : GT_BK);
            const int kv2 = (kv + 1) & ~1;
            const int b_cols2 = (pit.b_cols + 1) & ~1;                             // k-major: doubles per k-row
            char *sa = smem + (size_t)s * STAGE, *sb = sa + GT_A_BYTES;
            const uint32_t bytes = (uint32_t)kv * (uint32_t)pit.a_cols * 8u +
                                   (B_NMAJOR ? (uint32_t)pit.b_cols * (uint32_t)kv2 * 8u : (uint32_t)kv * (uint32_t)b_cols2 * 8u);
            if (lane == 0) mbar_arrive_expect_tx(&full[s], bytes);
            __syncwarp();
            if (lane < kv) {
                bulk_g2s(sa + (size_t)lane * GT_LD * 8, g.A + (k0 + lane) * g.lda + pit.m_base, (uint32_t)pit.a_cols * 8u, &full[s]);
                if (!B_NMAJOR)
                    bulk_g2s(sb + (size_t)lane * GT_LD * 8, g.B + (k0 + lane) * g.ldb + pit.n_base, (uint32_t)b_cols2 * 8u, &full[s]);
            }
            if (B_NMAJOR) {
#pragma unroll
                for (int i = 0; i < GM_BN / 32; ++i) {
                    const int n = lane + 32 * i;
                    if (n < pit.b_cols)
                        bulk_g2s(sb + (size_t)n * GT_LDBN * 8, g.B + (int64_t)(pit.n_base + n) * g.ldb + k0, (uint32_t)kv2 * 8u, &full[s]);
                }
            }
        }
        ++pkt; ++pgk;
        return true;
    };
    if (!SELF && warp == CW) {
        while (produce_next(true)) {}
        return;
    }
    if (SELF) {   // prologue: the first STAGES - 1 k-tiles of this CTA, by warps 0 .. STAGES - 2
#pragma unroll
        for (int i = 0; i < GT_STAGES - 1; ++i) produce_next(warp == i);
    }

    // ---- consumers ------------------------------------------------------------------------------------------------
    const int wm = (warp >> 2) * WM, wn = (warp & 3) * 32;
    const int fk = lane & 3, fr = lane >> 2;
    int gk = 0;                                   // k-tiles this CTA has consumed so far (all work items)
    for (int w = blockIdx.x; w < nwork; w += gridDim.x) {
        const int klen = gt_klen(g, w), nk = (klen + GT_BK - 1) / GT_BK;
        double acc[MI][4][2];
#pragma unroll
        for (int i = 0; i < MI; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }
        for (int kt = 0; kt < nk; ++kt, ++gk) {
            const int s = gk % GT_STAGES;
            if (SELF) produce_next(warp == (gk % CW));   // k-tile gk + STAGES - 1, into the stage freed by k-tile gk - 1
            mbar_wait(&full[s], (uint32_t)((gk / GT_STAGES) & 1));
            const double *as = reinterpret_cast<const double *>(smem + (size_t)s * STAGE) + fk * GT_LD + wm + fr;
            const double *bs = reinterpret_cast<const double *>(smem + (size_t)s * STAGE + GT_A_BYTES) +
                               (B_NMAJOR ? (wn + fr) * GT_LDBN + fk : fk * GT_LD + wn + fr);
            const int kleft = klen - kt * GT_BK;
            if (kleft >= GT_BK) {
#pragma unroll
                for (int kk = 0; kk < GT_BK / 4; ++kk) {
                    double a[MI], b[4];
#pragma unroll
                    for (int i = 0; i < MI; ++i) a[i] = as[kk * 4 * GT_LD + i * 8];
#pragma unroll
                    for (int j = 0; j < 4; ++j) b[j] = B_NMAJOR ? bs[j * 8 * GT_LDBN + kk * 4] : bs[kk * 4 * GT_LD + j * 8];
#pragma unroll
                    for (int i = 0; i < MI; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
                }
            } else {   // last k-tile of a split: rows past the end were not copied and hold stale shared memory
                const int kv = kleft;
#pragma unroll 1
                for (int kk = 0; kk * 4 < kv; ++kk) {
                    const bool live = kk * 4 + fk < kv;
                    double a[MI], b[4];
#pragma unroll
                    for (int i = 0; i < MI; ++i) a[i] = live ? as[kk * 4 * GT_LD + i * 8] : 0.0;
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        b[j] = live ? (B_NMAJOR ? bs[j * 8 * GT_LDBN + kk * 4] : bs[kk * 4 * GT_LD + j * 8]) : 0.0;
#pragma unroll
                    for (int i = 0; i < MI; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
                }
            }
            __syncwarp();
            if (lane == 0) gt_mbar_arrive(&empty[s]);
        }
        const GtItem it = gt_item(g, w);
        if (g.splitk == 1 && !g.syrk && !g.accum) {  // no partials needed: write the product directly
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int m = it.m_base + wm + i * 8 + fr, n = it.n_base + wn + j * 8 + fk * 2;
                    if (m >= g.M) continue;
                    if (!g.transC) {
                        if (n < g.N) g.C[(int64_t)m * g.ldc + n] = acc[i][j][0];
                        if (n + 1 < g.N) g.C[(int64_t)m * g.ldc + n + 1] = acc[i][j][1];
                    } else {
                        if (n < g.N) g.C[(int64_t)n * g.ldc + m] = acc[i][j][0];
                        if (n + 1 < g.N) g.C[(int64_t)(n + 1) * g.ldc + m] = acc[i][j][1];
                    }
                }
        } else {
            double *out = g.part + ((size_t)it.split * g.ntiles + it.tile) * (GM_BM * GM_BN);
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int row = wm + i * 8 + fr, col = wn + j * 8 + fk * 2;
                    *reinterpret_cast<double2 *>(out + row * GM_BN + col) = make_double2(acc[i][j][0], acc[i][j][1]);
                }
        }
    }
}

// C[m*ldc + n] = sum_s part[s][tile][...]; SYRK also writes the mirrored element
__global__ void __launch_bounds__(256) gemm_reduce_kernel(const GemmArgs g, double *__restrict__ C, int64_t ldc) {
    const int tile = blockIdx.x;
    int tm, tn;
    if (g.syrk) syrk_tile(tile, g.tiles_m, tm, tn);
    else { tm = tile / g.tiles_n; tn = tile % g.tiles_n; }
    for (int e = threadIdx.x; e < GM_BM * GM_BN; e += 256) {
        const int row = e / GM_BN, col = e % GM_BN;
        const int m = tm * GM_BM + row, n = tn * GM_BN + col;
        if (m >= g.M || n >= g.N) continue;
        double s = 0.0;
        if (g.accum) s = g.transC ? C[(int64_t)n * ldc + m] : C[(int64_t)m * ldc + n];
        for (int sp = 0; sp < g.splitk; ++sp) s += g.part[((size_t)sp * g.ntiles + tile) * (GM_BM * GM_BN) + e];
        if (g.transC) C[(int64_t)n * ldc + m] = s;
        else C[(int64_t)m * ldc + n] = s;
        if (g.syrk && tm != tn) C[(int64_t)n * ldc + m] = s;
    }
}

// C (M x N, row-major, ldc) = A' B.  syrk: A == B, M == N, C symmetric.  b_nmajor: B given as N x Kd.
// transC: store C transposed (N x M, row-major).
int launch_gemm_atb(plsb_ctx *c, const double *A, int64_t lda, const double *B, int64_t ldb, int64_t Kd, int M, int N,
                    double *C, int64_t ldc, bool syrk, bool b_nmajor, bool transC, int prof_slot, bool accumulate,
                    GemmLane *lane, bool no_split) {
    if (M == 0 || N == 0) return PLSB_OK;
    PLSB_ARG((lda % 2) == 0 && (ldb % 2) == 0, "gemm: leading dimensions must be even");
    cudaStream_t st = lane ? lane->stream : c->stream;
    GemmArgs g{};
    g.A = A; g.lda = lda; g.B = B; g.ldb = ldb; g.Kd = Kd; g.M = M; g.N = N; g.syrk = syrk ? 1 : 0;
    g.transC = transC ? 1 : 0; g.C = C; g.ldc = ldc; g.accum = accumulate ? 1 : 0;
    g.tiles_m = (int)ceil_div(M, GM_BM);
    g.tiles_n = (int)ceil_div(N, GM_BN);
    g.ntiles = syrk ? g.tiles_m * (g.tiles_m + 1) / 2 : g.tiles_m * g.tiles_n;
    // Split-K: at least 64 k-steps per split, at most ~12 waves of CTAs, and among those the smallest split count
    // whose last wave is (nearly) full — one CTA per SM, so a grid of 4.3 waves costs 5 (C4 X'X: 136 SYRK tiles
    // x 13 splits = 1768 CTAs = 11.95 waves instead of 5 splits = 4.59; a 2500-fold CV wave: 640 tiles x 2).
    int64_t maxsplit = ceil_div(Kd, (int64_t)GM_BK * 64);
    if (maxsplit > 64) maxsplit = 64;
    if (no_split) maxsplit = 1;   // the caller wants sums that do not depend on the batch width (cv.cu)
    int64_t want = 1;
    double best_eff = -1.0;
    for (int64_t sp = 1; sp <= maxsplit; ++sp) {
        const int64_t ctas = (int64_t)g.ntiles * sp;
        if (sp > 1 && ctas > (int64_t)12 * c->sm_count) break;
        const double eff = (double)ctas / (double)(ceil_div(ctas, c->sm_count) * c->sm_count);
        if (eff > best_eff + 0.02) { best_eff = eff; want = sp; }
    }
    g.k_per_split = round_up(ceil_div(Kd > 0 ? Kd : 1, want), GM_BK);
    g.splitk = (int)ceil_div(Kd > 0 ? Kd : 1, g.k_per_split);
    const bool direct = g.splitk == 1 && !syrk && !accumulate;
    if (!direct) {
        const size_t pbytes = (size_t)g.splitk * g.ntiles * GM_BM * GM_BN * sizeof(double);
        if (lane) {     // a lane of a multi-stream caller owns its scratch (the context's belongs to c->stream)
            if (pbytes > lane->part_bytes) {
                if (lane->part) cudaFreeAsync(lane->part, st);
                lane->part = nullptr; lane->part_bytes = 0;
                PLSB_CUDA(cudaMallocAsync((void **)&lane->part, pbytes, st));
                lane->part_bytes = pbytes;
            }
            g.part = lane->part;
        } else {
            PLSB_TRY(ensure_part(c, pbytes));
            g.part = c->part;
        }
    }
    // A/B switches: PLSB_GEMM_V1=1 the cp.async kernel, PLSB_GEMM_CW=16 sixteen 32x32 consumer warps, PLSB_GEMM_SELF=0/1
    // forces the producer-warp / self-fed pipeline (default: self-fed for lanes, i.e. when another stream may co-run)
    static const int variant = [] {
        const char *v1 = getenv("PLSB_GEMM_V1"), *cw = getenv("PLSB_GEMM_CW");
        if (v1 && *v1 == '1') return 1;
        return (cw && atoi(cw) == 16) ? 16 : 8;
    }();
    static const int self_env = [] { const char *e = getenv("PLSB_GEMM_SELF"); return e ? atoi(e) : -1; }();
    const bool self = self_env >= 0 ? self_env != 0 : lane != nullptr;
    static const int persist_env = [] { const char *e = getenv("PLSB_GEMM_PERSIST"); return e ? atoi(e) : -1; }();
    if (b_nmajor) PLSB_ARG(ldb >= round_up(Kd, 2), "gemm: an n-major B needs ldb >= Kd rounded up to even");
    if (variant == 1) {
        const size_t smem = (size_t)GM_STAGES * (GM_BK * GM_LD + (b_nmajor ? GM_BS_NMAJOR : GM_BS_KMAJOR)) * sizeof(double);
        auto kern = b_nmajor ? dmma_atb_kernel<true> : dmma_atb_kernel<false>;
        PLSB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ProfScope ps(c, lane ? -1 : prof_slot);
        kern<<<(unsigned)(g.ntiles * g.splitk), GM_NT, smem, st>>>(g);
    } else {
        const size_t smem = (size_t)GT_STAGES * gt_stage_bytes(b_nmajor) + 2 * GT_STAGES * sizeof(uint64_t) + 16;
        void (*kern)(const GemmArgs);
        unsigned threads;
        if (variant == 16) { kern = b_nmajor ? dmma_tma_kernel<true, 16, false> : dmma_tma_kernel<false, 16, false>; threads = 17 * 32; }
        else if (self) { kern = b_nmajor ? dmma_tma_kernel<true, 8, true> : dmma_tma_kernel<false, 8, true>; threads = 8 * 32; }
        else { kern = b_nmajor ? dmma_tma_kernel<true, 8, false> : dmma_tma_kernel<false, 8, false>; threads = 9 * 32; }
        PLSB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        // one CTA per SM looping over the work items when another stream is to co-run (see the kernel's header);
        // PLSB_GEMM_PERSIST=0/1 forces either shape (A/B)
        const bool persist = persist_env >= 0 ? persist_env != 0 : lane != nullptr;
        unsigned grid = (unsigned)(g.ntiles * g.splitk);
        if (persist && grid > (unsigned)c->sm_count) grid = (unsigned)c->sm_count;
        ProfScope ps(c, lane ? -1 : prof_slot);
        kern<<<grid, threads, smem, st>>>(g);
    }
    PLSB_CUDA(cudaGetLastError());
    c->launches += 1;
    if (!direct) {
        gemm_reduce_kernel<<<(unsigned)g.ntiles, 256, 0, st>>>(g, C, ldc);
        PLSB_CUDA(cudaGetLastError());
        c->launches += 1;
    }
    return PLSB_OK;
}

int launch_xtx(plsb_ctx *c, const plsb_dataset *ds, double *G, int64_t ldg) {
    if (ds->N == 0) {
        PLSB_CUDA(cudaMemsetAsync(G, 0, (size_t)ds->K * ldg * sizeof(double), c->stream));
        return PLSB_OK;
    }
    return launch_gemm_atb(c, ds->X, ds->ldk, ds->X, ds->ldk, ds->N, (int)ds->K, (int)ds->K, G, ldg, true,
                           false, false, PLSB_PROF_XTX);
}

}  // namespace plsb
