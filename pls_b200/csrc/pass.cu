// pls_b200/csrc/pass.cu — the streaming passes over X (HBM-bound).
//
// Replaces, on the device:
//   XY = X' * Y                      reference src/pls.cpp:396
//   t = X r ; tt = t't ; p = X' t    reference src/pls.cpp:419-421   (TWO reads of X in the reference)
//
// Data layout: X is kept ROW-major on the device (N x ldk, ldk even, zero padded), so a block of rows
// is one contiguous byte range.  Each CTA streams row tiles through a ring of shared-memory stages
// filled by the TMA engine (1-D cp.async.bulk + mbarrier complete_tx); compute threads pull their
// columns of the tile into registers, form the partial dot products with r, reduce them across the
// CTA (warp shuffles + one shared-memory exchange), and immediately accumulate x * t into per-thread
// column accumulators: X is read from HBM ONCE per component (the fused pass).
//
// Determinism: no atomics.  Every CTA writes its column partials to a scratch row; a second tiny
// kernel sums the rows in CTA order.  The grid depends only on the SM count, so results are
// reproducible run to run.
#include "internal.h"

namespace plsb {

// ------------------------------------------------------------------------------------------------
// layout transforms
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) cm_to_rm_kernel(const double *__restrict__ src, int64_t lds, int64_t rows,
                                                       int64_t cols, double *__restrict__ dst, int64_t ldd,
                                                       int64_t col0) {
    __shared__ double tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    const int64_t r0 = (int64_t)blockIdx.x * 32, c0 = (int64_t)blockIdx.y * 32;
#pragma unroll
    for (int j = ty; j < 32; j += 8) {
        int64_t r = r0 + tx, cidx = c0 + j;
        tile[j][tx] = (r < rows && cidx < cols) ? src[cidx * lds + r] : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int i = ty; i < 32; i += 8) {
        int64_t r = r0 + i, cidx = c0 + tx;
        if (r < rows && cidx < cols) dst[r * ldd + col0 + cidx] = tile[tx][i];
    }
}

__global__ void __launch_bounds__(256) rm_to_cm_kernel(const double *__restrict__ src, int64_t lds, int64_t rows,
                                                       int64_t cols, double *__restrict__ dst, int64_t ldd) {
    __shared__ double tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int64_t r0 = (int64_t)blockIdx.x * 32, c0 = (int64_t)blockIdx.y * 32;
#pragma unroll
    for (int i = ty; i < 32; i += 8) {
        int64_t r = r0 + i, cidx = c0 + tx;
        tile[i][tx] = (r < rows && cidx < cols) ? src[r * lds + cidx] : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int j = ty; j < 32; j += 8) {
        int64_t r = r0 + tx, cidx = c0 + j;
        if (r < rows && cidx < cols) dst[cidx * ldd + r] = tile[tx][j];
    }
}

int launch_cm_to_rm(plsb_ctx *c, const double *src, int64_t lds, int64_t rows, int64_t cols, double *dst,
                    int64_t ldd, int64_t col0) {
    if (rows == 0 || cols == 0) return PLSB_OK;
    dim3 grid((unsigned)ceil_div(rows, 32), (unsigned)ceil_div(cols, 32));
    cm_to_rm_kernel<<<grid, 256, 0, c->stream>>>(src, lds, rows, cols, dst, ldd, col0);
    c->launches++;
    PLSB_CUDA(cudaGetLastError());
    return PLSB_OK;
}
int launch_rm_to_cm(plsb_ctx *c, const double *src, int64_t lds, int64_t rows, int64_t cols, double *dst,
                    int64_t ldd) {
    if (rows == 0 || cols == 0) return PLSB_OK;
    dim3 grid((unsigned)ceil_div(rows, 32), (unsigned)ceil_div(cols, 32));
    rm_to_cm_kernel<<<grid, 256, 0, c->stream>>>(src, lds, rows, cols, dst, ldd);
    c->launches++;
    PLSB_CUDA(cudaGetLastError());
    return PLSB_OK;
}

// (the deterministic cross-CTA sum of the per-CTA partial rows lives in allreduce.cu: launch_reduce_partials)

// ------------------------------------------------------------------------------------------------
// the pass kernels
// ------------------------------------------------------------------------------------------------
struct PassArgs {
    const double *X;   // row-major N x ldk
    const double *Y;   // row-major N x ldm (xty only)
    const double *r;   // ldk (fused only)
    double *part;      // [grid][ldp]
    double *t_out;     // N or null (fused only)
    int64_t N, tiles;
    int ldk, ldm, ldp, stages;
    int m0, mc;        // xty: responses [m0, m0+mc)
    // fused pass: the tiles are dealt to `vgrid` VIRTUAL CTAs (virtual CTA p owns tiles p, p + vgrid, ... and partial
    // row p); a physical CTA runs virtual CTAs blockIdx.x, blockIdx.x + gridDim.x, ... one after the other.  The sums
    // therefore do not depend on how many CTAs share an SM (1 with a deep ring or 2 with a shallow one — tuned at run
    // time, see launch_fused_pass): results stay bit-identical across configurations.
    int vgrid;
};

// Thread layout: NT threads = G row groups x TPR threads per row.  A tile has ROWS = G*RB rows; group g
// owns rows [g*RB, g*RB+RB); thread c of a group owns columns {2*(c + TPR*v), +1 : v < VEC/2}.
template <int NT, int TPR, int VEC, int RB>
struct PassCfg {
    static constexpr int G = NT / TPR;
    static constexpr int ROWS = G * RB;
    static constexpr int NW = NT / 32;
    static constexpr int WPG = (TPR >= 32) ? TPR / 32 : 1;  // warps per row group
    static_assert(VEC % 2 == 0, "columns are owned in pairs (128-bit shared loads)");
    static_assert(TPR <= 32 ? (32 % TPR == 0) : (TPR % 32 == 0), "row group must tile warps");
};

__device__ __forceinline__ void issue_tile(const PassArgs &a, int64_t tile, int rows_per_tile, char *stage_x,
                                           char *stage_y, uint64_t *bar, bool with_y) {
    const int64_t row0 = tile * rows_per_tile;
    int64_t rows = a.N - row0;
    if (rows > rows_per_tile) rows = rows_per_tile;
    const uint32_t bx = (uint32_t)(rows * a.ldk * 8);
    const uint32_t by = with_y ? (uint32_t)(rows * a.ldm * 8) : 0u;
    mbar_arrive_expect_tx(bar, bx + by);
    bulk_g2s(stage_x, a.X + row0 * a.ldk, bx, bar);
    if (with_y) bulk_g2s(stage_y, a.Y + row0 * a.ldm, by, bar);
}

template <int NT, int TPR, int VEC, int RB>
__global__ void __launch_bounds__(NT) fused_pass_kernel(const PassArgs a) {
    using C = PassCfg<NT, TPR, VEC, RB>;
    extern __shared__ __align__(128) char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = tid / TPR, c = tid % TPR;
    const int tile_bytes = C::ROWS * a.ldk * 8;
    char *stage0 = smem;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)a.stages * tile_bytes);
    double *red = reinterpret_cast<double *>(full + 8);           // [2][NW][RB]
    double *gx = red;                                             // reused at the end: [G][ldk]

    if (tid == 0) {
        for (int s = 0; s < a.stages; ++s) mbar_init(&full[s], 1);
        mbar_fence_init();
    }
    __syncthreads();
    int j = 0;                 // tiles consumed so far by this CTA (all virtual CTAs): stage j % stages, parity from j / stages
    if (tid == 0) {
        for (int s = 0; s < a.stages; ++s) {
            int64_t tile = (int64_t)blockIdx.x + (int64_t)s * a.vgrid;
            if (tile < a.tiles) issue_tile(a, tile, C::ROWS, stage0 + (size_t)s * tile_bytes, nullptr, &full[s], false);
        }
    }

    // X does not change during a fit: the first tiles are already in flight while the kernel that produces r
    // (the step) is still running; everything below reads or overwrites data of the kernels before this one
    pdl_wait();
    pdl_launch();
    double rv[VEC], acc[VEC];
#pragma unroll
    for (int v = 0; v < VEC / 2; ++v) {
        const int col = 2 * (c + TPR * v);
        if (col < a.ldk) {
            rv[2 * v] = a.r[col];
            rv[2 * v + 1] = a.r[col + 1];
        } else {
            rv[2 * v] = 0.0;
            rv[2 * v + 1] = 0.0;
        }
    }

    for (int vb = blockIdx.x; vb < a.vgrid; vb += gridDim.x) {
    if (vb != (int)blockIdx.x) {
        __syncthreads();       // the previous virtual CTA is done with the ring and the reduction scratch
        if (tid == 0) {
            for (int s = 0; s < a.stages; ++s) {
                int64_t tile = (int64_t)vb + (int64_t)s * a.vgrid;
                const int slot = (j + s) % a.stages;
                if (tile < a.tiles) issue_tile(a, tile, C::ROWS, stage0 + (size_t)slot * tile_bytes, nullptr, &full[slot], false);
            }
        }
    }
#pragma unroll
    for (int v = 0; v < VEC; ++v) acc[v] = 0.0;
    double tt = 0.0;

    for (int64_t tile = vb; tile < a.tiles; tile += a.vgrid, ++j) {
        const int s = j % a.stages;
        const uint32_t parity = (uint32_t)((j / a.stages) & 1);
        const int64_t row0 = tile * C::ROWS + (int64_t)g * RB;
        mbar_wait(&full[s], parity);
        const double *xs = reinterpret_cast<const double *>(stage0 + (size_t)s * tile_bytes) + (size_t)g * RB * a.ldk;

        double x[RB][VEC], d[RB];
#pragma unroll
        for (int rb = 0; rb < RB; ++rb) {
            const bool live = row0 + rb < a.N;
            double s0 = 0.0, s1 = 0.0;
#pragma unroll
            for (int v = 0; v < VEC / 2; ++v) {
                const int col = 2 * (c + TPR * v);
                double2 xv = make_double2(0.0, 0.0);
                if (live && col < a.ldk) xv = *reinterpret_cast<const double2 *>(xs + (size_t)rb * a.ldk + col);
                x[rb][2 * v] = xv.x;
                x[rb][2 * v + 1] = xv.y;
                s0 = fma(xv.x, rv[2 * v], s0);
                s1 = fma(xv.y, rv[2 * v + 1], s1);
            }
            d[rb] = s0 + s1;
        }
        // reduce the RB partial dots across the TPR threads of the row group
        if (TPR <= 32) {
#pragma unroll
            for (int rb = 0; rb < RB; ++rb) d[rb] = group_sum<(TPR <= 32 ? TPR : 32)>(d[rb]);
            __syncthreads();  // every thread has its tile values in registers: the stage can be refilled
        } else {
            double *rbuf = red + (size_t)(j & 1) * C::NW * RB;
#pragma unroll
            for (int rb = 0; rb < RB; ++rb) d[rb] = warp_sum(d[rb]);
            if (lane == 0) {
#pragma unroll
                for (int rb = 0; rb < RB; ++rb) rbuf[warp * RB + rb] = d[rb];
            }
            __syncthreads();
#pragma unroll
            for (int rb = 0; rb < RB; ++rb) {
                double t = 0.0;
#pragma unroll
                for (int w = 0; w < C::WPG; ++w) t += rbuf[(g * C::WPG + w) * RB + rb];
                d[rb] = t;
            }
        }
        if (tid == 0) {
            const int64_t next = tile + (int64_t)a.stages * a.vgrid;
            if (next < a.tiles) issue_tile(a, next, C::ROWS, stage0 + (size_t)s * tile_bytes, nullptr, &full[s], false);
        }
        // g += x * t ; tt += t^2 ; T(:,a) = t
#pragma unroll
        for (int rb = 0; rb < RB; ++rb) {
            const double t = d[rb];  // zero for rows past N (their x are zero)
#pragma unroll
            for (int v = 0; v < VEC; ++v) acc[v] = fma(x[rb][v], t, acc[v]);
            tt = fma(t, t, tt);
        }
        if (c == 0 && a.t_out != nullptr) {
#pragma unroll
            for (int rb = 0; rb < RB; ++rb)
                if (row0 + rb < a.N) a.t_out[row0 + rb] = d[rb];
        }
    }

    // cross-group reduction inside the CTA, then one partial row per virtual CTA
    double *prow = a.part + (size_t)vb * a.ldp;
    if (C::G == 1) {
#pragma unroll
        for (int v = 0; v < VEC / 2; ++v) {
            const int col = 2 * (c + TPR * v);
            if (col < a.ldk) *reinterpret_cast<double2 *>(prow + col) = make_double2(acc[2 * v], acc[2 * v + 1]);
        }
        if (tid == 0) prow[a.ldk] = tt;
    } else {
        __syncthreads();  // red[] no longer in use
        double *tts = gx + (size_t)C::G * a.ldk;
#pragma unroll
        for (int v = 0; v < VEC / 2; ++v) {
            const int col = 2 * (c + TPR * v);
            if (col < a.ldk) {
                gx[(size_t)g * a.ldk + col] = acc[2 * v];
                gx[(size_t)g * a.ldk + col + 1] = acc[2 * v + 1];
            }
        }
        if (c == 0) tts[g] = tt;
        __syncthreads();
        for (int col = tid; col < a.ldk; col += NT) {
            double s = 0.0;
            for (int gg = 0; gg < C::G; ++gg) s += gx[(size_t)gg * a.ldk + col];
            prow[col] = s;
        }
        if (tid == 0) {
            double s = 0.0;
            for (int gg = 0; gg < C::G; ++gg) s += tts[gg];
            prow[a.ldk] = s;
        }
    }
    }   // virtual CTAs
}

// S partials: acc[m][v] += x[v] * y[m] for the mc (<= MC) responses starting at m0
template <int NT, int TPR, int VEC, int RB, int MC>
__global__ void __launch_bounds__(NT) xty_pass_kernel(const PassArgs a) {
    using C = PassCfg<NT, TPR, VEC, RB>;
    extern __shared__ __align__(128) char smem[];
    const int tid = threadIdx.x;
    const int g = tid / TPR, c = tid % TPR;
    const int tile_bytes = C::ROWS * a.ldk * 8;
    const int ytile_bytes = (C::ROWS * a.ldm * 8 + 127) / 128 * 128;
    const int stage_bytes = tile_bytes + ytile_bytes;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)a.stages * stage_bytes);
    double *gx = reinterpret_cast<double *>(full + 8);  // [G][ldk] for the final cross-group sum

    if (tid == 0) {
        for (int s = 0; s < a.stages; ++s) mbar_init(&full[s], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (tid == 0) {
        for (int s = 0; s < a.stages; ++s) {
            int64_t tile = (int64_t)blockIdx.x + (int64_t)s * gridDim.x;
            char *st = smem + (size_t)s * stage_bytes;
            if (tile < a.tiles) issue_tile(a, tile, C::ROWS, st, st + tile_bytes, &full[s], true);
        }
    }
    double acc[MC][VEC];
#pragma unroll
    for (int m = 0; m < MC; ++m)
#pragma unroll
        for (int v = 0; v < VEC; ++v) acc[m][v] = 0.0;

    int j = 0;
    for (int64_t tile = blockIdx.x; tile < a.tiles; tile += gridDim.x, ++j) {
        const int s = j % a.stages;
        const uint32_t parity = (uint32_t)((j / a.stages) & 1);
        const int64_t row0 = tile * C::ROWS + (int64_t)g * RB;
        mbar_wait(&full[s], parity);
        char *st = smem + (size_t)s * stage_bytes;
        const double *xs = reinterpret_cast<const double *>(st) + (size_t)g * RB * a.ldk;
        const double *ys = reinterpret_cast<const double *>(st + tile_bytes) + (size_t)g * RB * a.ldm + a.m0;
#pragma unroll
        for (int rb = 0; rb < RB; ++rb) {
            if (row0 + rb < a.N) {
                double y[MC];
#pragma unroll
                for (int m = 0; m < MC; ++m) y[m] = (m < a.mc) ? ys[(size_t)rb * a.ldm + m] : 0.0;
#pragma unroll
                for (int v = 0; v < VEC / 2; ++v) {
                    const int col = 2 * (c + TPR * v);
                    if (col < a.ldk) {
                        const double2 xv = *reinterpret_cast<const double2 *>(xs + (size_t)rb * a.ldk + col);
#pragma unroll
                        for (int m = 0; m < MC; ++m) {
                            acc[m][2 * v] = fma(xv.x, y[m], acc[m][2 * v]);
                            acc[m][2 * v + 1] = fma(xv.y, y[m], acc[m][2 * v + 1]);
                        }
                    }
                }
            }
        }
        __syncthreads();  // all reads of the stage are done
        if (tid == 0) {
            const int64_t next = tile + (int64_t)a.stages * gridDim.x;
            if (next < a.tiles) issue_tile(a, next, C::ROWS, st, st + tile_bytes, &full[s], true);
        }
    }
    // partial row layout: part[cta][m * ldk + col], m < mc
    double *prow = a.part + (size_t)blockIdx.x * a.ldp;
    for (int m = 0; m < MC; ++m) {
        if (m >= a.mc) break;
        if (C::G == 1) {
#pragma unroll
            for (int v = 0; v < VEC / 2; ++v) {
                const int col = 2 * (c + TPR * v);
                if (col < a.ldk) {
                    prow[(size_t)m * a.ldk + col] = acc[m][2 * v];
                    prow[(size_t)m * a.ldk + col + 1] = acc[m][2 * v + 1];
                }
            }
        } else {
            __syncthreads();
#pragma unroll
            for (int v = 0; v < VEC / 2; ++v) {
                const int col = 2 * (c + TPR * v);
                if (col < a.ldk) {
                    gx[(size_t)g * a.ldk + col] = acc[m][2 * v];
                    gx[(size_t)g * a.ldk + col + 1] = acc[m][2 * v + 1];
                }
            }
            __syncthreads();
            for (int col = tid; col < a.ldk; col += NT) {
                double s = 0.0;
                for (int gg = 0; gg < C::G; ++gg) s += gx[(size_t)gg * a.ldk + col];
                prow[(size_t)m * a.ldk + col] = s;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

// ------------------------------------------------------------------------------------------------
// Very wide matrices (K > 8192): a row no longer fits the shared-memory ring three times, so the component pass falls back
// to two plain streaming kernels — t = X r, then [g | tt] = [X't | t't] — and X'Y to the second one with Y's
// columns in the role of t.  Two reads of X per component (what the reference does, pls.cpp:419-421); every
// element is still read with coalesced 128-bit loads and summed without atomics.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) wide_xr_kernel(const double *__restrict__ X, int64_t N, int ldk,
                                                      const double *__restrict__ r, double *__restrict__ t) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t row = (int64_t)blockIdx.x * 8 + warp; row < N; row += (int64_t)gridDim.x * 8) {
        const double2 *x = reinterpret_cast<const double2 *>(X + row * ldk);
        const double2 *rr = reinterpret_cast<const double2 *>(r);
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        int k = lane;
        for (; k + 32 < ldk / 2; k += 64) {
            const double2 a = x[k], b = x[k + 32], ra = rr[k], rb = rr[k + 32];
            s0 = fma(a.x, ra.x, s0); s1 = fma(a.y, ra.y, s1);
            s2 = fma(b.x, rb.x, s2); s3 = fma(b.y, rb.y, s3);
        }
        for (; k < ldk / 2; k += 32) {
            const double2 a = x[k], ra = rr[k];
            s0 = fma(a.x, ra.x, s0); s1 = fma(a.y, ra.y, s1);
        }
        const double s = warp_sum((s0 + s1) + (s2 + s3));
        if (lane == 0) t[row] = s;
    }
}

// part[cta][v*ldk + k] = sum over the CTA's rows of X[i][k] * V[i*ldv + v0 + v], v < nv <= NV;
// with_tt: part[cta][ldk] = sum of V[i*ldv]^2 over the same rows (nv == 1)
template <int NV>
__global__ void __launch_bounds__(256) wide_xtv_kernel(const double *__restrict__ X, int64_t N, int ldk,
                                                       const double *__restrict__ V, int ldv, int v0, int nv,
                                                       double *__restrict__ part, int64_t ldp, int with_tt) {
    const int64_t rows_per = (N + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per, r1 = r0 + rows_per < N ? r0 + rows_per : N;
    double *prow = part + (size_t)blockIdx.x * ldp;
    for (int c0 = 2 * (int)threadIdx.x; c0 < ldk; c0 += 512) {
        double a0[NV], a1[NV];
#pragma unroll
        for (int v = 0; v < NV; ++v) { a0[v] = 0.0; a1[v] = 0.0; }
        int64_t i = r0;
        for (; i + 4 <= r1; i += 4) {      // four rows in flight per thread; the sums stay in ascending row order
            double2 x[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) x[u] = *reinterpret_cast<const double2 *>(X + (i + u) * ldk + c0);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
#pragma unroll
                for (int v = 0; v < NV; ++v) {
                    const double y = v < nv ? V[(i + u) * ldv + v0 + v] : 0.0;
                    a0[v] = fma(x[u].x, y, a0[v]);
                    a1[v] = fma(x[u].y, y, a1[v]);
                }
            }
        }
        for (; i < r1; ++i) {
            const double2 x = *reinterpret_cast<const double2 *>(X + i * ldk + c0);
#pragma unroll
            for (int v = 0; v < NV; ++v) {
                const double y = v < nv ? V[i * ldv + v0 + v] : 0.0;
                a0[v] = fma(x.x, y, a0[v]);
                a1[v] = fma(x.y, y, a1[v]);
            }
        }
#pragma unroll
        for (int v = 0; v < NV; ++v)
            if (v < nv) { prow[(size_t)v * ldk + c0] = a0[v]; prow[(size_t)v * ldk + c0 + 1] = a1[v]; }
    }
    if (with_tt) {
        __shared__ double red[8];
        double acc = 0.0;
        for (int64_t i = r0 + threadIdx.x; i < r1; i += 256) acc = fma(V[i * ldv], V[i * ldv], acc);
        acc = warp_sum(acc);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double tt = 0.0;
            for (int w = 0; w < 8; ++w) tt += red[w];
            prow[ldk] = tt;
        }
    }
}

static int run_wide_fused(plsb_ctx *c, const plsb_dataset *ds, const double *r, double *out, double *t_out, bool across_ranks) {
    const int ldk = (int)ds->ldk;
    double *t = t_out;
    if (!t) PLSB_TRY(dev_alloc(c, (void **)&t, (size_t)ds->N * sizeof(double)));
    const int64_t grid = (int64_t)c->sm_count * 8;
    const int64_t ldp = ldk + 2;
    PLSB_TRY(ensure_part(c, (size_t)grid * ldp * sizeof(double)));
    {
        ProfScope ps(c, PLSB_PROF_PASS_FUSED);
        wide_xr_kernel<<<(unsigned)grid, 256, 0, c->stream>>>(ds->X, ds->N, ldk, r, t);
        wide_xtv_kernel<1><<<(unsigned)grid, 256, 0, c->stream>>>(ds->X, ds->N, ldk, t, 1, 0, 1, c->part, ldp, 1);
    }
    PLSB_CUDA(cudaGetLastError());
    c->launches += 2;
    if (!t_out) dev_free(c, t);
    return launch_reduce_partials(c, c->part, ldp, (int)grid, ldk + 1, out, across_ranks);
}

static int run_wide_xty(plsb_ctx *c, const plsb_dataset *ds, double *S) {
    const int ldk = (int)ds->ldk;
    const int64_t grid = (int64_t)c->sm_count * 8;
    const int64_t ldp = (int64_t)8 * ldk;
    PLSB_TRY(ensure_part(c, (size_t)grid * ldp * sizeof(double)));
    for (int m0 = 0; m0 < ds->M; m0 += 8) {
        const int mc = (int)(ds->M - m0 < 8 ? ds->M - m0 : 8);
        {
            ProfScope ps(c, PLSB_PROF_PASS_XTY);
            wide_xtv_kernel<8><<<(unsigned)grid, 256, 0, c->stream>>>(ds->X, ds->N, ldk, ds->Y, (int)ds->ldm, m0, mc, c->part, ldp, 0);
        }
        PLSB_CUDA(cudaGetLastError());
        c->launches += 1;
        PLSB_TRY(launch_reduce_partials(c, c->part, ldp, (int)grid, (int64_t)mc * ldk, S + (size_t)m0 * ldk, false));
    }
    return PLSB_OK;
}

// One launch of the fused pass with `ctas_per_sm` resident CTAs per SM (0: the default for this shape).
template <int NT, int TPR, int VEC, int RB>
static int launch_fused_cfg(plsb_ctx *c, const plsb_dataset *ds, const double *r, double *t_out, int ctas_per_sm,
                            int *vgrid_out, int *ldp_out) {
    using C = PassCfg<NT, TPR, VEC, RB>;
    PassArgs a{};
    a.X = ds->X; a.r = r; a.t_out = t_out; a.N = ds->N; a.ldk = (int)ds->ldk; a.ldp = (int)ds->ldk + 2;
    a.tiles = ceil_div(ds->N, C::ROWS);
    const size_t tile_bytes = (size_t)C::ROWS * a.ldk * 8;
    const size_t tail = 64 + sizeof(double) * ((size_t)2 * C::NW * RB > (size_t)C::G * a.ldk + C::G
                                                   ? (size_t)2 * C::NW * RB : (size_t)C::G * a.ldk + C::G) + 128;
    const int vper = NT >= 512 ? 1 : 2;                       // virtual CTAs per SM: fixes the summation order
    if (ctas_per_sm <= 0) ctas_per_sm = env_int("PLSB_PASS_CTAS_PER_SM", vper);
    if (ctas_per_sm > vper) ctas_per_sm = vper;
    size_t budget = (c->smem_optin - 1024) / ctas_per_sm;
    if (budget > c->smem_optin) budget = c->smem_optin;
    int stages = budget > tail ? (int)((budget - tail) / tile_bytes) : 0;
    if (stages > env_int("PLSB_PASS_STAGES", 4)) stages = env_int("PLSB_PASS_STAGES", 4);
    if (stages > 8) stages = 8;
    if (stages < 1) {  // tile does not fit next to a second CTA: take the whole SM
        ctas_per_sm = 1;
        stages = c->smem_optin > tail ? (int)((c->smem_optin - tail) / tile_bytes) : 0;
        if (stages > 8) stages = 8;
        if (stages < 1) { set_error("fused pass: K=%lld does not fit shared memory", (long long)ds->K); return PLSB_ERR_UNSUPPORTED; }
    }
    a.stages = stages;
    const size_t smem = (size_t)stages * tile_bytes + tail;
    int64_t vgrid = (int64_t)c->sm_count * vper;
    if (vgrid > a.tiles) vgrid = a.tiles;
    if (vgrid < 1) vgrid = 1;
    a.vgrid = (int)vgrid;
    int64_t grid = (int64_t)c->sm_count * ctas_per_sm;
    if (grid > vgrid) grid = vgrid;
    PLSB_TRY(ensure_part(c, (size_t)vgrid * a.ldp * sizeof(double)));
    a.part = c->part;
    auto kern = fused_pass_kernel<NT, TPR, VEC, RB>;
    PLSB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    {
        ProfScope ps(c, PLSB_PROF_PASS_FUSED);
        PLSB_CUDA(launch_dep(kern, dim3((unsigned)grid), dim3(NT), smem, c->stream, a));
    }
    PLSB_CUDA(cudaGetLastError());
    c->launches += 1;
    *vgrid_out = (int)vgrid; *ldp_out = a.ldp;
    return PLSB_OK;
}

template <int NT, int TPR, int VEC, int RB>
static int run_fused(plsb_ctx *c, const plsb_dataset *ds, const double *r, double *out, double *t_out, bool across_ranks) {
    int vgrid = 0, ldp = 0;
    PLSB_TRY((launch_fused_cfg<NT, TPR, VEC, RB>(c, ds, r, t_out, 0, &vgrid, &ldp)));
    return launch_reduce_partials(c, c->part, ldp, vgrid, ds->ldk + 1, out, across_ranks);
}

int launch_fused_pass(plsb_ctx *c, const plsb_dataset *ds, const double *r, double *out, double *t_out, bool across_ranks) {
    const int64_t ldk = ds->ldk;
    if (ds->N == 0)  // empty shard: zero contribution, but it still takes part in the exchange
        return launch_reduce_partials(c, out, ldk + 2, 0, ldk + 1, out, across_ranks);
    const int rbsel = env_int("PLSB_PASS_RB", 0);
    if (ldk <= 64) return run_fused<256, 32, 2, 8>(c, ds, r, out, t_out, across_ranks);
    if (ldk <= 128) return run_fused<256, 32, 4, 4>(c, ds, r, out, t_out, across_ranks);
    if (ldk <= 256) return run_fused<256, 32, 8, 2>(c, ds, r, out, t_out, across_ranks);
    if (ldk <= 512) return run_fused<256, 256, 2, 8>(c, ds, r, out, t_out, across_ranks);
    if (ldk <= 1024) return run_fused<256, 256, 4, 4>(c, ds, r, out, t_out, across_ranks);
    if (ldk <= 2048) {
        // Streams of 2 GB and more: ONE 512-thread CTA per SM (two row groups of 2 rows each) behind a three-tile ring.
        // Smaller ones: two 256-thread CTAs per SM, 4 rows per thread, one tile each.  Measured under SUSTAINED load
        // (tools/devbench_shape.py, 3 s per cell, board at its 1000 W cap, profiles/r3p_shape.log), ms per 32-component fit,
        // one-CTA vs two-CTA: 2 GB 11.14 / 11.24, 4 GB 20.2 / 21.8, 8 GB 39.9 / 42.8, 16 GB 79.6 / 84.8 — the single CTA
        // re-arms its ring as soon as a tile is in registers, 16 warps share the arithmetic, and it draws less power per
        // byte, which is what counts once the cap sets the clock.  (Round 2 put the threshold at 6.4 GB after short runs
        // in which 2-4 GB shards had not yet reached the cap; the driver's 4-GPU run then lost 6 % on it.)
        const bool big = ds->N * ldk >= ((int64_t)240 << 20) && !getenv("PLSB_PASS_CTAS_PER_SM");
        if (rbsel == 2) return run_fused<256, 256, 8, 2>(c, ds, r, out, t_out, across_ranks);
        if (rbsel == 12 || (rbsel == 0 && big)) return run_fused<512, 256, 8, 2>(c, ds, r, out, t_out, across_ranks);
        return run_fused<256, 256, 8, 4>(c, ds, r, out, t_out, across_ranks);
    }
    if (ldk <= 4096) {
        if (rbsel == 4) return run_fused<512, 512, 8, 4>(c, ds, r, out, t_out, across_ranks);
        return run_fused<512, 512, 8, 2>(c, ds, r, out, t_out, across_ranks);
    }
    // 4096 < K <= 8192: one 64 KB row per tile, 16 columns per thread (x, the accumulators and r: 48 doubles = 96 of the
    // 128 registers a 512-thread CTA may use), three rows in flight per SM — still ONE read of X per component.  The
    // two-kernel fallback below (two reads) ran these shapes at 0.46 of this bound.  PLSB_PASS_WIDE_V1=1: A/B.
    if (ldk <= 8192 && !getenv("PLSB_PASS_WIDE_V1")) return run_fused<512, 512, 16, 1>(c, ds, r, out, t_out, across_ranks);
    return run_wide_fused(c, ds, r, out, t_out, across_ranks);
}

template <int NT, int TPR, int VEC, int RB, int MC>
static int run_xty(plsb_ctx *c, const plsb_dataset *ds, double *S) {
    using C = PassCfg<NT, TPR, VEC, RB>;
    PassArgs a{};
    a.X = ds->X; a.Y = ds->Y; a.N = ds->N; a.ldk = (int)ds->ldk; a.ldm = (int)ds->ldm;
    a.ldp = MC * (int)ds->ldk;
    a.tiles = ceil_div(ds->N, C::ROWS);
    const size_t tile_bytes = (size_t)C::ROWS * a.ldk * 8;
    const size_t ytile_bytes = ((size_t)C::ROWS * a.ldm * 8 + 127) / 128 * 128;
    const size_t stage_bytes = tile_bytes + ytile_bytes;
    const size_t tail = 64 + sizeof(double) * (size_t)C::G * a.ldk + 128;
    int ctas_per_sm = NT >= 512 ? 1 : 2;
    size_t budget = (c->smem_optin - 1024) / ctas_per_sm;
    int stages = budget > tail ? (int)((budget - tail) / stage_bytes) : 0;
    if (stages > 4) stages = 4;
    if (stages < 1) {
        ctas_per_sm = 1;
        stages = c->smem_optin > tail ? (int)((c->smem_optin - tail) / stage_bytes) : 0;
        if (stages > 8) stages = 8;
        if (stages < 1) { set_error("xty pass: K=%lld, M=%lld does not fit shared memory", (long long)ds->K, (long long)ds->M); return PLSB_ERR_UNSUPPORTED; }
    }
    a.stages = stages;
    const size_t smem = (size_t)stages * stage_bytes + tail;
    int64_t grid = (int64_t)c->sm_count * ctas_per_sm;
    if (grid > a.tiles) grid = a.tiles;
    if (grid < 1) grid = 1;
    PLSB_TRY(ensure_part(c, (size_t)grid * a.ldp * sizeof(double)));
    a.part = c->part;
    auto kern = xty_pass_kernel<NT, TPR, VEC, RB, MC>;
    PLSB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int m0 = 0; m0 < ds->M; m0 += MC) {
        a.m0 = m0;
        a.mc = (int)(ds->M - m0 < MC ? ds->M - m0 : MC);
        {
            ProfScope ps(c, PLSB_PROF_PASS_XTY);
            kern<<<(unsigned)grid, NT, smem, c->stream>>>(a);
        }
        PLSB_CUDA(cudaGetLastError());
        const int64_t len = (int64_t)a.mc * a.ldk;
        c->launches += 1;
        PLSB_TRY(launch_reduce_partials(c, c->part, a.ldp, (int)grid, len, S + (size_t)m0 * a.ldk, false));
    }
    return PLSB_OK;
}

int launch_xty_pass(plsb_ctx *c, const plsb_dataset *ds, double *S) {
    const int64_t ldk = ds->ldk;
    if (ds->N == 0) {
        PLSB_CUDA(cudaMemsetAsync(S, 0, (size_t)ds->M * ldk * sizeof(double), c->stream));
        return PLSB_OK;
    }
    // FP64 tensor-core path (xty.cu) for rows of 2-16 KB; PLSB_XTY_FMA=1 keeps the FMA kernels for A/B runs
    if (xty_dmma_supported(ds) && !env_int("PLSB_XTY_FMA", 0)) return launch_xty_dmma(c, ds, S);
    if (ldk <= 64) return run_xty<256, 32, 2, 8, 8>(c, ds, S);
    if (ldk <= 128) return run_xty<256, 32, 4, 4, 8>(c, ds, S);
    if (ldk <= 256) return run_xty<256, 32, 8, 2, 8>(c, ds, S);
    if (ldk <= 512) return run_xty<256, 256, 2, 8, 8>(c, ds, S);
    if (ldk <= 1024) return run_xty<256, 256, 4, 4, 8>(c, ds, S);
    if (ldk <= 2048) return run_xty<256, 256, 8, 2, 8>(c, ds, S);
    if (ldk <= 4096) return run_xty<512, 512, 8, 2, 4>(c, ds, S);
    return run_wide_xty(c, ds, S);
}

}  // namespace plsb
