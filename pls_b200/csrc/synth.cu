// pls_b200/csrc/synth.cu — deterministic synthetic latent-factor data, generated in HBM.
//
// Bit-identical to oracle/pls_oracle.c:orc_synth: integer hash -> Irwin-Hall(3) on
// 20-bit integers, X = (Z C) Bx + 0.5 Ex, Y = (Z C) By + 0.1 Ey with the latent sums accumulated in
// index order WITHOUT fused multiply-add (explicit __dmul_rn / __dadd_rn), so host and device agree
// to the last bit and parity can be checked on any row block of a data set that never leaves HBM.
#include "internal.h"

namespace plsb {

__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
__device__ __forceinline__ double synth_u(uint64_t seed_mixed, uint64_t idx) {
    const uint64_t h = splitmix64(seed_mixed + idx);
    const uint64_t s = (h & 0xFFFFF) + ((h >> 20) & 0xFFFFF) + ((h >> 40) & 0xFFFFF);
    return __dsub_rn(__ddiv_rn((double)s, 1048576.0), 1.5);
}

__global__ void synth_tables_kernel(uint64_t seed, int64_t K, int64_t M, int64_t L, double *Bx, double *By) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < L * K) Bx[i] = synth_u(splitmix64(seed + 1), (uint64_t)i);
    if (i < L * M) By[i] = synth_u(splitmix64(seed + 2), (uint64_t)i);
}

constexpr int SYN_ROWS = 32;
constexpr int SYN_MAXL = 32;

__global__ void __launch_bounds__(256) synth_rows_kernel(uint64_t seed, int64_t row0, int64_t N, int64_t K, int64_t M,
                                                         int L, const double *__restrict__ Bx,
                                                         const double *__restrict__ By, double *__restrict__ X,
                                                         int64_t ldk, double *__restrict__ Y, int64_t ldm) {
    __shared__ double z[SYN_ROWS][SYN_MAXL];
    const int tid = threadIdx.x;
    const int64_t base = (int64_t)blockIdx.x * SYN_ROWS;
    const uint64_t sz = splitmix64(seed), se = splitmix64(seed + 3), sy = splitmix64(seed + 4);
    for (int e = tid; e < SYN_ROWS * L; e += 256) {
        const int rr = e / L, l = e % L;
        const uint64_t i = (uint64_t)(row0 + base + rr);
        z[rr][l] = __dmul_rn(synth_u(sz, i * (uint64_t)L + (uint64_t)l), __ddiv_rn(4.0, __dadd_rn(4.0, (double)l)));
    }
    __syncthreads();
    const int nrows = (int)((N - base) < SYN_ROWS ? (N - base) : SYN_ROWS);
    for (int64_t j = tid; j < K; j += 256) {
        double bx[SYN_MAXL];
#pragma unroll
        for (int l = 0; l < SYN_MAXL; ++l) bx[l] = (l < L) ? Bx[(int64_t)l * K + j] : 0.0;
        for (int rr = 0; rr < nrows; ++rr) {
            const uint64_t i = (uint64_t)(row0 + base + rr);
            double acc = 0.0;
#pragma unroll
            for (int l = 0; l < SYN_MAXL; ++l)
                if (l < L) acc = __dadd_rn(acc, __dmul_rn(z[rr][l], bx[l]));
            X[(base + rr) * ldk + j] = __dadd_rn(acc, __dmul_rn(0.5, synth_u(se, i * (uint64_t)K + (uint64_t)j)));
        }
    }
    for (int e = tid; e < nrows * (int)M; e += 256) {
        const int rr = e / (int)M, m = e % (int)M;
        const uint64_t i = (uint64_t)(row0 + base + rr);
        double acc = 0.0;
        for (int l = 0; l < L; ++l) acc = __dadd_rn(acc, __dmul_rn(z[rr][l], By[(int64_t)l * M + m]));
        Y[(base + rr) * ldm + m] = __dadd_rn(acc, __dmul_rn(0.1, synth_u(sy, i * (uint64_t)M + (uint64_t)m)));
    }
}

int launch_synth(plsb_ctx *c, uint64_t seed, int64_t row0, int64_t N, int64_t K, int64_t M, int64_t L, double *X,
                 int64_t ldk, double *Y, int64_t ldm) {
    PLSB_ARG(L >= 1 && L <= SYN_MAXL, "synth: L must be in [1, %d]", SYN_MAXL);
    if (N == 0) return PLSB_OK;
    double *Bx = nullptr, *By = nullptr;
    PLSB_CUDA(cudaMallocAsync(&Bx, sizeof(double) * L * K, c->stream));
    PLSB_CUDA(cudaMallocAsync(&By, sizeof(double) * L * M, c->stream));
    const int64_t nt = L * (K > M ? K : M);
    synth_tables_kernel<<<(unsigned)ceil_div(nt, 256), 256, 0, c->stream>>>(seed, K, M, L, Bx, By);
    PLSB_CUDA(cudaGetLastError());
    synth_rows_kernel<<<(unsigned)ceil_div(N, SYN_ROWS), 256, 0, c->stream>>>(seed, row0, N, K, M, (int)L, Bx, By, X,
                                                                                ldk, Y, ldm);
    PLSB_CUDA(cudaGetLastError());
    c->launches += 2;
    PLSB_CUDA(cudaFreeAsync(Bx, c->stream));
    PLSB_CUDA(cudaFreeAsync(By, c->stream));
    return PLSB_OK;
}

}  // namespace plsb
