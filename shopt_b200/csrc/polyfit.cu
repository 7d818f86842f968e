// polyfit.cu -- per-pixel nth-degree (u,v) polynomial least squares over the training stars, the per-star
// reconstruction error, and the consumer-side evaluation p(u,v).
//
// Replaces, in the reference:
//   polynomial_optimizer(degree, u, v, z)            shopt/interpolate.jl:105-147   (called r*c times, shopt.jl:543-549)
//   objective_function / compute_mse                 shopt/interpolate.jl:89-102, shopt/shopt.jl:496-511, 551-555
//   p(u, v, polMatrix, degree)                        shopt/reader.jl:40-68
//
// The reference solves  A \ z  per pixel with the SAME design matrix A (n_train x K) for every pixel.  Here the
// normal equations are formed once:  G = V'V (K x K)  and  B = V'P (K x r*c)  in ONE streaming pass over the
// pixel-grid fits (the only O(N r c) traffic), then one Cholesky of G is shared by all pixels.
//
// Kernels (all FP64, all bandwidth-bound: K FMAs per 8-byte pixel loaded):
//   normal_kernel  grid (pixel tiles, star chunks): each thread owns PX pixels, keeps K*PX accumulators in
//                  registers and streams its chunk of stars with coalesced 8-byte loads (r*c is odd for the
//                  configs of interest, so 16-byte vectors would straddle stars); the basis values u^p v^q of the
//                  chunk sit in shared memory and are read as broadcasts.  Partials go to [chunk][K][r*c];
//   reduce_kernel  sums the chunk partials in chunk order (deterministic; no floating-point atomics anywhere);
//   solve_kernel   equilibrated Cholesky of G in shared memory, then one thread per pixel back-substitutes;
//   mse_kernel     same tiling as normal_kernel with the K*PX coefficients of the thread's pixels in registers;
//                  per-star partial sums per pixel tile, reduced in tile order by mse_reduce_kernel;
//   normal_ring_kernel, mse_ring_kernel   THE DEFAULT: the same two kernels (same arithmetic, same order, bit-identical
//                  results) with each lane prefetching its own pixels of the next stars into a private ring in shared
//                  memory by 8-byte cp.async, so 256 bytes per thread are in flight whatever the register budget:
//                  92 % / 84 % of the measured copy bandwidth at K = 10 against 61 % / 53 % with plain loads.
#include <algorithm>
#include <cstdlib>

#include "api_ctx.h"
#include "polyfit_factor.h"

namespace shopt {

__host__ __device__ constexpr int num_terms(int degree) { return (degree + 1) * (degree + 2) / 2; }

// basis[k] = u^p v^q in the reference's term order: p = 0..deg outer, q = 0..deg-p inner (interpolate.jl:106-114)
template <int DEG>
__device__ __forceinline__ void basis_eval(double u, double v, double *b) {
    double up = 1.0;
    int k = 0;
#pragma unroll
    for (int p = 0; p <= DEG; p++) {
        double vq = 1.0;
#pragma unroll
        for (int q = 0; q <= DEG - p; q++) {
            b[k++] = up * vq;
            vq *= v;
        }
        up *= u;
    }
}

// centred and scaled coordinates (polyfit_factor.h): ctr = {u0, 1/su, v0, 1/sv} on the device, or null = raw
__device__ __forceinline__ void centre_uv(const double *__restrict__ ctr, double &u, double &v) {
    if (ctr) { u = (u - ctr[0]) * ctr[1]; v = (v - ctr[2]) * ctr[3]; }
}

constexpr int POLY_THREADS = 128;
constexpr int POLY_CFG_DEFAULT = 2;   // 4 pixels per thread (K <= 10), star loop unrolled by 4: normal_kernel 2.98 -> 2.71 ms
constexpr int SCHUNK = 64;   // stars staged per shared-memory refill

template <int DEG> struct PolyCfg {
    static constexpr int K = num_terms(DEG);
    static constexpr int PX = (K <= 10) ? 4 : (K <= 21 ? 2 : 1);
    static constexpr int TILE = POLY_THREADS * PX;
};

// ---------------------------------------------------------------------------------------------
// G = V'V and B = V'P partials
// part layout: [chunk][K*K + K*rc]
// ---------------------------------------------------------------------------------------------
template <int DEG, int PX, int UNR>
__global__ void __launch_bounds__(POLY_THREADS) normal_kernel(const double *__restrict__ pgf, long long rc,
                                                              const double *__restrict__ u, const double *__restrict__ v,
                                                              const long long *__restrict__ idx, long long n_bound,
                                                              const long long *__restrict__ n_ptr,
                                                              const double *__restrict__ ctr, double *__restrict__ part) {
    constexpr int K = PolyCfg<DEG>::K;
    __shared__ double sV[SCHUNK][K];
    __shared__ long long sOff[SCHUNK];
    const long long n = n_ptr ? *n_ptr : n_bound;
    const long long per_chunk = (n + gridDim.y - 1) / gridDim.y;
    const long long t0 = (long long)blockIdx.y * per_chunk;
    const long long t1 = (t0 + per_chunk < n) ? t0 + per_chunk : n;
    const long long px0 = (long long)blockIdx.x * (POLY_THREADS * PX) + threadIdx.x;
    double acc[K][PX];
#pragma unroll
    for (int k = 0; k < K; k++)
#pragma unroll
        for (int j = 0; j < PX; j++) acc[k][j] = 0.0;
    for (long long base = t0; base < t1; base += SCHUNK) {
        const int m = (int)((t1 - base < SCHUNK) ? (t1 - base) : SCHUNK);
        __syncthreads();
        if (threadIdx.x < m) {
            const long long star = idx ? idx[base + threadIdx.x] : base + threadIdx.x;
            double b[K], cu = u[star], cv = v[star];
            centre_uv(ctr, cu, cv);
            basis_eval<DEG>(cu, cv, b);
#pragma unroll
            for (int k = 0; k < K; k++) sV[threadIdx.x][k] = b[k];
            sOff[threadIdx.x] = star * rc;
        }
        __syncthreads();
#pragma unroll UNR
        for (int s = 0; s < m; s++) {
            const double *row = pgf + sOff[s];
            double z[PX];
#pragma unroll
            for (int j = 0; j < PX; j++) {
                const long long px = px0 + (long long)j * POLY_THREADS;
                z[j] = (px < rc) ? __ldg(row + px) : 0.0;
            }
#pragma unroll
            for (int k = 0; k < K; k++) {
                const double vk = sV[s][k];
#pragma unroll
                for (int j = 0; j < PX; j++) acc[k][j] = fma(vk, z[j], acc[k][j]);
            }
        }
    }
    double *out = part + (size_t)blockIdx.y * ((size_t)K * K + (size_t)K * rc);
#pragma unroll
    for (int k = 0; k < K; k++)
#pragma unroll
        for (int j = 0; j < PX; j++) {
            const long long px = px0 + (long long)j * POLY_THREADS;
            if (px < rc) out[(size_t)K * K + (size_t)k * rc + px] = acc[k][j];
        }
}

// ---------------------------------------------------------------------------------------------
// The same V'P partials with every lane prefetching ITS OWN pixels RING_D - 1 stars ahead into a private slot ring in
// shared memory (8-byte cp.async: no alignment demands on the odd-length star rows, no registers held by loads in
// flight, and -- each lane reading back only what it copied -- no barrier besides cp.async.wait_group).  ncu on
// normal_kernel: 72 % of the warp samples wait on the long scoreboard with 16 loads per thread in flight; here
// RING_D - 1 stars x PX loads per thread are in flight all the time, across the 64-star basis refills as well.
// Same operations in the same order as normal_kernel: the partials are bit-identical.
// ---------------------------------------------------------------------------------------------
// resident CTAs per SM the ring kernels are compiled for, by accumulators per thread (K * PX doubles)
__host__ __device__ constexpr int ring_ctas(int acc) { return acc <= 42 ? 4 : (acc <= 60 ? 3 : 2); }
constexpr int RING_BYTES_PER_THREAD = 256;   // in flight per thread: RING_D = 32 / PX slots of PX doubles
#ifndef RING_U
#define RING_U 1
#endif
constexpr int RING_UNR = RING_U;
template <int DEG, int PX>
__global__ void __launch_bounds__(POLY_THREADS, ring_ctas(PolyCfg<DEG>::K * PX)) normal_ring_kernel(const double *__restrict__ pgf, long long rc,
                                                                      const double *__restrict__ u, const double *__restrict__ v,
                                                                      const long long *__restrict__ idx, long long n_bound,
                                                                      const long long *__restrict__ n_ptr,
                                                                      const double *__restrict__ ctr, double *__restrict__ part) {
    constexpr int K = PolyCfg<DEG>::K;
    constexpr int RING_D = (K <= 21) ? RING_BYTES_PER_THREAD / 8 / PX : 16;   // static shared memory stays under 48 KB
    static_assert(POLY_THREADS >= 2 * SCHUNK && (SCHUNK & (SCHUNK - 1)) == 0 && RING_D <= SCHUNK, "staging layout");
    __shared__ __align__(16) double sV[SCHUNK][K + (K & 1)];
    __shared__ long long sOff[2 * SCHUNK];                       // row offsets of this and of the next 64 stars
    __shared__ double ring[RING_D][PX][POLY_THREADS];            // [slot][j][thread]: conflict-free 8-byte rows
    const long long n = n_ptr ? *n_ptr : n_bound;
    const long long per_chunk = (n + gridDim.y - 1) / gridDim.y;
    const long long t0 = (long long)blockIdx.y * per_chunk;
    const long long t1 = (t0 + per_chunk < n) ? t0 + per_chunk : n;
    const int nloc = (t1 > t0) ? (int)(t1 - t0) : 0;             // stars of this chunk, addressed by local index below
    const long long px0 = (long long)blockIdx.x * (POLY_THREADS * PX) + threadIdx.x;
    int nlive = 0;                                               // this thread's pixels j < nlive exist (a prefix: px grows with j)
#pragma unroll
    for (int j = 0; j < PX; j++) nlive += (px0 + (long long)j * POLY_THREADS < rc) ? 1 : 0;
    double acc[K][PX];
#pragma unroll
    for (int k = 0; k < K; k++)
#pragma unroll
        for (int j = 0; j < PX; j++) acc[k][j] = 0.0;
    const double *mine = pgf + px0;
    double *my_slot = &ring[0][0][threadIdx.x];
    auto stage_off = [&](int i) {                                // local star i -> offset of its row
        if (i < nloc) sOff[i & (2 * SCHUNK - 1)] = (idx ? idx[t0 + i] : t0 + i) * rc;
    };
    auto issue = [&](int i) {                                    // local star i -> slot i mod RING_D; one group per star
        if (i < nloc) {
            const double *row = mine + sOff[i & (2 * SCHUNK - 1)];
            double *dst = my_slot + (i & (RING_D - 1)) * (PX * POLY_THREADS);
#pragma unroll
            for (int j = 0; j < PX; j++)
                if (j < nlive) cp_async8(dst + j * POLY_THREADS, row + (long long)j * POLY_THREADS);
        }
        cp_async_commit();
    };
    if (threadIdx.x < SCHUNK) stage_off(threadIdx.x);
    __syncthreads();
    for (int d = 0; d < RING_D - 1; d++) issue(d);
    for (int base = 0; base < nloc; base += SCHUNK) {
        const int m = (nloc - base < SCHUNK) ? (nloc - base) : SCHUNK;
        __syncthreads();
        if (threadIdx.x < m) {
            const long long star = idx ? idx[t0 + base + threadIdx.x] : t0 + base + threadIdx.x;
            double b[K], cu = u[star], cv = v[star];
            centre_uv(ctr, cu, cv);
            basis_eval<DEG>(cu, cv, b);
#pragma unroll
            for (int k = 0; k < K; k++) sV[threadIdx.x][k] = b[k];
        } else if (threadIdx.x >= SCHUNK && threadIdx.x < 2 * SCHUNK) {
            stage_off(base + threadIdx.x);                       // the NEXT 64 stars: the prefetch runs across the refill
        }
        __syncthreads();
#pragma unroll RING_UNR
        for (int s = 0; s < m; s++) {
            const int i = base + s;
            issue(i + RING_D - 1);                               // into the slot star i - 1 was read from
            cp_async_wait<RING_D - 1>();                         // all but the newest RING_D - 1 groups: star i has landed
            const double *src = my_slot + (i & (RING_D - 1)) * (PX * POLY_THREADS);
            double z[PX];
#pragma unroll
            for (int j = 0; j < PX; j++) z[j] = (j < nlive) ? src[j * POLY_THREADS] : 0.0;
#pragma unroll
            for (int k = 0; k < K; k++) {
                const double vk = sV[s][k];
#pragma unroll
                for (int j = 0; j < PX; j++) acc[k][j] = fma(vk, z[j], acc[k][j]);
            }
        }
    }
    cp_async_wait<0>();
    double *out = part + (size_t)blockIdx.y * ((size_t)K * K + (size_t)K * rc);
#pragma unroll
    for (int k = 0; k < K; k++)
#pragma unroll
        for (int j = 0; j < PX; j++) {
            const long long px = px0 + (long long)j * POLY_THREADS;
            if (px < rc) out[(size_t)K * K + (size_t)k * rc + px] = acc[k][j];
        }
}

// ---------------------------------------------------------------------------------------------
// The same V'P partials with the pixel-grid fits staged through shared memory by 1-D bulk (TMA) copies.
// normal_kernel keeps at most UNR x PX 8-byte loads per thread in flight and runs at 3 CTAs per SM (140 registers):
// 61 % of the measured copy bandwidth.  Here a CTA owns a tile of NB_TP pixels and walks its chunk of stars in stages
// of NB_S stars: one bulk copy per star row (NB_TP x 8 bytes) into a ring of NB_ST stages, completion on an mbarrier,
// so 2 x (NB_ST - 1) x NB_S x 4 KB are in flight per SM whatever the register budget, and the FMA loop reads its
// operands from shared memory (LDS.64 along the pixels, broadcast LDS of the basis values).
// r c is odd at the BASELINE sizes, so a star's tile starts on an odd multiple of 8 bytes every other star: the copy
// takes the enclosing 16-byte-aligned range (one leading double, recorded per row) and stops at the last even boundary
// inside the tile; a trailing odd element is fetched by a plain load.  No copy ever leaves the star's own pixels.
// Same operations in the same order as normal_kernel: the partials are bit-identical.
// ---------------------------------------------------------------------------------------------
constexpr int NB_TP = 512, NB_S = 8, NB_ST = 3, NB_PITCH = NB_TP + 2;
template <int DEG>
__global__ void __launch_bounds__(POLY_THREADS) normal_bulk_kernel(const double *__restrict__ pgf, long long rc,
                                                                   const double *__restrict__ u, const double *__restrict__ v,
                                                                   const long long *__restrict__ idx, long long n_bound,
                                                                   const long long *__restrict__ n_ptr,
                                                                   const double *__restrict__ ctr, double *__restrict__ part) {
    constexpr int K = PolyCfg<DEG>::K, PX = NB_TP / POLY_THREADS;
    extern __shared__ __align__(16) unsigned char nb_smem[];
    double *ring = reinterpret_cast<double *>(nb_smem);                        // [NB_ST][NB_S][NB_PITCH]
    double *sV = ring + (size_t)NB_ST * NB_S * NB_PITCH;                         // [2][NB_S][K]
    uint64_t *bars = reinterpret_cast<uint64_t *>(sV + 2 * NB_S * K);            // [NB_ST]
    int *sM = reinterpret_cast<int *>(bars + NB_ST);                             // [NB_ST][NB_S] leading doubles of each row
    const int tid = threadIdx.x, lane = tid & 31;
    const long long n = n_ptr ? *n_ptr : n_bound;
    const long long per_chunk = (n + gridDim.y - 1) / gridDim.y;
    const long long t0 = (long long)blockIdx.y * per_chunk;
    const long long t1 = (t0 + per_chunk < n) ? t0 + per_chunk : n;
    const long long px0 = (long long)blockIdx.x * NB_TP;
    const int cnt = (int)((rc - px0 < NB_TP) ? (rc - px0) : NB_TP);             // pixels of this tile
    const long long nstage = (t1 > t0) ? (t1 - t0 + NB_S - 1) / NB_S : 0;
    if (tid == 0) {
        for (int i = 0; i < NB_ST; i++) mbar_init(&bars[i], 1);
        mbar_fence_init();
    }
    __syncthreads();
    // warp 0 issues stage `stg` (rows = stars t0 + stg NB_S ...): lane r < rows copies row r
    auto issue = [&](long long stg) {
        const int slot = (int)(stg % NB_ST);
        const long long base = t0 + stg * NB_S;
        const int rows = (int)((t1 - base < NB_S) ? (t1 - base) : NB_S);
        uint32_t bytes = 0;
        const double *src = nullptr;
        double *dst = ring + ((size_t)slot * NB_S + lane) * NB_PITCH;
        long long off = 0;
        int m = 0, tail = 0;
        if (lane < rows) {
            const long long star = idx ? idx[base + lane] : base + lane;
            off = star * rc + px0;
            m = (int)(((reinterpret_cast<uintptr_t>(pgf) >> 3) + (unsigned long long)off) & 1ULL);   // doubles before the tile
            const long long start = off - m, end = off + cnt;                   // needed: [off, end)
            const long long cend = end - (((reinterpret_cast<uintptr_t>(pgf) >> 3) + (unsigned long long)end) & 1ULL);
            tail = (cend < end);
            bytes = (uint32_t)((cend - start) * 8);
            src = pgf + start;
            sM[slot * NB_S + lane] = m;
        }
        uint32_t total = bytes;
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
        if (lane == 0) mbar_arrive_expect_tx(&bars[slot], total);
        __syncwarp();
        if (lane < rows) {
            if (bytes) bulk_g2s(dst, src, bytes, &bars[slot]);
            if (tail) dst[m + cnt - 1] = __ldg(pgf + off + cnt - 1);            // visible after the next __syncthreads
        }
    };
    auto stage_basis = [&](long long stg) {
        const long long base = t0 + stg * NB_S;
        const int rows = (int)((t1 - base < NB_S) ? (t1 - base) : NB_S);
        if (tid < rows) {
            const long long star = idx ? idx[base + tid] : base + tid;
            double b[K], cu = u[star], cv = v[star];
            centre_uv(ctr, cu, cv);
            basis_eval<DEG>(cu, cv, b);
            double *o = sV + ((size_t)(stg & 1) * NB_S + tid) * K;
#pragma unroll
            for (int k = 0; k < K; k++) o[k] = b[k];
        }
    };
    double acc[K][PX];
#pragma unroll
    for (int k = 0; k < K; k++)
#pragma unroll
        for (int j = 0; j < PX; j++) acc[k][j] = 0.0;
    if (tid < 32) for (long long g = 0; g < NB_ST - 1 && g < nstage; g++) issue(g);
    if (nstage > 0) stage_basis(0);
    for (long long it = 0; it < nstage; it++) {
        __syncthreads();                                         // stage it-1 consumed by everyone: its slot is free
        if (tid < 32 && it + NB_ST - 1 < nstage) {
            fence_proxy_async();                                 // generic reads of that slot before the async writes
            issue(it + NB_ST - 1);
        }
        if (it + 1 < nstage) stage_basis(it + 1);
        const int slot = (int)(it % NB_ST);
        mbar_wait(&bars[slot], (uint32_t)((it / NB_ST) & 1));
        const long long base = t0 + it * NB_S;
        const int rows = (int)((t1 - base < NB_S) ? (t1 - base) : NB_S);
        const double *bv = sV + (size_t)(it & 1) * NB_S * K;
#pragma unroll 2
        for (int s = 0; s < rows; s++) {
            const double *row = ring + ((size_t)slot * NB_S + s) * NB_PITCH + sM[slot * NB_S + s];
            double z[PX];
#pragma unroll
            for (int j = 0; j < PX; j++) {
                const int p = tid + j * POLY_THREADS;
                z[j] = (p < cnt) ? row[p] : 0.0;
            }
#pragma unroll
            for (int k = 0; k < K; k++) {
                const double vk = bv[s * K + k];
#pragma unroll
                for (int j = 0; j < PX; j++) acc[k][j] = fma(vk, z[j], acc[k][j]);
            }
        }
    }
    double *out = part + (size_t)blockIdx.y * ((size_t)K * K + (size_t)K * rc);
#pragma unroll
    for (int k = 0; k < K; k++)
#pragma unroll
        for (int j = 0; j < PX; j++) {
            const long long px = px0 + tid + (long long)j * POLY_THREADS;
            if (px < rc) out[(size_t)K * K + (size_t)k * rc + px] = acc[k][j];
        }
}

// G = V'V partials: one block per chunk, the K*K entries strided over the threads
template <int DEG>
__global__ void __launch_bounds__(POLY_THREADS) gram_kernel(const double *__restrict__ u, const double *__restrict__ v,
                                                            const long long *__restrict__ idx, long long n_bound,
                                                            const long long *__restrict__ n_ptr, const double *__restrict__ ctr,
                                                            long long rc, double *__restrict__ part) {
    constexpr int K = PolyCfg<DEG>::K;
    __shared__ double sV[SCHUNK][K];
    const long long n = n_ptr ? *n_ptr : n_bound;
    const long long per_chunk = (n + gridDim.x - 1) / gridDim.x;
    const long long t0 = (long long)blockIdx.x * per_chunk;
    const long long t1 = (t0 + per_chunk < n) ? t0 + per_chunk : n;
    constexpr int NE = (K * K + POLY_THREADS - 1) / POLY_THREADS;
    double g[NE];
#pragma unroll
    for (int e = 0; e < NE; e++) g[e] = 0.0;
    for (long long base = t0; base < t1; base += SCHUNK) {
        const int m = (int)((t1 - base < SCHUNK) ? (t1 - base) : SCHUNK);
        __syncthreads();
        if (threadIdx.x < m) {
            const long long star = idx ? idx[base + threadIdx.x] : base + threadIdx.x;
            double b[K], cu = u[star], cv = v[star];
            centre_uv(ctr, cu, cv);
            basis_eval<DEG>(cu, cv, b);
#pragma unroll
            for (int k = 0; k < K; k++) sV[threadIdx.x][k] = b[k];
        }
        __syncthreads();
#pragma unroll
        for (int e = 0; e < NE; e++) {
            const int ent = threadIdx.x + e * POLY_THREADS;
            if (ent < K * K) {
                const int i = ent / K, j = ent % K;
                for (int s = 0; s < m; s++) g[e] = fma(sV[s][i], sV[s][j], g[e]);
            }
        }
    }
    double *out = part + (size_t)blockIdx.x * ((size_t)K * K + (size_t)K * rc);
#pragma unroll
    for (int e = 0; e < NE; e++) {
        const int ent = threadIdx.x + e * POLY_THREADS;
        if (ent < K * K) out[ent] = g[e];
    }
}

// acc[e] = sum over chunks, in chunk order
__global__ void reduce_kernel(const double *__restrict__ part, long long len, int nchunk, double *__restrict__ acc) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= len) return;
    double s = 0.0;
    for (int c = 0; c < nchunk; c++) s += part[(size_t)c * len + e];
    acc[e] = s;
}

// ---------------------------------------------------------------------------------------------
// solve: factor_kernel turns G = Vc'Vc (K x K) into the solve operator of polyfit_factor.h (one thread: K <= 45, a few
// thousand dependent flops in shared memory); solve_kernel applies it to every pixel's right-hand side:
// forward substitution with the pivoted Cholesky factor, then one K x rc matrix-vector product.
// median_constraint (interpolate.jl:118-124, 139-141): the first design column is the constant m = median(z) instead of
// 1, so for m != 0 the solution is the unconstrained one with coefficient 1 divided by m -- and coefficient 1 is then
// overwritten by m; for m == 0 the column vanishes, LAPACK's rank-revealing solve drops it, and the other coefficients
// are those of the fit WITHOUT a constant term (the `red` factorisation); m = NaN makes every coefficient NaN.
// ---------------------------------------------------------------------------------------------
__global__ void factor_kernel(int degree, const double *__restrict__ G, const double *__restrict__ ctr,
                              const long long *__restrict__ nrows_ptr, long long nrows, int drop_const,
                              PolyFactor *__restrict__ F) {
    extern __shared__ __align__(16) double fac_work[];
    polyfit_factor(degree, G, ctr, nrows_ptr ? *nrows_ptr : nrows, drop_const, fac_work, F, (int)threadIdx.x, 32);
}

template <int DEG>
__global__ void __launch_bounds__(128) solve_kernel(const double *__restrict__ acc, long long rc,
                                                    const PolyFactor *__restrict__ F, const PolyFactor *__restrict__ Fred,
                                                    const double *__restrict__ med, double *__restrict__ poly) {
    constexpr int K = PolyCfg<DEG>::K;
    extern __shared__ __align__(16) double solve_smem[];           // (Fred ? 4 : 2) matrices of K x (K+1)
    typedef double (*Mat)[K + 1];
    Mat L = reinterpret_cast<Mat>(solve_smem), W = reinterpret_cast<Mat>(solve_smem + K * (K + 1));
    Mat Lr = reinterpret_cast<Mat>(solve_smem + 2 * K * (K + 1)), Wr = reinterpret_cast<Mat>(solve_smem + 3 * K * (K + 1));
    __shared__ double dsc[2][K];
    __shared__ int piv[2][K];
    for (int e = threadIdx.x; e < K * K; e += blockDim.x) {
        L[e / K][e % K] = F->L[e]; W[e / K][e % K] = F->W2[e];
        if (Fred) { Lr[e / K][e % K] = Fred->L[e]; Wr[e / K][e % K] = Fred->W2[e]; }
    }
    for (int e = threadIdx.x; e < K; e += blockDim.x) {
        dsc[0][e] = F->dsc[e]; piv[0][e] = F->piv[e];
        if (Fred) { dsc[1][e] = Fred->dsc[e]; piv[1][e] = Fred->piv[e]; }
    }
    __syncthreads();
    const long long px = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (px >= rc) return;
    double b[K], y[K];
#pragma unroll
    for (int k = 0; k < K; k++) b[k] = acc[(size_t)K * K + (size_t)k * rc + px];
    const double m = med ? med[px] : 1.0;
    const bool red = med && (m == 0.0);
    const int rcr = red ? Fred->rc_rank : F->rc_rank;
    if (!red) {
#pragma unroll
        for (int i = 0; i < K; i++) {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < K; k++) if (k == piv[0][i]) s = b[k] * dsc[0][k];
#pragma unroll
            for (int k = 0; k < i; k++) s -= L[i][k] * y[k];
            y[i] = (i < rcr) ? s / L[i][i] : 0.0;
        }
#pragma unroll
        for (int j = 0; j < K; j++) {
            double s = 0.0;
#pragma unroll
            for (int i = 0; i < K; i++) s = fma(W[j][i], y[i], s);
            poly[px + (size_t)j * rc] = (med && j == 0) ? m : (med && isnan(m)) ? m : s;
        }
    } else {
#pragma unroll
        for (int i = 0; i < K; i++) {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < K; k++) if (k == piv[1][i]) s = b[k] * dsc[1][k];
#pragma unroll
            for (int k = 0; k < i; k++) s -= Lr[i][k] * y[k];
            y[i] = (i < rcr) ? s / Lr[i][i] : 0.0;
        }
#pragma unroll
        for (int j = 0; j < K; j++) {
            double s = 0.0;
#pragma unroll
            for (int i = 0; i < K; i++) s = fma(Wr[j][i], y[i], s);
            poly[px + (size_t)j * rc] = (j == 0) ? m : s;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// per-star reconstruction error: part[tile][t] = sum over the tile's pixels of (sum_k poly[px,k] V[t,k] - pgf[t,px])^2
// ---------------------------------------------------------------------------------------------
template <int DEG, int PX, int UNR>
__global__ void __launch_bounds__(POLY_THREADS) mse_kernel(const double *__restrict__ pgf, long long rc,
                                                           const double *__restrict__ u, const double *__restrict__ v,
                                                           const long long *__restrict__ idx, long long n_bound,
                                                           const long long *__restrict__ n_ptr, const double *__restrict__ poly,
                                                           double *__restrict__ part) {
    constexpr int K = PolyCfg<DEG>::K, NW = POLY_THREADS / 32;
    __shared__ double sV[SCHUNK][K];
    __shared__ long long sOff[SCHUNK];
    __shared__ double sE[SCHUNK][NW];
    const long long n = n_ptr ? *n_ptr : n_bound;
    const long long per_chunk = ((n + gridDim.y - 1) / gridDim.y + SCHUNK - 1) / SCHUNK * SCHUNK;
    const long long t0 = (long long)blockIdx.y * per_chunk;
    const long long t1 = (t0 + per_chunk < n) ? t0 + per_chunk : n;
    const long long px0 = (long long)blockIdx.x * (POLY_THREADS * PX) + threadIdx.x;
    double cf[K][PX];
    bool live[PX];
#pragma unroll
    for (int j = 0; j < PX; j++) {
        const long long px = px0 + (long long)j * POLY_THREADS;
        live[j] = px < rc;
#pragma unroll
        for (int k = 0; k < K; k++) cf[k][j] = live[j] ? poly[px + (size_t)k * rc] : 0.0;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (long long base = t0; base < t1; base += SCHUNK) {
        const int m = (int)((t1 - base < SCHUNK) ? (t1 - base) : SCHUNK);
        __syncthreads();
        if (threadIdx.x < m) {
            const long long star = idx ? idx[base + threadIdx.x] : base + threadIdx.x;
            double b[K];
            basis_eval<DEG>(u[star], v[star], b);
#pragma unroll
            for (int k = 0; k < K; k++) sV[threadIdx.x][k] = b[k];
            sOff[threadIdx.x] = star * rc;
        }
        __syncthreads();
        // eight stars at a time: their per-thread partial errors are reduced across the warp by ONE halving tree
        // (9 shuffles per 8 stars instead of 40), value q landing on lane 4 q
        for (int s0 = 0; s0 < m; s0 += 8) {
            double e8[8];
#pragma unroll
            for (int q = 0; q < 8; q++) {
                const int s = (s0 + q < m) ? s0 + q : m - 1;   // a ragged tail re-reads the last star (result unused)
                double e = 0.0;
                {
                    const double *row = pgf + sOff[s];
                    double val[PX];
#pragma unroll
                    for (int j = 0; j < PX; j++) {
                        const long long px = px0 + (long long)j * POLY_THREADS;
                        val[j] = live[j] ? -__ldg(row + px) : 0.0;
                    }
#pragma unroll
                    for (int k = 0; k < K; k++) {
                        const double vk = sV[s][k];
#pragma unroll
                        for (int j = 0; j < PX; j++) val[j] = fma(cf[k][j], vk, val[j]);
                    }
#pragma unroll
                    for (int j = 0; j < PX; j++) e = fma(val[j], val[j], e);
                }
                e8[q] = e;
            }
            const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
            double k4[4], k2[2];
#pragma unroll
            for (int i = 0; i < 4; i++)
                k4[i] = (b4 ? e8[i + 4] : e8[i]) + __shfl_xor_sync(0xffffffffu, b4 ? e8[i] : e8[i + 4], 16);
#pragma unroll
            for (int i = 0; i < 2; i++)
                k2[i] = (b3 ? k4[i + 2] : k4[i]) + __shfl_xor_sync(0xffffffffu, b3 ? k4[i] : k4[i + 2], 8);
            double tot = (b2 ? k2[1] : k2[0]) + __shfl_xor_sync(0xffffffffu, b2 ? k2[0] : k2[1], 4);
            tot += __shfl_xor_sync(0xffffffffu, tot, 2);
            tot += __shfl_xor_sync(0xffffffffu, tot, 1);
            if ((lane & 3) == 0 && s0 + (lane >> 2) < m) sE[s0 + (lane >> 2)][warp] = tot;
        }
        __syncthreads();
        if (threadIdx.x < m) {
            double e = sE[threadIdx.x][0];
#pragma unroll
            for (int w = 1; w < NW; w++) e += sE[threadIdx.x][w];
            part[(size_t)blockIdx.x * n + base + threadIdx.x] = e;
        }
    }
}
// mse_kernel with the per-lane cp.async ring of normal_ring_kernel (see there): same arithmetic in the same order, the
// loads of the next RING_D - 1 stars in flight while a star is evaluated.
template <int DEG, int PX>
__global__ void __launch_bounds__(POLY_THREADS, ring_ctas(PolyCfg<DEG>::K * PX)) mse_ring_kernel(const double *__restrict__ pgf, long long rc,
                                                                   const double *__restrict__ u, const double *__restrict__ v,
                                                                   const long long *__restrict__ idx, long long n_bound,
                                                                   const long long *__restrict__ n_ptr, const double *__restrict__ poly,
                                                                   double *__restrict__ part) {
    constexpr int K = PolyCfg<DEG>::K;
    constexpr int RING_D = (K <= 21) ? RING_BYTES_PER_THREAD / 8 / PX : 16;   // static shared memory stays under 48 KB
    constexpr int NW = POLY_THREADS / 32;
    static_assert(RING_D % 8 == 0 && RING_D <= SCHUNK, "groups of eight stars stay inside the ring");
    __shared__ __align__(16) double sV[SCHUNK][K + (K & 1)];
    __shared__ long long sOff[2 * SCHUNK];
    __shared__ double sE[SCHUNK][NW];
    __shared__ double ring[RING_D][PX][POLY_THREADS];
    const long long n = n_ptr ? *n_ptr : n_bound;
    const long long per_chunk = ((n + gridDim.y - 1) / gridDim.y + SCHUNK - 1) / SCHUNK * SCHUNK;
    const long long t0 = (long long)blockIdx.y * per_chunk;
    const long long t1 = (t0 + per_chunk < n) ? t0 + per_chunk : n;
    const int nloc = (t1 > t0) ? (int)(t1 - t0) : 0;
    const long long px0 = (long long)blockIdx.x * (POLY_THREADS * PX) + threadIdx.x;
    double cf[K][PX];
    int nlive = 0;
#pragma unroll
    for (int j = 0; j < PX; j++) {
        const long long px = px0 + (long long)j * POLY_THREADS;
        const bool lv = px < rc;
        nlive += lv ? 1 : 0;
#pragma unroll
        for (int k = 0; k < K; k++) cf[k][j] = lv ? poly[px + (size_t)k * rc] : 0.0;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const double *mine = pgf + px0;
    double *my_slot = &ring[0][0][threadIdx.x];
    auto stage_off = [&](int i) {
        if (i < nloc) sOff[i & (2 * SCHUNK - 1)] = (idx ? idx[t0 + i] : t0 + i) * rc;
    };
    auto issue = [&](int i) {
        if (i < nloc) {
            const double *row = mine + sOff[i & (2 * SCHUNK - 1)];
            double *dst = my_slot + (i & (RING_D - 1)) * (PX * POLY_THREADS);
#pragma unroll
            for (int j = 0; j < PX; j++)
                if (j < nlive) cp_async8(dst + j * POLY_THREADS, row + (long long)j * POLY_THREADS);
        }
        cp_async_commit();
    };
    if (threadIdx.x < SCHUNK) stage_off(threadIdx.x);
    __syncthreads();
    for (int d = 0; d < RING_D - 1; d++) issue(d);
    for (int base = 0; base < nloc; base += SCHUNK) {
        const int m = (nloc - base < SCHUNK) ? (nloc - base) : SCHUNK;
        __syncthreads();
        if (threadIdx.x < m) {
            const long long star = idx ? idx[t0 + base + threadIdx.x] : t0 + base + threadIdx.x;
            double b[K];
            basis_eval<DEG>(u[star], v[star], b);
#pragma unroll
            for (int k = 0; k < K; k++) sV[threadIdx.x][k] = b[k];
        } else if (threadIdx.x >= SCHUNK && threadIdx.x < 2 * SCHUNK) {
            stage_off(base + threadIdx.x);
        }
        __syncthreads();
        for (int s0 = 0; s0 < m; s0 += 8) {
            double e8[8];
#pragma unroll
            for (int q = 0; q < 8; q++) {
                const int i = base + s0 + q;
                issue(i + RING_D - 1);                           // empty group past the end: the count stays in step
                cp_async_wait<RING_D - 1>();
                double e = 0.0;
                if (s0 + q < m) {
                    const double *src = my_slot + (i & (RING_D - 1)) * (PX * POLY_THREADS);
                    double val[PX];
#pragma unroll
                    for (int j = 0; j < PX; j++) val[j] = (j < nlive) ? -src[j * POLY_THREADS] : 0.0;
#pragma unroll
                    for (int k = 0; k < K; k++) {
                        const double vk = sV[s0 + q][k];
#pragma unroll
                        for (int j = 0; j < PX; j++) val[j] = fma(cf[k][j], vk, val[j]);
                    }
#pragma unroll
                    for (int j = 0; j < PX; j++) e = fma(val[j], val[j], e);
                }
                e8[q] = e;
            }
            const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
            double k4[4], k2[2];
#pragma unroll
            for (int i = 0; i < 4; i++)
                k4[i] = (b4 ? e8[i + 4] : e8[i]) + __shfl_xor_sync(0xffffffffu, b4 ? e8[i] : e8[i + 4], 16);
#pragma unroll
            for (int i = 0; i < 2; i++)
                k2[i] = (b3 ? k4[i + 2] : k4[i]) + __shfl_xor_sync(0xffffffffu, b3 ? k4[i] : k4[i + 2], 8);
            double tot = (b2 ? k2[1] : k2[0]) + __shfl_xor_sync(0xffffffffu, b2 ? k2[0] : k2[1], 4);
            tot += __shfl_xor_sync(0xffffffffu, tot, 2);
            tot += __shfl_xor_sync(0xffffffffu, tot, 1);
            if ((lane & 3) == 0 && s0 + (lane >> 2) < m) sE[s0 + (lane >> 2)][warp] = tot;
        }
        __syncthreads();
        if (threadIdx.x < m) {
            double e = sE[threadIdx.x][0];
#pragma unroll
            for (int w = 1; w < NW; w++) e += sE[threadIdx.x][w];
            part[(size_t)blockIdx.x * n + t0 + base + threadIdx.x] = e;
        }
    }
    cp_async_wait<0>();
}
__global__ void mse_reduce_kernel(const double *__restrict__ part, long long n_bound, const long long *__restrict__ n_ptr,
                                  int ntile, double inv_rc, double *__restrict__ mse) {
    const long long n = n_ptr ? *n_ptr : n_bound;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    double s = 0.0;
    for (int b = 0; b < ntile; b++) s += part[(size_t)b * n + t];
    mse[t] = s * inv_rc;
}

// ---------------------------------------------------------------------------------------------
// reader.jl:40-68  p(u, v, polMatrix, degree): one CTA per position; evaluate, sum, divide
// ---------------------------------------------------------------------------------------------
template <int DEG>
__global__ void __launch_bounds__(256) psf_at_kernel(const double *__restrict__ poly, long long rc,
                                                     const double *__restrict__ u, const double *__restrict__ v,
                                                     double *__restrict__ out) {
    constexpr int K = PolyCfg<DEG>::K;
    __shared__ double scratch[2 * 8 * 1];
    const long long m = blockIdx.x;
    double b[K];
    basis_eval<DEG>(u[m], v[m], b);
    double *star = out + (size_t)m * rc;
    double sum[1] = {0.0};
    for (long long px = threadIdx.x; px < rc; px += 256) {
        double val = 0.0;
#pragma unroll
        for (int k = 0; k < K; k++) val = fma(__ldg(poly + px + (size_t)k * rc), b[k], val);
        star[px] = val;
        sum[0] += val;
    }
    int buf = 0;
    cta_sum<1, 8>(sum, scratch, buf);
    const double s = sum[0];
    for (long long px = threadIdx.x; px < rc; px += 256) star[px] = star[px] / s;   // same thread wrote it
}

// ---------------------------------------------------------------------------------------------
// fewer than 5 training stars (interpolate.jl:142-145): coefficients = [median(z), 0, ...]; and the per-pixel
// median of the non-NaN values of a SMALL training set (n <= 64) used for that override.
// ---------------------------------------------------------------------------------------------
// rows: (n, rc) the training stars' pixel-grid fits gathered into one small buffer (see small_gather_kernel)
__global__ void small_median_kernel(const double *__restrict__ rows, long long rc, int n, int K, double *__restrict__ poly) {
    const long long px = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (px >= rc) return;
    double z[4];
    int m = 0;
    for (int t = 0; t < n && t < 4; t++) {
        const double val = rows[(size_t)t * rc + px];
        if (!isnan(val)) {   // insertion sort
            int i = m++;
            while (i > 0 && z[i - 1] > val) { z[i] = z[i - 1]; i--; }
            z[i] = val;
        }
    }
    double med = nan("");
    if (m > 0) med = (m & 1) ? z[m / 2] : (z[m / 2 - 1] / 2 + z[m / 2] / 2);   // Statistics.middle
    poly[px] = med;
    for (int k = 1; k < K; k++) poly[px + (size_t)k * rc] = 0.0;
}
// The (at most four) training stars of the whole catalogue may live on different shards: every shard marks the
// permutation positions of its own ones in a bit set (all-reduced by the caller), a star's slot is its rank among the
// marked positions, and the shards' rows (zero elsewhere) add up to the gathered (n, rc) block.
__global__ void alive_bits_kernel(const unsigned *__restrict__ pos, const long long *__restrict__ n_dev, unsigned *__restrict__ bits) {
    const long long n = *n_dev;
    for (long long t = threadIdx.x; t < n; t += blockDim.x) atomicOr(&bits[pos[t] >> 5], 1u << (pos[t] & 31));
}
__global__ void __launch_bounds__(256) small_gather_kernel(const double *__restrict__ pgf, long long rc,
                                                           const long long *__restrict__ idx, const unsigned *__restrict__ pos,
                                                           const long long *__restrict__ n_dev, const unsigned *__restrict__ bits,
                                                           double *__restrict__ rows) {
    __shared__ int s_part[8];
    const long long n = *n_dev;
    const int t = blockIdx.y;
    if (t >= n) return;
    const unsigned p = pos[t];
    int cnt = 0;
    for (unsigned w = threadIdx.x; w <= (p >> 5); w += blockDim.x) {
        unsigned b = bits[w];
        if (w == (p >> 5)) b &= (1u << (p & 31)) - 1u;
        cnt += __popc(b);
    }
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, m);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = cnt;
    __syncthreads();
    int slot = 0;
    for (int w = 0; w < 8; w++) slot += s_part[w];
    if (slot >= 4) return;
    const long long px = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (px < rc) rows[(size_t)slot * rc + px] = pgf[idx[t] * rc + px];
}

// ---------------------------------------------------------------------------------------------
// Per-pixel median over the training stars, median(filter(!isnan, z)) of interpolate.jl:119,128,140 -- the two places
// the reference needs it at full training-set size:
//   mode 1  median_constraint = true: coefficient 1 of every pixel is replaced by the median (:139-141).  With the
//           first design column set to the constant m (:119) the least-squares solution is the standard one with
//           coefficient 1 divided by m, so the other coefficients are those of the unconstrained fit (m != 0).
//   mode 2  median_constraint = false and a pixel has NaN values (:126-130): the NaNs are replaced by the median
//           before the solve.  Only pixels whose accumulated V'P is NaN are touched: their median is selected and
//           their column of V'P is re-accumulated with the replacement, in place in `acc`.
// Selection: MSB-first radix select on the order-preserving 64-bit key, 8 bits per pass, histograms in shared memory
// (256 bins x 64 pixels), 64 pixels x 8 star lanes per CTA; coalesced 8-byte loads along the pixels.  Integer
// arithmetic only, so the result is exact and deterministic.  An even count takes Statistics.middle = a/2 + b/2.
// ---------------------------------------------------------------------------------------------
constexpr int MED_PX = 64, MED_LANES = 8, MED_THREADS = MED_PX * MED_LANES;
__device__ __forceinline__ unsigned long long order_key(double z) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(z);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}
__device__ __forceinline__ double key_value(unsigned long long k) {
    const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffULL) : ~k;
    return __longlong_as_double((long long)b);
}
template <int DEG>
__global__ void __launch_bounds__(MED_THREADS) median_fix_kernel(const double *__restrict__ pgf, long long rc,
                                                                 const double *__restrict__ u, const double *__restrict__ v,
                                                                 const long long *__restrict__ idx, long long n_bound,
                                                                 const long long *__restrict__ n_ptr,
                                                                 const double *__restrict__ ctr, int mode,
                                                                 double *__restrict__ acc, double *__restrict__ med_out) {
    constexpr int K = PolyCfg<DEG>::K;
    const long long n = n_ptr ? *n_ptr : n_bound;
    extern __shared__ __align__(16) unsigned char med_smem[];
    unsigned int *hist = reinterpret_cast<unsigned int *>(med_smem);            // [256][MED_PX]
    __shared__ unsigned long long s_prefix[MED_PX], s_next[MED_PX];
    __shared__ unsigned int s_k[MED_PX], s_nvalid[MED_PX], s_cnt[MED_PX];
    __shared__ double s_med[MED_PX];
    __shared__ int s_need[MED_PX];
    const int pl = threadIdx.x % MED_PX, lane = threadIdx.x / MED_PX;
    const long long px = (long long)blockIdx.x * MED_PX + pl;
    if (lane == 0) {
        int need = px < rc;
        if (need && mode == 2) need = isnan(acc[(size_t)K * K + px]);          // V'P of the constant term: NaN iff z has one
        s_need[pl] = need; s_prefix[pl] = 0ULL; s_k[pl] = 0u; s_nvalid[pl] = 0u; s_cnt[pl] = 0u;
        s_next[pl] = ~0ULL; s_med[pl] = nan("");
    }
    __syncthreads();
    {
        int any = 0;
        for (int i = 0; i < MED_PX; i++) any |= s_need[i];
        if (!any) return;
    }
    for (int pass = 0; pass < 8; pass++) {
        const int shift = 56 - 8 * pass;
        for (int e = threadIdx.x; e < 256 * MED_PX; e += MED_THREADS) hist[e] = 0u;
        __syncthreads();
        if (s_need[pl]) {
            const unsigned long long prefix = s_prefix[pl];
            for (long long t = lane; t < n; t += MED_LANES) {
                const long long star = idx ? idx[t] : t;
                const double z = __ldg(pgf + star * rc + px);
                if (isnan(z)) continue;
                const unsigned long long key = order_key(z);
                if (pass == 0 || (key >> (shift + 8)) == (prefix >> (shift + 8)))
                    atomicAdd(&hist[(unsigned)((key >> shift) & 255ULL) * MED_PX + pl], 1u);
            }
        }
        __syncthreads();
        if (lane == 0 && s_need[pl]) {
            unsigned int k = s_k[pl];
            if (pass == 0) {
                unsigned int total = 0;
                for (int b = 0; b < 256; b++) total += hist[b * MED_PX + pl];
                s_nvalid[pl] = total;
                k = total ? (total - 1) / 2 : 0;
                if (!total) s_need[pl] = (mode == 2) ? 2 : 0;     // no finite value at all: median is NaN
            }
            unsigned int cum = 0;
            for (int b = 0; b < 256; b++) {
                const unsigned int h = hist[b * MED_PX + pl];
                if (cum + h > k) { s_prefix[pl] |= (unsigned long long)b << shift; k -= cum; break; }
                cum += h;
            }
            s_k[pl] = k;
        }
        __syncthreads();
    }
    // even count: the upper middle value is v1 again when enough copies of it exist, else the smallest value above it
    const bool sel = s_need[pl] == 1;
    if (sel && (s_nvalid[pl] & 1u) == 0u) {
        const unsigned long long k1 = s_prefix[pl];
        unsigned int cnt = 0;
        unsigned long long nxt = ~0ULL;
        for (long long t = lane; t < n; t += MED_LANES) {
            const long long star = idx ? idx[t] : t;
            const double z = __ldg(pgf + star * rc + px);
            if (isnan(z)) continue;
            const unsigned long long key = order_key(z);
            if (key <= k1) cnt++;
            else if (key < nxt) nxt = key;
        }
        atomicAdd(&s_cnt[pl], cnt);
        atomicMin(&s_next[pl], nxt);
    }
    __syncthreads();
    if (lane == 0 && sel) {
        const double v1 = key_value(s_prefix[pl]);
        double m = v1;
        if ((s_nvalid[pl] & 1u) == 0u) {
            const double v2 = (s_cnt[pl] >= s_nvalid[pl] / 2 + 1) ? v1 : key_value(s_next[pl]);
            m = v1 / 2 + v2 / 2;
        }
        s_med[pl] = m;
    }
    __syncthreads();
    if (mode == 1) {
        if (lane == 0 && px < rc) med_out[px] = s_med[pl];
        return;
    }
    // mode 2: V'P of the flagged pixels with NaN -> median
    double a[K];
#pragma unroll
    for (int k = 0; k < K; k++) a[k] = 0.0;
    const bool fix = s_need[pl] != 0;
    if (fix) {
        const double m = s_med[pl];
        for (long long t = lane; t < n; t += MED_LANES) {
            const long long star = idx ? idx[t] : t;
            double z = __ldg(pgf + star * rc + px);
            if (isnan(z)) z = m;
            double b[K], cu = u[star], cv = v[star];
            centre_uv(ctr, cu, cv);
            basis_eval<DEG>(cu, cv, b);
#pragma unroll
            for (int k = 0; k < K; k++) a[k] = fma(b[k], z, a[k]);
        }
    }
    double *red = reinterpret_cast<double *>(med_smem);                         // [MED_LANES][MED_PX]
#pragma unroll
    for (int k = 0; k < K; k++) {
        __syncthreads();
        red[lane * MED_PX + pl] = a[k];
        __syncthreads();
        if (lane == 0 && fix) {
            double sum = red[pl];
            for (int l = 1; l < MED_LANES; l++) sum += red[l * MED_PX + pl];
            acc[(size_t)K * K + (size_t)k * rc + px] = sum;
        }
    }
}
// ---------------------------------------------------------------------------------------------
// The same per-pixel median when the training stars are SHARDED over ranks (one process per GPU): the radix select
// runs on every rank with the same prefix; the only thing ranks exchange is the 256-bin histogram of each pass (one
// int32 sum-all-reduce per pass), after which every rank takes the same bin.  Two selections per pixel -- the lower
// and the upper middle order statistic -- so that an even count gives Statistics.middle(v1, v2) = v1/2 + v2/2 without a
// min-reduction.  Integer arithmetic only: exact, and identical on every rank.  Pixels are processed in tiles so that
// the histogram buffer stays bounded (SHARD_MED_TILE pixels x 256 bins).
// ---------------------------------------------------------------------------------------------
__global__ void shard_med_init_kernel(const double *__restrict__ b0, long long rc, int mode, int *__restrict__ need,
                                      unsigned long long *__restrict__ prefix, unsigned *__restrict__ kleft,
                                      unsigned *__restrict__ total) {
    const long long px = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (px >= rc) return;
    need[px] = (mode == 1) ? 1 : (isnan(b0[px]) ? 1 : 0);       // mode 2: V'P of the constant term is NaN iff z has a NaN
    prefix[px] = 0ULL; prefix[rc + px] = 0ULL; kleft[px] = 0u; kleft[rc + px] = 0u; total[px] = 0u;
}
__global__ void __launch_bounds__(MED_THREADS) shard_med_hist_kernel(const double *__restrict__ pgf, long long rc,
                                                                     const long long *__restrict__ idx, long long n_bound,
                                                                     const long long *__restrict__ n_ptr, long long px0,
                                                                     long long npx, int pass, const int *__restrict__ need,
                                                                     const unsigned long long *__restrict__ prefix,
                                                                     unsigned *__restrict__ hist) {
    const long long n = n_ptr ? *n_ptr : n_bound;
    extern __shared__ __align__(16) unsigned char med_smem[];
    unsigned int *sh = reinterpret_cast<unsigned int *>(med_smem);              // [256][MED_PX]
    __shared__ int s_need[MED_PX];
    const int pl = threadIdx.x % MED_PX, lane = threadIdx.x / MED_PX;
    const long long loc = (long long)blockIdx.x * MED_PX + pl, px = px0 + loc;
    if (lane == 0) s_need[pl] = (loc < npx && px < rc) ? need[px] : 0;
    for (int e = threadIdx.x; e < 256 * MED_PX; e += MED_THREADS) sh[e] = 0u;
    __syncthreads();
    int any = 0;
    for (int i = 0; i < MED_PX; i++) any |= s_need[i];
    if (!any) return;                                                           // (hist was zeroed by the host)
    const int shift = 56 - 8 * pass;
    if (s_need[pl]) {
        const unsigned long long pre = prefix[px];
        for (long long t = lane; t < n; t += MED_LANES) {
            const long long star = idx ? idx[t] : t;
            const double z = __ldg(pgf + star * rc + px);
            if (isnan(z)) continue;
            const unsigned long long key = order_key(z);
            if (pass == 0 || (key >> (shift + 8)) == (pre >> (shift + 8)))
                atomicAdd(&sh[(unsigned)((key >> shift) & 255ULL) * MED_PX + pl], 1u);
        }
    }
    __syncthreads();
    for (int e = threadIdx.x; e < 256 * MED_PX; e += MED_THREADS) {
        const int b = e / MED_PX, q = e % MED_PX;
        const long long l2 = (long long)blockIdx.x * MED_PX + q;
        if (l2 < npx && s_need[q]) hist[(size_t)l2 * 256 + b] = sh[e];
    }
}
// which = 0: order statistic (total - 1) / 2, which = 1: total / 2 (0-based, over the whole catalogue's finite values)
__global__ void shard_med_pick_kernel(const unsigned *__restrict__ hist, long long rc, long long px0, long long npx, int pass,
                                      int which, const int *__restrict__ need, unsigned long long *__restrict__ prefix,
                                      unsigned *__restrict__ kleft, unsigned *__restrict__ total) {
    const long long loc = (long long)blockIdx.x * blockDim.x + threadIdx.x, px = px0 + loc;
    if (loc >= npx || px >= rc || !need[px]) return;
    const unsigned *h = hist + (size_t)loc * 256;
    unsigned k = kleft[(size_t)which * rc + px];
    if (pass == 0) {
        unsigned tot = 0;
        for (int b = 0; b < 256; b++) tot += h[b];
        total[px] = tot;
        k = tot ? (which ? tot / 2 : (tot - 1) / 2) : 0;
    }
    const int shift = 56 - 8 * pass;
    unsigned cum = 0;
    for (int b = 0; b < 256; b++) {
        const unsigned hb = h[b];
        if (cum + hb > k) { prefix[(size_t)which * rc + px] |= (unsigned long long)b << shift; k -= cum; break; }
        cum += hb;
    }
    kleft[(size_t)which * rc + px] = k;
}
__global__ void shard_med_final_kernel(long long rc, const int *__restrict__ need, const unsigned long long *__restrict__ prefix,
                                       const unsigned *__restrict__ total, double *__restrict__ med) {
    const long long px = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (px >= rc) return;
    double m = nan("");
    if (need[px] && total[px]) {
        const double v1 = key_value(prefix[px]), v2 = key_value(prefix[rc + px]);
        m = (total[px] & 1u) ? v1 : v1 / 2 + v2 / 2;
    }
    med[px] = m;
}
// mode 2: this shard's share of V'P for the flagged pixels, NaN values replaced by the catalogue-wide median
template <int DEG>
__global__ void __launch_bounds__(MED_THREADS) shard_refill_kernel(const double *__restrict__ pgf, long long rc,
                                                                   const double *__restrict__ u, const double *__restrict__ v,
                                                                   const long long *__restrict__ idx, long long n_bound,
                                                                   const long long *__restrict__ n_ptr,
                                                                   const double *__restrict__ ctr, const int *__restrict__ need,
                                                                   const double *__restrict__ med, double *__restrict__ fix) {
    constexpr int K = PolyCfg<DEG>::K;
    const long long n = n_ptr ? *n_ptr : n_bound;
    __shared__ double red[MED_LANES][MED_PX];
    const int pl = threadIdx.x % MED_PX, lane = threadIdx.x / MED_PX;
    const long long px = (long long)blockIdx.x * MED_PX + pl;
    const bool fixp = px < rc && need[px];
    if (!__syncthreads_or(fixp)) return;
    double a[K];
#pragma unroll
    for (int k = 0; k < K; k++) a[k] = 0.0;
    if (fixp) {
        const double m = med[px];
        for (long long t = lane; t < n; t += MED_LANES) {
            const long long star = idx ? idx[t] : t;
            double z = __ldg(pgf + star * rc + px);
            if (isnan(z)) z = m;
            double b[K], cu = u[star], cv = v[star];
            centre_uv(ctr, cu, cv);
            basis_eval<DEG>(cu, cv, b);
#pragma unroll
            for (int k = 0; k < K; k++) a[k] = fma(b[k], z, a[k]);
        }
    }
#pragma unroll
    for (int k = 0; k < K; k++) {
        __syncthreads();
        red[lane][pl] = a[k];
        __syncthreads();
        if (lane == 0 && fixp) {
            double sum = red[0][pl];
            for (int l = 1; l < MED_LANES; l++) sum += red[l][pl];
            fix[(size_t)k * rc + px] = sum;
        }
    }
}
__global__ void shard_apply_fix_kernel(long long rc, int K, const int *__restrict__ need, const double *__restrict__ fix,
                                       double *__restrict__ vtp) {
    const long long px = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (px >= rc || !need[px]) return;
    for (int k = 0; k < K; k++) vtp[(size_t)k * rc + px] = fix[(size_t)k * rc + px];
}

// ---- launch -----------------------------------------------------------------------------------
static int pick_chunks(long long n, int tiles) {
    long long want = ((long long)sm_count() * 16 + tiles - 1) / tiles;
    const long long most = (n + SCHUNK - 1) / SCHUNK;
    if (want > most) want = most;
    if (want < 1) want = 1;
    return (int)want;
}

// Tiling variants of the two streaming kernels: pixels per thread x star-loop unroll.  SHOPT_POLY_CFG (0..3) selects one
// for measurements; the default is the fastest measured on B200 (profiles/README.md).
static int poly_cfg() {
    static int cfg = -1;
    if (cfg < 0) { const char *e = getenv("SHOPT_POLY_CFG"); cfg = e ? atoi(e) & 3 : POLY_CFG_DEFAULT; }
    return cfg;
}
// SHOPT_POLY_RING: bit 0 = normal_ring_kernel, bit 1 = mse_ring_kernel (default 3: both; 0 = the plain-load kernels).
// Measured on the bench workload (87k x 131 x 131, K = 10): normal 3.01 -> 1.99 ms (61 % -> 92 % of the measured copy
// bandwidth), mse 3.43 -> 2.17 ms (53 % -> 84 %); K = 21 (5.4k x 301 x 301): 1.55 -> 1.21 and 1.91 -> 1.69 ms.
static int poly_ring() {           // read per launch: the parity tests compare both forms in one process
    const char *e = getenv("SHOPT_POLY_RING");
    return e ? atoi(e) : 3;
}
// Four pixels per thread for 10 < K <= 21 as well (84 accumulators at K = 21, two CTAs per SM: the ring keeps the loads in
// flight whatever the occupancy, and a basis value read from shared memory feeds four FMAs instead of two).  Measured at
// K = 21 (5.4k x 301 x 301): normal 1.21 -> 1.13 ms, mse 1.69 -> 1.31 ms.  SHOPT_POLY_PX4=0 goes back to two pixels.
static int poly_px4() {
    const char *e = getenv("SHOPT_POLY_PX4");
    return e ? atoi(e) : 1;
}
static int poly_px(int K, int cfg) {
    const int base = (K <= 10) ? 4 : (K <= 21 ? 2 : 1);
    return (cfg & 1) ? (base > 1 ? base / 2 : 1) : base;
}
// The list length may live on the device (sl.n_dev): grids are sized from the bound sl.n, the kernels split the ACTUAL
// count over the chunks they were launched with (empty chunks write zeros).
template <int DEG, int PX, int UNR>
static int normal_launch_v(const double *pgf, long long rc, const double *u, const double *v, StarList sl, const double *ctr,
                           double *acc, double *scratch, size_t scratch_bytes, cudaStream_t st) {
    constexpr int K = PolyCfg<DEG>::K;
    const long long len = (long long)K * K + (long long)K * rc;
    if (sl.n == 0) { SHOPT_CUDA_TRY(cudaMemsetAsync(acc, 0, (size_t)len * 8, st)); return SHOPT_OK; }
    const int tiles = (int)((rc + POLY_THREADS * PX - 1) / (POLY_THREADS * PX));
    int nchunk = pick_chunks(sl.n, tiles);
    while (nchunk > 1 && (size_t)nchunk * len * 8 > scratch_bytes) nchunk--;
    if ((size_t)nchunk * len * 8 > scratch_bytes) return SHOPT_E_NOMEM;
    if (poly_ring() & 1)
        normal_ring_kernel<DEG, PX><<<dim3(tiles, nchunk), POLY_THREADS, 0, st>>>(pgf, rc, u, v, sl.idx, sl.n, sl.n_dev, ctr, scratch);
    else
        normal_kernel<DEG, PX, UNR><<<dim3(tiles, nchunk), POLY_THREADS, 0, st>>>(pgf, rc, u, v, sl.idx, sl.n, sl.n_dev, ctr, scratch);
    gram_kernel<DEG><<<nchunk, POLY_THREADS, 0, st>>>(u, v, sl.idx, sl.n, sl.n_dev, ctr, rc, scratch);
    reduce_kernel<<<(unsigned)((len + 255) / 256), 256, 0, st>>>(scratch, len, nchunk, acc);
    count_launches(3);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
template <int DEG>
static int normal_bulk_launch(const double *pgf, long long rc, const double *u, const double *v, StarList sl, const double *ctr,
                              double *acc, double *scratch, size_t scratch_bytes, cudaStream_t st) {
    constexpr int K = PolyCfg<DEG>::K;
    const long long len = (long long)K * K + (long long)K * rc;
    if (sl.n == 0) { SHOPT_CUDA_TRY(cudaMemsetAsync(acc, 0, (size_t)len * 8, st)); return SHOPT_OK; }
    const int tiles = (int)((rc + NB_TP - 1) / NB_TP);
    int nchunk = pick_chunks(sl.n, tiles);
    while (nchunk > 1 && (size_t)nchunk * len * 8 > scratch_bytes) nchunk--;
    if ((size_t)nchunk * len * 8 > scratch_bytes) return SHOPT_E_NOMEM;
    const size_t smem = ((size_t)NB_ST * NB_S * NB_PITCH + (size_t)2 * NB_S * K) * 8 + NB_ST * 8 + (size_t)NB_ST * NB_S * 4;
    auto kern = normal_bulk_kernel<DEG>;
    SHOPT_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<dim3(tiles, nchunk), POLY_THREADS, smem, st>>>(pgf, rc, u, v, sl.idx, sl.n, sl.n_dev, ctr, scratch);
    gram_kernel<DEG><<<nchunk, POLY_THREADS, 0, st>>>(u, v, sl.idx, sl.n, sl.n_dev, ctr, rc, scratch);
    reduce_kernel<<<(unsigned)((len + 255) / 256), 256, 0, st>>>(scratch, len, nchunk, acc);
    count_launches(3);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
// SHOPT_POLY_BULK=1 selects the bulk-copy staged kernel where it applies (4 pixels per thread: K <= 10).  Measured on the
// bench workload (87k x 131 x 131, K = 10): 3.54 ms against 3.00 ms for plain loads (52 % vs 61 % of the copy bandwidth),
// with 128 KB in flight per SM instead of ~49 KB: bytes in flight are NOT what holds normal_kernel back -- its FMA loop
// (40 DFMA + 14 LDS per star and thread on 8-12 resident warps) is; see DESIGN.md 4.3.  Kept as a measurement switch.
static int poly_bulk() {
    static int v = -1;
    if (v < 0) { const char *e = getenv("SHOPT_POLY_BULK"); v = e ? atoi(e) : 0; }
    return v;
}
template <int DEG>
static int normal_launch_t(const double *pgf, long long rc, const double *u, const double *v, StarList sl, const double *ctr,
                           double *acc, double *scratch, size_t scratch_bytes, cudaStream_t st) {
    constexpr int P0 = PolyCfg<DEG>::PX, P1 = (P0 > 1 ? P0 / 2 : 1);
    if constexpr (PolyCfg<DEG>::K <= 10) {
        if (poly_bulk() && (reinterpret_cast<uintptr_t>(pgf) & 7) == 0)
            return normal_bulk_launch<DEG>(pgf, rc, u, v, sl, ctr, acc, scratch, scratch_bytes, st);
    }
    if constexpr (PolyCfg<DEG>::K > 10 && PolyCfg<DEG>::K <= 21) {
        if (poly_px4() && poly_ring()) return normal_launch_v<DEG, 4, 4>(pgf, rc, u, v, sl, ctr, acc, scratch, scratch_bytes, st);
    }
    switch (poly_cfg()) {
        case 0: return normal_launch_v<DEG, P0, 2>(pgf, rc, u, v, sl, ctr, acc, scratch, scratch_bytes, st);
        case 1: return normal_launch_v<DEG, P1, 2>(pgf, rc, u, v, sl, ctr, acc, scratch, scratch_bytes, st);
        case 2: return normal_launch_v<DEG, P0, 4>(pgf, rc, u, v, sl, ctr, acc, scratch, scratch_bytes, st);
        default: return normal_launch_v<DEG, P1, 4>(pgf, rc, u, v, sl, ctr, acc, scratch, scratch_bytes, st);
    }
}
template <int DEG>
static int solve_launch_t(const double *acc, long long rc, const PolyFactor *F, const PolyFactor *Fred, const double *med,
                          double *poly, cudaStream_t st) {
    constexpr int K = PolyCfg<DEG>::K;
    const size_t smem = (size_t)(Fred ? 4 : 2) * K * (K + 1) * sizeof(double);
    auto kern = solve_kernel<DEG>;
    if (smem > 48 * 1024) SHOPT_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)((rc + 127) / 128), 128, smem, st>>>(acc, rc, F, Fred, med, poly);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
template <int DEG, int PX, int UNR>
static int mse_launch_v(const double *pgf, long long rc, const double *u, const double *v, StarList sl, const double *poly,
                        double *mse, double *scratch, size_t scratch_bytes, cudaStream_t st) {
    if (sl.n == 0) return SHOPT_OK;
    const int tiles = (int)((rc + POLY_THREADS * PX - 1) / (POLY_THREADS * PX));
    if ((size_t)tiles * sl.n * 8 > scratch_bytes) return SHOPT_E_NOMEM;
    const int nchunk = pick_chunks(sl.n, tiles);
    if (poly_ring() & 2)
        mse_ring_kernel<DEG, PX><<<dim3(tiles, nchunk), POLY_THREADS, 0, st>>>(pgf, rc, u, v, sl.idx, sl.n, sl.n_dev, poly, scratch);
    else
        mse_kernel<DEG, PX, UNR><<<dim3(tiles, nchunk), POLY_THREADS, 0, st>>>(pgf, rc, u, v, sl.idx, sl.n, sl.n_dev, poly, scratch);
    mse_reduce_kernel<<<(unsigned)((sl.n + 255) / 256), 256, 0, st>>>(scratch, sl.n, sl.n_dev, tiles, 1.0 / (double)rc, mse);
    count_launches(2);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
template <int DEG>
static int mse_launch_t(const double *pgf, long long rc, const double *u, const double *v, StarList sl, const double *poly,
                        double *mse, double *scratch, size_t scratch_bytes, cudaStream_t st) {
    constexpr int P0 = PolyCfg<DEG>::PX, P1 = (P0 > 1 ? P0 / 2 : 1);
    if constexpr (PolyCfg<DEG>::K > 10 && PolyCfg<DEG>::K <= 21) {
        if (poly_px4() && poly_ring()) return mse_launch_v<DEG, 4, 4>(pgf, rc, u, v, sl, poly, mse, scratch, scratch_bytes, st);
    }
    switch (poly_cfg()) {
        case 0: return mse_launch_v<DEG, P0, 2>(pgf, rc, u, v, sl, poly, mse, scratch, scratch_bytes, st);
        case 1: return mse_launch_v<DEG, P1, 2>(pgf, rc, u, v, sl, poly, mse, scratch, scratch_bytes, st);
        case 2: return mse_launch_v<DEG, P0, 4>(pgf, rc, u, v, sl, poly, mse, scratch, scratch_bytes, st);
        default: return mse_launch_v<DEG, P1, 4>(pgf, rc, u, v, sl, poly, mse, scratch, scratch_bytes, st);
    }
}
template <int DEG>
static int psf_launch_t(const double *poly, long long rc, const double *u, const double *v, long long M, double *out,
                        cudaStream_t st) {
    if (M == 0) return SHOPT_OK;
    psf_at_kernel<DEG><<<(unsigned)M, 256, 0, st>>>(poly, rc, u, v, out);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}

#define DEG_SWITCH(degree, CALL)                     \
    switch (degree) {                                \
        case 0: return CALL(0);                      \
        case 1: return CALL(1);                      \
        case 2: return CALL(2);                      \
        case 3: return CALL(3);                      \
        case 4: return CALL(4);                      \
        case 5: return CALL(5);                      \
        case 6: return CALL(6);                      \
        case 7: return CALL(7);                      \
        case 8: return CALL(8);                      \
        default: return SHOPT_E_BADARG;              \
    }

size_t polyfit_scratch_bytes(long long rc, long long n, int degree) {
    const long long K = num_terms(degree);
    const long long len = K * K + K * rc;
    const int px = poly_px((int)K, 1);                      // the smaller tile of the variants: the larger scratch
    const int tiles = (int)((rc + POLY_THREADS * px - 1) / (POLY_THREADS * px));
    const size_t a = (size_t)pick_chunks(n > 0 ? n : 1, tiles) * len * 8;
    const size_t b = (size_t)tiles * (n > 0 ? n : 1) * 8;
    return a > b ? a : b;
}
// doubles of per-tile partial sums that polyfit_mse_launch needs for n stars (the smaller pixel tile of its variants)
size_t polyfit_mse_part_len(long long rc, long long n, int degree) {
    const int px = poly_px(num_terms(degree), 1);
    const long long tiles = (rc + POLY_THREADS * px - 1) / (POLY_THREADS * px);
    return (size_t)tiles * (size_t)(n > 0 ? n : 0);
}
int polyfit_normal_launch(const double *pgf, long long rc, const double *u, const double *v, StarList sl, int degree,
                          const double *ctr, double *acc, double *scratch, size_t scratch_bytes, cudaStream_t st) {
#define CALL(D) normal_launch_t<D>(pgf, rc, u, v, sl, ctr, acc, scratch, scratch_bytes, st)
    DEG_SWITCH(degree, CALL)
#undef CALL
}
int polyfit_factor_launch(const double *G, int degree, const double *ctr, const long long *nrows_dev, long long nrows,
                          int drop_const, PolyFactor *F, cudaStream_t st) {
    if (degree < 0 || degree > 8) return SHOPT_E_BADARG;
    const int K = num_terms(degree);
    const size_t smem = ((size_t)7 * K * K + 2 * K) * sizeof(double);
    if (smem > 48 * 1024)
        SHOPT_CUDA_TRY(cudaFuncSetAttribute(factor_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    factor_kernel<<<1, 32, smem, st>>>(degree, G, ctr, nrows_dev, nrows, drop_const, F);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
int polyfit_solve_launch(const double *acc, long long rc, int degree, const PolyFactor *F, const PolyFactor *Fred,
                         const double *med, double *poly, cudaStream_t st) {
#define CALL(D) solve_launch_t<D>(acc, rc, F, Fred, med, poly, st)
    DEG_SWITCH(degree, CALL)
#undef CALL
}
int polyfit_mse_launch(const double *pgf, long long rc, const double *u, const double *v, StarList sl, int degree,
                       const double *poly, double *mse, double *scratch, size_t scratch_bytes, cudaStream_t st) {
#define CALL(D) mse_launch_t<D>(pgf, rc, u, v, sl, poly, mse, scratch, scratch_bytes, st)
    DEG_SWITCH(degree, CALL)
#undef CALL
}
int psf_at_launch(const double *poly, long long rc, int degree, const double *u, const double *v, long long M,
                  double *out, cudaStream_t st) {
#define CALL(D) psf_launch_t<D>(poly, rc, u, v, M, out, st)
    DEG_SWITCH(degree, CALL)
#undef CALL
}
int small_median_launch(const double *rows, long long rc, int n, int degree, double *poly, cudaStream_t st) {
    small_median_kernel<<<(unsigned)((rc + 127) / 128), 128, 0, st>>>(rows, rc, n, num_terms(degree), poly);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
// fewer than five training stars in the whole catalogue: gather their rows (see small_gather_kernel); `bits` holds
// NG bits, rows 4 * rc doubles
int small_gather_launch(const double *pgf, long long rc, const long long *idx, const unsigned *pos, const long long *n_dev,
                        long long NG, unsigned *bits, double *rows, const Exchange &ex, cudaStream_t st) {
    const long long W = (NG + 31) >> 5;
    SHOPT_CUDA_TRY(cudaMemsetAsync(bits, 0, (size_t)(W > 0 ? W : 1) * 4, st));
    SHOPT_CUDA_TRY(cudaMemsetAsync(rows, 0, (size_t)4 * rc * 8, st));
    alive_bits_kernel<<<1, 32, 0, st>>>(pos, n_dev, bits);
    TRY(ex.sum(bits, W, 1));
    small_gather_kernel<<<dim3((unsigned)((rc + 255) / 256), 4), 256, 0, st>>>(pgf, rc, idx, pos, n_dev, bits, rows);
    TRY(ex.sum(rows, 4 * rc, 0));
    count_launches(2);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}

template <int DEG>
static int median_launch_t(const double *pgf, long long rc, const double *u, const double *v, StarList sl, const double *ctr,
                           int mode, double *acc, double *med, cudaStream_t st) {
    if (sl.n == 0) return SHOPT_OK;
    const size_t smem = (size_t)256 * MED_PX * sizeof(unsigned int);
    auto kern = median_fix_kernel<DEG>;
    SHOPT_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)((rc + MED_PX - 1) / MED_PX), MED_THREADS, smem, st>>>(pgf, rc, u, v, sl.idx, sl.n, sl.n_dev, ctr, mode, acc, med);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
int polyfit_median_launch(const double *pgf, long long rc, const double *u, const double *v, StarList sl, int degree,
                          const double *ctr, int mode, double *acc, double *med, cudaStream_t st) {
    if ((mode == 1 && !med) || (mode == 2 && !acc) || mode < 1 || mode > 2) return SHOPT_E_BADARG;
#define CALL(D) median_launch_t<D>(pgf, rc, u, v, sl, ctr, mode, acc, med, st)
    DEG_SWITCH(degree, CALL)
#undef CALL
}

// acc -= sub (the downdate of the normal equations by the stars a round rejected)
__global__ void sub_kernel(double *__restrict__ acc, const double *__restrict__ sub, long long len) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e < len) acc[e] -= sub[e];
}
int polyfit_sub_launch(double *acc, const double *sub, long long len, cudaStream_t st) {
    sub_kernel<<<(unsigned)((len + 255) / 256), 256, 0, st>>>(acc, sub, len);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}

// ---- sharded median: buffers and driver ----------------------------------------------------------------------
constexpr long long SHARD_MED_TILE = 16384;
size_t polyfit_shard_median_bytes(long long rc, int degree) {
    const size_t tile = (size_t)std::min<long long>(rc, SHARD_MED_TILE);
    return (size_t)rc * (2 * 8 + 2 * 4 + 4 + 4) + tile * 256 * 4 + (size_t)num_terms(degree) * rc * 8 + 8 * 256;   // Bump pads every block to 256 bytes
}
template <int DEG>
static int shard_refill_launch_t(const double *pgf, long long rc, const double *u, const double *v, StarList sl,
                                 const double *ctr, const int *need, const double *med, double *fix, cudaStream_t st) {
    shard_refill_kernel<DEG><<<(unsigned)((rc + MED_PX - 1) / MED_PX), MED_THREADS, 0, st>>>(pgf, rc, u, v, sl.idx, sl.n, sl.n_dev,
                                                                                             ctr, need, med, fix);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
static int shard_refill_launch(const double *pgf, long long rc, const double *u, const double *v, StarList sl, int degree,
                               const double *ctr, const int *need, const double *med, double *fix, cudaStream_t st) {
#define CALL(D) shard_refill_launch_t<D>(pgf, rc, u, v, sl, ctr, need, med, fix, st)
    DEG_SWITCH(degree, CALL)
#undef CALL
}
// mode 1: med (rc) = per-pixel median over the training stars of EVERY rank; mode 2: the NaN -> median repair of acc.
// work: polyfit_shard_median_bytes(rc, degree) bytes of device scratch.  Every rank must call it (collectives inside).
int polyfit_median_sharded_launch(const double *pgf, long long rc, const double *u, const double *v, StarList sl, int degree,
                                  const double *ctr, int mode, double *acc, double *med, void *work, const Exchange &ex,
                                  cudaStream_t st) {
    if (mode < 1 || mode > 2 || !work || !med || (mode == 2 && !acc) || degree < 0 || degree > 8) return SHOPT_E_BADARG;
    const int K = num_terms(degree);
    const long long tile = std::min<long long>(rc, SHARD_MED_TILE);
    Bump a(work);
    unsigned long long *prefix = a.take<unsigned long long>(2 * (size_t)rc);
    unsigned *kleft = a.take<unsigned>(2 * (size_t)rc), *total = a.take<unsigned>((size_t)rc);
    int *need = a.take<int>((size_t)rc);
    unsigned *hist = a.take<unsigned>((size_t)tile * 256);
    double *fix = a.take<double>((size_t)K * rc);
    const unsigned g256 = (unsigned)((rc + 255) / 256);
    shard_med_init_kernel<<<g256, 256, 0, st>>>(acc ? acc + (size_t)K * K : nullptr, rc, mode, need, prefix, kleft, total);
    count_launches(1);
    const size_t smem = (size_t)256 * MED_PX * sizeof(unsigned int);
    SHOPT_CUDA_TRY(cudaFuncSetAttribute(shard_med_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (long long px0 = 0; px0 < rc; px0 += tile) {
        const long long npx = std::min<long long>(tile, rc - px0);
        for (int which = 0; which < 2; which++)
            for (int pass = 0; pass < 8; pass++) {
                SHOPT_CUDA_TRY(cudaMemsetAsync(hist, 0, (size_t)npx * 256 * 4, st));
                shard_med_hist_kernel<<<(unsigned)((npx + MED_PX - 1) / MED_PX), MED_THREADS, smem, st>>>(
                    pgf, rc, sl.idx, sl.n, sl.n_dev, px0, npx, pass, need, prefix + (size_t)which * rc, hist);
                TRY(ex.sum(hist, npx * 256, 1));
                shard_med_pick_kernel<<<(unsigned)((npx + 255) / 256), 256, 0, st>>>(hist, rc, px0, npx, pass, which, need, prefix,
                                                                                     kleft, total);
                count_launches(2);
            }
    }
    shard_med_final_kernel<<<g256, 256, 0, st>>>(rc, need, prefix, total, med);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    if (mode == 1) return SHOPT_OK;
    SHOPT_CUDA_TRY(cudaMemsetAsync(fix, 0, (size_t)K * rc * 8, st));
    TRY(shard_refill_launch(pgf, rc, u, v, sl, degree, ctr, need, med, fix, st));
    TRY(ex.sum(fix, (long long)K * rc, 0));
    shard_apply_fix_kernel<<<g256, 256, 0, st>>>(rc, K, need, fix, acc + (size_t)K * K);
    count_launches(1);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}

}  // namespace shopt
