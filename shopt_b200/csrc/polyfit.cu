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
//                  per-star partial sums per pixel tile, reduced in tile order by mse_reduce_kernel.
#include "common.cuh"
#include "kernels.h"

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

constexpr int POLY_THREADS = 128;
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
template <int DEG>
__global__ void __launch_bounds__(POLY_THREADS) normal_kernel(const double *__restrict__ pgf, long long rc,
                                                              const double *__restrict__ u, const double *__restrict__ v,
                                                              const long long *__restrict__ idx, long long n,
                                                              long long per_chunk, double *__restrict__ part) {
    constexpr int K = PolyCfg<DEG>::K, PX = PolyCfg<DEG>::PX;
    __shared__ double sV[SCHUNK][K];
    __shared__ long long sOff[SCHUNK];
    const long long t0 = (long long)blockIdx.y * per_chunk;
    const long long t1 = (t0 + per_chunk < n) ? t0 + per_chunk : n;
    const long long px0 = (long long)blockIdx.x * PolyCfg<DEG>::TILE + threadIdx.x;
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
            double b[K];
            basis_eval<DEG>(u[star], v[star], b);
#pragma unroll
            for (int k = 0; k < K; k++) sV[threadIdx.x][k] = b[k];
            sOff[threadIdx.x] = star * rc;
        }
        __syncthreads();
#pragma unroll 2
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

// G = V'V partials: one block per chunk, the K*K entries strided over the threads
template <int DEG>
__global__ void __launch_bounds__(POLY_THREADS) gram_kernel(const double *__restrict__ u, const double *__restrict__ v,
                                                            const long long *__restrict__ idx, long long n,
                                                            long long per_chunk, long long rc, double *__restrict__ part) {
    constexpr int K = PolyCfg<DEG>::K;
    __shared__ double sV[SCHUNK][K];
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
            double b[K];
            basis_eval<DEG>(u[star], v[star], b);
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
// solve: G = L L' after symmetric diagonal equilibration; poly[px + k*rc] = (G^-1 B)[k][px]
// ---------------------------------------------------------------------------------------------
template <int DEG>
__global__ void __launch_bounds__(128) solve_kernel(const double *__restrict__ acc, long long rc, double *__restrict__ poly,
                                                    int *__restrict__ info) {
    constexpr int K = PolyCfg<DEG>::K;
    __shared__ double L[K][K + 1];
    __shared__ double dsc[K];
    __shared__ int bad;
    if (threadIdx.x == 0) {
        bad = 0;
        for (int i = 0; i < K; i++) {
            const double d = acc[i * K + i];
            if (!(d > 0.0) || !isfinite(d)) bad = 1;
            dsc[i] = bad ? 1.0 : 1.0 / sqrt(d);
        }
        for (int i = 0; i < K; i++)
            for (int j = 0; j < K; j++) L[i][j] = acc[i * K + j] * dsc[i] * dsc[j];
        for (int j = 0; j < K && !bad; j++) {
            double d = L[j][j];
            for (int k = 0; k < j; k++) d -= L[j][k] * L[j][k];
            if (!(d > 0.0) || !isfinite(d)) { bad = 1; break; }
            const double lj = sqrt(d);
            L[j][j] = lj;
            for (int i = j + 1; i < K; i++) {
                double s = L[i][j];
                for (int k = 0; k < j; k++) s -= L[i][k] * L[j][k];
                L[i][j] = s / lj;
            }
        }
        if (bad && blockIdx.x == 0) *info = 1;
    }
    __syncthreads();
    const long long px = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (px >= rc) return;
    double y[K];
    if (bad) {
#pragma unroll
        for (int k = 0; k < K; k++) poly[px + (size_t)k * rc] = nan("");
        return;
    }
#pragma unroll
    for (int i = 0; i < K; i++) {
        double s = acc[(size_t)K * K + (size_t)i * rc + px] * dsc[i];
#pragma unroll
        for (int k = 0; k < i; k++) s -= L[i][k] * y[k];
        y[i] = s / L[i][i];
    }
#pragma unroll
    for (int i = K - 1; i >= 0; i--) {
        double s = y[i];
#pragma unroll
        for (int k = i + 1; k < K; k++) s -= L[k][i] * y[k];
        y[i] = s / L[i][i];
    }
#pragma unroll
    for (int k = 0; k < K; k++) poly[px + (size_t)k * rc] = y[k] * dsc[k];
}

// ---------------------------------------------------------------------------------------------
// per-star reconstruction error: part[tile][t] = sum over the tile's pixels of (sum_k poly[px,k] V[t,k] - pgf[t,px])^2
// ---------------------------------------------------------------------------------------------
template <int DEG>
__global__ void __launch_bounds__(POLY_THREADS) mse_kernel(const double *__restrict__ pgf, long long rc,
                                                           const double *__restrict__ u, const double *__restrict__ v,
                                                           const long long *__restrict__ idx, long long n,
                                                           long long per_chunk, const double *__restrict__ poly,
                                                           double *__restrict__ part) {
    constexpr int K = PolyCfg<DEG>::K, PX = PolyCfg<DEG>::PX, NW = POLY_THREADS / 32;
    __shared__ double sV[SCHUNK][K];
    __shared__ long long sOff[SCHUNK];
    __shared__ double sE[SCHUNK][NW];
    const long long t0 = (long long)blockIdx.y * per_chunk;
    const long long t1 = (t0 + per_chunk < n) ? t0 + per_chunk : n;
    const long long px0 = (long long)blockIdx.x * PolyCfg<DEG>::TILE + threadIdx.x;
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
#pragma unroll 2
        for (int s = 0; s < m; s++) {
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
            double e = 0.0;
#pragma unroll
            for (int j = 0; j < PX; j++) e = fma(val[j], val[j], e);
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
            if (lane == 0) sE[s][warp] = e;
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
__global__ void mse_reduce_kernel(const double *__restrict__ part, long long n, int ntile, double inv_rc,
                                  double *__restrict__ mse) {
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
__global__ void small_median_kernel(const double *__restrict__ pgf, long long rc, const long long *__restrict__ idx,
                                    int n, int K, double *__restrict__ poly) {
    const long long px = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (px >= rc) return;
    double z[4];
    int m = 0;
    for (int t = 0; t < n && t < 4; t++) {
        const long long star = idx ? idx[t] : t;
        const double val = pgf[star * rc + px];
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

// ---- launch -----------------------------------------------------------------------------------
static int pick_chunks(long long n, int tiles) {
    long long want = ((long long)sm_count() * 16 + tiles - 1) / tiles;
    const long long most = (n + SCHUNK - 1) / SCHUNK;
    if (want > most) want = most;
    if (want < 1) want = 1;
    return (int)want;
}

template <int DEG>
static int normal_launch_t(const double *pgf, long long rc, const double *u, const double *v, const long long *idx,
                           long long n, double *acc, double *scratch, size_t scratch_bytes, cudaStream_t st) {
    constexpr int K = PolyCfg<DEG>::K;
    const long long len = (long long)K * K + (long long)K * rc;
    if (n == 0) { SHOPT_CUDA_TRY(cudaMemsetAsync(acc, 0, (size_t)len * 8, st)); return SHOPT_OK; }
    const int tiles = (int)((rc + PolyCfg<DEG>::TILE - 1) / PolyCfg<DEG>::TILE);
    int nchunk = pick_chunks(n, tiles);
    while (nchunk > 1 && (size_t)nchunk * len * 8 > scratch_bytes) nchunk--;
    if ((size_t)nchunk * len * 8 > scratch_bytes) return SHOPT_E_NOMEM;
    const long long per_chunk = (n + nchunk - 1) / nchunk;
    nchunk = (int)((n + per_chunk - 1) / per_chunk);
    normal_kernel<DEG><<<dim3(tiles, nchunk), POLY_THREADS, 0, st>>>(pgf, rc, u, v, idx, n, per_chunk, scratch);
    gram_kernel<DEG><<<nchunk, POLY_THREADS, 0, st>>>(u, v, idx, n, per_chunk, rc, scratch);
    reduce_kernel<<<(unsigned)((len + 255) / 256), 256, 0, st>>>(scratch, len, nchunk, acc);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
template <int DEG>
static int solve_launch_t(const double *acc, long long rc, double *poly, int *info, cudaStream_t st) {
    SHOPT_CUDA_TRY(cudaMemsetAsync(info, 0, sizeof(int), st));
    solve_kernel<DEG><<<(unsigned)((rc + 127) / 128), 128, 0, st>>>(acc, rc, poly, info);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
template <int DEG>
static int mse_launch_t(const double *pgf, long long rc, const double *u, const double *v, const long long *idx,
                        long long n, const double *poly, double *mse, double *scratch, size_t scratch_bytes,
                        cudaStream_t st) {
    if (n == 0) return SHOPT_OK;
    const int tiles = (int)((rc + PolyCfg<DEG>::TILE - 1) / PolyCfg<DEG>::TILE);
    if ((size_t)tiles * n * 8 > scratch_bytes) return SHOPT_E_NOMEM;
    int nchunk = pick_chunks(n, tiles);
    const long long per_chunk = ((n + nchunk - 1) / nchunk + SCHUNK - 1) / SCHUNK * SCHUNK;
    nchunk = (int)((n + per_chunk - 1) / per_chunk);
    mse_kernel<DEG><<<dim3(tiles, nchunk), POLY_THREADS, 0, st>>>(pgf, rc, u, v, idx, n, per_chunk, poly, scratch);
    mse_reduce_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(scratch, n, tiles, 1.0 / (double)rc, mse);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
template <int DEG>
static int psf_launch_t(const double *poly, long long rc, const double *u, const double *v, long long M, double *out,
                        cudaStream_t st) {
    if (M == 0) return SHOPT_OK;
    psf_at_kernel<DEG><<<(unsigned)M, 256, 0, st>>>(poly, rc, u, v, out);
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
    const int px = (K <= 10) ? 4 : (K <= 21 ? 2 : 1);
    const int tiles = (int)((rc + POLY_THREADS * px - 1) / (POLY_THREADS * px));
    const size_t a = (size_t)pick_chunks(n > 0 ? n : 1, tiles) * len * 8;
    const size_t b = (size_t)tiles * (n > 0 ? n : 1) * 8;
    return a > b ? a : b;
}
int polyfit_normal_launch(const double *pgf, long long rc, const double *u, const double *v, const long long *idx,
                          long long n, int degree, double *acc, double *scratch, size_t scratch_bytes, cudaStream_t st) {
#define CALL(D) normal_launch_t<D>(pgf, rc, u, v, idx, n, acc, scratch, scratch_bytes, st)
    DEG_SWITCH(degree, CALL)
#undef CALL
}
int polyfit_solve_launch(const double *acc, long long rc, int degree, double *poly, int *info, cudaStream_t st) {
#define CALL(D) solve_launch_t<D>(acc, rc, poly, info, st)
    DEG_SWITCH(degree, CALL)
#undef CALL
}
int polyfit_mse_launch(const double *pgf, long long rc, const double *u, const double *v, const long long *idx,
                       long long n, int degree, const double *poly, double *mse, double *scratch, size_t scratch_bytes,
                       cudaStream_t st) {
#define CALL(D) mse_launch_t<D>(pgf, rc, u, v, idx, n, poly, mse, scratch, scratch_bytes, st)
    DEG_SWITCH(degree, CALL)
#undef CALL
}
int psf_at_launch(const double *poly, long long rc, int degree, const double *u, const double *v, long long M,
                  double *out, cudaStream_t st) {
#define CALL(D) psf_launch_t<D>(poly, rc, u, v, M, out, st)
    DEG_SWITCH(degree, CALL)
#undef CALL
}
int small_median_launch(const double *pgf, long long rc, const long long *idx, int n, int degree, double *poly,
                        cudaStream_t st) {
    small_median_kernel<<<(unsigned)((rc + 127) / 128), 128, 0, st>>>(pgf, rc, idx, n, num_terms(degree), poly);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}

}  // namespace shopt
