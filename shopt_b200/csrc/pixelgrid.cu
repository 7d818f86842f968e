// pixelgrid.cu -- per-star pixel-grid fit in "PCA" mode, and the standalone smooth().
//
// Replaces the loop shopt/shopt.jl:281-314 of the reference:
//     data = nanToZero(star)                       shopt/masks.jl:38-50      (NaN -> 0.001)
//     pca_image(data, PCAterms)                    shopt/pca.jl:1-18         (MultivariateStats fit/transform/reconstruct,
//                                                                             rows = variables, columns = observations)
//     smooth(., lanczos)  or  ./ sum(.)            shopt/dataPreprocessing.jl:94-139, shopt.jl:292-302
// MultivariateStats 0.10.2 is not vendored; the algorithm is the one recorded in SURVEY.md A.3: mean over the columns,
// SVD of the centred matrix Z, variances sigma_i^2/(c-1), k = first count reaching pratio of the total (capped by
// ncomp and min(r,c)), output mean + U_k U_k' Z.
//
// Design (B200): one persistent CTA per star.  The centred stamp W = Z (r x c, column-major) lives in shared memory
// (137 KB at 131 x 131) and is diagonalised in place by ONE-SIDED (Hestenes) JACOBI rotations: a sweep is c-1 (or c)
// steps of a round-robin tournament, each step rotating floor(c/2) disjoint column pairs, one warp per pair (the three
// dot products and the rotation keep the two columns in registers).  On convergence the columns are sigma_j u_j, which
// is all PCA needs: variances are the squared column norms, and the projection  Y = U_k' Z  re-reads the stamp from
// global memory (L2-hot), so no second r x c buffer and no V accumulation is needed.  The reconstruction is written to
// the output, pulled back into the (now free) shared buffer and smoothed from there.  Stamps too large for shared
// memory (301 x 301) run the same code with W in a per-CTA global scratch block that stays L2-resident.
// The stage is FP64-compute-bound (O(r c^2) flops per sweep on 8 r c bytes); see DESIGN.md for the roofline.
#include "common.cuh"
#include "kernels.h"

namespace shopt {

struct PcaArgs {
    const double *stamps;
    int r, c;
    long long N;
    int ncomp;
    double pratio;
    int lanczos_a, unity_sum;
    double *pgf;
    int *k_out, *status;
    unsigned long long *counter;
    double *wglobal;      // non-null: per-CTA scratch of wstride doubles replaces the shared-memory W / Y buffers
    size_t wstride;
    int kmax;             // min(ncomp, r, c)
    double kern[15 * 15]; // normalised lanczos kernel (2a+1)^2, a <= 7, column-major like the reference's
};

__device__ __forceinline__ double nan_to_zero(double x) { return isnan(x) ? 0.001 : x; }   // masks.jl:43

template <int NW>
__device__ __forceinline__ double block_sum1(double v, double *scratch, int &buf) {
    double a[1] = {v};
    cta_sum<1, NW>(a, scratch, buf);
    return a[0];
}

// rows handled by one lane: i = lane + 32 t, t < RT
template <int NW, int RT>
__global__ void __launch_bounds__(NW * 32, 1) pca_kernel(const PcaArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int r = a.r, c = a.c;
    const long long rc = (long long)r * c;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nthr = NW * 32;
    // shared layout: [ misc 64 B | scratch 2*NW doubles | mu r | sv c | ord c ints | (W rc | Y kmax*c) unless global ]
    long long *s_next = reinterpret_cast<long long *>(smem_raw);
    int *s_flag = reinterpret_cast<int *>(smem_raw + 16);
    double *scratch = reinterpret_cast<double *>(smem_raw + 64);
    double *mu = scratch + 2 * NW;
    double *sv = mu + r;
    int *ord = reinterpret_cast<int *>(sv + c);
    double *tail = reinterpret_cast<double *>(ord + ((c + 1) & ~1));
    double *W = a.wglobal ? a.wglobal + (size_t)blockIdx.x * a.wstride : tail;
    double *Y = W + rc;
    int buf = 0;
    const int npos = (c & 1) ? c + 1 : c;     // tournament positions (one dummy when c is odd)
    const int nm1 = npos - 1, npairs = npos / 2;

    for (;;) {
        __syncthreads();
        if (tid == 0) *s_next = (long long)atomicAdd(a.counter, 1ULL);
        __syncthreads();
        const long long star = *s_next;
        if (star >= a.N) break;
        const double *X = a.stamps + star * rc;
        double *out = a.pgf + star * rc;

        // ---- load, nanToZero, finiteness (LAPACK raises on Inf -> shopt.jl:305-309 zeros + failedStars) ----
        int bad = 0;
        for (long long p = tid; p < rc; p += nthr) {
            const double x = nan_to_zero(__ldg(X + p));
            if (!isfinite(x)) bad = 1;
            W[p] = x;
        }
        bad = __syncthreads_or(bad);
        if (bad) {
            for (long long p = tid; p < rc; p += nthr) out[p] = 0.0;
            if (tid == 0) { if (a.k_out) a.k_out[star] = 0; if (a.status) a.status[star] = SHOPT_PCA_FAILED; }
            continue;
        }
        // ---- mean over the columns (sequential in j like mean(X, dims=2)), centre ----
        for (int i = tid; i < r; i += nthr) {
            double s = 0.0;
            for (int j = 0; j < c; j++) s += W[i + (size_t)j * r];
            mu[i] = s / c;
        }
        __syncthreads();
        double fro = 0.0;
        for (long long p = tid; p < rc; p += nthr) {
            const double z = W[p] - mu[(int)(p % r)];
            W[p] = z;
            fro = fma(z, z, fro);
        }
        // columns whose squared norm falls below eps^2 |Z|_F^2 are numerically zero (always c - r + 1 of them when
        // c > r): rotating them only shuffles round-off, so they are left alone and do not hold up convergence
        const double tiny = block_sum1<NW>(fro, scratch, buf) * (DBL_EPSILON * DBL_EPSILON);

        // ---- one-sided Jacobi sweeps ----
        int sweeps = 0, rotated = 1;
        while (rotated && sweeps < 60) {
            rotated = 0;
            for (int step = 0; step < nm1; step++) {
                for (int pr = warp; pr < npairs; pr += NW) {
                    int p, q;
                    if (pr == 0) { p = nm1; q = step; }
                    else { p = step + pr; if (p >= nm1) p -= nm1; q = step - pr; if (q < 0) q += nm1; }
                    if (p > q) { const int t = p; p = q; q = t; }
                    if (q >= c) continue;                       // the dummy position
                    double *cp = W + (size_t)p * r, *cq = W + (size_t)q * r;
                    double xa[RT], xb[RT];
                    double al = 0.0, be = 0.0, ga = 0.0;
#pragma unroll
                    for (int t = 0; t < RT; t++) {
                        const int i = lane + 32 * t;
                        xa[t] = (i < r) ? cp[i] : 0.0;
                        xb[t] = (i < r) ? cq[i] : 0.0;
                        al = fma(xa[t], xa[t], al);
                        be = fma(xb[t], xb[t], be);
                        ga = fma(xa[t], xb[t], ga);
                    }
#pragma unroll
                    for (int o = 16; o >= 1; o >>= 1) {
                        al += __shfl_xor_sync(0xffffffffu, al, o);
                        be += __shfl_xor_sync(0xffffffffu, be, o);
                        ga += __shfl_xor_sync(0xffffffffu, ga, o);
                    }
                    if (al <= tiny || be <= tiny) continue;
                    if (fabs(ga) <= DBL_EPSILON * sqrt(al) * sqrt(be)) continue;
                    rotated = 1;
                    const double zeta = (be - al) / (2.0 * ga);
                    const double tt = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                    const double cs = 1.0 / sqrt(1.0 + tt * tt), sn = cs * tt;
#pragma unroll
                    for (int t = 0; t < RT; t++) {
                        const int i = lane + 32 * t;
                        if (i < r) {
                            cp[i] = cs * xa[t] - sn * xb[t];
                            cq[i] = sn * xa[t] + cs * xb[t];
                        }
                    }
                }
                __syncthreads();
            }
            rotated = __syncthreads_or(rotated);
            sweeps++;
        }

        // ---- principal variances, stable descending rank, choose k (choose_pcadim) ----
        for (int j = tid; j < c; j += nthr) {
            double s = 0.0;
            const double *w = W + (size_t)j * r;
            for (int i = 0; i < r; i++) s = fma(w[i], w[i], s);
            sv[j] = s;
        }
        __syncthreads();
        for (int j = tid; j < c; j += nthr) {
            const double x = sv[j];
            int rank = 0;
            for (int i = 0; i < c; i++) rank += (sv[i] > x) || (sv[i] == x && i < j);
            ord[rank] = j;
        }
        __syncthreads();
        if (tid == 0) {
            const int m = r < c ? r : c;
            const double den = (double)(c - 1);
            double vsum = 0.0;
            for (int j = 0; j < m; j++) vsum += sv[ord[j]] / den;
            int k = 1;
            double acc = sv[ord[0]] / den;
            const double thres = vsum * a.pratio;
            while (k < m && k < a.ncomp && acc < thres) { acc += sv[ord[k]] / den; k++; }
            *s_flag = k;
        }
        __syncthreads();
        const int k = *s_flag;
        // ---- normalise the k leading columns: u_a = w_a / sigma_a ----
        for (int e = tid; e < k * r; e += nthr) {
            const int aa = e / r, i = e - aa * r;
            const int col = ord[aa];
            const double nrm = sqrt(sv[col]);
            if (nrm > 0.0) W[i + (size_t)col * r] /= nrm;
        }
        __syncthreads();
        // ---- Y = U_k' Z, Z re-read from the stamp (transform) ----
        for (int j = warp; j < c; j += NW) {
            double z[RT];
#pragma unroll
            for (int t = 0; t < RT; t++) {
                const int i = lane + 32 * t;
                z[t] = (i < r) ? nan_to_zero(__ldg(X + i + (size_t)j * r)) - mu[i] : 0.0;
            }
            for (int aa = 0; aa < k; aa++) {
                const double *ua = W + (size_t)ord[aa] * r;
                double s = 0.0;
#pragma unroll
                for (int t = 0; t < RT; t++) {
                    const int i = lane + 32 * t;
                    if (i < r) s = fma(ua[i], z[t], s);
                }
#pragma unroll
                for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
                if (lane == 0) Y[aa + (size_t)j * k] = s;
            }
        }
        __syncthreads();
        // ---- reconstruct: out = U_k Y + mean ----
        double part = 0.0;
        for (long long p = tid; p < rc; p += nthr) {
            const int j = (int)(p / r), i = (int)(p - (long long)j * r);
            double s = 0.0;
            for (int aa = 0; aa < k; aa++) s = fma(W[i + (size_t)ord[aa] * r], Y[aa + (size_t)j * k], s);
            s += mu[i];
            out[p] = s;
            part += s;
        }
        if (a.unity_sum) {                                   // shopt.jl:295: pca ./ sum(pca), no smoothing
            const double tot = block_sum1<NW>(part, scratch, buf);
            for (long long p = tid; p < rc; p += nthr) out[p] = out[p] / tot;   // same thread wrote it
        } else if (a.lanczos_a > 0) {                        // shopt.jl:299: smooth(pca, lanczos)
            __syncthreads();
            for (long long p = tid; p < rc; p += nthr) W[p] = out[p];            // same thread wrote it
            __syncthreads();
            const int la = a.lanczos_a, kw = 2 * la + 1;
            for (long long p = tid; p < rc; p += nthr) {
                const int j = (int)(p / r), i = (int)(p - (long long)j * r);
                if (isnan(W[p])) continue;
                const int dlo = i - la < 0 ? 0 : i - la, dhi = i + la > r - 1 ? r - 1 : i + la;
                const int elo = j - la < 0 ? 0 : j - la, ehi = j + la > c - 1 ? c - 1 : j + la;
                double norm = 0.0, val = 0.0;
                for (int di = dlo; di <= dhi; di++)
                    for (int dj = elo; dj <= ehi; dj++) {
                        const double d = W[di + (size_t)dj * r];
                        if (!isnan(d)) {
                            const double kk = a.kern[(di - i + la) + (dj - j + la) * kw];
                            val += kk * d;
                            norm += kk;
                        }
                    }
                out[p] = val / norm;
            }
        }
        if (tid == 0) {
            if (a.k_out) a.k_out[star] = k;
            if (a.status) a.status[star] = (rotated && sweeps >= 60) ? SHOPT_PCA_NOCONV : SHOPT_PCA_OK;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// smooth(data, a) alone (dataPreprocessing.jl:112-139); mode "smooth" applies it to nanToZero(star) (shopt.jl:316-348).
// One CTA per star; taps come straight from global memory (the 8 r c bytes of a stamp stay in L1/L2).
// ---------------------------------------------------------------------------------------------
struct SmoothArgs {
    const double *in;
    double *out;
    int r, c, la, nan_to_zero;
    double kern[15 * 15];
};
__global__ void __launch_bounds__(256) smooth_kernel(const SmoothArgs a) {
    const long long rc = (long long)a.r * a.c;
    const double *X = a.in + (size_t)blockIdx.x * rc;
    double *out = a.out + (size_t)blockIdx.x * rc;
    const int r = a.r, c = a.c, la = a.la, kw = 2 * la + 1;
    for (long long p = threadIdx.x; p < rc; p += 256) {
        const int j = (int)(p / r), i = (int)(p - (long long)j * r);
        double x = __ldg(X + p);
        if (a.nan_to_zero) x = nan_to_zero(x);
        if (isnan(x) || la <= 0) { out[p] = x; continue; }
        const int dlo = i - la < 0 ? 0 : i - la, dhi = i + la > r - 1 ? r - 1 : i + la;
        const int elo = j - la < 0 ? 0 : j - la, ehi = j + la > c - 1 ? c - 1 : j + la;
        double norm = 0.0, val = 0.0;
        for (int di = dlo; di <= dhi; di++)
            for (int dj = elo; dj <= ehi; dj++) {
                double d = __ldg(X + di + (size_t)dj * r);
                if (a.nan_to_zero) d = nan_to_zero(d);
                if (!isnan(d)) {
                    const double kk = a.kern[(di - i + la) + (dj - j + la) * kw];
                    val += kk * d;
                    norm += kk;
                }
            }
        out[p] = val / norm;
    }
}

// dataPreprocessing.jl:94-110: a*sinc(pi x)*sinc(pi x / a) at integer offsets, outer product, normalised by the sum
static void lanczos_kernel2d(int a, double *kernel) {
    const int w = 2 * a + 1;
    auto sinc = [](double x) { return x == 0 ? 1.0 : sin(x) / x; };
    auto lz = [&](double x) { return (fabs(x) >= a) ? 0.0 : (x == 0 ? 1.0 : a * sinc(M_PI * x) * sinc(M_PI * x / a)); };
    double sum = 0.0;
    for (int j = 0; j < w; j++)
        for (int i = 0; i < w; i++) kernel[i + j * w] = lz(i - a) * lz(j - a);
    for (int i = 0; i < w * w; i++) sum += kernel[i];
    for (int i = 0; i < w * w; i++) kernel[i] /= sum;
}

// ---- launch -----------------------------------------------------------------------------------
static size_t pca_fixed_smem(int nw, int r, int c) {
    return 64 + (size_t)2 * nw * 8 + (size_t)r * 8 + (size_t)c * 8 + (size_t)((c + 1) & ~1) * 4;
}
size_t pca_scratch_bytes(int r, int c, int ncomp) {
    int kmax = ncomp < r ? ncomp : r; if (c < kmax) kmax = c; if (kmax < 1) kmax = 1;
    const size_t wdoubles = (size_t)r * c + (size_t)kmax * c;
    if (pca_fixed_smem(32, r, c) + wdoubles * 8 <= 227 * 1024) return 0;
    return (size_t)sm_count() * wdoubles * 8;
}

template <int NW, int RT>
static int pca_launch_t(PcaArgs a, double *scratch, size_t scratch_bytes, cudaStream_t st) {
    const size_t wdoubles = (size_t)a.r * a.c + (size_t)a.kmax * a.c;
    size_t smem = pca_fixed_smem(NW, a.r, a.c) + wdoubles * 8;
    long long grid = sm_count();
    a.wglobal = nullptr; a.wstride = wdoubles;
    if (smem > 227 * 1024) {
        smem = pca_fixed_smem(NW, a.r, a.c);
        if ((size_t)grid * wdoubles * 8 > scratch_bytes) return SHOPT_E_NOMEM;
        a.wglobal = scratch;
    }
    auto kern = pca_kernel<NW, RT>;
    SHOPT_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (grid > a.N) grid = a.N;
    SHOPT_CUDA_TRY(cudaMemsetAsync(a.counter, 0, sizeof(unsigned long long), st));
    kern<<<(unsigned)grid, NW * 32, smem, st>>>(a);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}
template <int RT>
static int pca_launch_nw(const PcaArgs &a, double *scratch, size_t scratch_bytes, cudaStream_t st) {
    const int npairs = (a.c + 1) / 2;
    if (RT > 5) return pca_launch_t<16, RT>(a, scratch, scratch_bytes, st);
    // fewest (rounds x warps): 16, 22 or 32 warps
    const int cost16 = ((npairs + 15) / 16) * 16, cost22 = ((npairs + 21) / 22) * 22, cost32 = ((npairs + 31) / 32) * 32;
    if (RT <= 3 && cost32 < cost16 && cost32 < cost22) return pca_launch_t<32, RT>(a, scratch, scratch_bytes, st);
    if (cost22 < cost16) return pca_launch_t<22, RT>(a, scratch, scratch_bytes, st);
    return pca_launch_t<16, RT>(a, scratch, scratch_bytes, st);
}

int pca_launch(const double *d_stamps, int r, int c, long long N, int ncomp, double pratio, int lanczos_a, int unity_sum,
               double *d_pgf, int *d_k, int *d_status, unsigned long long *d_counter, double *scratch,
               size_t scratch_bytes, cudaStream_t st) {
    if (!d_stamps || !d_pgf || !d_counter || r < 1 || c < 2 || N < 0 || ncomp < 1 || lanczos_a > 7) return SHOPT_E_BADARG;
    if (N == 0) return SHOPT_OK;
    if (r > 320) return SHOPT_E_UNSUPPORTED;
    PcaArgs a;
    a.stamps = d_stamps; a.r = r; a.c = c; a.N = N; a.ncomp = ncomp; a.pratio = pratio;
    a.lanczos_a = lanczos_a; a.unity_sum = unity_sum; a.pgf = d_pgf; a.k_out = d_k; a.status = d_status;
    a.counter = d_counter; a.wglobal = nullptr; a.wstride = 0;
    a.kmax = ncomp < r ? ncomp : r; if (c < a.kmax) a.kmax = c;
    if (lanczos_a > 0) lanczos_kernel2d(lanczos_a, a.kern);
    const int rt = (r + 31) / 32;
    switch (rt) {
        case 1: return pca_launch_nw<1>(a, scratch, scratch_bytes, st);
        case 2: return pca_launch_nw<2>(a, scratch, scratch_bytes, st);
        case 3: return pca_launch_nw<3>(a, scratch, scratch_bytes, st);
        case 4: return pca_launch_nw<4>(a, scratch, scratch_bytes, st);
        case 5: return pca_launch_nw<5>(a, scratch, scratch_bytes, st);
        case 6: case 7: return pca_launch_nw<7>(a, scratch, scratch_bytes, st);
        default: return pca_launch_nw<10>(a, scratch, scratch_bytes, st);
    }
}

int smooth_launch(const double *d_in, int r, int c, long long N, int lanczos_a, int nan_to_zero_flag, double *d_out,
                  cudaStream_t st) {
    if (!d_in || !d_out || r < 1 || c < 1 || N < 0 || lanczos_a > 7) return SHOPT_E_BADARG;
    if (N == 0) return SHOPT_OK;
    SmoothArgs a;
    a.in = d_in; a.out = d_out; a.r = r; a.c = c; a.la = lanczos_a; a.nan_to_zero = nan_to_zero_flag;
    if (lanczos_a > 0) lanczos_kernel2d(lanczos_a, a.kern);
    smooth_kernel<<<(unsigned)N, 256, 0, st>>>(a);
    SHOPT_CUDA_TRY(cudaGetLastError());
    return SHOPT_OK;
}

}  // namespace shopt
