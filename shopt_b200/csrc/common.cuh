// common.cuh -- shared device helpers for libshopt_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <stdint.h>
#include <math.h>
#include <float.h>

#include "../../include/shopt_b200.h"

namespace shopt {

// ---- host-side error plumbing ------------------------------------------------------------
void set_last_error(const char *what, cudaError_t e);
#define SHOPT_CUDA_TRY(expr)                                                         \
    do {                                                                             \
        cudaError_t _e = (expr);                                                     \
        if (_e != cudaSuccess) {                                                     \
            ::shopt::set_last_error(#expr, _e);                                      \
            return (_e == cudaErrorMemoryAllocation) ? SHOPT_E_NOMEM : SHOPT_E_CUDA; \
        }                                                                            \
    } while (0)

int sm_count();   // SMs of the current device (cached per device)

// kernels launched by this process (every *_launch adds what it enqueues): bench.py's `gpu_launches` is a count, not a
// constant
extern std::atomic<long long> g_launches;
inline void count_launches(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// ---- deterministic CTA-wide sum of NV doubles ----------------------------------------------
// Every thread returns with the same bits in v[]: xor-butterfly inside the warp (commutative, so all lanes agree), one
// partial per warp in shared memory, then every warp adds the warp partials with a second butterfly (lane w holds
// warp w's partial; NWARPS <= 32) -- a fixed tree, independent of timing.  `buf` alternates between two scratch areas so
// that one __syncthreads per call is enough (a thread can only reach call n+2 after every thread has left call n).
template <int NV, int NWARPS>
__device__ __forceinline__ void cta_sum(double (&v)[NV], double *scratch /* 2*NWARPS*NV */, int &buf) {
    static_assert(NWARPS <= 32, "one lane per warp partial");
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) {
#pragma unroll
        for (int i = 0; i < NV; i++) v[i] += __shfl_xor_sync(0xffffffffu, v[i], m);
    }
    double *s = scratch + buf * (NWARPS * NV);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < NV; i++) s[warp * NV + i] = v[i];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NV; i++) {
        double t = (lane < NWARPS) ? s[lane * NV + i] : 0.0;
#pragma unroll
        for (int m = 16; m >= 1; m >>= 1) t += __shfl_xor_sync(0xffffffffu, t, m);
        v[i] = t;
    }
    buf ^= 1;
}

// ---- mbarrier + 1-D bulk (TMA) copy, global -> shared ---------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// size and both addresses must be multiples of 16 bytes
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
// make generic-proxy reads of a buffer visible before the async proxy overwrites it
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// 8-byte asynchronous copy global -> shared (LDGSTS): no register holds the data in flight.  A lane that reads back only
// what it copied itself needs no barrier besides cp_async_wait.
__device__ __forceinline__ void cp_async8(double *smem_dst, const double *gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ bool finite_(double x) { return isfinite(x); }

}  // namespace shopt
