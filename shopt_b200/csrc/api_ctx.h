// api_ctx.h -- host-side context shared by the api_*.cu files: per-thread/per-device stream, timing events, work
// counters and a grow-only device scratch block; self-freeing device buffers; error plumbing.
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace shopt {

struct Ctx {
    int dev = -1;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;      // H2D of the pipelined host entry points
    static constexpr int NCHUNK_EV = 16;
    cudaEvent_t chunk_ev[NCHUNK_EV] = {};
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    unsigned long long *counter = nullptr;   // 8 slots, rotated per launch so overlapping launches never share one
    unsigned slot = 0;
    double ms[3] = {0, 0, 0};
    void *scratch = nullptr;                 // grow-only device scratch of the _dev building blocks
    size_t scratch_bytes = 0;
    // grow-only device blocks of the host entry points (stamp cube, pixel-grid cube, ...): allocating and freeing tens
    // of GB per call costs tens of milliseconds; shopt_release_buffers() gives them back
    static constexpr int NCACHE = 4;
    void *cache[NCACHE] = {};
    size_t cache_bytes[NCACHE] = {};
    int need_cache(int i, size_t bytes, void **out);
    void release_cache();
    unsigned long long *next_counter() { return counter + (slot++ & 7u); }
    // Work on a context is stream-ordered by the caller; growing frees the old block only after the device is idle.
    int need_scratch(size_t bytes);
};
int get_ctx(Ctx **out);
int finish_timing(Ctx *c);   // synchronise the context's stream and read ev[0..3] into ms[0..2]
void set_last_error_text(const char *text);

// device allocation that frees itself
struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    int alloc(size_t bytes) {
        if (bytes == 0) bytes = 16;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) { p = nullptr; set_last_error("cudaMalloc", e); (void)cudaGetLastError(); return SHOPT_E_NOMEM; }
        return SHOPT_OK;
    }
    template <class T> T *as() const { return static_cast<T *>(p); }
};
// non-owning view with DevBuf's accessor (blocks that live in the context's cache)
struct RawBuf {
    void *p = nullptr;
    template <class T> T *as() const { return static_cast<T *>(p); }
};
#define TRY(expr) do { int _rc = (expr); if (_rc != SHOPT_OK) return _rc; } while (0)

}  // namespace shopt
