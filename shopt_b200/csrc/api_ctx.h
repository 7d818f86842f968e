// api_ctx.h -- host-side context shared by the api_*.cu files: per-thread/per-device stream, timing events, work
// counters, grow-only device blocks (scratch, stage arena, stamp caches) and a grow-only pinned host block;
// self-freeing device buffers for the one-shot host entry points; error plumbing; the communicator of a sharded run.
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
    // grow-only device arena of the stage (indices, keys, histograms, normal equations, ...) and pinned host staging:
    // a stage call allocates and frees NOTHING once the first call of a given size has run (cudaFree synchronises the
    // device and cudaMalloc takes driver locks -- both showed up as 100+ ms of idle GPU in round 1)
    void *arena = nullptr;
    size_t arena_bytes = 0;
    void *arena2 = nullptr;                  // the host-buffer entry points' own block (small per-star arrays)
    size_t arena2_bytes = 0;
    void *pinned = nullptr;
    size_t pinned_bytes = 0;
    int need_cache(int i, size_t bytes, void **out);
    int need_arena(size_t bytes);
    int need_arena2(size_t bytes);
    int need_pinned(size_t bytes);
    void release_cache();
    unsigned long long *next_counter() { return counter + (slot++ & 7u); }
    // Work on a context is stream-ordered by the caller; growing frees the old block only after the device is idle.
    int need_scratch(size_t bytes);
};
int get_ctx(Ctx **out);
int finish_timing(Ctx *c);   // synchronise the context's stream and read ev[0..3] into ms[0..2]
void set_last_error_text(const char *text);

// bump allocator over a block that is already large enough (sizes are computed twice: once to size the block)
struct Bump {
    char *base = nullptr;
    size_t off = 0;
    explicit Bump(void *b = nullptr) : base(static_cast<char *>(b)) {}
    template <class T> T *take(size_t count) {
        const size_t bytes = (count * sizeof(T) + 255) & ~(size_t)255;
        T *p = base ? reinterpret_cast<T *>(base + off) : nullptr;
        off += bytes;
        return p;
    }
};

// device allocation that frees itself (one-shot host entry points only; the stage uses the arena)
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

// The exchange of a sharded run: a sum-all-reduce on a device buffer, in stream order.  Either the caller's callback
// (shopt_allreduce_fn) or the library's own NCCL communicator (comm.cu; shopt_comm_allreduce is such a callback).
struct Exchange {
    shopt_allreduce_fn fn = nullptr;
    void *user = nullptr;
    void *stream = nullptr;
    bool on() const { return fn != nullptr; }
    int sum(void *d_buf, long long count, int dtype) const {
        if (!fn) return SHOPT_OK;
        if (fn(d_buf, (int64_t)count, dtype, stream, user) != 0) { set_last_error_text("all-reduce failed"); return SHOPT_E_CUDA; }
        return SHOPT_OK;
    }
};

}  // namespace shopt
