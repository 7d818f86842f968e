// comm.cu -- the library's own communicator for sharded runs (include/shopt_b200.h, shopt_comm_*): NCCL over
// NVLink 5 / NVSwitch, one rank per GPU.  The reference has no distributed mode; the only exchange
// steps of the stage are sum-all-reduces of small device buffers ([V'V | V'P], selection histograms, survivor bits),
// all enqueued on the stage's stream from C -- no host round trip, no Python in the loop.
//
// NCCL is bound at run time (dlopen of libnccl.so.2, reusing the copy the host program has already loaded, e.g.
// torch's) so that single-GPU users need no NCCL at all and the .so links only against cudart.
#include <dlfcn.h>
#include <nccl.h>

#include <cstdio>
#include <cstring>
#include <mutex>

#include "api_ctx.h"

struct shopt_comm {
    ncclComm_t comm;
    int rank, world, dev;
};

namespace shopt {

struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
};
static NcclApi g_nccl;
static std::mutex g_nccl_mutex;

static int nccl_load() {
    std::lock_guard<std::mutex> lock(g_nccl_mutex);
    if (g_nccl.handle) return SHOPT_OK;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);          // the host program's copy, if it has one
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { set_last_error_text("libnccl.so.2 not found (needed only for sharded runs)"); return SHOPT_E_UNSUPPORTED; }
#define BIND(field, name)                                                                  \
    *(void **)(&g_nccl.field) = dlsym(h, name);                                            \
    if (!g_nccl.field) { set_last_error_text("libnccl: missing symbol " name); return SHOPT_E_UNSUPPORTED; }
    BIND(GetUniqueId, "ncclGetUniqueId")
    BIND(CommInitRank, "ncclCommInitRank")
    BIND(CommDestroy, "ncclCommDestroy")
    BIND(AllReduce, "ncclAllReduce")
    BIND(GetErrorString, "ncclGetErrorString")
    BIND(GetVersion, "ncclGetVersion")
#undef BIND
    g_nccl.handle = h;
    return SHOPT_OK;
}
static int nccl_fail(const char *what, ncclResult_t r) {
    char msg[256];
    snprintf(msg, sizeof msg, "%s: %s", what, g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
    set_last_error_text(msg);
    return SHOPT_E_CUDA;
}

}  // namespace shopt

using namespace shopt;

extern "C" {

int shopt_comm_unique_id(void *id_out) {
    if (!id_out) return SHOPT_E_BADARG;
    TRY(nccl_load());
    static_assert(sizeof(ncclUniqueId) == SHOPT_UNIQUE_ID_BYTES, "ncclUniqueId size");
    ncclUniqueId id;
    ncclResult_t r = g_nccl.GetUniqueId(&id);
    if (r != ncclSuccess) return nccl_fail("ncclGetUniqueId", r);
    memcpy(id_out, &id, sizeof id);
    return SHOPT_OK;
}

int shopt_comm_init(const void *id_in, int rank, int world, shopt_comm **out) {
    if (!id_in || !out || world < 1 || rank < 0 || rank >= world) return SHOPT_E_BADARG;
    TRY(nccl_load());
    Ctx *ctx; TRY(get_ctx(&ctx));                                          // binds to the CURRENT device
    ncclUniqueId id;
    memcpy(&id, id_in, sizeof id);
    ncclComm_t c;
    ncclResult_t r = g_nccl.CommInitRank(&c, world, id, rank);
    if (r != ncclSuccess) return nccl_fail("ncclCommInitRank", r);
    *out = new shopt_comm{c, rank, world, ctx->dev};
    return SHOPT_OK;
}

int shopt_comm_destroy(shopt_comm *c) {
    if (!c) return SHOPT_OK;
    if (g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    delete c;
    return SHOPT_OK;
}

int shopt_comm_rank(const shopt_comm *c) { return c ? c->rank : SHOPT_E_BADARG; }
int shopt_comm_world(const shopt_comm *c) { return c ? c->world : SHOPT_E_BADARG; }

// a shopt_allreduce_fn: user = the shopt_comm*.  In place, sum, enqueued on `stream`.
int shopt_comm_allreduce(void *d_buf, int64_t count, int dtype, void *stream, void *user) {
    shopt_comm *c = static_cast<shopt_comm *>(user);
    if (!c || !d_buf || count < 0 || !g_nccl.AllReduce) return 1;
    if (count == 0 || c->world == 1) return 0;
    ncclResult_t r = g_nccl.AllReduce(d_buf, d_buf, (size_t)count, dtype == 0 ? ncclDouble : ncclInt32, ncclSum, c->comm,
                                      (cudaStream_t)stream);
    if (r != ncclSuccess) { nccl_fail("ncclAllReduce", r); return 1; }
    return 0;
}

}  // extern "C"
