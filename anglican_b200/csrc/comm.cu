// comm.cu — inter-rank plumbing.  The reference has no multi-process code at
// all (the only cross-"node" exchange is the Clojure vector of
// log-Zs in ipmcmc.cljc:183-187); this layer is new.  NCCL is bound at run time
// with dlopen so that libanglican_pop.so loads (and its symbol table can be
// checked) on machines without NCCL; a world > 1 context without NCCL fails
// loudly with APOP_ERR_NCCL.
#include <dlfcn.h>
#include <stdlib.h>
#include <string.h>

#include "apop_internal.cuh"

namespace apop {

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclInt8 = 0, ncclChar = 0, ncclInt32 = 2, ncclInt64 = 4, ncclFloat64 = 8 };

struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool ok = false;
};

static NcclApi g_nccl;

static bool nccl_load() {
    if (g_nccl.ok) return true;
    if (g_nccl.handle) return false;
    const char* env = getenv("APOP_NCCL_LIB");
    const char* names[] = {env, "libnccl.so.2", "libnccl.so", nullptr};
    void* h = nullptr;
    // prefer an already-loaded copy (e.g. the one bundled with torch)
    for (int i = 0; i < 3 && !h; ++i)
        if (names[i]) h = dlopen(names[i], RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    for (int i = 0; i < 3 && !h; ++i)
        if (names[i]) h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
    if (!h) return false;
    g_nccl.handle = h;
#define LOAD(sym) *(void**)(&g_nccl.sym) = dlsym(h, "nccl" #sym)
    LOAD(GetUniqueId); LOAD(CommInitRank); LOAD(CommDestroy); LOAD(AllGather);
    LOAD(Send); LOAD(Recv); LOAD(GroupStart); LOAD(GroupEnd); LOAD(GetErrorString);
#undef LOAD
    g_nccl.ok = g_nccl.GetUniqueId && g_nccl.CommInitRank && g_nccl.CommDestroy &&
                g_nccl.AllGather && g_nccl.Send && g_nccl.Recv && g_nccl.GroupStart &&
                g_nccl.GroupEnd;
    return g_nccl.ok;
}

#define APOP_NCCL(ctx, call)                                                               \
    do {                                                                                   \
        int r__ = (call);                                                                  \
        if (r__ != ncclSuccess)                                                            \
            return set_err(ctx, APOP_ERR_NCCL, "%s failed: %s", #call,                     \
                           g_nccl.GetErrorString ? g_nccl.GetErrorString(r__) : "?");      \
    } while (0)

int comm_get_unique_id(void* out) {
    if (!out) return APOP_ERR_INVALID;
    if (!nccl_load()) return APOP_ERR_NCCL;
    ncclUniqueId id;
    if (g_nccl.GetUniqueId(&id) != ncclSuccess) return APOP_ERR_NCCL;
    memcpy(out, &id, sizeof id);
    return APOP_OK;
}

int comm_init(apop_ctx* c, const void* nccl_id) {
    if (!nccl_id) return set_err(c, APOP_ERR_INVALID, "world > 1 needs an NCCL unique id");
    if (!nccl_load()) return set_err(c, APOP_ERR_NCCL, "libnccl.so.2 could not be loaded");
    ncclUniqueId id;
    memcpy(&id, nccl_id, sizeof id);
    ncclComm_t comm = nullptr;
    APOP_NCCL(c, g_nccl.CommInitRank(&comm, c->world, id, c->rank));
    c->nccl = comm;
    return APOP_OK;
}

void comm_free(apop_ctx* c) {
    if (c->nccl && g_nccl.ok) g_nccl.CommDestroy((ncclComm_t)c->nccl);
    c->nccl = nullptr;
}

// device-buffer allgather of `bytes` per rank on the ctx stream
int comm_allgather_dev(apop_ctx* c, const void* send, void* recv, size_t bytes) {
    if (c->world == 1) {
        if (send != recv)
            APOP_CUDA(c, cudaMemcpyAsync(recv, send, bytes, cudaMemcpyDeviceToDevice, c->stream));
        return APOP_OK;
    }
    APOP_NCCL(c, g_nccl.AllGather(send, recv, bytes, ncclChar, (ncclComm_t)c->nccl, c->stream));
    return APOP_OK;
}

// grouped point-to-point exchange: for each peer, send_bytes[p] from send_ptr[p]
// and recv_bytes[p] into recv_ptr[p] (zero-byte entries skipped)
int comm_exchange_dev(apop_ctx* c, void* const* send_ptr, const size_t* send_bytes,
                      void* const* recv_ptr, const size_t* recv_bytes) {
    if (c->world == 1) return APOP_OK;
    APOP_NCCL(c, g_nccl.GroupStart());
    for (int p = 0; p < c->world; ++p) {
        if (p == c->rank) continue;
        if (send_bytes[p])
            APOP_NCCL(c, g_nccl.Send(send_ptr[p], send_bytes[p], ncclChar, p, (ncclComm_t)c->nccl,
                                     c->stream));
        if (recv_bytes[p])
            APOP_NCCL(c, g_nccl.Recv(recv_ptr[p], recv_bytes[p], ncclChar, p, (ncclComm_t)c->nccl,
                                     c->stream));
    }
    APOP_NCCL(c, g_nccl.GroupEnd());
    return APOP_OK;
}

int comm_allgather_f64(apop_ctx* c, const double* send_host, int count, double* recv_host) {
    const size_t bytes = sizeof(double) * (size_t)count;
    APOP_TRY(ensure_scratch(c, bytes * (size_t)(c->world + 1)));
    char* base = (char*)c->scratch;
    APOP_CUDA(c, cudaMemcpyAsync(base, send_host, bytes, cudaMemcpyHostToDevice, c->stream));
    APOP_TRY(comm_allgather_dev(c, base, base + bytes, bytes));
    APOP_CUDA(c, cudaMemcpyAsync(recv_host, base + bytes, bytes * (size_t)c->world,
                                 cudaMemcpyDeviceToHost, c->stream));
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    return APOP_OK;
}

} // namespace apop
