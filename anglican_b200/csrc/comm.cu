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

int comm_allgather_dev(apop_ctx* c, const void* send, void* recv, size_t bytes);

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
    LOAD(GetErrorString);
#undef LOAD
    g_nccl.ok = g_nccl.GetUniqueId && g_nccl.CommInitRank && g_nccl.CommDestroy && g_nccl.AllGather;
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

// ------------------------------------------------------------------ fabric
Fabric fabric_view(const apop_ctx* c) {
    Fabric F;
    memset(&F, 0, sizeof F);
    F.rank = c->rank;
    F.world = c->world;
    F.n_local = c->n;
    {
        // read once, thread-safely (function-local static initialiser): contexts are driven
        // from several host threads
        static const long long cyc = [] {
            const char* e = getenv("APOP_SPIN_SECONDS");
            const double sec = (e && atof(e) > 0.0) ? atof(e) : 3.0;
            return (long long)(sec * 2.0e9);
        }();
        F.spin_cycles = cyc;
    }
    if (c->world > 1 && c->fab.linked)
        for (int q = 0; q < c->world; ++q) F.mbox[q] = (Mailbox*)c->fab.heap[q];
    return F;
}

// every rank raises its flag in every peer's mailbox, then waits for all flags
__global__ void k_fabric_barrier(Fabric F, unsigned long long seq, int* err) {
    __threadfence_system();
    const int q = threadIdx.x;
    if (q < F.world) {
        st_release_sys(&F.mbox[q]->bar[F.rank], seq);
        if (!spin_until(&F.mbox[F.rank]->bar[q], seq, F.spin_cycles) && err)
            atomicOr(err, EB_COMM | (CW_BARRIER << 16) | (q << 24));
    }
}

void preload_comm_kernels() {
    cudaFuncAttributes a;
    cudaFuncGetAttributes(&a, k_fabric_barrier);
}

int fabric_barrier_enqueue(apop_ctx* c) {
    if (c->world == 1) return APOP_OK;
    if (!c->fab.linked) return set_err(c, APOP_ERR_INVALID, "fabric not linked");
    k_fabric_barrier<<<1, 32, 0, c->stream>>>(fabric_view(c), ++c->fab.bar_epoch,
                                              c->ws.info ? &c->ws.info->err : nullptr);
    c->launches += 1;
    APOP_CUDA(c, cudaGetLastError());
    return APOP_OK;
}

static void fabric_unmap(apop_ctx* c) {
    FabricHost& f = c->fab;
    for (int q = 0; q < kMaxWorld; ++q) {
        if (q != c->rank && f.ipc_open[q] && f.heap[q]) cudaIpcCloseMemHandle(f.heap[q]);
        if (q != c->rank) { f.heap[q] = nullptr; f.ipc_open[q] = false; }
    }
    f.linked = false;
}

void fabric_free(apop_ctx* c) {
    FabricHost& f = c->fab;
    fabric_unmap(c);
    if (f.heap[c->rank]) cudaFree(f.heap[c->rank]);
    f.heap[c->rank] = nullptr;
    f.heap_bytes = 0;
    f.cap_n = 0;
}

// allocate this rank's symmetric heap for n_cap particles (layout identical on
// every rank: mailbox | xa | xb | anc)
static int fabric_alloc_own(apop_ctx* c, size_t n_cap) {
    FabricHost& f = c->fab;
    const size_t col = ((n_cap + 8) * 8 + 255) & ~(size_t)255;
    const size_t acol = ((n_cap + 8) * 4 + 255) & ~(size_t)255;
    f.off_xa = kMailboxBytes;
    f.off_xb = f.off_xa + col;
    f.off_anc = f.off_xb + col;
    const size_t bytes = f.off_anc + acol;
    cudaSetDevice(c->device);
    if (f.heap[c->rank]) {
        APOP_CUDA(c, cudaStreamSynchronize(c->stream));
        cudaFree(f.heap[c->rank]);
        f.heap[c->rank] = nullptr;
    }
    if (cudaMalloc((void**)&f.heap[c->rank], bytes) != cudaSuccess)
        return set_err(c, APOP_ERR_OOM, "fabric: heap allocation of %zu bytes failed", bytes);
    APOP_CUDA(c, cudaMemset(f.heap[c->rank], 0, kMailboxBytes));
    APOP_CUDA(c, cudaDeviceSynchronize());       // ctx streams do not order with the NULL stream
    f.heap_bytes = bytes;
    f.cap_n = n_cap;
    return APOP_OK;
}

// Collective over all ranks (same n_cap everywhere).  Multi-process: IPC handles
// are exchanged with one NCCL allgather and mapped with cudaIpcOpenMemHandle.
// In-process group (apop_link_group): pointers are shared directly, with peer
// access enabled between distinct devices.
int fabric_prepare(apop_ctx* c, size_t n_cap) {
    FabricHost& f = c->fab;
    if (c->world == 1) return APOP_OK;
    if (c->world > kMaxWorld) return set_err(c, APOP_ERR_UNSUPPORTED, "world > %d", kMaxWorld);
    if (f.linked && f.cap_n >= n_cap) return APOP_OK;
    if (f.in_process) {
        if (c->rank != 0) return APOP_OK;          // rank 0 of the group (re)builds everything
        for (apop_ctx* g : f.group) {
            fabric_unmap(g);
            APOP_TRY(fabric_alloc_own(g, n_cap));
        }
        for (apop_ctx* g : f.group) {
            cudaSetDevice(g->device);
            for (apop_ctx* h : f.group) {
                if (h == g) continue;
                if (h->device != g->device) {
                    cudaError_t e = cudaDeviceEnablePeerAccess(h->device, 0);
                    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                        return set_err(c, APOP_ERR_CUDA, "peer access %d -> %d: %s", g->device,
                                       h->device, cudaGetErrorString(e));
                    cudaGetLastError();
                }
                g->fab.heap[h->rank] = h->fab.heap[h->rank];
            }
            g->fab.linked = true;
            APOP_CUDA(g, cudaDeviceSynchronize());
        }
        return APOP_OK;
    }
    // ---- one process per GPU
    if (!c->nccl) return set_err(c, APOP_ERR_NCCL, "fabric: no communicator");
    fabric_unmap(c);
    APOP_TRY(fabric_alloc_own(c, n_cap));
    cudaIpcMemHandle_t mine;
    APOP_CUDA(c, cudaIpcGetMemHandle(&mine, f.heap[c->rank]));
    const size_t hb = sizeof(cudaIpcMemHandle_t);
    APOP_TRY(ensure_scratch(c, hb * (size_t)(c->world + 1)));
    char* base = (char*)c->scratch;
    APOP_CUDA(c, cudaMemcpyAsync(base, &mine, hb, cudaMemcpyHostToDevice, c->stream));
    APOP_TRY(comm_allgather_dev(c, base, base + hb, hb));
    std::vector<cudaIpcMemHandle_t> all((size_t)c->world);
    APOP_CUDA(c, cudaMemcpyAsync(all.data(), base + hb, hb * (size_t)c->world,
                                 cudaMemcpyDeviceToHost, c->stream));
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    for (int q = 0; q < c->world; ++q) {
        if (q == c->rank) continue;
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, all[(size_t)q], cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess)
            return set_err(c, APOP_ERR_CUDA, "cudaIpcOpenMemHandle(rank %d): %s", q,
                           cudaGetErrorString(e));
        f.heap[q] = (char*)p;
        f.ipc_open[q] = true;
    }
    f.linked = true;
    // nobody may touch a peer heap before every rank has mapped and zeroed its own
    APOP_TRY(comm_allgather_dev(c, base, base + hb, hb));
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
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
