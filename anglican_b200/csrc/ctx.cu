// ctx.cu — context lifecycle, columns, and the weight / resample / gather
// entry points of the C ABI (include/anglican_pop.h).
#include <dlfcn.h>
#include <math.h>
#include <stdarg.h>
#include <string.h>

#include "apop_internal.cuh"
#include "rng.cuh"

namespace apop {

int set_err(apop_ctx* c, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c) c->err = buf;
    return code;
}

void kt_begin(apop_ctx* c, int tag) {
    if (!c->ktimers) return;
    Timer& t = c->timers[tag];
    if (t.open) return;
    if (cudaEventCreate(&t.a) != cudaSuccess) return;
    cudaEventRecord(t.a, c->stream);
    t.open = true;
}
void kt_end(apop_ctx* c, int tag) {
    if (!c->ktimers) return;
    Timer& t = c->timers[tag];
    if (!t.open) return;
    cudaEvent_t b;
    if (cudaEventCreate(&b) != cudaSuccess) return;
    cudaEventRecord(b, c->stream);
    t.pending.push_back({t.a, b});
    t.a = nullptr;
    t.open = false;
}

Column* get_col(apop_ctx* c, apop_col h, int dtype, int width) {
    if (h < 0 || (size_t)h >= c->cols.size() || !c->cols[h].used) {
        set_err(c, APOP_ERR_INVALID, "bad column handle %d", (int)h);
        return nullptr;
    }
    Column* col = &c->cols[h];
    if (dtype >= 0 && col->dtype != dtype) {
        set_err(c, APOP_ERR_INVALID, "column %d has dtype %d, expected %d", (int)h, col->dtype,
                dtype);
        return nullptr;
    }
    if (width > 0 && col->width != width) {
        set_err(c, APOP_ERR_INVALID, "column %d has width %d, expected %d", (int)h, col->width,
                width);
        return nullptr;
    }
    return col;
}

int ensure_scratch(apop_ctx* c, size_t bytes) {
    if (c->scratch_bytes >= bytes) return APOP_OK;
    if (c->scratch) {
        APOP_CUDA(c, cudaStreamSynchronize(c->stream));
        APOP_CUDA(c, cudaFree(c->scratch));
        c->scratch = nullptr;
        c->scratch_bytes = 0;
    }
    cudaError_t e = cudaMalloc(&c->scratch, bytes);
    if (e != cudaSuccess) return set_err(c, APOP_ERR_OOM, "scratch alloc of %zu bytes failed", bytes);
    c->scratch_bytes = bytes;
    return APOP_OK;
}

// small parameter arena: host arrays (distribution parameters) are staged
// through pinned memory into a device ring so calls stay asynchronous.
int param_upload(apop_ctx* c, const void* host, size_t bytes, void** dev) {
    ParamArena* a = &c->arena;
    const size_t cap = 4u << 20;
    if (!a->dev) {
        APOP_CUDA(c, cudaMalloc((void**)&a->dev, cap));
        APOP_CUDA(c, cudaMallocHost((void**)&a->host, cap));
        a->cap = cap;
        a->off = 0;
    }
    const size_t need = (bytes + 255) & ~(size_t)255;
    if (need > a->cap) return set_err(c, APOP_ERR_INVALID, "parameter block too large");
    if (a->off + need > a->cap) {
        APOP_CUDA(c, cudaStreamSynchronize(c->stream));
        a->off = 0;
    }
    memcpy(a->host + a->off, host, bytes);
    APOP_CUDA(c, cudaMemcpyAsync(a->dev + a->off, a->host + a->off, bytes, cudaMemcpyHostToDevice,
                                 c->stream));
    *dev = a->dev + a->off;
    a->off += need;
    return APOP_OK;
}

static void arena_free(apop_ctx* c) {
    if (c->arena.dev) cudaFree(c->arena.dev);
    if (c->arena.host) cudaFreeHost(c->arena.host);
}

void sweep_free(apop_ctx* c);          // ssm.cu
int comm_init(apop_ctx* c, const void* nccl_id);   // comm.cu
void comm_free(apop_ctx* c);
int comm_get_unique_id(void* out);
int comm_allgather_f64(apop_ctx* c, const double* send_host, int count, double* recv_host);

} // namespace apop

using namespace apop;

#define CHECK_CTX(c) do { if (!(c)) return APOP_ERR_INVALID; } while (0)

extern "C" {

int apop_version(void) { return 100; }

int apop_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int apop_create(apop_ctx** out, int64_t n_local, uint64_t seed, int device, int rank, int world,
                const void* nccl_id) {
    if (!out || n_local <= 0 || world < 1 || rank < 0 || rank >= world) return APOP_ERR_INVALID;
    if (n_local >= (1LL << 31)) return APOP_ERR_INVALID;
    // Every population owns a stream, and the kernels of a sharded population wait on
    // peers: two populations whose streams share a hardware queue can block each other
    // across ranks (A behind B on one rank, B behind A on the next).  Ask for one queue
    // per stream (32 instead of the default 8); only effective if this call creates the
    // CUDA context - a host that initialises CUDA first sets the variable itself.
    setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) return APOP_ERR_CUDA;
    if (device < 0 || device >= ndev) return APOP_ERR_INVALID;
    if (cudaSetDevice(device) != cudaSuccess) return APOP_ERR_CUDA;
    apop_ctx* c = new apop_ctx();
    c->device = device;
    {
        int sms = 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && sms > 0)
            c->sm_count = sms;
    }
    c->rank = rank;
    c->world = world;
    c->n = n_local;
    c->n_global = n_local * world;
    c->g0 = n_local * rank;
    c->seed = seed;
    // the engine stream takes the greatest priority
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    if (cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prio_hi) != cudaSuccess) {
        delete c;
        return APOP_ERR_CUDA;
    }
    if (world > 1 && nccl_id) {
        int rc = comm_init(c, nccl_id);
        if (rc != APOP_OK) {
            cudaStreamDestroy(c->stream);
            delete c;
            return rc;
        }
    }
    // world > 1 without an NCCL id: a rank of an in-process group, to be wired
    // with apop_link_group before its first sweep
    *out = c;
    return APOP_OK;
}

int apop_link_group(apop_ctx* const* ctxs, int32_t world) {
    if (!ctxs || world < 2 || world > kMaxWorld) return APOP_ERR_INVALID;
    for (int i = 0; i < world; ++i) {
        apop_ctx* c = ctxs[i];
        if (!c) return APOP_ERR_INVALID;
        if (c->world != world || c->rank != i || c->n != ctxs[0]->n || c->nccl)
            return set_err(ctxs[0], APOP_ERR_INVALID,
                           "link_group: need ranks 0..world-1 of equal size, created without an "
                           "NCCL id, in rank order");
    }
    for (int i = 0; i < world; ++i) {
        FabricHost& f = ctxs[i]->fab;
        f.in_process = true;
        f.linked = false;
        f.group.assign(ctxs, ctxs + world);
    }
    return APOP_OK;
}

int apop_destroy(apop_ctx* c) {
    CHECK_CTX(c);
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    sweep_free(c);
    fabric_free(c);
    comm_free(c);
    for (auto& col : c->cols)
        if (col.used && col.ptr) { cudaFree(col.ptr); if (col.alt) cudaFree(col.alt); }
    ResampleWS& w = c->ws;
    if (w.slab) cudaFree(w.slab);
    if (w.edges) cudaFree(w.edges);
    if (w.info_host) cudaFreeHost(w.info_host);
    if (c->scratch) cudaFree(c->scratch);
    for (auto& t : c->timers) {
        for (auto& pr : t.pending) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
        if (t.a) cudaEventDestroy(t.a);
    }
    arena_free(c);
    cudaStreamDestroy(c->stream);
    delete c;
    return APOP_OK;
}

const char* apop_last_error(apop_ctx* c) { return c ? c->err.c_str() : "null context"; }

int apop_get_unique_id(void* out) { return comm_get_unique_id(out); }

int apop_synchronize(apop_ctx* c) {
    CHECK_CTX(c);
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    return APOP_OK;
}

int apop_host_alloc(apop_ctx* c, void** out, uint64_t bytes) {
    CHECK_CTX(c);
    if (!out || bytes == 0) return set_err(c, APOP_ERR_INVALID, "host_alloc: bad arguments");
    cudaSetDevice(c->device);
    if (cudaHostAlloc(out, (size_t)bytes, cudaHostAllocDefault) != cudaSuccess) {
        *out = nullptr;
        cudaGetLastError();
        return set_err(c, APOP_ERR_OOM, "host_alloc: %llu pinned bytes unavailable",
                       (unsigned long long)bytes);
    }
    return APOP_OK;
}

int apop_host_free(apop_ctx* c, void* p) {
    CHECK_CTX(c);
    if (p) APOP_CUDA(c, cudaFreeHost(p));
    return APOP_OK;
}

int64_t apop_num_particles(apop_ctx* c) { return c ? c->n : 0; }
int64_t apop_launch_count(apop_ctx* c) { return c ? c->launches : 0; }
void* apop_stream(apop_ctx* c) { return c ? (void*)c->stream : nullptr; }

int apop_enable_kernel_timers(apop_ctx* c, int on) {
    CHECK_CTX(c);
    c->ktimers = on != 0;
    return APOP_OK;
}

int apop_set_seed(apop_ctx* c, uint64_t seed) {
    CHECK_CTX(c);
    c->seed = seed;
    return APOP_OK;
}

int apop_set_step(apop_ctx* c, uint32_t step) {
    CHECK_CTX(c);
    c->step = step;
    return APOP_OK;
}

// ---- timers: pairs of events recorded on the ctx stream -------------------
int apop_timer_begin(apop_ctx* c, int tag) {
    CHECK_CTX(c);
    if (tag < 0 || tag >= 16) return APOP_ERR_INVALID;
    Timer& t = c->timers[tag];
    if (t.open) return set_err(c, APOP_ERR_INVALID, "timer %d already open", tag);
    APOP_CUDA(c, cudaEventCreate(&t.a));
    APOP_CUDA(c, cudaEventRecord(t.a, c->stream));
    t.open = true;
    return APOP_OK;
}
int apop_timer_end(apop_ctx* c, int tag) {
    CHECK_CTX(c);
    if (tag < 0 || tag >= 16) return APOP_ERR_INVALID;
    Timer& t = c->timers[tag];
    if (!t.open) return set_err(c, APOP_ERR_INVALID, "timer %d not open", tag);
    cudaEvent_t b;
    APOP_CUDA(c, cudaEventCreate(&b));
    APOP_CUDA(c, cudaEventRecord(b, c->stream));
    t.pending.push_back({t.a, b});
    t.a = nullptr;
    t.open = false;
    return APOP_OK;
}
int apop_timer_elapsed_ms(apop_ctx* c, int tag, double* total_ms, int64_t* count) {
    CHECK_CTX(c);
    if (tag < 0 || tag >= 16) return APOP_ERR_INVALID;
    Timer& t = c->timers[tag];
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    for (auto& pr : t.pending) {
        float ms = 0.f;
        APOP_CUDA(c, cudaEventElapsedTime(&ms, pr.first, pr.second));
        t.total_ms += ms;
        t.count += 1;
        cudaEventDestroy(pr.first);
        cudaEventDestroy(pr.second);
    }
    t.pending.clear();
    if (total_ms) *total_ms = t.total_ms;
    if (count) *count = t.count;
    return APOP_OK;
}
int apop_timer_reset(apop_ctx* c) {
    CHECK_CTX(c);
    for (int i = 0; i < 16; ++i) {
        double d; int64_t n;
        apop_timer_elapsed_ms(c, i, &d, &n);
        c->timers[i].total_ms = 0;
        c->timers[i].count = 0;
    }
    return APOP_OK;
}

// ---- columns ------------------------------------------------------------
int apop_column_alloc(apop_ctx* c, int dtype, int width, apop_col* out) {
    CHECK_CTX(c);
    if (!out || width <= 0 || (dtype != APOP_F64 && dtype != APOP_I32))
        return set_err(c, APOP_ERR_INVALID, "column_alloc: bad arguments");
    cudaSetDevice(c->device);
    const size_t esz = dtype == APOP_F64 ? 8 : 4;
    // one spare row: conditional SMC resamples N-1 and re-inserts the retained particle
    const size_t bytes = (size_t)(c->n + 1) * width * esz;
    void* p = nullptr;
    if (cudaMalloc(&p, bytes) != cudaSuccess)
        return set_err(c, APOP_ERR_OOM, "column_alloc: %zu bytes", bytes);
    APOP_CUDA(c, cudaMemsetAsync(p, 0, bytes, c->stream));
    size_t h = 0;
    for (; h < c->cols.size(); ++h)
        if (!c->cols[h].used) break;
    if (h == c->cols.size()) c->cols.push_back(Column());
    c->cols[h].ptr = p;
    c->cols[h].dtype = dtype;
    c->cols[h].width = width;
    c->cols[h].used = true;
    *out = (apop_col)h;
    return APOP_OK;
}

int apop_column_free(apop_ctx* c, apop_col h) {
    CHECK_CTX(c);
    Column* col = get_col(c, h, -1, 0);
    if (!col) return APOP_ERR_INVALID;
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    cudaFree(col->ptr);
    if (col->alt) cudaFree(col->alt);
    *col = Column();
    return APOP_OK;
}

int apop_column_upload(apop_ctx* c, apop_col h, const void* host, int64_t n_values) {
    CHECK_CTX(c);
    Column* col = get_col(c, h, -1, 0);
    if (!col) return APOP_ERR_INVALID;
    if (n_values < 0 || n_values > c->n * col->width)
        return set_err(c, APOP_ERR_INVALID, "upload: n_values out of range");
    const size_t esz = col->dtype == APOP_F64 ? 8 : 4;
    APOP_CUDA(c, cudaMemcpyAsync(col->ptr, host, (size_t)n_values * esz, cudaMemcpyHostToDevice,
                                 c->stream));
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    return APOP_OK;
}

int apop_column_download(apop_ctx* c, apop_col h, void* host, int64_t n_values) {
    CHECK_CTX(c);
    Column* col = get_col(c, h, -1, 0);
    if (!col) return APOP_ERR_INVALID;
    if (n_values < 0 || n_values > c->n * col->width)
        return set_err(c, APOP_ERR_INVALID, "download: n_values out of range");
    const size_t esz = col->dtype == APOP_F64 ? 8 : 4;
    APOP_CUDA(c, cudaMemcpyAsync(host, col->ptr, (size_t)n_values * esz, cudaMemcpyDeviceToHost,
                                 c->stream));
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    return APOP_OK;
}

} // extern "C"

namespace apop {
__global__ void k_fill_f64(double* p, int64_t n, double v) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x)
        p[i] = v;
}
__global__ void k_fill_i32(int32_t* p, int64_t n, int32_t v) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x)
        p[i] = v;
}
} // namespace apop

extern "C" {

int apop_column_fill(apop_ctx* c, apop_col h, double value) {
    CHECK_CTX(c);
    Column* col = get_col(c, h, -1, 0);
    if (!col) return APOP_ERR_INVALID;
    const int64_t total = c->n * col->width;
    int blocks = (int)((total + 255) / 256);
    if (blocks > c->sm_count * 32) blocks = c->sm_count * 32;
    if (col->dtype == APOP_F64)
        k_fill_f64<<<blocks, 256, 0, c->stream>>>((double*)col->ptr, total, value);
    else
        k_fill_i32<<<blocks, 256, 0, c->stream>>>((int32_t*)col->ptr, total, (int32_t)value);
    c->launches += 1;
    APOP_CUDA(c, cudaGetLastError());
    return APOP_OK;
}

int apop_column_device_ptr(apop_ctx* c, apop_col h, void** out) {
    CHECK_CTX(c);
    Column* col = get_col(c, h, -1, 0);
    if (!col || !out) return APOP_ERR_INVALID;
    *out = col->ptr;
    return APOP_OK;
}

// ---- weights ------------------------------------------------------------
int apop_normalize(apop_ctx* c, apop_col logw, double* max_lw, double* log_mean_w, double* ess) {
    CHECK_CTX(c);
    Column* l = get_col(c, logw, APOP_F64, 1);
    if (!l) return APOP_ERR_INVALID;
    if (c->world > 1) return set_err(c, APOP_ERR_UNSUPPORTED, "normalize: use the sweep API when sharded");
    ResampleArgs a{};
    a.lw = (const double*)l->ptr; a.w = nullptr; a.n = c->n; a.n_new = c->n; a.u = 0.0;
    a.order = APOP_ORDER_SORTED; a.anc_out = nullptr; a.out_offset = 0; a.have_partials = false;
    APOP_TRY(resample_enqueue(c, a));
    ResampleInfo I;
    APOP_TRY(resample_fetch_info(c, &I));
    if (I.err & EB_NAN) return info_to_status(c, I);
    if (max_lw) *max_lw = I.m;
    if (log_mean_w) *log_mean_w = I.log_mean_approx;
    if (ess) *ess = I.ess;
    return APOP_OK;
}

static int do_resample(apop_ctx* c, const double* lw, const double* w, int scheme, double u,
                       int64_t n_new, int order, apop_col anc_out, ResampleInfo* I) {
    if (c->world > 1)
        return set_err(c, APOP_ERR_UNSUPPORTED, "resample: use the sweep API when sharded");
    if (n_new <= 0 || n_new > c->n)
        return set_err(c, APOP_ERR_INVALID, "resample: n_new must be in [1, N]");
    Column* a = get_col(c, anc_out, APOP_I32, 1);
    if (!a) return APOP_ERR_INVALID;
    if (scheme == APOP_SYSTEMATIC) {
        ResampleArgs r{};
        r.lw = lw; r.w = w; r.n = c->n; r.n_new = n_new; r.u = u; r.order = order;
        r.anc_out = (int32_t*)a->ptr; r.out_offset = 0; r.have_partials = false;
        APOP_TRY(resample_enqueue(c, r));
    } else if (scheme == APOP_MULTINOMIAL) {
        APOP_TRY(multinomial_enqueue(c, lw, w, nullptr, n_new, (int32_t*)a->ptr));
    } else {
        return set_err(c, APOP_ERR_INVALID, "resample: unknown scheme %d", scheme);
    }
    APOP_TRY(resample_fetch_info(c, I));
    return info_to_status(c, *I);
}

int apop_resample(apop_ctx* c, apop_col logw, int scheme, double u, int64_t n_new, int order,
                  apop_col anc_out, double* new_logw, int64_t* n_first_lap) {
    CHECK_CTX(c);
    Column* l = get_col(c, logw, APOP_F64, 1);
    if (!l) return APOP_ERR_INVALID;
    ResampleInfo I;
    APOP_TRY(do_resample(c, (const double*)l->ptr, nullptr, scheme, u, n_new, order, anc_out, &I));
    if (new_logw) *new_logw = I.new_lw;
    if (n_first_lap) *n_first_lap = I.J;
    return APOP_OK;
}

int apop_resample_weights(apop_ctx* c, apop_col weights, int scheme, double u, int64_t n_new,
                          int order, apop_col anc_out, double* total, int64_t* n_first_lap) {
    CHECK_CTX(c);
    Column* w = get_col(c, weights, APOP_F64, 1);
    if (!w) return APOP_ERR_INVALID;
    ResampleInfo I;
    APOP_TRY(do_resample(c, nullptr, (const double*)w->ptr, scheme, u, n_new, order, anc_out, &I));
    if (total) *total = I.total;
    if (n_first_lap) *n_first_lap = I.J;
    return APOP_OK;
}

int apop_resample_multinomial_u(apop_ctx* c, apop_col weights, apop_col uniforms, int64_t n_new,
                                apop_col anc_out) {
    CHECK_CTX(c);
    Column* w = get_col(c, weights, APOP_F64, 1);
    Column* uu = get_col(c, uniforms, APOP_F64, 1);
    Column* a = get_col(c, anc_out, APOP_I32, 1);
    if (!w || !uu || !a) return APOP_ERR_INVALID;
    if (n_new <= 0 || n_new > c->n) return set_err(c, APOP_ERR_INVALID, "n_new out of range");
    APOP_TRY(multinomial_enqueue(c, nullptr, (const double*)w->ptr, (const double*)uu->ptr, n_new,
                                 (int32_t*)a->ptr));
    ResampleInfo I;
    APOP_TRY(resample_fetch_info(c, &I));
    return info_to_status(c, I);
}

int apop_debug_chain_trace(apop_ctx* c, double* out512) {
    CHECK_CTX(c);
    if (!out512 || !c->ws.trace) return APOP_ERR_INVALID;
    APOP_CUDA(c, cudaMemcpyAsync(out512, c->ws.trace, 512 * sizeof(double), cudaMemcpyDeviceToHost,
                                 c->stream));
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    return APOP_OK;
}

int apop_debug_resolve_phases(apop_ctx* c, double* out8) {
    CHECK_CTX(c);
    if (!out8 || !c->ws.info) return APOP_ERR_INVALID;
    ResampleInfo I;
    APOP_TRY(resample_fetch_info(c, &I));
    for (int i = 0; i < 4; ++i) out8[i] = I.wx_sum[i];
    out8[4] = I.nh0; out8[5] = I.nh1; out8[6] = I.first_skip1; out8[7] = I.asq;
    return APOP_OK;
}

int apop_gather(apop_ctx* c, apop_col anc, int64_t n_new, const apop_col* cols, int ncols) {
    CHECK_CTX(c);
    Column* a = get_col(c, anc, APOP_I32, 1);
    if (!a || !cols || ncols <= 0) return APOP_ERR_INVALID;
    if (n_new <= 0 || n_new > c->n) return set_err(c, APOP_ERR_INVALID, "gather: n_new out of range");
    for (int i = 0; i < ncols; ++i) {
        Column* col = get_col(c, cols[i], -1, 0);
        if (!col) return APOP_ERR_INVALID;
        if (col == a) return set_err(c, APOP_ERR_INVALID, "gather: cannot gather the index column");
        APOP_TRY(gather_enqueue(c, (const int32_t*)a->ptr, n_new, col));
    }
    return APOP_OK;
}

// ---- PIMH / IPMCMC host arithmetic ----------------------------------------
// accept?, lmh.cljc:146-151
int apop_pimh_accept(double next_log_z, double prev_log_z, double uniform) {
    return (prev_log_z == -INFINITY) || ((next_log_z - prev_log_z) > log(uniform));
}

// gibbs-update-csmc-indices, ipmcmc.cljc:106-146, with norm-exp (85-104).
// Sampling rule: ipmcmc.cljc:133 tests `(> 0.0 (apply max ps))`, which is never
// true for probabilities, so as written the reference always draws from
// uniform-discrete.  ipmcmc_test.clj:33-45 pins the INTENDED rule (draw
// proportional to ps when any ps > 0), which is what is implemented here.
int apop_ipmcmc_gibbs(const double* log_zs, int m, int p, const double* uniforms,
                      int32_t* csmc_idx, double* zeta_sums) {
    if (!log_zs || !uniforms || !csmc_idx || !zeta_sums || m < 2 || p < 1 || p >= m)
        return APOP_ERR_INVALID;
    const int n_smc = m - p, np = n_smc + 1;
    std::vector<int32_t> smc(n_smc);
    std::vector<double> ps(np);
    for (int i = 0; i < p; ++i) csmc_idx[i] = i;
    for (int i = 0; i < n_smc; ++i) smc[i] = p + i;
    for (int i = 0; i < m; ++i) zeta_sums[i] = 0.0;
    for (int i = 0; i < p; ++i) {
        auto lz = [&](int j) { return j < n_smc ? log_zs[smc[j]] : log_zs[i]; };
        double mx = lz(0);
        for (int j = 1; j < np; ++j) mx = lz(j) > mx ? lz(j) : mx;
        double total = 0.0;
        if (mx == -INFINITY) {
            for (int j = 0; j < np; ++j) ps[j] = 0.0;
        } else {
            for (int j = 0; j < np; ++j) { ps[j] = exp(lz(j) - mx); total = total + ps[j]; }
            for (int j = 0; j < np; ++j) ps[j] = ps[j] / total;
        }
        for (int j = 0; j < n_smc; ++j) zeta_sums[smc[j]] += ps[j];
        zeta_sums[i] += ps[n_smc];
        double pmax = ps[0];
        for (int j = 1; j < np; ++j) pmax = ps[j] > pmax ? ps[j] : pmax;
        int k;
        if (pmax > 0.0) {                         // (discrete ps): runtime.cljc:281-287
            double tot = 0.0;
            for (int j = 0; j < np; ++j) tot = tot + ps[j];
            double x = uniforms[i] * tot, acc = 0.0;
            k = np - 1;
            for (int j = 0; j < np; ++j) { acc = acc + ps[j]; if (x < acc) { k = j; break; } }
        } else {                                  // (uniform-discrete 0 (count ps))
            k = (int)(uniforms[i] * np);
            if (k >= np) k = np - 1;
        }
        if (k != n_smc) {
            int32_t tmp = smc[k];
            smc[k] = csmc_idx[i];
            csmc_idx[i] = tmp;
        }
    }
    return APOP_OK;
}

int apop_ipmcmc_exchange_logz(apop_ctx* c, const double* local, int n_local, double* out_all) {
    CHECK_CTX(c);
    if (!local || !out_all || n_local <= 0) return APOP_ERR_INVALID;
    if (c->world == 1) {
        memcpy(out_all, local, sizeof(double) * (size_t)n_local);
        return APOP_OK;
    }
    return comm_allgather_f64(c, local, n_local, out_all);
}

} // extern "C"
