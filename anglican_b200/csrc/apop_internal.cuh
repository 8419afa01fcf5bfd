// apop_internal.cuh — context, workspace and small device helpers shared by
// the translation units of libanglican_pop.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "../../include/anglican_pop.h"
#include "exact_fold.cuh"
#include "fabric.cuh"

struct apop_ctx;

namespace apop {

constexpr int TB = 512;          // threads per tile CTA
constexpr int IPT = 8;           // elements per thread (blocked arrangement)
constexpr int TILE = TB * IPT;   // particles per tile
constexpr int NWARP = TB / 32;

// error bits raised by kernels
enum {
    EB_NAN = 1,          // NaN log-weight
    EB_ZERO = 2,         // total weight not > 0 / not finite
    EB_CTX = 4,          // classification / chain context mismatch (internal)
    EB_COMB = 8,         // comb table overflow / stuck comb
    EB_COUNT = 16,       // emitted offspring count mismatch (internal)
    EB_COMM = 32,        // a peer rank never arrived at an exchange (spin timeout)
    EB_INF = 64          // +inf log-weight in a fused sweep (infinite observe density)
};

struct ResampleInfo {
    double m;            // max log-weight (0 in explicit-weights mode)
    double atot;         // approximate total weight (tree sum)
    double asq;          // approximate sum of squared weights
    double ess;
    double log_mean_approx;
    double u;            // resampling uniform
    double total;        // exact sequential-fold total (smc.cljc:94)
    double step;         // total / n_new
    double new_lw;       // log(step) + m  (smc.cljc:98-103)
    double wx_sum[4];    // spare accumulators (filtering mean numerator, ...)
    long long J;         // emissions of the first lap
    long long n_new;
    int err;
    int nh0, nh1;        // hard tile counts per lap
    int first_skip1;     // first tile skipped in lap 1
    int weights_mode;
    int skip;            // adaptive resampling: this observe keeps the particles as they are
};

struct SweepDev {           // device-resident scalars of a fused sweep
    double base;            // common accumulated log-weight (state.cljc:27-31)
    double log_z;           // sum_t log-mean-weight (ipmcmc.cljc:67-74)
    int err;
    int carry;              // 1: the last observe did not resample (ESS above the threshold) -
                            //    every particle carries its own accumulated log-weight
    unsigned long long seed; // graph-replayed sweeps read the RNG key from here (see run_group)
};

// per-step outputs of a fused sweep, written by k_chain
struct StepOut {
    SweepDev* sd;           // nullptr outside sweeps
    double* log_mean_w;
    double* ess;
    double* filt;
    int32_t* anc;
    int t, reset, cond, has_x;
    double ess_thr;         // > 0: resample only when ESS < ess_thr * N (adaptive resampling)
};

struct Column {
    void* ptr = nullptr;
    void* alt = nullptr;     // second buffer, allocated by the first gather (result swapped in)
    int dtype = 0;
    int width = 0;
    bool used = false;
};

struct Timer {
    cudaEvent_t a = nullptr, b = nullptr;
    double total_ms = 0.0;
    int64_t count = 0;
    bool open = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending;
};

struct ParamArena {
    char* dev = nullptr;
    char* host = nullptr;
    size_t cap = 0, off = 0;
};

// host side of the peer-memory fabric (fabric.cuh)
struct FabricHost {
    bool linked = false;
    bool in_process = false;               // all ranks live in this process (apop_link_group)
    char* heap[kMaxWorld] = {};            // heap[q]: rank q's symmetric heap as mapped here
    bool ipc_open[kMaxWorld] = {};
    size_t heap_bytes = 0;
    size_t off_xa = 0, off_xb = 0, off_anc = 0;
    size_t cap_n = 0;
    unsigned long long epoch = 0;          // per-step exchange sequence number
    unsigned long long bar_epoch = 0;      // barrier sequence number
    std::vector<apop_ctx*> group;          // in-process group, rank order
};

struct RunRec {
    long long k0;                // first pointer index of the run
    int count;                   // offspring
    int val;                     // ancestor id | lap << 31
};
// Everything the streaming path of k_ancestors needs to know about one tile, built
// by k_expand (one 64-byte record, read with four 128-bit loads).  Fast mode: the
// tile emits in exactly one lap, all its edges and the pointers below them live in
// one binade / one comb segment, and k_fold_agg saw no tie and no saturated
// element; anything else takes the general path.
struct TileRec {
    double sD0;                  // (units of the exact running sum entering the tile) - X0 of the segment
    double qD, invq;             // comb stride of the segment in units of the binade, 1 / stride
    double inv_u;                // 2^-e of the binade
    double units;                // >= 0: exact units the tile must add; < 0: -(bound it must stay below)
    double m;                    // max log-weight (smc.cljc:75)
    int32_t* dst;                // where window entry 0 goes (local ancestor array or a peer's)
    int k0;                      // index, relative to the segment, of the first pointer of the window
    unsigned wm;                 // window length | mode << 28
};
enum { TM_NONE = 0, TM_FAST = 1, TM_GENERAL = 2, TM_IDENT = 3 };   // IDENT: no resample, anc[k] = k
static_assert(sizeof(TileRec) == 64, "TileRec is read as four 16-byte words");

constexpr int HEAVY_W = 1 << 18; // offspring of one (tile, lap) above which emission is deferred
constexpr int BIG_RUN = 1 << 16; // runs split over all CTAs of the fill kernel
constexpr int SEGB = 512;        // tiles per block of the segmented tile scan (k_seg_scan)
constexpr int HCAP = 128;        // hard tiles per lap with precomputed prefix / suffix tables

// Device workspace of the resampling pipeline (one slab).  Arrays indexed [lap]:
// lap 0 = first pass over the wheel, lap 1 = after the wrap (smc.cljc:133-136).
struct ResampleWS {
    void* slab = nullptr;
    int64_t nt = 0;              // tiles
    int nb = 0;                  // blocks of SEGB tiles
    double *tmax = nullptr, *tsum = nullptr, *tsum2 = nullptr, *ain = nullptr;  // tsum2: sum w x
    double *tacc = nullptr, *tacc2 = nullptr;   // accurate per-tile sum w, sum w^2
    int* cls[2] = {};            // class of every tile: binade context, kCtxHard or kCtxSkip
    Xd* agg[2] = {};             // easy tile: transducer of the whole tile in its context
    Xd* ex[2] = {};              // easy tile: composite of the tiles since the last hard tile /
                                 //   block start (exclusive)
    int* hix[2] = {};            // hard tile: its index in hl, else -1
    int* hp[2] = {};             // easy tile: index of the last hard tile before it in its block, or -1
    int* hl[2] = {};             // sorted list of hard tiles
    int* hctx[2] = {};           // per hard tile: context eA of its tables, or kCtxHard (general fold)
    Xd* bc[2] = {};              // per block: composite from the last hard tile (or block start) to
                                 //   the block end
    double* sblk[2] = {};        // exact running sum entering each block
    double* sin[2] = {};         // exact running sum entering every tile (k_expand), [nt + 1]
    long long* cnt[2] = {};      // pointers of the comb strictly below sin[lap][t], [nt + 1]
    int* gseg[2] = {};           // comb segment that holds pointer cnt[lap][t]
    TileRec* rec = nullptr;      // per tile: what k_ancestors does with it (k_expand)
    int* tflag = nullptr;        // per tile: bit lap set = integer fast path valid in that lap (k_fold_agg)
    int* glist = nullptr;        // tiles listed for k_ancestors_general, ngen[0] of them
    int* ngen = nullptr;
    // finalize (k_fin_a / k_fin_b)
    long long* gmax = nullptr;   // ordered-integer key of the max log-weight (atomicMax by the
                                 //   producer of the tile partials), reset by k_fin_b
    int* ticket = nullptr;       // [2] CTAs done in k_fin_a / k_fin_b
    int* hcnt = nullptr;         // [2] hard tiles staged per lap
    int* skipmin = nullptr;      // first tile skipped in lap 1
    double* fsc = nullptr;       // [8] scalars between the two finalize kernels
    double *csum = nullptr, *coff = nullptr;   // per chunk of 1024 tiles: sum, exclusive offset
    int* hst[2] = {};            // staged hard tiles (unsorted) and their table context
    int* hea[2] = {};
    double* shin[2] = {};        // exact running sum entering / leaving each hard tile
    double* shout[2] = {};
    Xd* hpa[2] = {};             // [HCAP][TILE] per-element inclusive prefix transducers, context eA
    Xd* hsb[2] = {};             // [HCAP][TILE] per-element suffix transducers, context eA + 1
    double* hw[2] = {};          // [HCAP][TILE] weights of the hard tiles
    double *bacc = nullptr, *bacc2 = nullptr, *baccx = nullptr;   // per block: sums of tacc, tacc2, tsum2
    double* send = nullptr;      // [2] exact running sum at the end of each lap (this rank)
    double* trace = nullptr;     // [2][256] chain diagnostics: ns per hard-tile event (lap 0 / 1)
    int hcap = 0;                // table slots allocated per lap (min(HCAP, nt))
    // deferred emission of heavy tiles (a tile that owns more than HEAVY_W offspring
    // would serialise the step): k_ancestors records one run per element and
    // k_fill_runs writes them with the whole GPU
    RunRec* runs = nullptr;      // [runs_cap], blocks of TILE records per heavy (tile, lap)
    RunRec* bigruns = nullptr;   // [big_cap] copies of the runs of >= BIG_RUN offspring
    int* nruns = nullptr;        // [0]: records reserved, [1]: big runs
    int runs_cap = 0, big_cap = 0;
    int heavy_w = HEAVY_W;       // APOP_HEAVY_W overrides (tests force the deferred path at small n)
    Comb* comb = nullptr;
    ResampleInfo* info = nullptr;      // device
    ResampleInfo* info_host = nullptr; // pinned
    double* edges = nullptr;           // multinomial only, lazily allocated
};

} // namespace apop

struct apop_ctx {
    int device = 0;
    int sm_count = 148;       // multiprocessors of the device (grid sizing)
    int rank = 0, world = 1;
    int64_t n = 0;            // local particles
    int64_t n_global = 0;
    int64_t g0 = 0;           // global id of local particle 0
    uint64_t seed = 0;
    uint32_t step = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    std::vector<apop::Column> cols;
    int64_t launches = 0;
    apop::ResampleWS ws;
    void* scratch = nullptr;      // gather scratch
    size_t scratch_bytes = 0;
    apop::Timer timers[16];
    void* nccl = nullptr;         // ncclComm_t
    apop::ParamArena arena;       // staged distribution parameters
    bool ktimers = false;         // per-kernel CUDA-event timers (bench roofline)
    apop::FabricHost fab;         // world > 1: peer heaps and mailboxes
    // last fused sweep
    struct SweepState* sweep = nullptr;
};

namespace apop {

int set_err(apop_ctx* c, int code, const char* fmt, ...);
#define APOP_CUDA(ctx, call)                                                        \
    do {                                                                            \
        cudaError_t e__ = (call);                                                   \
        if (e__ != cudaSuccess)                                                     \
            return apop::set_err(ctx, APOP_ERR_CUDA, "%s failed: %s (%s:%d)", #call, \
                                 cudaGetErrorString(e__), __FILE__, __LINE__);      \
    } while (0)
#define APOP_TRY(expr)               \
    do {                             \
        int rc__ = (expr);           \
        if (rc__ != APOP_OK) return rc__; \
    } while (0)

// kernel-class timer tags (apop_timer_elapsed_ms)
enum { KT_SWEEP = 0, KT_PROPAGATE = 1, KT_TILE_LSE = 2, KT_FINALIZE = 3, KT_FOLD_AGG = 4,
       KT_RESOLVE = 5, KT_ANCESTORS = 6, KT_STATS = 7, KT_GATHER = 8, KT_MIGRATE = 9 };
void kt_begin(apop_ctx* c, int tag);
void kt_end(apop_ctx* c, int tag);

Column* get_col(apop_ctx* c, apop_col h, int dtype, int width);
int ensure_scratch(apop_ctx* c, size_t bytes);
int ensure_resample_ws(apop_ctx* c);
int param_upload(apop_ctx* c, const void* host, size_t bytes, void** dev);

// resample pipeline entry points (resample.cu)
struct ResampleArgs {
    const double* lw;        // log-weights (or NULL in weights mode)
    const double* w;         // explicit weights (weights mode)
    int64_t n;
    int64_t n_new;
    double u;                // NaN -> draw from stream
    int order;
    int32_t* anc_out;        // n_new entries
    int64_t out_offset;      // added to every written slot (CSMC keeps slot 0)
    bool have_partials;      // tile partials (tmax/tsum/tsum2) already in ws
    const double* xfilt;     // optional state column: per-step filtering mean sum w x / sum w
    double rel_eps;          // accuracy of the supplied tile partials (0: fp64 partials)
    StepOut step;            // optional per-step outputs (sd == nullptr: none)
    bool sharded;            // world > 1: exchange through the fabric, offspring stored to peers
    PeerI32 anc_peer;        // sharded: anc_peer.p[q] = rank q's ancestor array
    bool keep_err;           // second resample inside a sweep step: error bits stay sticky
    const unsigned long long* seedp;   // non-null: the RNG key lives in device memory (graph replay)
};
int resample_enqueue(apop_ctx* c, const ResampleArgs& a);
int resample_fetch_info(apop_ctx* c, ResampleInfo* out);   // syncs
int info_to_status(apop_ctx* c, const ResampleInfo& info);
int multinomial_enqueue(apop_ctx* c, const double* lw, const double* w, const double* uni,
                        int64_t n_new, int32_t* anc);
int stat_lse_enqueue(apop_ctx* c, const double* lw, int64_t n, double* out2);   // stats.cu
int stat_lse_prepare(apop_ctx* c);
int gather_enqueue(apop_ctx* c, const int32_t* anc, int64_t n_new, Column* col);
// mvn.cu: fp64 tensor-core (DMMA) kernels
struct MvnSweepArgs {
    const double* A;        // [d*d] dense transition
    const double* Lq;       // [d*d] lower factor of the process covariance
    const double* iL;       // [d*d] inverse lower factor of the observation covariance
    double Z;               // 0.5 d log(2 pi) + log det L   (runtime.cljc:355-358)
    const double* xp;       // parent states [n*d]
    const int32_t* anc;     // ancestors (sorted physical order), unused on the first step
    int first_step;
    double* xn;             // new states [n*d]
    double* lw;             // log-weights [n]
    const double* y;        // observation of this step [d] (device)
    const double* retained; // conditional sweep: state of slot 0 at this step, or NULL
    const SweepDev* sd;
    int64_t n, g0;
    uint64_t seed;
    uint32_t t;
    // sharded population: parents are global ids; rows of a peer's shard are read
    // from its heap over NVLink (world == 1: unused)
    PeerF64 xpeer;
    int world;
};

bool mvn_dim_supported(int d);
void mvn_preload_kernels(int d);
int mvn_propagate_enqueue(apop_ctx* c, int d, const MvnSweepArgs& a);
int mvn_init_enqueue(apop_ctx* c, double* x, int d, double s0, const double* retained);
int mvn_observe_dmma_enqueue(apop_ctx* c, int d, const double* iL, double Z, const double* mean,
                             int ms, const double* val, int vs, double* lw);
int mvn_prepare_host(const double* cov, int d, std::vector<double>& L, std::vector<double>& iL,
                     double* Z);
Fabric fabric_view(const apop_ctx* c);                      // by-value kernel argument
int fabric_barrier_enqueue(apop_ctx* c);                    // all ranks: flags over peer memory
int fabric_prepare(apop_ctx* c, size_t n_cap);              // (re)allocate + map the heaps (comm.cu)
void fabric_free(apop_ctx* c);

// ---------------------------------------------------------------- device
#if defined(__CUDACC__)
} // namespace apop
#include "fastmath.cuh"
namespace apop {

// ordered-integer key of a double (no NaN): integer max == floating-point max
__device__ __forceinline__ long long dkey(double x) {
    const long long b = __double_as_longlong(x);
    return b >= 0 ? b : (b ^ 0x7fffffffffffffffLL);
}
__device__ __forceinline__ double dkey_inv(long long k) {
    return __longlong_as_double(k >= 0 ? k : (k ^ 0x7fffffffffffffffLL));
}

__device__ __forceinline__ double weight_from_lw(double lw, double m) {
    // smc.cljc:78-82
    if (lw == m) return 1.0;
    if (!(lw > -INFINITY)) return 0.0;
    return fm::exp_any(lw - m);
}

// blocked load of IPT consecutive doubles per thread; out-of-range -> fill
__device__ __forceinline__ void load_blocked(const double* __restrict__ p, int64_t base,
                                             int64_t n, double (&v)[IPT], double fill) {
    const int64_t i0 = base + (int64_t)threadIdx.x * IPT;
    if (i0 + IPT <= n && ((((uintptr_t)(p + i0)) & 15) == 0)) {
        const double2* q = reinterpret_cast<const double2*>(p + i0);
#pragma unroll
        for (int j = 0; j < IPT / 2; ++j) {
            double2 t = __ldg(q + j);
            v[2 * j] = t.x;
            v[2 * j + 1] = t.y;
        }
    } else {
#pragma unroll
        for (int j = 0; j < IPT; ++j) v[j] = (i0 + j < n) ? p[i0 + j] : fill;
    }
}

__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ i64 shfl_up_i64(i64 v, int d) {
    int lo = __shfl_up_sync(0xffffffffu, (int)(v & 0xffffffffLL), d);
    int hi = __shfl_up_sync(0xffffffffu, (int)(v >> 32), d);
    return ((i64)hi << 32) | (u64)(unsigned)lo;
}
__device__ __forceinline__ i64 shfl_i64(i64 v, int src) {
    int lo = __shfl_sync(0xffffffffu, (int)(v & 0xffffffffLL), src);
    int hi = __shfl_sync(0xffffffffu, (int)(v >> 32), src);
    return ((i64)hi << 32) | (u64)(unsigned)lo;
}

#endif // __CUDACC__

} // namespace apop
