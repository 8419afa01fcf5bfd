// resample.cu — systematic / multinomial resampling, bit-exact with the
// sequential fp64 wheel of anglican.smc/resample (reference
// src/anglican/smc.cljc:86-137) and particle-weights (smc.cljc:70-84).
//
// Pipeline (all on ctx->stream, no host round trip):
//   K2  k_tile_lse      per-tile (max, sum exp, sum exp^2) of the log-weights
//                       (inside a fused sweep k_propagate produces these)
//       k_fin_a/b       global max m (atomicMax of the producer), approximate tile prefix,
//                       per-tile binade classification for both laps of the wheel
//   K3a k_fold_agg      per-tile aggregate of the exact fold (integer units of
//                       the tile's binade) + accurate sum w, sum w^2
//       k_seg_scan, k_chain, k_expand ("resolve")
//                       chain the aggregates -> exact running sum at every
//                       tile boundary, exact total, pointer-comb table, J,
//                       ESS / log-mean-weight, per-step sweep outputs
//   K3b/K4 k_ancestors  per-element exact edges merged against the pointer
//                       comb, writes the ancestor index array
//   K5  k_gather        x_new[k] = x[a[k]]
// The arithmetic (why this reproduces a sequential fold exactly) is in
// exact_fold.cuh.  "Easy" tiles (the fold provably stays in one binade, no
// exact ties) take a fast path: fp64 magic-number quantisation, plain int64
// prefix sums, and a merge of edges and pointers in integer units.  Everything
// else takes the general parity-transducer path.
#include <limits.h>
#include <math.h>
#include <stdlib.h>
#include <stddef.h>

#include "apop_internal.cuh"
#include "rng.cuh"

namespace apop {

using WSView = ResampleWS;      // passed to kernels by value (device pointers only)
#ifndef APOP_FOLD_CTAS
#define APOP_FOLD_CTAS 3
#endif
#ifndef APOP_ANC_CTAS
#define APOP_ANC_CTAS 3
#endif
static WSView view_of(const ResampleWS& w) { return w; }

static constexpr int kTieCap = 32;      // ties one fast iteration of cta_tile_fold resolves (one warp)
struct FoldSmem {
    Xd warp_tot[NWARP];
    i64 wsum[NWARP];
    double bcast_S;
    int cross;
    int imin;
    double dsum[NWARP];          // cta_tile_fold fast iteration: per-warp sums of integer-valued doubles
    int ntie;                    //   its ties in front of the crossing: element, parity of the
    int tie_idx[kTieCap];        //   uncorrected sum, order by element, corrections so far
    int tie_par[kTieCap];
    int tie_ord[kTieCap];
    int tie_sidx[kTieCap];
    int tie_cum[kTieCap];
    double last_edge[TB];
    double head[1024];           // generic_emit: literal fold of the first elements of a tile that
                                 //   starts at zero (a binade crossing at every doubling of the sum)
};

__device__ __forceinline__ long long gtimer_ns() {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// ------------------------------------------------------------------ helpers
__device__ __forceinline__ Xd cta_excl_scan_xd(Xd t, Xd* warp_tot) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    Xd inc = t;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        Xd o;
        o.a0 = shfl_up_i64(inc.a0, d);
        o.a1 = shfl_up_i64(inc.a1, d);
        if (lane >= d) inc = xd_compose(o, inc);
    }
    Xd ex;
    ex.a0 = shfl_up_i64(inc.a0, 1);
    ex.a1 = shfl_up_i64(inc.a1, 1);
    if (lane == 0) ex = xd_identity();
    if (lane == 31) warp_tot[wid] = inc;
    __syncthreads();
    Xd pre = xd_identity();
    for (int i = 0; i < wid; ++i) pre = xd_compose(pre, warp_tot[i]);
    return xd_compose(pre, ex);
}

__device__ __forceinline__ Xd cta_total_xd(const Xd* warp_tot) {
    Xd tot = xd_identity();
#pragma unroll
    for (int i = 0; i < NWARP; ++i) tot = xd_compose(tot, warp_tot[i]);
    return tot;
}

__device__ __forceinline__ i64 shfl_down_i64(i64 v, int d) {
    int lo = __shfl_down_sync(0xffffffffu, (int)(v & 0xffffffffLL), d);
    int hi = __shfl_down_sync(0xffffffffu, (int)(v >> 32), d);
    return ((i64)hi << 32) | (u64)(unsigned)lo;
}

// exclusive SUFFIX scan over the CTA: returns the composite of the transducers of
// all later threads (thread order), identity for the last thread
__device__ __forceinline__ Xd cta_excl_suffix_scan_xd(Xd t, Xd* warp_tot) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    Xd inc = t;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        Xd o;
        o.a0 = shfl_down_i64(inc.a0, d);
        o.a1 = shfl_down_i64(inc.a1, d);
        if (lane + d < 32) inc = xd_compose(inc, o);
    }
    Xd ex;
    ex.a0 = shfl_down_i64(inc.a0, 1);
    ex.a1 = shfl_down_i64(inc.a1, 1);
    if (lane == 31) ex = xd_identity();
    __syncthreads();
    if (lane == 0) warp_tot[wid] = inc;
    __syncthreads();
    Xd post = xd_identity();
    for (int i = NWARP - 1; i > wid; --i) post = xd_compose(warp_tot[i], post);
    return xd_compose(ex, post);
}

__device__ __forceinline__ void quantize_ctx(double w, int e, bool fast, double inv_u, i64& f,
                                             int& up, int& tie) {
    if (fast) { up = 0; quantize_fast(w, inv_u, f, tie); }
    else quantize(w, e, f, up, tie);
}

// General exact fold of one tile (IPT consecutive elements per thread) starting
// from the exact running sum S_in: parity-transducer scan, one iteration per
// binade crossing inside the tile.  edge[j] = exact inclusive running sum.
__device__ __noinline__ double cta_tile_fold(double S_in, const double (&w)[IPT], int nvalid,
                                             double (&edge)[IPT], FoldSmem& sm, int pos_start = 0) {
    if (nvalid <= pos_start) return S_in;
    int pos = pos_start;
    double S = S_in;
    const int i0 = threadIdx.x * IPT;
    for (;;) {
        const int e = ctx_of(S);
        const i64 s_in = units_of(S);
        const bool fast = ctx_fast_ok(e);
        const double inv_u = fast ? pow2(-e) : 0.0;
        int mycross = INT_MAX;
        i64 sprev = 0;
        if (threadIdx.x == 0) { sm.cross = INT_MAX; sm.imin += 1; sm.ntie = 0; }
        bool done = false;
        if (fast) {
            // ---- fast iteration: increments are RN(w / ulp) by the 2^52 magic add and the
            // prefix sums are plain fp64 additions of integer-valued doubles.  Only additions
            // are used: a prefix below 2^53 is exact, one at or above 2^53 can only round to a
            // value >= 2^53, so "s_in + prefix >= 2^53" (the crossing) is decided exactly.
            // An element of 2^52 ulps or more crosses wherever it stands: it counts as 2^53.
            // A tie (remainder exactly half an ulp) adds floor(w / ulp) and one more when the
            // sum before it is odd (round half to even, step_units): the few ties in front of
            // the crossing are listed with the parity of their uncorrected sum, one lane walks
            // the list in element order and every thread takes the corrections made before its
            // first element.  Ties and saturating elements are the rule where this function
            // runs - at the head of the fold an element is comparable to the running sum -, so
            // without them the iteration fell back to the transducer scan three times out of
            // four (chain trace at 2^27).  ~5x fewer instructions than that scan; 7-9 k instead of
            // 9-10 k cycles per iteration (the rest is barrier and pipeline latency, see APOP_HEAD).
            double incD[IPT];
            double loc = 0.0;
            unsigned tmask = 0;
#pragma unroll
            for (int j = 0; j < IPT; ++j) {
                const int idx = i0 + j;
                double rn = 0.0;
                if (idx >= pos && idx < nvalid) {
                    const double scaled = w[j] * inv_u;
                    if (!(scaled < 4503599627370496.0)) rn = 9007199254740992.0;
                    else {
                        rn = (scaled + 4503599627370496.0) - 4503599627370496.0;
                        const double diff = scaled - rn;
                        if (diff == 0.5) tmask |= 1u << j;
                        else if (diff == -0.5) { tmask |= 1u << j; rn -= 1.0; }
                    }
                }
                incD[j] = rn;
                loc += rn;
            }
            const int any_tie = __syncthreads_or(tmask != 0);
            const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
            double inc = loc;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const double o = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d) inc += o;
            }
            double exl = __shfl_up_sync(0xffffffffu, inc, 1);
            if (lane == 0) exl = 0.0;
            if (lane == 31) sm.dsum[wid] = inc;
            __syncthreads();
            double pre = 0.0;
            for (int i = 0; i < wid; ++i) pre += sm.dsum[i];
            pre += exl;
            const double L = (double)(kTwo53 - s_in), sD0 = (double)s_in, ulp = pow2(e);
            done = true;
            if (any_tie) {
                if (tmask) {
                    double b = pre;
#pragma unroll
                    for (int j = 0; j < IPT; ++j) {
                        b += incD[j];
                        // a tie whose uncorrected sum already crosses lies behind the crossing
                        if (((tmask >> j) & 1u) && b < L) {
                            const int k = atomicAdd(&sm.ntie, 1);
                            if (k < kTieCap) {
                                sm.tie_idx[k] = i0 + j;
                                sm.tie_par[k] = (int)((s_in + (i64)b) & 1);
                            }
                        }
                    }
                }
                __syncthreads();
                const int K = sm.ntie;
                if (K > kTieCap) done = false;              // block-uniform: transducer scan instead
                else if (K > 0) {
                    if (wid == 0) {
                        if (lane < K) {
                            const int mine = sm.tie_idx[lane];
                            int r = 0;
                            for (int q = 0; q < K; ++q) r += sm.tie_idx[q] < mine;
                            sm.tie_ord[r] = lane;
                        }
                        __syncwarp();
                        if (lane == 0) {
                            int C = 0;
                            for (int r = 0; r < K; ++r) {
                                const int q = sm.tie_ord[r];
                                C += (sm.tie_par[q] + C) & 1;
                                sm.tie_sidx[r] = sm.tie_idx[q];
                                sm.tie_cum[r] = C;
                            }
                        }
                    }
                    __syncthreads();
                    int corr = 0;
                    for (int r = 0; r < K; ++r)
                        if (sm.tie_sidx[r] < i0) corr = sm.tie_cum[r];
                    pre += (double)corr;
                }
            }
            if (done) {
                double sD = pre;
#pragma unroll
                for (int j = 0; j < IPT; ++j) {
                    const int idx = i0 + j;
                    if (idx >= pos && idx < nvalid && mycross == INT_MAX) {
                        double d = incD[j];
                        if ((tmask >> j) & 1u) d += (double)((s_in + (i64)(sD + d)) & 1);
                        const double nx = sD + d;
                        if (nx >= L) { mycross = idx; sprev = s_in + (i64)sD; }
                        else { sD = nx; edge[j] = (sD0 + sD) * ulp; }
                    }
                }
            }
        }
        if (!done) {
        i64 qf[IPT];
        int qfl[IPT];
        Xd t = xd_identity();
#pragma unroll
        for (int j = 0; j < IPT; ++j) {
            const int idx = i0 + j;
            qfl[j] = 0;
            qf[j] = 0;
            if (idx >= pos && idx < nvalid) {
                i64 f; int up, tie;
                quantize_ctx(w[j], e, fast, inv_u, f, up, tie);
                qf[j] = f;
                qfl[j] = up | (tie << 1) | 4;
                xd_push(t, f, up, tie);
            }
        }
        Xd ex = cta_excl_scan_xd(t, sm.warp_tot);      // has a __syncthreads
        i64 s = xd_apply(s_in, ex);
#pragma unroll
        for (int j = 0; j < IPT; ++j) {
            if ((qfl[j] & 4) && mycross == INT_MAX) {
                i64 s2 = step_units(s, qf[j], qfl[j] & 1, (qfl[j] >> 1) & 1);
                if (s2 >= kTwo53) { mycross = i0 + j; sprev = s; }
                else { s = s2; edge[j] = from_units(s, e); }
            }
        }
        }
        {
            // every thread behind the crossing "crosses" at its first element: reduce inside the
            // warp first (up to 512 atomics on one shared word cost ~8 k cycles per iteration)
            const int wmin = __reduce_min_sync(0xffffffffu, mycross);
            if ((threadIdx.x & 31) == 0 && wmin != INT_MAX) atomicMin(&sm.cross, wmin);
        }
        __syncthreads();
        const int cross = sm.cross;
        if (cross == INT_MAX) {
            if ((nvalid - 1) / IPT == (int)threadIdx.x) {
#pragma unroll
                for (int j = 0; j < IPT; ++j)
                    if (i0 + j == nvalid - 1) sm.bcast_S = edge[j];
            }
            __syncthreads();
            return sm.bcast_S;
        }
        if (cross / IPT == (int)threadIdx.x) {
#pragma unroll
            for (int j = 0; j < IPT; ++j)
                if (i0 + j == cross) {
                    // the crossing add itself: one hardware fp64 add (smc.cljc:120)
                    double Sn = from_units(sprev, e) + w[j];
                    edge[j] = Sn;
                    sm.bcast_S = Sn;
                }
        }
        __syncthreads();
        S = sm.bcast_S;
        pos = cross + 1;
        if (pos >= nvalid) return S;
    }
}

// Length of the literal head fold of a tile whose running sum starts at zero (chain general
// path, generic_emit).  Measured with the chain trace (tools/dbg_resolve.py, cycles of tile 0
// at 2^27 particles): a literal add costs 13 cycles; one parallel crossing iteration of
// cta_tile_fold costs 7-9 k cycles (9-10 k before ties were resolved inside the fast
// iteration) and that cost is latency, not work: per iteration thread 0 sees ~0.7-1.1 k cycles
// for the 8 quantisations of a thread, ~0.9 k waiting for the slowest warp, ~0.6 k for the
// scan, ~1.6-2.3 k for the tie list when the tile has a tie, ~1.1-1.5 k for the replay and
// ~0.9 k for the crossing minimum - a dozen short phases of fp64 work (half rate per
// scheduler, four warps each) separated by block barriers, whether sixteen warps or one take
// part.  Shorter heads or window-by-window folding therefore only add iterations:
//   head 1024: 13.1 k + 3-4 iterations 30-36 k   | head 256: 3.5 k + 5-7 iterations 41-54 k
//   head 64: 1.1 k + 7-9 iterations 53-65 k       (gpurun_out/r03tie; before: 41 / 66 / 90 k)
#ifndef APOP_HEAD
#define APOP_HEAD 1024
#endif

// raw tile values: the weights themselves (weights mode) or the log-weights
__device__ __forceinline__ void load_raw(const double* __restrict__ lw,
                                         const double* __restrict__ win, int64_t base,
                                         int64_t n, double (&v)[IPT]) {
    if (win) load_blocked(win, base, n, v, 0.0);
    else load_blocked(lw, base, n, v, -INFINITY);      // out of range: -inf -> weight 0
}
// particle-weights (smc.cljc:70-84) from the raw tile values
__device__ __forceinline__ void weights_from_raw(bool is_w, const double (&v)[IPT], int64_t base,
                                                 int64_t n, double m, double (&w)[IPT]) {
    if (is_w) {
#pragma unroll
        for (int j = 0; j < IPT; ++j) w[j] = v[j];
    } else if (m > -INFINITY) {
        double a[4], b[4];
#pragma unroll
        for (int h = 0; h < IPT; h += 4) {
#pragma unroll
            for (int j = 0; j < 4; ++j) a[j] = v[h + j];
            fm::weights_v<4>(a, m, b);
#pragma unroll
            for (int j = 0; j < 4; ++j) w[h + j] = b[j];
        }
    } else {                                           // all -inf: every weight is 1 (smc.cljc:80)
        const int64_t i0 = base + (int64_t)threadIdx.x * IPT;
#pragma unroll
        for (int j = 0; j < IPT; ++j) w[j] = (i0 + j < n) ? weight_from_lw(v[j], m) : 0.0;
    }
}
__device__ __forceinline__ void load_weights(const double* __restrict__ lw,
                                             const double* __restrict__ win, int64_t base,
                                             int64_t n, double m, double (&w)[IPT]) {
    double v[IPT];
    load_raw(lw, win, base, n, v);
    weights_from_raw(win != nullptr, v, base, n, m, w);
}

// ------------------------------------------------------------------ K2
// per-tile partials of particle-weights (smc.cljc:70-84): max, sum of
// exp(lw - tilemax), sum of squares.  In weights mode: plain sums.
__global__ void __launch_bounds__(TB) k_tile_lse(const double* __restrict__ lw,
                                                 const double* __restrict__ win, int64_t n,
                                                 WSView v) {
    __shared__ double red[NWARP];
    __shared__ double red2[NWARP];
    const int64_t base = (int64_t)blockIdx.x * TILE;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    double x[IPT];
    if (win) {
        load_blocked(win, base, n, x, 0.0);
        double s = 0.0, s2 = 0.0;
        bool bad = false;
#pragma unroll
        for (int j = 0; j < IPT; ++j) { s += x[j]; s2 += x[j] * x[j]; bad |= !(x[j] >= 0.0); }
        if (bad) atomicOr(&v.info->err, EB_NAN);
        s = warp_sum(s); s2 = warp_sum(s2);
        if (lane == 0) { red[wid] = s; red2[wid] = s2; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double a = 0, b = 0;
            for (int i = 0; i < NWARP; ++i) { a += red[i]; b += red2[i]; }
            v.tmax[blockIdx.x] = 0.0; v.tsum[blockIdx.x] = a; (void)b;
        }
        return;
    }
    load_blocked(lw, base, n, x, -INFINITY);
    double mx = -INFINITY;
    bool nan = false;
#pragma unroll
    for (int j = 0; j < IPT; ++j) { nan |= (x[j] != x[j]); mx = fmax(mx, x[j]); }
    if (nan) atomicOr(&v.info->err, EB_NAN);
    mx = warp_max(mx);
    if (lane == 0) red[wid] = mx;
    __syncthreads();
    mx = red[0];
#pragma unroll
    for (int i = 1; i < NWARP; ++i) mx = fmax(mx, red[i]);
    __syncthreads();
    const int64_t i0 = base + (int64_t)threadIdx.x * IPT;
    double s = 0.0, s2 = 0.0;
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
        if (i0 + j < n) {
            double wv = (x[j] == mx) ? 1.0 : ((x[j] > -INFINITY) ? exp(x[j] - mx) : 0.0);
            s += wv; s2 += wv * wv;
        }
    }
    s = warp_sum(s); s2 = warp_sum(s2);
    if (lane == 0) { red[wid] = s; red2[wid] = s2; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0, b = 0;
        for (int i = 0; i < NWARP; ++i) { a += red[i]; b += red2[i]; }
        v.tmax[blockIdx.x] = mx; v.tsum[blockIdx.x] = a; (void)b;
        atomicMax(v.gmax, dkey(mx));
    }
}

// ------------------------------------------------------------------ finalize
constexpr int FB = 1024;
constexpr int FCH = 4;                 // values per thread per scan chunk

// Exclusive prefix sums of non-negative values.  Exclusive values are taken from
// the neighbour's inclusive one (never "inclusive - own": one dominant value
// would cancel the whole prefix and break the relative error bound that
// classify_tile relies on).
__device__ __forceinline__ double block_excl_scan_f64(double v, double* sm /*[33]*/,
                                                      double& total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    double inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        double o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    double ex = __shfl_up_sync(0xffffffffu, inc, 1);
    if (lane == 0) ex = 0.0;
    if (lane == 31) sm[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        double winc = (lane < (int)(blockDim.x >> 5)) ? sm[lane] : 0.0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            double o = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += o;
        }
        double wex = __shfl_up_sync(0xffffffffu, winc, 1);
        if (lane == 0) wex = 0.0;
        sm[lane] = wex;                             // exclusive over warps
        if (lane == 31) sm[32] = winc;
    }
    __syncthreads();
    double r = sm[wid] + ex;
    total = sm[32];
    __syncthreads();
    return r;
}

__device__ __forceinline__ double block_sum_f64(double v, double* sm /*[33]*/) {
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    double r = 0.0;
    if (threadIdx.x < 32) {
        r = (threadIdx.x < (blockDim.x >> 5)) ? sm[threadIdx.x] : 0.0;
        r = warp_sum(r);
        if (threadIdx.x == 0) sm[32] = r;
    }
    __syncthreads();
    r = sm[32];
    __syncthreads();
    return r;
}

// Finalize, in two launches of ceil(nt / FB) CTAs (one tile per thread).  The global
// max arrives as an ordered-integer atomicMax from the kernel that produced the tile
// partials (k_propagate / k_tile_lse), so nothing here walks the tile arrays serially:
//   k_fin_a  rescale the tile sums to the max, per-chunk sums; the LAST CTA to finish
//            turns them into chunk offsets (exclusive prefix), runs the sharded
//            exchange and writes the scalars;
//   k_fin_b  approximate exclusive prefix ain[t], classification of every tile for
//            both laps, hard tiles appended to a staging list; the LAST CTA ranks
//            the list (hard tiles are rare) and resets the counters.
// fsc[]: 0 gfac (local max -> global max), 1 atot, 2 upper bound of the last pointer
__global__ void __launch_bounds__(FB) k_fin_a(WSView v, int64_t n, int64_t n_new, double u_in,
                                              uint64_t seed, uint32_t step, int weights_mode,
                                              double rel_eps, Fabric F, unsigned long long seq,
                                              int64_t n_global,
                                              const unsigned long long* __restrict__ seedp) {
    __shared__ double smd[33];
    __shared__ int s_last;
    if (seedp) seed = *seedp;
    __shared__ double pm[kMaxWorld], ps[kMaxWorld];
    __shared__ int perr[kMaxWorld];
    const int64_t nt = v.nt;
    const int tid = threadIdx.x, G = gridDim.x;
    const int64_t t = (int64_t)blockIdx.x * FB + tid;
    // ---- global max (smc.cljc:75) and tile sums relative to it
    double m = weights_mode ? 0.0 : dkey_inv(v.gmax[0]);
    double val = 0.0;
    if (t < nt) {
        val = v.tsum[t];
        if (!weights_mode) {
            const double tm = v.tmax[t];
            val *= (tm == m) ? 1.0 : fm::exp_any(tm - m);
            v.tsum[t] = val;
        }
    }
    const double cs = block_sum_f64(val, smd);
    if (tid == 0) {
        v.csum[blockIdx.x] = cs;
        __threadfence();
        s_last = (atomicAdd(v.ticket, 1) == G - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // ---- tail, one CTA: chunk offsets and scalars (G <= FB: nt < 2^31 / TILE)
    const double cv = (tid < G) ? __ldcg(v.csum + tid) : 0.0;
    double gfac = 1.0, offset = 0.0, gtot = 0.0;
    if (F.world > 1) {
        // exchange (local max, local sum, error bits) with every peer through the
        // mailboxes; afterwards m is the global max, `offset` the approximate weight
        // of all lower ranks and gtot the global total
        const double lsum = block_sum_f64(cv, smd);
        const int par = (int)(seq & 1);
        if (tid < F.world) {
            mail_send(&F.mbox[tid]->fin[par][F.rank], m, lsum, (double)v.info->err, seq);
            double o[3];
            const bool ok = mail_wait(&F.mbox[F.rank]->fin[par][tid], seq, o, F.spin_cycles);
            pm[tid] = o[0]; ps[tid] = o[1];
            perr[tid] = (int)o[2] | (ok ? 0 : (EB_COMM | (CW_FIN << 16) | (tid << 24)));
        }
        __syncthreads();
        const double m_loc = m;
        double gm = pm[0];
        int gerr = 0;
        for (int q = 1; q < F.world; ++q) gm = fmax(gm, pm[q]);
        if (weights_mode) gm = 0.0;
        for (int q = 0; q < F.world; ++q) {
            const double sq = ps[q] * ((pm[q] == gm || weights_mode) ? 1.0 : fm::exp_any(pm[q] - gm));
            if (q < F.rank) offset += sq;
            gtot += sq;
            gerr |= perr[q];
        }
        m = gm;
        gfac = (m_loc == gm || weights_mode) ? 1.0 : fm::exp_any(m_loc - gm);
        if (tid == 0 && gerr) atomicOr(&v.info->err, gerr);
        __syncthreads();
    }
    double ctot;
    const double ex = offset + block_excl_scan_f64(cv * gfac, smd, ctot);
    if (tid < G) v.coff[tid] = ex;
    if (tid == 0) {
        const double carry = offset + ctot;
        const double atot = (F.world > 1) ? gtot : carry;
        v.coff[G] = carry;
        v.ain[nt] = carry;
        const double u = (u_in != u_in) ? uniform_at(seed, 0, step, STREAM_RESAMPLE, 0) : u_in;
        // x_last < x_0 + total <= total*(u+1): generous upper bound for the skip test
        // (the comb is n_new repeated additions: its rounding error grows with n_new, not n)
        const double eps_tot = ((double)(2 * n_global) + (double)n_new + 8192.0) * 4.44089209850062616e-16 + rel_eps;
        v.fsc[0] = gfac;
        v.fsc[1] = atot;
        v.fsc[2] = atot * (u + 1.0) * (1.0 + 4.0 * eps_tot);
        ResampleInfo* I = v.info;
        I->m = m;
        I->atot = atot;
        I->asq = 0.0;
        I->ess = 0.0;
        I->log_mean_approx = (m > -INFINITY && m < INFINITY) ? log(atot / (double)n_global) + m
                                                               : -INFINITY;
        I->u = u;
        I->n_new = n_new;
        I->weights_mode = weights_mode;
        if (!weights_mode && m != m) I->err |= EB_NAN;
        v.ngen[0] = 0;
        v.ticket[0] = 0;
        v.hcnt[0] = 0; v.hcnt[1] = 0;
        v.skipmin[0] = (nt < INT_MAX) ? (int)nt : INT_MAX;
    }
}

constexpr int HB = 4096;                   // staged hard tiles per lap

__global__ void __launch_bounds__(FB) k_fin_b(WSView v, int64_t n, double rel_eps, int64_t g0,
                                              int64_t n_global) {
    __shared__ double smd[33];
    __shared__ double sA[FB + 1];
    __shared__ int s_last;
    const int64_t nt = v.nt;
    const int tid = threadIdx.x, G = gridDim.x;
    const int64_t t = (int64_t)blockIdx.x * FB + tid;
    const double gfac = v.fsc[0], atot = v.fsc[1], xlast_ub = v.fsc[2];
    // ---- approximate exclusive prefix of the tile sums.  Exclusive values come from
    // the scan itself, never as "inclusive - own" (one dominant value would cancel the
    // prefix and break the relative error bound classify_tile relies on).
    const double val = (t < nt) ? v.tsum[t] * gfac : 0.0;
    double ctot;
    const double a_in = v.coff[blockIdx.x] + block_excl_scan_f64(val, smd, ctot);
    sA[tid] = a_in;
    if (tid == 0) sA[FB] = v.coff[blockIdx.x + 1];
    __syncthreads();
    int myskip = INT_MAX;
    if (t < nt) {
        const double ao = (t + 1 == nt) ? v.coff[G] : sA[tid + 1];
        v.ain[t] = a_in;
        const int64_t hi = ((t + 1) * TILE < n) ? (t + 1) * TILE : n;
        int ea0, eb0, ea1 = kCtxHard, eb1 = kCtxHard;
        const int c0 = classify_tile2(a_in, ao, g0 + hi, rel_eps, &ea0, &eb0);
        v.cls[0][t] = c0;
        int c1;
        const double eps1 = ((double)(n_global + g0 + hi) + 4096.0) * 4.44089209850062616e-16 + rel_eps;
        if ((atot + a_in) * (1.0 - eps1) > xlast_ub) {
            c1 = kCtxSkip;
            myskip = (int)t;
        } else {
            c1 = classify_tile2(atot + a_in, atot + ao, n_global + g0 + hi, rel_eps, &ea1, &eb1);
        }
        v.cls[1][t] = c1;
        v.hix[0][t] = -1;
        v.hix[1][t] = -1;
        // a hard tile whose bounds span exactly two binades gets prefix / suffix tables
        if (c0 == kCtxHard) {
            const int k = atomicAdd(&v.hcnt[0], 1);
            if (k < HB) {
                v.hst[0][k] = (int)t;
                v.hea[0][k] = (ea0 != kCtxHard && eb0 == ea0 + 1 && ea0 > -1000 && ea0 < 900) ? ea0 : kCtxHard;
            }
        }
        if (c1 == kCtxHard) {
            const int k = atomicAdd(&v.hcnt[1], 1);
            if (k < HB) {
                v.hst[1][k] = (int)t;
                v.hea[1][k] = (ea1 != kCtxHard && eb1 == ea1 + 1 && ea1 > -1000 && ea1 < 900) ? ea1 : kCtxHard;
            }
        }
    }
    {
        // first skipped tile: one global atomic per warp (beyond the last pointer every tile skips)
        const int wmin = __reduce_min_sync(0xffffffffu, myskip);
        if ((tid & 31) == 0 && wmin != INT_MAX) atomicMin(v.skipmin, wmin);
    }
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        s_last = (atomicAdd(v.ticket + 1, 1) == G - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // ---- tail, one CTA: the sorted hard lists
    const int skip1 = __ldcg(v.skipmin);
    const int h0tot = __ldcg(v.hcnt), h1tot = __ldcg(v.hcnt + 1);
    if (h0tot <= HB && h1tot <= HB) {
#pragma unroll 1
        for (int lap = 0; lap < 2; ++lap) {
            const int cnt = lap ? h1tot : h0tot;
            const int* list = v.hst[lap];
            for (int i = tid; i < cnt; i += FB) {
                const int tt = __ldcg(list + i);
                int rank = 0;
                for (int k = 0; k < cnt; ++k) rank += (__ldcg(list + k) < tt);
                // tiles at or beyond the first skipped tile are never folded in lap 1
                const bool usable = (lap == 0 || tt < skip1);
                v.hl[lap][rank] = tt;
                v.hctx[lap][rank] = usable ? __ldcg(v.hea[lap] + i) : kCtxHard;
                v.hix[lap][tt] = rank;
            }
        }
    } else {
        // overflow (pathological weights): contiguous walk, each thread owns a range of tiles
        auto emit_hard = [&](int lap, int rank, int64_t tt) {
            const int64_t hi = ((tt + 1) * TILE < n) ? (tt + 1) * TILE : n;
            const double ai = __ldcg(v.ain + tt), ao = __ldcg(v.ain + tt + 1);
            int ea, eb;
            if (lap == 0) classify_tile2(ai, ao, g0 + hi, rel_eps, &ea, &eb);
            else classify_tile2(atot + ai, atot + ao, n_global + g0 + hi, rel_eps, &ea, &eb);
            const bool usable = (lap == 0 || tt < skip1) && ea != kCtxHard && eb == ea + 1 &&
                                ea > -1000 && ea < 900;
            v.hl[lap][rank] = (int)tt;
            v.hctx[lap][rank] = usable ? ea : kCtxHard;
            v.hix[lap][tt] = rank;
        };
        const int64_t per = (nt + FB - 1) / FB;
        const int64_t tb = (int64_t)tid * per;
        const int64_t te = (tb + per < nt) ? tb + per : nt;
        int nh0 = 0, nh1 = 0;
        for (int64_t tt = tb; tt < te; ++tt) {
            nh0 += (__ldcg(v.cls[0] + tt) == kCtxHard);
            nh1 += (__ldcg(v.cls[1] + tt) == kCtxHard);
        }
        double d0, d1;
        __syncthreads();
        int o0 = (int)block_excl_scan_f64((double)nh0, smd, d0);
        int o1 = (int)block_excl_scan_f64((double)nh1, smd, d1);
        for (int64_t tt = tb; tt < te; ++tt) {
            if (__ldcg(v.cls[0] + tt) == kCtxHard) emit_hard(0, o0++, tt);
            if (__ldcg(v.cls[1] + tt) == kCtxHard) emit_hard(1, o1++, tt);
        }
    }
    if (tid == 0) {
        ResampleInfo* I = v.info;
        I->nh0 = h0tot;
        I->nh1 = h1tot;
        I->first_skip1 = skip1;
        v.ticket[1] = 0;
        v.gmax[0] = dkey(-INFINITY);        // for the next producer of tile partials
    }
}

// ------------------------------------------------------------------ K3a
// Tables of one hard tile whose fold crosses (at most) the boundary between the
// binades eA and eA + 1: per-element inclusive prefix transducers in eA, suffix
// transducers in eA + 1, and the weights.  With them the chain kernel resolves
// the tile for ANY exact running sum entering it in O(1) serial steps: find the
// crossing element in the prefix table, do the crossing add in hardware fp64
// (smc.cljc:120), finish with one suffix transducer.
__device__ __noinline__ void hard_tile_tables(const double (&w)[IPT], int eA, Xd* __restrict__ pa,
                                              Xd* __restrict__ sb, double* __restrict__ hw,
                                              Xd* warp_tot) {
    const int i0 = threadIdx.x * IPT;
    Xd lp[IPT];
    Xd run = xd_identity();
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
        i64 f; int up, tie;
        quantize(w[j], eA, f, up, tie);
        xd_push(run, f, up, tie);
        lp[j] = run;
    }
    const Xd exl = cta_excl_scan_xd(run, warp_tot);
#pragma unroll
    for (int j = 0; j < IPT; ++j) pa[i0 + j] = xd_compose(exl, lp[j]);
    __syncthreads();
    Xd suf = xd_identity();
#pragma unroll
    for (int j = IPT - 1; j >= 0; --j) {
        i64 f; int up, tie;
        quantize(w[j], eA + 1, f, up, tie);
        Xd one = xd_identity();
        xd_push(one, f, up, tie);
        suf = xd_compose(one, suf);
        lp[j] = suf;
    }
    const Xd post = cta_excl_suffix_scan_xd(suf, warp_tot);
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
        sb[i0 + j] = xd_compose(lp[j], post);
        hw[i0 + j] = w[j];
    }
    __syncthreads();
}

// rare path of k_fold_agg, kept out of line (and loading the weights itself) so
// that it costs the streaming path neither registers nor stack
__device__ __noinline__ void hard_tile_prework(const double* __restrict__ lw,
                                               const double* __restrict__ win, int64_t n,
                                               double m, int t, int eA, Xd* __restrict__ pa,
                                               Xd* __restrict__ sb, double* __restrict__ hw,
                                               Xd* warp_tot) {
    double w[IPT];
    load_weights(lw, win, (int64_t)t * TILE, n, m, w);
    hard_tile_tables(w, eA, pa, sb, hw, warp_tot);
}

// (Writing the weights exp(lw - m) out here for the emission kernels was measured:
// k_ancestors 617 -> 530 us, k_fold_agg 348 -> 461 us at 2^27 particles - the extra
// 8 B per particle cost more than the second exponential, so the kernels recompute.)
__global__ void __launch_bounds__(TB, APOP_FOLD_CTAS) k_fold_agg(const double* __restrict__ lw,
                                                    const double* __restrict__ win, int64_t n,
                                                    WSView v, const double* __restrict__ xf) {
    __shared__ Xd warp_tot[NWARP];
    __shared__ i64 ws0[NWARP], ws1[NWARP];
    __shared__ double wa[NWARP], wb[NWARP];
    const int t = blockIdx.x;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t base = (int64_t)t * TILE;
    // every global read of the streaming path is issued up front so that the
    // tile data, the classes and the max arrive in ONE memory round trip
    double raw[IPT];
    load_raw(lw, win, base, n, raw);
    const int c0 = v.cls[0][t], c1 = v.cls[1][t];
    const double m_glob = v.info->m;
    const bool do0 = (c0 != kCtxHard), do1 = (c1 != kCtxHard && c1 != kCtxSkip);
    double w[IPT];
    weights_from_raw(win != nullptr, raw, base, n, m_glob, w);
    // accurate sum w, sum w^2 of the tile (ESS, log-mean-weight)
    double s = 0.0, s2 = 0.0;
#pragma unroll
    for (int j = 0; j < IPT; ++j) { s += w[j]; s2 += w[j] * w[j]; }
    if (xf) {                  // filtering-mean numerator sum w x (uniform branch)
        __shared__ double wx[NWARP];
        double xv[IPT], sx = 0.0;
        load_blocked(xf, base, n, xv, 0.0);
#pragma unroll
        for (int j = 0; j < IPT; ++j) sx += w[j] * xv[j];
        sx = warp_sum(sx);
        if (lane == 0) wx[wid] = sx;
        __syncthreads();
        if (threadIdx.x == 0) {
            double a = 0.0;
            for (int i = 0; i < NWARP; ++i) a += wx[i];
            v.tsum2[t] = a;
        }
    }
    // fast integer aggregates (valid when no element ties / saturates)
    const bool f0 = do0 && ctx_fast_ok(c0), f1 = do1 && ctx_fast_ok(c1);
    const double inv0 = f0 ? pow2(-c0) : 0.0, inv1 = f1 ? pow2(-c1) : 0.0;
    // integer-valued doubles: RN(w / ulp) by the 2^52 magic add; exact sums below 2^53
    double sum0 = 0.0, sum1 = 0.0;
    int bad = (do0 && !f0) | (do1 && !f1);
    const double k52 = 4503599627370496.0;
    if (f0) {
#pragma unroll
        for (int j = 0; j < IPT; ++j) {
            const double scaled = w[j] * inv0;
            const double rn = (scaled + k52) - k52;
            bad |= !(scaled < k52) | (fabs(scaled - rn) == 0.5);
            sum0 += rn;
        }
    }
    if (f1) {
#pragma unroll
        for (int j = 0; j < IPT; ++j) {
            const double scaled = w[j] * inv1;
            const double rn = (scaled + k52) - k52;
            bad |= !(scaled < k52) | (fabs(scaled - rn) == 0.5);
            sum1 += rn;
        }
    }
    s = warp_sum(s); s2 = warp_sum(s2);
    sum0 = warp_sum(sum0); sum1 = warp_sum(sum1);
    bad |= !(sum0 < 9007199254740992.0) | !(sum1 < 9007199254740992.0);
    if (lane == 0) { wa[wid] = s; wb[wid] = s2; ws0[wid] = (i64)sum0; ws1[wid] = (i64)sum1; }
    const int any_bad = __syncthreads_or(bad);
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        i64 t0 = 0, t1 = 0;
        for (int i = 0; i < NWARP; ++i) { a += wa[i]; b += wb[i]; t0 += ws0[i]; t1 += ws1[i]; }
        v.tacc[t] = a;
        v.tacc2[t] = b;
        v.tflag[t] = any_bad ? 0 : ((f0 ? 1 : 0) | (f1 ? 2 : 0));
        if (!any_bad) {
            Xd x0; x0.a0 = t0; x0.a1 = t0;
            Xd x1; x1.a0 = t1; x1.a1 = t1;
            if (do0) v.agg[0][t] = x0;
            if (do1) v.agg[1][t] = x1;
        }
    }
    if (any_bad) {
        // ---- general path: parity transducers (ties) / bit-level quantisation
        Xd a = xd_identity(), b = xd_identity();
#pragma unroll
        for (int j = 0; j < IPT; ++j) {
            i64 f; int up, tie;
            if (do0) { quantize(w[j], c0, f, up, tie); xd_push(a, f, up, tie); }
            if (do1) { quantize(w[j], c1, f, up, tie); xd_push(b, f, up, tie); }
        }
        if (do0) {
            (void)cta_excl_scan_xd(a, warp_tot);
            if (threadIdx.x == 0) v.agg[0][t] = cta_total_xd(warp_tot);
            __syncthreads();
        }
        if (do1) {
            (void)cta_excl_scan_xd(b, warp_tot);
            if (threadIdx.x == 0) v.agg[1][t] = cta_total_xd(warp_tot);
        }
    }
    // ---- hard tiles (rare, uniform per CTA): tables for the chain kernel.  Last,
    // so that the streaming path keeps no registers live across the call.
    if (c0 == kCtxHard || c1 == kCtxHard) {
        __syncthreads();
#pragma unroll 1
        for (int lap = 0; lap < 2; ++lap) {
            if ((lap ? c1 : c0) != kCtxHard) continue;
            const int h = v.hix[lap][t];
            if (h < 0 || h >= v.hcap) continue;
            const int eA = v.hctx[lap][h];
            if (eA == kCtxHard) continue;
            hard_tile_prework(lw, win, n, m_glob, t, eA, v.hpa[lap] + (size_t)h * TILE,
                              v.hsb[lap] + (size_t)h * TILE, v.hw[lap] + (size_t)h * TILE,
                              warp_tot);
        }
    }
}

// ------------------------------------------------------------------ tile scan
// Segmented scan of the tile transducers, SEGB tiles per CTA.  Hard tiles cut
// the segments (their fold depends on the exact running sum entering them), so
//   ex[t]  = composite of the easy tiles between the last hard tile before t in
//            this block (or the block start) and t, exclusive;
//   hp[t]  = index (in hl) of that hard tile, or -1;
//   bc[b]  = composite from the last hard tile of the block (or its start) to
//            the block end.
// None of this depends on the running sum, so it runs at full width; the chain
// kernel then only touches one value per block and per hard tile.
__global__ void __launch_bounds__(SEGB) k_seg_scan(WSView v, int has_x) {
    __shared__ Xd wv[SEGB / 32];
    __shared__ int wf[SEGB / 32], wh[SEGB / 32];
    __shared__ double ra[SEGB / 32], rb[SEGB / 32], rx[SEGB / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t t = (int64_t)blockIdx.x * SEGB + threadIdx.x;
    const int64_t nt = v.nt;
    {   // per-block sums of the accurate tile sums (ESS, log-mean-weight, filtering mean)
        double a = (t < nt) ? v.tacc[t] : 0.0, b2 = (t < nt) ? v.tacc2[t] : 0.0;
        double x = (has_x && t < nt) ? v.tsum2[t] : 0.0;
        a = warp_sum(a); b2 = warp_sum(b2); x = warp_sum(x);
        if (lane == 0) { ra[wid] = a; rb[wid] = b2; rx[wid] = x; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double sa = 0.0, sb = 0.0, sx = 0.0;
            for (int i = 0; i < SEGB / 32; ++i) { sa += ra[i]; sb += rb[i]; sx += rx[i]; }
            v.bacc[blockIdx.x] = sa; v.bacc2[blockIdx.x] = sb; v.baccx[blockIdx.x] = sx;
        }
    }
#pragma unroll 1
    for (int lap = 0; lap < 2; ++lap) {
        int64_t stop = nt;
        if (lap == 1 && v.info->first_skip1 < stop) stop = v.info->first_skip1;
        const int c = (t < stop) ? v.cls[lap][t] : kCtxSkip;
        const bool hard = (c == kCtxHard);
        const bool act = (c != kCtxHard) && (c != kCtxSkip);
        Xd val = act ? v.agg[lap][t] : xd_identity();
        int f = hard ? 1 : 0;
        int hm = hard ? v.hix[lap][t] : -1;
        // inclusive segmented scan inside the warp
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            Xd o;
            o.a0 = shfl_up_i64(val.a0, d);
            o.a1 = shfl_up_i64(val.a1, d);
            const int of = __shfl_up_sync(0xffffffffu, f, d);
            const int oh = __shfl_up_sync(0xffffffffu, hm, d);
            if (lane >= d) {
                if (!f) val = xd_compose(o, val);
                f |= of;
                hm = max(hm, oh);
            }
        }
        if (lane == 31) { wv[wid] = val; wf[wid] = f; wh[wid] = hm; }
        __syncthreads();
        // prefix of the preceding warps
        Xd pv = xd_identity();
        int pf = 0, ph = -1;
        for (int i = 0; i < wid; ++i) {
            pv = wf[i] ? wv[i] : xd_compose(pv, wv[i]);
            pf |= wf[i];
            ph = max(ph, wh[i]);
        }
        // inclusive value of this tile, then shift by one for the exclusive result
        Xd inc = f ? val : xd_compose(pv, val);
        const int ih = max(ph, hm);
        Xd exv;
        exv.a0 = shfl_up_i64(inc.a0, 1);
        exv.a1 = shfl_up_i64(inc.a1, 1);
        int exh = __shfl_up_sync(0xffffffffu, ih, 1);
        if (lane == 0) { exv = pv; exh = ph; }
        if (t < nt) { v.ex[lap][t] = exv; v.hp[lap][t] = exh; }
        if (threadIdx.x == SEGB - 1) v.bc[lap][blockIdx.x] = inc;
        __syncthreads();
    }
}

// ------------------------------------------------------------------ chain
constexpr int CH_BC = 1024;      // blocks of the tile scan staged in shared memory per lap

struct ChainSmem {
    FoldSmem fold;
    double bcast;
    int cross, fallback;
    i64 last[TB];                // running units after each thread's elements (hard-tile search)
    Xd bc[CH_BC];                // staged per-block composites / class of the block's last tile
    int cend[CH_BC];
    int hl[HCAP];                // staged head of the hard list: tile, ex[tile], class of the
    Xd hex[HCAP];                //   run before it, table context
    int hcr[HCAP];
    int hctx[HCAP];
};

// S' = S + (transducer of a run of easy tiles); the run must live in the binade of S
__device__ __forceinline__ double apply_run(double S, const Xd& x, int cls_run, int* err) {
    if ((x.a0 | x.a1) == 0) return S;
    const int e = ctx_of(S);
    const i64 s = xd_apply(units_of(S), x);
    if (s >= kTwo53 || e != cls_run) atomicOr(err, EB_CTX | 0x100);      // high bits: which check
    return from_units(s, e);
}

// coarse level of the prefix table of one hard tile: the entry after this
// thread's IPT elements.  Prefetched while the previous event of the chain is
// still in flight (the tables do not depend on S).
struct HardPre {
    Xd c;
    int h;                       // hard index the entry belongs to, or -1
};

__device__ __forceinline__ void hard_prefetch(const WSView& v, int lap, int h, int nh, int eA,
                                              HardPre& P) {
    P.h = -1;
    if (h >= nh || h >= v.hcap || eA == kCtxHard) return;
    P.c = v.hpa[lap][(size_t)h * TILE + threadIdx.x * IPT + (IPT - 1)];
    P.h = h;
}

// exact fold of hard tile t (index h of the lap's hard list) from the exact
// running sum S entering it.  Uniform over the CTA.  eA: context of the tile's
// tables (kCtxHard: none).  P holds this thread's coarse entry; on return it
// holds the entry of hard tile h + 1 (context eA_next).
__device__ __forceinline__ double chain_hard_tile(const WSView& v, int lap, int h, int nh,
                                                  int64_t t, double S, int eA, int eA_next,
                                                  HardPre& P, const double* __restrict__ lw,
                                                  const double* __restrict__ win, int64_t n,
                                                  double m, ChainSmem& sm) {
    const int e = ctx_of(S);
    bool general = true;
    if (P.h == h && eA != kCtxHard && (e == eA || e == eA + 1)) {
        const Xd* pa = v.hpa[lap] + (size_t)h * TILE;
        const Xd* sb = v.hsb[lap] + (size_t)h * TILE;
        const double* hw = v.hw[lap] + (size_t)h * TILE;
        const i64 s_in = units_of(S);
        if (threadIdx.x == 0) { sm.cross = INT_MAX; sm.fallback = 0; }
        __syncthreads();
        if (e == eA + 1) {                         // already in the upper binade: one suffix step
            hard_prefetch(v, lap, h + 1, nh, eA_next, P);
            if (threadIdx.x == 0) {
                const i64 s2 = xd_apply(s_in, sb[0]);
                if (s2 >= kTwo53) sm.fallback = 1;
                sm.bcast = from_units(s2, e);
            }
        } else {
            // coarse search: first thread whose elements take the sum across 2^53 units
            const i64 mine = xd_apply(s_in, P.c);
            sm.last[threadIdx.x] = mine;
            {
                const int wmin = __reduce_min_sync(0xffffffffu, (mine >= kTwo53) ? (int)threadIdx.x : INT_MAX);
                if ((threadIdx.x & 31) == 0 && wmin != INT_MAX) atomicMin(&sm.cross, wmin);
            }
            hard_prefetch(v, lap, h + 1, nh, eA_next, P);
            __syncthreads();
            const int cc = sm.cross;
            if (cc == INT_MAX) {                   // stays in eA over the whole tile
                if (threadIdx.x == TB - 1) sm.bcast = from_units(mine, eA);
            } else if ((int)threadIdx.x == cc) {   // fine search inside the chunk, one thread
                const int i0 = cc * IPT;
                Xd f[IPT], sf[IPT];
                double wj[IPT];
#pragma unroll
                for (int j = 0; j < IPT; ++j) {
                    f[j] = pa[i0 + j];
                    wj[j] = hw[i0 + j];
                    sf[j] = (i0 + j + 1 < TILE) ? sb[i0 + j + 1] : xd_identity();
                }
                i64 run = cc ? sm.last[cc - 1] : s_in, sprev = 0;
                double wc = 0.0;
                Xd sfx = xd_identity();
                bool found = false;
#pragma unroll
                for (int j = 0; j < IPT; ++j) {
                    const i64 nx = xd_apply(s_in, f[j]);
                    if (!found && nx >= kTwo53) { found = true; sprev = run; wc = wj[j]; sfx = sf[j]; }
                    run = nx;
                }
                const double Sn = from_units(sprev, eA) + wc;          // the crossing add (smc.cljc:120)
                double out = Sn;
                if (!found || ctx_of(Sn) != eA + 1) {
                    sm.fallback = 1;
                } else {
                    const i64 s2 = xd_apply(units_of(Sn), sfx);
                    if (s2 >= kTwo53) sm.fallback = 1;
                    out = from_units(s2, eA + 1);
                }
                sm.bcast = out;
            }
        }
        __syncthreads();
        general = sm.fallback != 0;
        const double r = sm.bcast;
        __syncthreads();
        if (!general) return r;
    }
    P.h = -1;              // not live across the call below; re-fetched afterwards
    // general path: parity-transducer fold, one iteration per binade crossing
    if (threadIdx.x == 0) v.info->asq += 1.0;          // diagnostics: general folds in the chain
    double w[IPT], edge[IPT];
    const int64_t base = t * TILE;
    const int nvalid = (int)((n - base < TILE) ? (n - base) : TILE);
    long long tq0 = 0, tq1 = 0, tq2 = 0;
    if (threadIdx.x == 0) tq0 = clock64();
    load_weights(lw, win, base, n, m, w);
    if (threadIdx.x == 0) tq1 = clock64();
    // A fold that starts near zero crosses a binade at every doubling of the sum
    // (elements 1, 2, 4, 8, ...): one thread adds the first elements literally,
    // the parallel fold takes over from there.
    constexpr int HEAD = APOP_HEAD;
    int pos = 0;
    if (S == 0.0 && nvalid > HEAD) {
        if (threadIdx.x < HEAD / IPT) {
#pragma unroll
            for (int j = 0; j < IPT; ++j) sm.fold.head[threadIdx.x * IPT + j] = w[j];
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            double acc = S;
#pragma unroll 8
            for (int i = 0; i < HEAD; ++i) acc = acc + sm.fold.head[i];         // smc.cljc:94
            sm.bcast = acc;
        }
        __syncthreads();
        S = sm.bcast;
        pos = HEAD;
        __syncthreads();
    }
    if (threadIdx.x == 0) { tq2 = clock64(); sm.fold.imin = 0; }
    const double r = cta_tile_fold(S, w, nvalid, edge, sm.fold, pos);
    __syncthreads();
    if (threadIdx.x == 0 && lap == 0 && h == 0) {      // diagnostics: phases of the general fold
        v.trace[480] = (double)(tq1 - tq0);            // weights (exp)
        v.trace[481] = (double)(tq2 - tq1);            // literal head
        v.trace[482] = (double)(clock64() - tq2);      // parallel fold
        v.trace[483] = (double)sm.fold.imin;           // its iterations
    }
    hard_prefetch(v, lap, h + 1, nh, eA_next, P);
    return r;
}

// One lap of the exact chain over this rank's tiles [0, stop): walks the blocks
// of the tile scan and the hard tiles in order.  Records the exact running sum
// entering every block (sblk) and entering / leaving every hard tile (shin,
// shout); the running sum entering any other tile follows from ex[] (tile_sin).
// Stops once S exceeds `limit` (second lap: no pointer lies beyond xlast);
// everything not reached is marked +inf.  *reached_stop: the walk got to `stop`.
__device__ double chain_lap(const WSView& v, int lap, double S, int64_t stop, double limit,
                            int nh, const double* __restrict__ lw, const double* __restrict__ win,
                            int64_t n, double m, ChainSmem& sm, int* err, bool* reached_stop) {
    const int nb = v.nb;
    const int nb_stop = (int)((stop + SEGB - 1) / SEGB);
    // ---- stage what the walk reads into shared memory (none of it depends on S)
    __syncthreads();
    for (int b = threadIdx.x; b < nb_stop && b < CH_BC; b += TB) {
        const int64_t be = ((int64_t)(b + 1) * SEGB < stop) ? (int64_t)(b + 1) * SEGB : stop;
        sm.bc[b] = v.bc[lap][b];
        sm.cend[b] = v.cls[lap][be - 1];
    }
    for (int i = threadIdx.x; i < nh && i < HCAP; i += TB) {
        const int t = v.hl[lap][i];
        sm.hl[i] = t;
        sm.hex[i] = v.ex[lap][t];
        sm.hcr[i] = (t > 0) ? v.cls[lap][t - 1] : kCtxHard;
        sm.hctx[i] = (i < v.hcap) ? v.hctx[lap][i] : kCtxHard;
    }
    __syncthreads();
    auto hctx_of = [&](int i) { return (i < nh && i < HCAP) ? sm.hctx[i] : kCtxHard; };
    HardPre P;
    hard_prefetch(v, lap, 0, nh, hctx_of(0), P);
    int h = 0, b = 0;
    int shin_done = 0;             // hard entries whose shin has been written
    bool over = false;
    // the running sum as (integer significand, binade): easy runs are integer adds
    i64 su = units_of(S);
    int se = ctx_of(S);
    int bad = 0;
    auto apply_units = [&](const Xd& x, int cls_run) {
        if ((x.a0 | x.a1) == 0) return;
        su = xd_apply(su, x);
        bad |= (su >= kTwo53) | (se != cls_run);
    };
    for (; b < nb_stop; ++b) {
        S = from_units(su, se);
        if (S > limit) { over = true; break; }
        if (threadIdx.x == 0) v.sblk[lap][b] = S;
        const int64_t blk_end = ((int64_t)(b + 1) * SEGB < stop) ? (int64_t)(b + 1) * SEGB : stop;
        while (h < nh) {
            const int64_t t = (h < HCAP) ? sm.hl[h] : v.hl[lap][h];
            if (t >= blk_end) break;
            const Xd x = (h < HCAP) ? sm.hex[h] : v.ex[lap][t];
            const int crun = (h < HCAP) ? sm.hcr[h] : ((t > 0) ? v.cls[lap][t - 1] : kCtxHard);
            apply_units(x, crun);
            const double S_at = from_units(su, se);
            if (threadIdx.x == 0) v.shin[lap][h] = S_at;
            shin_done = h + 1;
            if (S_at > limit) { over = true; break; }
            long long tc0 = 0;
            if (threadIdx.x == 0) tc0 = clock64();
            S = chain_hard_tile(v, lap, h, nh, t, S_at, hctx_of(h), hctx_of(h + 1), P, lw, win, n,
                                m, sm);
            if (threadIdx.x == 0 && h < 128) {       // diagnostics (apop_debug_chain_trace): cycles
                v.trace[lap * 256 + 2 * h] = (double)(clock64() - tc0);
                v.trace[lap * 256 + 2 * h + 1] = (double)tc0;
            }
            su = units_of(S);
            se = ctx_of(S);
            if (threadIdx.x == 0) v.shout[lap][h] = S;
            ++h;
        }
        if (over) { ++b; break; }
        const Xd xb = (b < CH_BC) ? sm.bc[b] : v.bc[lap][b];
        const int ce = (b < CH_BC) ? sm.cend[b] : v.cls[lap][blk_end - 1];
        apply_units(xb, ce);
    }
    S = from_units(su, se);
    if (bad && threadIdx.x == 0) atomicOr(err, EB_CTX | 0x200);
    *reached_stop = !over;
    // everything the walk did not reach
    for (int i = b + threadIdx.x; i <= nb; i += TB) v.sblk[lap][i] = INFINITY;
    for (int i = h + threadIdx.x; i < nh; i += TB) {
        if (i >= shin_done) v.shin[lap][i] = INFINITY;
        v.shout[lap][i] = INFINITY;
    }
    return S;
}

// exact running sum entering tile t in `lap` (t == nt: at the end of the lap)
__device__ __forceinline__ double tile_sin(const WSView& v, int lap, int64_t t, int* err) {
    if (t >= v.nt) return v.send[lap];
    if (lap == 1 && t >= v.info->first_skip1) return INFINITY;
    const int c = v.cls[lap][t];
    if (c == kCtxSkip) return INFINITY;
    if (c == kCtxHard) return v.shin[lap][v.hix[lap][t]];
    const int h = v.hp[lap][t];
    const double base = (h >= 0) ? v.shout[lap][h] : v.sblk[lap][t / SEGB];
    if (!(base < INFINITY)) return INFINITY;
    const Xd x = v.ex[lap][t];
    if ((x.a0 | x.a1) == 0) return base;
    const i64 s = xd_apply(units_of(base), x);
    if (s >= kTwo53 || ctx_of(base) != c) { atomicOr(err, EB_CTX | 0x400); return INFINITY; }
    return from_units(s, c);
}

// ---- sharded fold: what one shard contributes to a lap of the sequential fold ----
// While the running sum stays inside one binade a whole shard is ONE parity
// transducer (exact_fold.cuh), and transducers compose: a shard without a
// binade-crossing ("hard") tile publishes its composite right away, without knowing
// the sum entering it (CK_EASY), and every rank applies the composites of the lower
// ranks itself.  Only a shard that owns a crossing has to wait for its exact input;
// it publishes the exact sum leaving it (CK_EXACT).  The fold of smc.cljc:94,118-137
// crosses about log2(N) binades in all, nearly all of them in rank 0, so a lap
// costs a handful of dependent hops instead of one per rank.
enum { CK_EASY = 1, CK_EXACT = 2 };

// all ranks receive this rank's part of lap `lap` (threads 0 .. world-1 send)
__device__ __forceinline__ void comp_publish(const Fabric& F, int par, int lap, int kind,
                                             long long a, long long b, int cls,
                                             unsigned long long seq) {
    if ((int)threadIdx.x < F.world)
        mail_send_bits(&F.mbox[threadIdx.x]->comp[par][lap][F.rank], a, b,
                       ((long long)kind << 32) | (long long)(unsigned)cls, seq);
}
// the running sum after the parts of ranks [q0, q1) have been applied to S, in rank
// order (one thread)
__device__ double comp_walk(const Fabric& F, int par, int lap, double S, int q0, int q1,
                            unsigned long long seq, int* err, int* comm) {
    for (int q = q0; q < q1; ++q) {
        long long o[3];
        if (!mail_wait_bits(&F.mbox[F.rank]->comp[par][lap][q], seq, o, F.spin_cycles)) {
            *comm = EB_COMM | ((lap ? CW_COMP1 : CW_COMP0) << 16) | (q << 24);
            return S;
        }
        const int kind = (int)(o[2] >> 32), cls = (int)(unsigned)(o[2] & 0xffffffffLL);
        if (kind == CK_EXACT) {
            S = __longlong_as_double(o[0]);
        } else {
            Xd x; x.a0 = o[0]; x.a1 = o[1];
            S = apply_run(S, x, cls, err);
        }
    }
    return S;
}

__global__ void __launch_bounds__(TB) k_chain(const double* __restrict__ lw,
                                                const double* __restrict__ win, int64_t n,
                                                int64_t n_total, WSView v, StepOut so, Fabric F,
                                                unsigned long long seq) {
    __shared__ ChainSmem sm;
    __shared__ double smd[33];
    __shared__ double sh_in;
    __shared__ double pa[kMaxWorld], pb[kMaxWorld], px[kMaxWorld];
    __shared__ int sh_comm;
    __shared__ Xd sh_x;
    const int64_t nt = v.nt;
    ResampleInfo* I = v.info;
    const double m = I->m;
    const bool sharded = F.world > 1;
    const int par = (int)(seq & 1);
    if (threadIdx.x == 0) sh_comm = 0;
    long long tg0 = 0, tg1 = 0, tg2 = 0;
    if (threadIdx.x == 0) tg0 = gtimer_ns();
    // composite of this rank's blocks [0, nbk) of lap `lap` (no hard tile among them)
    auto shard_composite = [&](int lap, int nbk) {
        if (threadIdx.x == 0) {
            Xd x = xd_identity();
            for (int b = 0; b < nbk; ++b) x = xd_compose(x, v.bc[lap][b]);
            sh_x = x;
        }
        __syncthreads();
    };

    // ---------------- lap 0: exact total (smc.cljc:94)
    double S0 = 0.0;
    const bool easy0 = sharded && I->nh0 == 0;
    if (sharded) {
        if (easy0) {
            shard_composite(0, v.nb);
            comp_publish(F, par, 0, CK_EASY, sh_x.a0, sh_x.a1, v.cls[0][0], seq);
        }
        if (threadIdx.x == 0)
            sh_in = comp_walk(F, par, 0, 0.0, 0, F.rank, seq, &I->err, &sh_comm);
        __syncthreads();
        S0 = sh_in;
    }
    bool reached;
    const double Send0 = chain_lap(v, 0, S0, nt, INFINITY, I->nh0, lw, win, n, m, sm, &I->err,
                                   &reached);
    double total = Send0;
    if (sharded) {
        if (!easy0) comp_publish(F, par, 0, CK_EXACT, __double_as_longlong(Send0), 0, 0, seq);
        if (threadIdx.x == 0)
            sh_in = comp_walk(F, par, 0, Send0, F.rank + 1, F.world, seq, &I->err, &sh_comm);
        __syncthreads();
        total = sh_in;
    }
    if (threadIdx.x == 0) tg1 = gtimer_ns();
    // accurate sum w / sum w^2 (ESS = (sum w)^2 / sum w^2, dist_test.cljc:83-94)
    double la = 0.0, lb = 0.0, lx = 0.0;
    for (int b = threadIdx.x; b < v.nb; b += TB) { la += v.bacc[b]; lb += v.bacc2[b]; lx += v.baccx[b]; }
    double acc = block_sum_f64(la, smd);
    double acc2 = block_sum_f64(lb, smd);
    double accx = block_sum_f64(lx, smd);
    if (sharded) {
        if (threadIdx.x < F.world) {
            const int q = threadIdx.x;
            mail_send(&F.mbox[q]->acc[par][F.rank], acc, acc2, accx, seq);
            double o[3];
            if (!mail_wait(&F.mbox[F.rank]->acc[par][q], seq, o, F.spin_cycles))
                sh_comm = EB_COMM | (CW_ACC << 16) | (q << 24);
            pa[q] = o[0]; pb[q] = o[1]; px[q] = o[2];
        }
        __syncthreads();
        acc = 0.0; acc2 = 0.0; accx = 0.0;
        for (int q = 0; q < F.world; ++q) { acc += pa[q]; acc2 += pb[q]; accx += px[q]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        v.send[0] = Send0;
        const i64 n_new = I->n_new;
        int err = sh_comm;
        if (!(total > 0.0) || !(total < INFINITY)) err |= EB_ZERO;
        Comb& c = *v.comb;
        if (!(err & EB_ZERO)) {
            comb_build(c, total, I->u, n_new);
            if (c.err) err |= EB_COMB;
        } else {
            c.nseg = 0; c.n_new = n_new; c.x0 = 0; c.xlast = 0; c.step = 0; c.err = 1;
        }
        int hint = 0;
        I->total = total;
        I->step = total / (double)n_new;                                  // smc.cljc:95
        I->J = (err & (EB_ZERO | EB_COMB)) ? 0 : comb_count_below(c, total, hint);
        const bool mfin = (m > -INFINITY && m < INFINITY);
        I->new_lw = mfin ? (log(I->step) + m) : -INFINITY;                // smc.cljc:98-103
        I->ess = (acc2 > 0.0) ? acc * acc / acc2 : 0.0;
        I->log_mean_approx = mfin ? log(acc / (double)n_total) + m : -INFINITY;
        if (err) atomicOr(&I->err, err);
        I->skip = 0;
        if (so.sd) {                       // per-step outputs of a fused sweep
            SweepDev* sd = so.sd;
            so.ess[so.t - 1] = I->ess;
            if (so.has_x) so.filt[so.t - 1] = accx / acc;
            sd->err |= I->err | err;
            if (so.cond) so.anc[0] = 0;          // retained particle back at position 0
            // adaptive resampling (the reference resamples at every
            // observe): keep the weighted particles while ESS >= threshold * N
            const bool skip = so.ess_thr > 0.0 && !(err | I->err) &&
                              I->ess >= so.ess_thr * (double)n_total;
            if (skip) {
                I->skip = 1;
                sd->carry = 1;
                so.log_mean_w[so.t - 1] = 0.0;   // deferred to the next resample
            } else {
                const double nl = I->new_lw;
                const double inc = nl - sd->base;
                so.log_mean_w[so.t - 1] = inc;
                sd->log_z += inc;
                sd->base = so.reset ? 0.0 : nl;  // pgibbs.cljc:88 / smc.cljc:98-103
                sd->carry = 0;
            }
        }
        tg2 = gtimer_ns();
    }
    __syncthreads();
    if (I->err & (EB_ZERO | EB_COMB | EB_NAN | EB_INF)) {
        for (int i = threadIdx.x; i <= v.nb; i += TB) v.sblk[1][i] = INFINITY;
        for (int i = threadIdx.x; i < I->nh1; i += TB) { v.shin[1][i] = INFINITY; v.shout[1][i] = INFINITY; }
        if (threadIdx.x == 0) v.send[1] = INFINITY;
        // keep the fold moving so that no peer waits for this rank
        if (sharded) comp_publish(F, par, 1, CK_EXACT, __double_as_longlong(INFINITY), 0, 0, seq);
        return;
    }
    const double xlast = v.comb->xlast;

    // ---------------- lap 1: the wheel wraps with acc NOT reset (smc.cljc:133-136).
    // The lap ends at the first tile that no pointer can reach (first_skip1): a shard
    // that is folded completely and owns no crossing is one transducer again, a shard
    // that is skipped completely is the identity.
    int64_t stop = I->first_skip1;
    if (stop > nt) stop = nt;
    double S1 = total;
    const bool easy1 = sharded && ((stop == nt && I->nh1 == 0) || stop == 0);
    if (sharded) {
        if (easy1) {
            shard_composite(1, stop == 0 ? 0 : v.nb);
            comp_publish(F, par, 1, CK_EASY, sh_x.a0, sh_x.a1, stop == 0 ? 0 : v.cls[1][0], seq);
        }
        if (threadIdx.x == 0) {
            int comm = 0;
            sh_in = comp_walk(F, par, 1, total, 0, F.rank, seq, &I->err, &comm);
            if (comm) {
                atomicOr(&I->err, comm);
                if (so.sd) so.sd->err |= comm;
            }
        }
        __syncthreads();
        S1 = sh_in;
    }
    const double S = chain_lap(v, 1, S1, stop, xlast, I->nh1, lw, win, n, m, sm, &I->err, &reached);
    if (sharded && !easy1) comp_publish(F, par, 1, CK_EXACT, __double_as_longlong(S), 0, 0, seq);
    // the walk may only end at a skipped tile if no pointer of the second lap lies beyond
    if (reached && stop < nt && !(S > xlast) && threadIdx.x == 0) {
        atomicOr(&I->err, EB_CTX | 0x800);
        if (so.sd) so.sd->err |= EB_CTX;
    }
    if (threadIdx.x == 0) v.send[1] = S;
    if (threadIdx.x == 0) {          // phase durations in ns (apop_debug_resolve_phases)
        const long long tg3 = gtimer_ns();
        I->wx_sum[0] = (double)(tg1 - tg0);      // lap 0 (incl. waiting for the lower ranks)
        I->wx_sum[1] = (double)(tg2 - tg1);      // sums, exchange, comb table
        I->wx_sum[2] = (double)(tg3 - tg2);      // lap 1
        I->wx_sum[3] = (double)v.comb->nseg;
    }
}

// ------------------------------------------------------------------ K3b/K4
__device__ __forceinline__ int64_t slot_of(i64 k, int lap, int order, i64 J, i64 n_new) {
    if (order == APOP_ORDER_EMISSION) return k;
    if (order == APOP_ORDER_REFERENCE) return n_new - 1 - k;      // conj onto nil: smc.cljc:127
    return lap == 0 ? k + (n_new - J) : k - J;                    // sorted physical order
}

// where offspring indices go: the local ancestor array, or (sharded) the array
// of whichever rank owns the output slot, written through peer memory
struct AncOut {
    RunRec* runs; RunRec* bigruns; int* nruns; int runs_cap, big_cap, heavy_w;
    int32_t* local;
    const PeerI32* peerp;        // points at the kernel's __grid_constant__ parameter
    int world;
    i64 nl;                      // particles per rank
    __device__ __forceinline__ int32_t* at(i64 slot) const {
        if (world == 1) return local + slot;
        const int q = (int)(slot / nl);
        return peerp->p[q] + (slot - (i64)q * nl);
    }
};

struct EmitCtx {
    AncOut out;
    int64_t out_offset;
    i64 J, n_new;
    int order, lap;
    __device__ __forceinline__ void put(i64 k, int32_t val) const {
        *out.at(out_offset + slot_of(k, lap, order, J, n_new)) = val;
    }
};

// ---- fast path of one lap of an easy tile ---------------------------------
// Every edge of the tile and every pointer that can fall below them live in one
// binade (context e) and in ONE comb segment g, strictly inside it
// (C0 > g.j0, C1 < g.j0 + g.len).  Then, in integer units of 2^e,
//   pointer k of the segment = X0 + k q,      edge_j = s_j,
//   #pointers below edge_j   = ceil((s_j - X0) / q)            (strict <, smc.cljc:122)
// and (s_j - X0) is an integer below 2^53, so the whole computation runs in
// fp64 on integer-valued doubles, exactly: est = RN(sD * (1/q)) is within 1/2
// of the true ratio, r = fma(-est, q, sD) is the exact remainder, and
// ceil = est + (r > 0).  Offspring are staged in shared memory as "head" marks
// (element j writes j+1 at its first slot), a max-scan fills the runs, and the
// window is copied out with coalesced stores.
constexpr int STAGE_CAP = 8192;            // staged slots per window (int32)

struct StageSmem {
    int buf[STAGE_CAP];
    int wmax[NWARP];
    double dsum[NWARP];
    i64 C0, C1;
    int gi, ok;
};

// exclusive block scan of integer-valued doubles (exact below 2^53)
__device__ __forceinline__ double cta_excl_scan_f64(double v, double* wsum, double& total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    double inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        double o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    // every warp scans the NWARP (= 16) warp totals with shuffles
    double ws = (lane < NWARP) ? wsum[lane] : 0.0;
#pragma unroll
    for (int d = 1; d < NWARP; d <<= 1) {
        double o = __shfl_up_sync(0xffffffffu, ws, d);
        if (lane >= d) ws += o;
    }
    total = __shfl_sync(0xffffffffu, ws, NWARP - 1);
    double pre = __shfl_sync(0xffffffffu, ws, (wid > 0) ? wid - 1 : 0);
    if (wid == 0) pre = 0.0;
    __syncthreads();
    return pre + (inc - v);
}

// Quantise the tile in context e with the fp64 magic-number trick; incD[j] is
// the exact integer increment of element j as a double.  Returns (uniformly)
// false if any element ties or saturates.
__device__ __forceinline__ bool tile_quantize_f64(int e, const double (&w)[IPT],
                                                  double (&incD)[IPT]) {
    const double inv_u = pow2(-e);
    int bad = 0;
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
        const double scaled = w[j] * inv_u;
        const double t = scaled + 4503599627370496.0;
        const double rn = t - 4503599627370496.0;
        const double diff = scaled - rn;                          // exact
        bad |= !(scaled < 4503599627370496.0) | (fabs(diff) == 0.5);
        incD[j] = rn;
    }
    return !__syncthreads_or(bad);
}

__device__ __forceinline__ int find_seg_of_index(const Comb& c, i64 k) {
    int lo = 0, hi = c.nseg - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (c.seg[mid].j0 <= k) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// returns false (uniformly, nothing written) when the tile does not qualify
__device__ __forceinline__ bool staged_emit(const Comb& comb, int e, double Sin, double Sout,
                                            i64 C0in, i64 C1in, int gsel,
                                            const double (&incD)[IPT], int32_t val_tile0,
                                            const AncOut& A, int64_t out_offset, i64 J,
                                            i64 n_new, int order, int lap, StageSmem& st,
                                            int* err) {
    // the window [C0, C1) of pointer indices and its comb segment were computed
    // per tile boundary by k_expand (uniform over the CTA)
    const CombSeg& gs = comb.seg[gsel];
    const int ok = (C1in == C0in) ? 2
                 : (gs.e == e && gs.q > 0 && C0in > gs.j0 && C1in < gs.j0 + gs.len) ? 1 : 0;
    if (ok != 1) return ok == 2;                                   // 2: no offspring in this lap
    double loc = 0.0;
#pragma unroll
    for (int j = 0; j < IPT; ++j) loc += incD[j];
    double tile_units;
    const double pre = cta_excl_scan_f64(loc, st.dsum, tile_units);
    const i64 s_in = units_of(Sin);
    const i64 s_out = s_in + (i64)tile_units;
    // consistency: the tile's integer sum must land on the chain's exact edge
    if ((s_out >= kTwo53 || (Sout < INFINITY && from_units(s_out, e) != Sout)) && threadIdx.x == 0)
        atomicOr(err, EB_CTX | 0x1000);
    const i64 C0 = C0in, C1 = C1in;
    const CombSeg& g = gs;
    const double qD = (double)g.q;
    const double invq = fm::div_fast(1.0, qD);
    const double magic = 6755399441055744.0;                      // 1.5 * 2^52
    // relative pointer index k(s) = ceil((s - X0)/q); k0 = k(S_in) = C0 - g.j0
    const int k0 = (int)(C0 - g.j0);
    double sD = (double)(s_in - g.X0) + pre;                      // exact
    int kprev;
    {
        const double tm = sD * invq + magic;
        const double est = tm - magic;
        const double r = fma(-est, qD, sD);
        kprev = __double2loint(tm) + (r > 0.0 ? 1 : 0);
    }
    int ks[IPT], kst[IPT];
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
        sD += incD[j];
        const double tm = sD * invq + magic;
        const double est = tm - magic;
        const double r = fma(-est, qD, sD);
        const int k = __double2loint(tm) + (r > 0.0 ? 1 : 0);
        kst[j] = kprev - k0;                                      // first slot (window relative)
        ks[j] = k - k0;                                           // one past the last slot
        kprev = k;
    }
    const int W = (int)(C1 - C0);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (W > A.heavy_w) {
        // This tile owns so many offspring that emitting them here would serialise
        // the step (degenerate weights: one particle with millions of copies).
        // Record one run per element; k_fill_runs writes them with the whole GPU.
        if (tid == 0) {
            const int b = atomicAdd(A.nruns, TILE);
            st.gi = (b + TILE <= A.runs_cap) ? b : -1;
        }
        __syncthreads();
        const int rb = st.gi;
        __syncthreads();
        if (rb >= 0) {
#pragma unroll
            for (int j = 0; j < IPT; ++j) {
                RunRec r;
                r.k0 = C0 + kst[j];
                r.count = ks[j] - kst[j];
                r.val = (int)((unsigned)(val_tile0 + tid * IPT + j) | ((unsigned)lap << 31));
                A.runs[rb + tid * IPT + j] = r;
                if (r.count >= BIG_RUN) {
                    const int bi = atomicAdd(A.nruns + 1, 1);
                    if (bi < A.big_cap) { A.bigruns[bi] = r; A.runs[rb + tid * IPT + j].count = 0; }
                }
            }
            return true;
        }
    }
    // windows of at most STAGE_CAP - 4 slots; inside shared memory the window is
    // shifted by a0 = (destination address / 4) mod 4 so that 16-byte groups of
    // the staging buffer line up with 16-byte groups of the output array and
    // the copy-out runs on 128-bit loads and stores
    constexpr int WCAP = STAGE_CAP - 4;
    for (int wb = 0; wb < W; wb += WCAP) {
        const int wn = (W - wb < WCAP) ? (W - wb) : WCAP;
        const i64 slot0 = out_offset + slot_of(C0 + wb, lap, order, J, n_new);
        const bool vec = (A.world == 1) && (order != APOP_ORDER_REFERENCE);
        const int a0 = vec ? (int)((reinterpret_cast<uintptr_t>(A.local + slot0) >> 2) & 3) : 0;
        const int tot = a0 + wn;
        const int wn8 = (tot + 7) & ~7;
        for (int i = tid * 4; i < wn8; i += TB * 4)
            *reinterpret_cast<int4*>(&st.buf[i]) = make_int4(0, 0, 0, 0);
        __syncthreads();
        if (W <= WCAP) {
#pragma unroll
            for (int j = 0; j < IPT; ++j)
                if (kst[j] < ks[j]) st.buf[a0 + kst[j]] = tid * IPT + j + 1;
        } else {
#pragma unroll
            for (int j = 0; j < IPT; ++j) {
                const int a = kst[j] > wb ? kst[j] : wb;
                const int b = ks[j] < wb + wn ? ks[j] : wb + wn;
                if (a < b) st.buf[a0 + a - wb] = tid * IPT + j + 1;
            }
        }
        __syncthreads();
        // max-scan over the window, 8 consecutive entries per thread per round
        int carry = 0;
        for (int rb = 0; rb < wn8; rb += TB * 8) {
            const int i = rb + tid * 8;
            int v[8];
            if (i < wn8) {
                const int4 a = *reinterpret_cast<const int4*>(&st.buf[i]);
                const int4 b = *reinterpret_cast<const int4*>(&st.buf[i + 4]);
                v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
                v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
            } else {
#pragma unroll
                for (int q = 0; q < 8; ++q) v[q] = 0;
            }
#pragma unroll
            for (int q = 1; q < 8; ++q) v[q] = max(v[q], v[q - 1]);
            int inc = v[7];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int o = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d) inc = max(inc, o);
            }
            int ex = __shfl_up_sync(0xffffffffu, inc, 1);
            if (lane == 0) ex = 0;
            if (lane == 31) st.wmax[wid] = inc;
            __syncthreads();
            int wm = (lane < NWARP) ? st.wmax[lane] : 0;
#pragma unroll
            for (int d = 1; d < NWARP; d <<= 1) {
                const int o = __shfl_up_sync(0xffffffffu, wm, d);
                if (lane >= d) wm = max(wm, o);
            }
            const int all = max(carry, __shfl_sync(0xffffffffu, wm, NWARP - 1));
            int pre_w = __shfl_sync(0xffffffffu, wm, (wid > 0) ? wid - 1 : 0);
            pre_w = (wid > 0) ? max(pre_w, carry) : carry;
            ex = max(ex, pre_w);
            if (i < wn8) {
#pragma unroll
                for (int q = 0; q < 8; ++q) v[q] = max(v[q], ex) - 1 + val_tile0;
                *reinterpret_cast<int4*>(&st.buf[i]) = make_int4(v[0], v[1], v[2], v[3]);
                *reinterpret_cast<int4*>(&st.buf[i + 4]) = make_int4(v[4], v[5], v[6], v[7]);
            }
            carry = all;
            __syncthreads();
        }
        // coalesced copy-out (the three orders are contiguous runs, ascending or descending)
        if (vec) {
            // entries [a0, tot) of the buffer go to dst[0 .. wn); dst - a0 is 16-byte aligned
            int32_t* dst = A.local + slot0 - a0;
            const int g_lo = (a0 + 3) >> 2, g_hi = tot >> 2;          // full groups [g_lo, g_hi)
            for (int g = g_lo + tid; g < g_hi; g += TB)
                *reinterpret_cast<int4*>(dst + 4 * g) = *reinterpret_cast<const int4*>(&st.buf[4 * g]);
            if (tid < 8) {
                // ragged ends: the first group's [a0, 4) and the last group's [4 g_hi, tot)
                const int i = (tid < 4) ? tid : 4 * g_hi + (tid - 4);
                const bool head = (tid < 4) && i >= a0 && i < tot && g_lo > 0;
                const bool tail = (tid >= 4) && i < tot && i >= a0 && (g_hi >= g_lo);
                if (head || tail) dst[i] = st.buf[i];
            }
        } else if (A.world == 1) {
            int32_t* dst = A.local + slot0;
            for (int i = tid; i < wn; i += TB) dst[-i] = st.buf[i];
        } else {
            // the window may straddle shard boundaries: the part beyond this
            // rank's slots is stored into the owner's array over NVLink
            const int dir = (order == APOP_ORDER_REFERENCE) ? -1 : 1;
            const i64 lo = (dir > 0) ? slot0 : slot0 - (wn - 1);
            const int q_lo = (int)(lo / A.nl);
            const i64 off_lo = lo - (i64)q_lo * A.nl;
            for (int i = tid; i < wn; i += TB) {
                i64 li = off_lo + ((dir > 0) ? i : (wn - 1 - i));
                int q = q_lo;
                while (li >= A.nl) { li -= A.nl; ++q; }
                A.peerp->p[q][li] = st.buf[i];
            }
        }
        __syncthreads();
    }
    return true;
}

// general path of one lap: exact edges as doubles, one count per element
__device__ __noinline__ void generic_emit(const Comb& comb, double Sin, const double (&w)[IPT],
                                          int nvalid, int32_t val0, const EmitCtx& E,
                                          FoldSmem& sm) {
    const int lane = threadIdx.x & 31;
    const int i0 = threadIdx.x * IPT;
    double edge[IPT];
    {
        // A fold that starts at zero crosses a binade at every doubling of the sum
        // (elements 1, 2, 4, 8, ...): one thread adds the first elements literally
        // (smc.cljc:120), the parallel fold takes over from there (as in k_chain).
        constexpr int HEAD = 1024;           // measured (chain trace, cycles of tile 0): 1024 -> 58.8k, 128 -> 78.6k
        int pos = 0;
        double S = Sin;
        if (Sin == 0.0 && nvalid > HEAD) {
            if (threadIdx.x < HEAD / IPT) {
#pragma unroll
                for (int j = 0; j < IPT; ++j) sm.head[threadIdx.x * IPT + j] = w[j];
            }
            __syncthreads();
            if (threadIdx.x == 0) {
                double acc = 0.0;
#pragma unroll 8
                for (int i = 0; i < HEAD; ++i) { acc = acc + sm.head[i]; sm.head[i] = acc; }
            }
            __syncthreads();
            if (threadIdx.x < HEAD / IPT) {
#pragma unroll
                for (int j = 0; j < IPT; ++j) edge[j] = sm.head[threadIdx.x * IPT + j];
            }
            S = sm.head[HEAD - 1];
            pos = HEAD;
            __syncthreads();
        }
        (void)cta_tile_fold(S, w, nvalid, edge, sm, pos);
    }
    sm.last_edge[threadIdx.x] = edge[IPT - 1];
    __syncthreads();
    const double pred = (threadIdx.x == 0) ? Sin : sm.last_edge[threadIdx.x - 1];
    int hint = comb.nseg - 1;
    i64 c_prev = (i0 < nvalid) ? comb_count_below(comb, pred, hint) : 0;
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
        const bool valid = (i0 + j < nvalid);
        i64 c = c_prev;
        if (valid) c = comb_count_below(comb, edge[j], hint);
        i64 cnt = c - c_prev;
        const int32_t val = val0 + j;
        if (cnt >= BIG_RUN) {
            // one element with a huge number of offspring: hand the run to k_fill_runs
            const int bi = atomicAdd(E.out.nruns + 1, 1);
            if (bi < E.out.big_cap) {
                RunRec r;
                r.k0 = c_prev; r.count = (int)cnt;
                r.val = (int)((unsigned)val | ((unsigned)E.lap << 31));
                E.out.bigruns[bi] = r;
                cnt = 0;
            }
        }
        const bool heavy = cnt > 8;
        if (!heavy)
            for (i64 k = c_prev; k < c_prev + cnt; ++k) E.put(k, val);
        unsigned hm = __ballot_sync(0xffffffffu, heavy);
        while (hm) {
            const int src = __ffs(hm) - 1;
            hm &= hm - 1;
            const i64 hs = shfl_i64(c_prev, src), he = shfl_i64(c_prev + cnt, src);
            const int32_t hv = __shfl_sync(0xffffffffu, val, src);
            for (i64 k = hs + lane; k < he; k += 32) E.put(k, hv);
        }
        c_prev = c;
    }
    __syncthreads();
}

// exact running sum entering every tile, both laps, from the chain's records, the
// pointer counts at the tile boundaries and the comb segment they fall in, and from
// those the per-tile record of k_ancestors: the emission kernel needs no serial
// search and no decision of its own.  A block covers 255 tiles (256 boundaries).
constexpr int XB = 256;
__global__ void __launch_bounds__(XB) k_expand(WSView v, int order, int32_t* __restrict__ anc,
                                               int64_t out_offset,
                                               const __grid_constant__ PeerI32 ancp, int world,
                                               int64_t nl) {
    __shared__ double sS[2][XB];
    __shared__ i64 sC[2][XB];
    __shared__ int sG[2][XB];
    const int64_t t = (int64_t)blockIdx.x * (XB - 1) + threadIdx.x;
    if (t == 0) { v.nruns[0] = 0; v.nruns[1] = 0; }           // run lists of this resample
    const ResampleInfo* I = v.info;
    const Comb& comb = *v.comb;
    const bool dead = comb.err || comb.nseg <= 0;
    if (t <= v.nt) {
        int* err = &v.info->err;
        const double s0 = tile_sin(v, 0, t, err), s1 = tile_sin(v, 1, t, err);
        i64 c0 = 0, c1 = 0;
        int g0 = 0, g1 = 0;
        if (!dead) {
            int h0 = comb.nseg - 1, h1 = comb.nseg - 1;
            c0 = comb_count_below(comb, s0, h0);
            c1 = comb_count_below(comb, s1, h1);
            g0 = find_seg_of_index(comb, c0);
            g1 = find_seg_of_index(comb, c1);
        }
        sS[0][threadIdx.x] = s0; sS[1][threadIdx.x] = s1;
        sC[0][threadIdx.x] = c0; sC[1][threadIdx.x] = c1;
        sG[0][threadIdx.x] = g0; sG[1][threadIdx.x] = g1;
        if (threadIdx.x < XB - 1 || t == v.nt) {               // each boundary written once
            v.sin[0][t] = s0; v.sin[1][t] = s1;
            v.cnt[0][t] = c0; v.cnt[1][t] = c1;
            v.gseg[0][t] = g0; v.gseg[1][t] = g1;
        }
    }
    __syncthreads();
    if (threadIdx.x >= XB - 1 || t >= v.nt) return;
    const int i = threadIdx.x;
    TileRec R;
    R.sD0 = 0.0; R.qD = 1.0; R.invq = 1.0; R.inv_u = 1.0; R.units = 0.0; R.m = I->m;
    R.dst = nullptr; R.k0 = 0; R.wm = (unsigned)TM_NONE << 28;
    if (I->skip) {
        // no resample at this observe: every particle is its own ancestor (slot k <- k)
        R.dst = anc + t * TILE;
        R.wm = (unsigned)TM_IDENT << 28;
    } else if (!dead && !I->err) {
        const double x0 = comb.x0, xlast = comb.xlast;
        // a lap emits if some pointer falls into this tile (x0 is uniform on the wheel:
        // all tiles but the one holding x0 emit in exactly one lap)
        const bool do0 = (sS[0][i + 1] > x0) && !(sS[0][i] > xlast) && (sC[0][i] != sC[0][i + 1]);
        const bool do1 = !(sS[1][i] > xlast) && (sC[1][i] != sC[1][i + 1]);
        if (do0 || do1) {
            R.wm = (unsigned)TM_GENERAL << 28;
            const int lap = do0 ? 0 : 1;
            const int cls = v.cls[lap][t];
            const double Sin = sS[lap][i], Sout = sS[lap][i + 1];
            const i64 C0 = sC[lap][i], C1 = sC[lap][i + 1];
            const i64 W = C1 - C0;
            const CombSeg& gs = comb.seg[sG[lap][i]];
            constexpr int WCAP = 8192 - 4;
            if (!(do0 && do1) && ((v.tflag[t] >> lap) & 1) && cls != kCtxHard && cls != kCtxSkip &&
                ctx_fast_ok(cls) && ctx_of(Sin) == cls && gs.e == cls && gs.q > 0 && C0 > gs.j0 &&
                C1 < gs.j0 + gs.len && W <= WCAP && W <= v.heavy_w &&
                order != APOP_ORDER_REFERENCE) {
                const i64 slot0 = out_offset + slot_of(C0, lap, order, I->J, I->n_new);
                int32_t* dst = nullptr;
                if (world == 1) {
                    dst = anc + slot0;
                } else {
                    const int q0 = (int)(slot0 / nl), q1 = (int)((slot0 + W - 1) / nl);
                    if (q0 == q1) dst = ancp.p[q0] + (slot0 - (i64)q0 * nl);   // one owner: peer stores
                }
                if (dst) {
                    const i64 s_in = units_of(Sin);
                    R.sD0 = (double)(s_in - gs.X0);                            // exact
                    R.qD = (double)gs.q;
                    R.invq = fm::div_fast(1.0, R.qD);
                    R.inv_u = pow2(-cls);
                    // the tile's integer sum must land on the chain's exact edge
                    if (Sout < INFINITY)
                        R.units = (ctx_of(Sout) == cls) ? (double)(units_of(Sout) - s_in)
                                                        : 1152921504606846976.0;   // 2^60: never
                    else
                        R.units = -(double)(kTwo53 - s_in);
                    R.dst = dst;
                    R.k0 = (int)(C0 - gs.j0);
                    R.wm = (unsigned)W | ((unsigned)TM_FAST << 28);
                }
            }
        }
    }
    v.rec[t] = R;
    if ((R.wm >> 28) == TM_GENERAL) v.glist[atomicAdd(v.ngen, 1)] = (int)t;
}

union AncSmem {
    FoldSmem fold;
    StageSmem stage;
};

// general path of one tile: any class, any order, both laps, deferred runs.  The
// tiles that need it (hard tiles, the tile that holds x0, ties, windows that straddle
// a shard boundary, reference order) were listed by k_expand.
__device__ __forceinline__ void ancestors_general(const int t, const double* __restrict__ lw,
                                               const double* __restrict__ win, int64_t n,
                                               int64_t g0, const WSView& v, int order,
                                               int32_t* __restrict__ anc, int64_t out_offset,
                                               const PeerI32& ancp, int world, AncSmem& sm) {
    const Comb& comb = *v.comb;          // global (L1/L2 resident); the fast path reads one segment
    const ResampleInfo* I = v.info;
    const int64_t base = (int64_t)t * TILE;
    // Every global read of the common path is issued here, unconditionally and
    // before the first branch, so that the tile data and its metadata arrive in
    // one memory round trip instead of a chain of dependent ones.
    double raw[IPT];
    load_raw(lw, win, base, n, raw);
    const int ierr = I->err;
    const double x0 = comb.x0, xlast = comb.xlast;
    const double si0 = v.sin[0][t], so0 = v.sin[0][t + 1];
    const double si1 = v.sin[1][t], so1 = v.sin[1][t + 1];
    const i64 c00 = v.cnt[0][t], c01 = v.cnt[0][t + 1];
    const i64 c10 = v.cnt[1][t], c11 = v.cnt[1][t + 1];
    const int cls0 = v.cls[0][t], cls1 = v.cls[1][t];
    const int gs0 = v.gseg[0][t], gs1 = v.gseg[1][t];
    const double m = I->m;
    const i64 J = I->J, n_new = I->n_new;
    if (ierr) return;
    // a lap emits if some pointer falls into this tile
    const bool do0 = (so0 > x0) && !(si0 > xlast) && (c00 != c01);
    const bool do1 = !(si1 > xlast) && (c10 != c11);
    if (!do0 && !do1) return;
    const int nvalid = (int)((n - base < TILE) ? (n - base) : TILE);
    const int i0 = threadIdx.x * IPT;
    AncOut A;
    A.local = anc; A.peerp = &ancp; A.world = world; A.nl = n;
    A.runs = v.runs; A.bigruns = v.bigruns; A.nruns = v.nruns; A.runs_cap = v.runs_cap;
    A.big_cap = v.big_cap; A.heavy_w = v.heavy_w;
    // x0 is uniform on the wheel, so all tiles but the one holding x0 emit in
    // exactly one lap: that lap consumes the preloaded tile ...
    const int lap = do0 ? 0 : 1;
    {
        const double Sin = lap ? si1 : si0;
        const int cls = lap ? cls1 : cls0;
        bool fast = (cls != kCtxHard) && (cls != kCtxSkip) && ctx_fast_ok(cls) &&
                    (ctx_of(Sin) == cls);
        if (fast) {
            double incD[IPT];
            {
                double w[IPT];
                weights_from_raw(win != nullptr, raw, base, n, m, w);
                fast = tile_quantize_f64(cls, w, incD);
            }
            if (fast)
                fast = staged_emit(comb, cls, Sin, lap ? so1 : so0, lap ? c10 : c00,
                                   lap ? c11 : c01, lap ? gs1 : gs0, incD, (int32_t)(g0 + base),
                                   A, out_offset, J, n_new, order, lap, sm.stage, &v.info->err);
        }
        if (!fast) {
            EmitCtx E;
            E.out = A; E.out_offset = out_offset; E.J = J; E.n_new = n_new; E.order = order;
            E.lap = lap;
            double w[IPT];
            load_weights(lw, win, base, n, m, w);      // rare path: reload instead of keeping w live
            __syncthreads();
            generic_emit(comb, Sin, w, nvalid, (int32_t)(g0 + base + i0), E, sm.fold);
        }
    }
    // ... and the tile that straddles x0 emits its second lap on the general path
    if (do0 && do1) {
        EmitCtx E;
        E.out = A; E.out_offset = out_offset; E.J = J; E.n_new = n_new; E.order = order;
        E.lap = 1;
        double w[IPT];
        load_weights(lw, win, base, n, m, w);
        __syncthreads();
        generic_emit(comb, si1, w, nvalid, (int32_t)(g0 + base + i0), E, sm.fold);
    }
}


// ---- streaming path of one tile (TileRec mode TM_FAST) ----------------------
// Integer increments by the 2^52 magic add (k_fold_agg has checked that no element
// ties or saturates), exact prefix in integer-valued doubles, pointer index per
// edge k(s) = ceil((s - X0) / q) in exact fp64, then the offspring are staged as
// "head" marks (element j writes j+1 at its first slot), a max-scan fills the runs
// and every thread stores its E consecutive window entries with 128-bit stores.
template <int E>
__device__ __forceinline__ void scan_store(StageSmem& st, int a0, int tot, int32_t* dstA,
                                           int32_t val_tile0) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int i = tid * E;
    int x[E];
#pragma unroll
    for (int g = 0; g < E / 4; ++g) {
        const int4 a = *reinterpret_cast<const int4*>(&st.buf[i + 4 * g]);
        x[4 * g] = a.x; x[4 * g + 1] = a.y; x[4 * g + 2] = a.z; x[4 * g + 3] = a.w;
    }
#pragma unroll
    for (int q = 1; q < E; ++q) x[q] = max(x[q], x[q - 1]);
    int inc = x[E - 1];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc = max(inc, o);
    }
    int ex = __shfl_up_sync(0xffffffffu, inc, 1);
    if (lane == 0) ex = 0;
    if (lane == 31) st.wmax[wid] = inc;
    __syncthreads();
    int wm = (lane < NWARP) ? st.wmax[lane] : 0;
#pragma unroll
    for (int d = 1; d < NWARP; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, wm, d);
        if (lane >= d) wm = max(wm, o);
    }
    const int pre_w = __shfl_sync(0xffffffffu, wm, (wid > 0) ? wid - 1 : 0);
    if (wid > 0) ex = max(ex, pre_w);
    const int add = val_tile0 - 1;
#pragma unroll
    for (int q = 0; q < E; ++q) x[q] = max(x[q], ex) + add;
#pragma unroll
    for (int g = 0; g < E / 4; ++g) {
        const int idx = i + 4 * g;
        if (idx >= a0 && idx + 4 <= tot) {
            *reinterpret_cast<int4*>(dstA + idx) = make_int4(x[4 * g], x[4 * g + 1], x[4 * g + 2], x[4 * g + 3]);
        } else {
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if (idx + c >= a0 && idx + c < tot) dstA[idx + c] = x[4 * g + c];
        }
    }
}

__device__ __forceinline__ void ancestors_fast(const TileRec& R, const double (&raw)[IPT], bool is_w,
                                               int64_t base, int64_t n, int32_t val_tile0,
                                               StageSmem& st, int* err) {
    const int tid = threadIdx.x;
    const int W = (int)(R.wm & 0x0fffffffu);
    // the window is shifted by a0 = (destination address / 4) mod 4 inside the staging
    // buffer so that 16-byte groups of the buffer line up with those of the output
    const int a0 = (int)((reinterpret_cast<uintptr_t>(R.dst) >> 2) & 3);
    const int tot = a0 + W;
    const int E = ((tot + TB * 4 - 1) / (TB * 4)) * 4;             // entries per thread: 4 .. 16
    for (int e = tid * 4; e < TB * E; e += TB * 4)
        *reinterpret_cast<int4*>(&st.buf[e]) = make_int4(0, 0, 0, 0);
    double inc[IPT], loc = 0.0;
    {
        double w[IPT];
        weights_from_raw(is_w, raw, base, n, R.m, w);
        const double inv_u = R.inv_u;
#pragma unroll
        for (int j = 0; j < IPT; ++j) {
            inc[j] = (w[j] * inv_u + 4503599627370496.0) - 4503599627370496.0;
            loc += inc[j];
        }
    }
    double tile_units;
    const double pre = cta_excl_scan_f64(loc, st.dsum, tile_units);     // two barriers: zeroing done
    if (tid == 0) {
        const double u = R.units;
        if ((u >= 0.0) ? (tile_units != u) : !(tile_units < -u)) atomicOr(err, EB_CTX | 0x1000);
    }
    const double qD = R.qD, invq = R.invq;
    const double magic = 6755399441055744.0;                           // 1.5 * 2^52
    const int k0 = R.k0 - a0;
    double sD = R.sD0 + pre;                                           // exact
    int kprev;
    {
        const double tm = sD * invq + magic;
        const double r = fma(-(tm - magic), qD, sD);
        kprev = __double2loint(tm) + (r > 0.0 ? 1 : 0);
    }
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
        sD += inc[j];
        const double tm = sD * invq + magic;
        const double r = fma(-(tm - magic), qD, sD);
        const int k = __double2loint(tm) + (r > 0.0 ? 1 : 0);
        if (kprev < k) st.buf[kprev - k0] = tid * IPT + j + 1;          // head of element j's run
        kprev = k;
    }
    __syncthreads();
    int32_t* dstA = R.dst - a0;                                        // 16-byte aligned
    if (E == 8) scan_store<8>(st, a0, tot, dstA, val_tile0);
    else if (E == 12) scan_store<12>(st, a0, tot, dstA, val_tile0);
    else if (E == 4) scan_store<4>(st, a0, tot, dstA, val_tile0);
    else scan_store<16>(st, a0, tot, dstA, val_tile0);
}

__global__ void __launch_bounds__(TB) k_ancestors_general(const double* __restrict__ lw,
                                                     const double* __restrict__ win, int64_t n,
                                                     int64_t g0, const __grid_constant__ WSView v,
                                                     int order, int32_t* __restrict__ anc,
                                                     int64_t out_offset,
                                                     const __grid_constant__ PeerI32 ancp,
                                                     int world) {
    __shared__ AncSmem sm;
    const int ng = v.ngen[0];
    for (int i = blockIdx.x; i < ng; i += gridDim.x) {
        ancestors_general(v.glist[i], lw, win, n, g0, v, order, anc, out_offset, ancp, world, sm);
        __syncthreads();
    }
}

__global__ void __launch_bounds__(TB, APOP_ANC_CTAS) k_ancestors(const double* __restrict__ lw,
                                                     const double* __restrict__ win, int64_t n,
                                                     int64_t g0, const TileRec* __restrict__ rec,
                                                     int* err) {
    __shared__ StageSmem sm;
    const int t = blockIdx.x;
    const int64_t base = (int64_t)t * TILE;
    // the tile and its record: all global reads of the streaming path, issued up front
    double raw[IPT];
    load_raw(lw, win, base, n, raw);
    TileRec R;
    {
        const int4* rp = reinterpret_cast<const int4*>(rec + t);
        int4* rq = reinterpret_cast<int4*>(&R);
#pragma unroll
        for (int k = 0; k < 4; ++k) rq[k] = __ldg(rp + k);
    }
    if ((R.wm >> 28) == TM_IDENT) {
        const int64_t i0 = base + (int64_t)threadIdx.x * IPT;
#pragma unroll
        for (int j = 0; j < IPT; ++j)
            if (i0 + j < n) R.dst[threadIdx.x * IPT + j] = (int32_t)(g0 + i0 + j);
        return;
    }
    if ((R.wm >> 28) != TM_FAST) return;           // nothing to emit, or listed for the general kernel
    ancestors_fast(R, raw, win != nullptr, base, n, (int32_t)(g0 + base), sm, err);
}

// Deferred emission (see RunRec): every warp takes run records in turn; the
// runs of >= BIG_RUN offspring are split in chunks over all CTAs.
constexpr int FILL_TB = 256;
constexpr int FILL_CHUNK = 8192;
__global__ void __launch_bounds__(FILL_TB) k_fill_runs(WSView v, int order, int32_t* __restrict__ anc,
                                                      int64_t out_offset, int64_t nl,
                                                      const __grid_constant__ PeerI32 ancp,
                                                      int world) {
    const ResampleInfo* I = v.info;
    int nr = v.nruns[0], nbig = v.nruns[1];
    if (nr == 0 && nbig == 0) return;
    if (I->err) return;
    if (nr > v.runs_cap) nr = (v.runs_cap / TILE) * TILE;      // blocks that did not fit fell back
    if (nbig > v.big_cap) nbig = v.big_cap;
    const i64 J = I->J, n_new = I->n_new;
    AncOut A;
    A.local = anc; A.peerp = &ancp; A.world = world; A.nl = nl;
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * FILL_TB + threadIdx.x) >> 5, nwarps = (gridDim.x * FILL_TB) >> 5;
    for (int r = warp; r < nr; r += nwarps) {
        const RunRec rec = v.runs[r];
        if (rec.count <= 0) continue;
        const int lap = (int)((unsigned)rec.val >> 31);
        const int32_t val = rec.val & 0x7fffffff;
        for (i64 k = rec.k0 + lane; k < rec.k0 + rec.count; k += 32)
            *A.at(out_offset + slot_of(k, lap, order, J, n_new)) = val;
    }
    for (int r = 0; r < nbig; ++r) {
        const RunRec rec = v.bigruns[r];
        const int lap = (int)((unsigned)rec.val >> 31);
        const int32_t val = rec.val & 0x7fffffff;
        const int chunks = (rec.count + FILL_CHUNK - 1) / FILL_CHUNK;
        for (int ch = blockIdx.x; ch < chunks; ch += gridDim.x) {
            const i64 kb = rec.k0 + (i64)ch * FILL_CHUNK;
            const i64 ke = (kb + FILL_CHUNK < rec.k0 + rec.count) ? kb + FILL_CHUNK : rec.k0 + rec.count;
            for (i64 k = kb + threadIdx.x; k < ke; k += FILL_TB)
                *A.at(out_offset + slot_of(k, lap, order, J, n_new)) = val;
        }
    }
}

// ------------------------------------------------------------------ K5
// x_new[k] = x[anc[k]] (the per-offspring copy of smc.cljc:127-129).  Ancestors come
// out of the resampler in runs (sorted order: monotone), so the reads are nearly
// sequential.  width 1: four offspring per thread, one 128-bit index load, 128-bit
// stores.  Rows (width > 1): 16-byte chunks, row index by shift when the chunk
// count per row is a power of two.  The result goes to the column's second buffer
// and the two are swapped (no copy-back).
template <typename T>
__global__ void __launch_bounds__(256) k_gather1(const T* __restrict__ src, T* __restrict__ dst,
                                                 const int32_t* __restrict__ anc, int64_t n_new) {
    const int64_t k0 = ((int64_t)blockIdx.x * 256 + threadIdx.x) * 4;
    if (k0 >= n_new) return;
    if (k0 + 4 <= n_new) {
        const int4 a = __ldg(reinterpret_cast<const int4*>(anc + k0));
        const T v0 = __ldg(src + a.x), v1 = __ldg(src + a.y), v2 = __ldg(src + a.z), v3 = __ldg(src + a.w);
        if (sizeof(T) == 8) {
            double2* q = reinterpret_cast<double2*>(dst + k0);
            q[0] = make_double2((double)v0, (double)v1);
            q[1] = make_double2((double)v2, (double)v3);
        } else {
            *reinterpret_cast<int4*>(dst + k0) = make_int4((int)v0, (int)v1, (int)v2, (int)v3);
        }
    } else {
        for (int64_t k = k0; k < n_new; ++k) dst[k] = src[anc[k]];
    }
}
// rows of `chunks` 16-byte words (row bytes = 16 * chunks); shift >= 0: chunks == 1 << shift
__global__ void __launch_bounds__(256) k_gather_rows16(const int4* __restrict__ src,
                                                       int4* __restrict__ dst,
                                                       const int32_t* __restrict__ anc,
                                                       int64_t n_new, int chunks, int shift) {
    const int64_t total = n_new * (int64_t)chunks;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * 256) {
        const int64_t k = (shift >= 0) ? (i >> shift) : (i / chunks);
        const int c = (int)(i - k * chunks);
        dst[i] = __ldg(src + (int64_t)__ldg(anc + k) * chunks + c);
    }
}
// any width / element size (rows that are not a multiple of 16 bytes)
template <typename T>
__global__ void __launch_bounds__(256) k_gather_any(const T* __restrict__ src, T* __restrict__ dst,
                                                    const int32_t* __restrict__ anc, int64_t n_new,
                                                    int width) {
    const int64_t total = n_new * (int64_t)width;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * 256) {
        const int64_t k = i / width;
        dst[i] = src[(int64_t)anc[k] * width + (i - k * width)];
    }
}

// ------------------------------------------------------------------ multinomial
// exact inclusive edges of the left fold (runtime.cljc:281-287 `acc`)
__global__ void __launch_bounds__(TB) k_edges(const double* __restrict__ lw,
                                              const double* __restrict__ win, int64_t n, WSView v,
                                              double* __restrict__ edges) {
    __shared__ FoldSmem sm;
    const int t = blockIdx.x;
    const int64_t base = (int64_t)t * TILE;
    const int nvalid = (int)((n - base < TILE) ? (n - base) : TILE);
    double w[IPT], edge[IPT];
    load_weights(lw, win, base, n, v.info->m, w);
    (void)cta_tile_fold(v.sin[0][t], w, nvalid, edge, sm);
    const int64_t i0 = base + (int64_t)threadIdx.x * IPT;
#pragma unroll
    for (int j = 0; j < IPT; ++j)
        if (i0 + j < n) edges[i0 + j] = edge[j];
}

// first i with x < edge[i]  (strict, runtime.cljc:285)
__global__ void k_multinomial_search(const double* __restrict__ edges, int64_t n,
                                     const double* __restrict__ uni, uint64_t seed, uint32_t step,
                                     int64_t n_new, const ResampleInfo* I,
                                     int32_t* __restrict__ anc, int64_t g0) {
    const double total = I->total;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n_new;
         k += (int64_t)gridDim.x * blockDim.x) {
        const double u = uni ? uni[k] : uniform_at(seed, (uint64_t)k, step, STREAM_RESAMPLE, 1);
        const double x = u * total;
        int64_t lo = 0, hi = n;
        while (lo < hi) {
            int64_t mid = (lo + hi) >> 1;
            if (x < edges[mid]) hi = mid; else lo = mid + 1;
        }
        anc[k] = (int32_t)(g0 + (lo < n ? lo : n - 1));
    }
}

__global__ void k_info_reset(ResampleInfo* I, WSView v) {
    v.gmax[0] = dkey(-INFINITY);
    v.ticket[0] = 0; v.ticket[1] = 0;
    I->err = 0;
    I->J = 0;
    I->total = 0;
    I->step = 0;
    I->new_lw = 0;
}

// ------------------------------------------------------------------ host side
void preload_resample_kernels() {
    cudaFuncAttributes a;
    cudaFuncGetAttributes(&a, k_tile_lse);
    cudaFuncGetAttributes(&a, k_fin_a);
    cudaFuncGetAttributes(&a, k_fin_b);
    cudaFuncGetAttributes(&a, k_fold_agg);
    cudaFuncGetAttributes(&a, k_seg_scan);
    cudaFuncGetAttributes(&a, k_chain);
    cudaFuncGetAttributes(&a, k_expand);
    cudaFuncGetAttributes(&a, k_ancestors);
    cudaFuncGetAttributes(&a, k_ancestors_general);
    cudaFuncGetAttributes(&a, k_fill_runs);
    cudaFuncGetAttributes(&a, k_info_reset);
}

int ensure_resample_ws(apop_ctx* c) {
    ResampleWS& w = c->ws;
    if (w.info) return APOP_OK;
    const int64_t nt = (c->n + TILE - 1) / TILE;
    w.nt = nt;
    w.nb = (int)((nt + SEGB - 1) / SEGB);
    w.hcap = (int)((nt < HCAP) ? nt : HCAP);
    const size_t nd = (size_t)(nt + 1), nbd = (size_t)(w.nb + 1);
    // one slab, carved in two passes (measure, then assign)
    char* base = nullptr;
    size_t off = 0;
    auto carve = [&](size_t bytes) -> void* {
        void* p = base ? base + off : nullptr;
        off += (bytes + 255) & ~(size_t)255;
        return p;
    };
    for (int pass = 0; pass < 2; ++pass) {
        off = 0;
        w.tmax = (double*)carve(nd * 8); w.tsum = (double*)carve(nd * 8);
        w.tsum2 = (double*)carve(nd * 8); w.ain = (double*)carve(nd * 8);
        w.tacc = (double*)carve(nd * 8); w.tacc2 = (double*)carve(nd * 8);
        w.rec = (TileRec*)carve(nd * sizeof(TileRec)); w.tflag = (int*)carve(nd * 4);
        w.glist = (int*)carve(nd * 4); w.ngen = (int*)carve(4);
        w.gmax = (long long*)carve(8); w.ticket = (int*)carve(8); w.hcnt = (int*)carve(8);
        w.skipmin = (int*)carve(4); w.fsc = (double*)carve(8 * 8);
        w.csum = (double*)carve((nd / 1024 + 2) * 8); w.coff = (double*)carve((nd / 1024 + 3) * 8);
        for (int l = 0; l < 2; ++l) { w.hst[l] = (int*)carve(4096 * 4); w.hea[l] = (int*)carve(4096 * 4); }
        for (int l = 0; l < 2; ++l) {
            w.cls[l] = (int*)carve(nd * 4); w.agg[l] = (Xd*)carve(nd * sizeof(Xd));
            w.ex[l] = (Xd*)carve(nd * sizeof(Xd)); w.hix[l] = (int*)carve(nd * 4);
            w.hp[l] = (int*)carve(nd * 4); w.hl[l] = (int*)carve(nd * 4);
            w.hctx[l] = (int*)carve(nd * 4); w.bc[l] = (Xd*)carve(nbd * sizeof(Xd));
            w.sblk[l] = (double*)carve(nbd * 8); w.shin[l] = (double*)carve(nd * 8);
            w.sin[l] = (double*)carve(nd * 8);
            w.cnt[l] = (long long*)carve(nd * 8); w.gseg[l] = (int*)carve(nd * 4);
            w.shout[l] = (double*)carve(nd * 8);
            w.hpa[l] = (Xd*)carve((size_t)w.hcap * TILE * sizeof(Xd));
            w.hsb[l] = (Xd*)carve((size_t)w.hcap * TILE * sizeof(Xd));
            w.hw[l] = (double*)carve((size_t)w.hcap * TILE * 8);
        }
        w.bacc = (double*)carve(nbd * 8); w.bacc2 = (double*)carve(nbd * 8);
        w.baccx = (double*)carve(nbd * 8);
        w.send = (double*)carve(2 * 8);
        w.trace = (double*)carve(2 * 256 * 8);
        w.comb = (Comb*)carve(sizeof(Comb));
        {
            // a (tile, lap) is heavy when it owns > HEAVY_W of the n_new pointers; n_new is
            // the global population in a sharded sweep.  A list that still overflows
            // (callers asking for n_new >> n) falls back to in-tile emission.
            const int64_t nmax = (c->n_global > c->n) ? c->n_global : c->n;
            if (pass == 0) {
                w.heavy_w = HEAVY_W;
                if (const char* e = getenv("APOP_HEAVY_W")) w.heavy_w = atoi(e) > 0 ? atoi(e) : HEAVY_W;
            }
            int64_t blocks = nmax / w.heavy_w + 2;
            if (blocks > nt + 2) blocks = nt + 2;
            w.runs_cap = (int)(blocks * TILE);
            w.big_cap = (int)(nmax / BIG_RUN + 8);
            w.runs = (RunRec*)carve((size_t)w.runs_cap * sizeof(RunRec));
            w.bigruns = (RunRec*)carve((size_t)w.big_cap * sizeof(RunRec));
            w.nruns = (int*)carve(2 * sizeof(int));
        }
        ResampleInfo* info = (ResampleInfo*)carve(sizeof(ResampleInfo));
        if (pass == 0) {
            if (cudaMalloc(&w.slab, off) != cudaSuccess) {
                w.slab = nullptr;
                return set_err(c, APOP_ERR_OOM, "resample workspace: %zu bytes", off);
            }
            base = (char*)w.slab;
        } else {
            w.info = info;
        }
    }
    APOP_CUDA(c, cudaMemsetAsync(w.info, 0, sizeof(ResampleInfo), c->stream));
    k_info_reset<<<1, 1, 0, c->stream>>>(w.info, view_of(w));      // max cell, tickets
    APOP_CUDA(c, cudaMallocHost(&w.info_host, sizeof(ResampleInfo)));
    return APOP_OK;
}

// Enqueue the resampling pipeline.  If a.have_partials the per-tile partials
// were already produced by a fused propagate kernel.
int resample_enqueue(apop_ctx* c, const ResampleArgs& a) {
    APOP_TRY(ensure_resample_ws(c));
    if (a.n != c->n) return set_err(c, APOP_ERR_INVALID, "resample: n mismatch");
    if (a.n_new <= 0) return set_err(c, APOP_ERR_INVALID, "resample: n_new must be > 0");
    WSView v = view_of(c->ws);
    const int nt = (int)v.nt;
    const int wm = a.w ? 1 : 0;
    if (!a.have_partials) {
        if (!a.keep_err) k_info_reset<<<1, 1, 0, c->stream>>>(v.info, v);
        kt_begin(c, KT_TILE_LSE);
        k_tile_lse<<<nt, TB, 0, c->stream>>>(a.lw, a.w, a.n, v);
        kt_end(c, KT_TILE_LSE);
        c->launches += a.keep_err ? 1 : 2;
    }
    Fabric F = fabric_view(c);
    if (!a.sharded) { F.world = 1; F.rank = 0; }
    const int64_t g0 = a.sharded ? c->g0 : 0;
    const int64_t n_glob = a.sharded ? c->n_global : a.n;
    const unsigned long long seq = a.sharded ? ++c->fab.epoch : 0ULL;
    kt_begin(c, KT_FINALIZE);
    const int fg = (int)((v.nt + FB - 1) / FB);
    k_fin_a<<<fg, FB, 0, c->stream>>>(v, a.n, a.n_new, a.u, c->seed, c->step, wm, a.rel_eps, F, seq,
                                      n_glob, a.seedp);
    k_fin_b<<<fg, FB, 0, c->stream>>>(v, a.n, a.rel_eps, g0, n_glob);
    kt_end(c, KT_FINALIZE);
    kt_begin(c, KT_FOLD_AGG);
    k_fold_agg<<<nt, TB, 0, c->stream>>>(a.lw, a.w, a.n, v, a.xfilt);
    kt_end(c, KT_FOLD_AGG);
    kt_begin(c, KT_RESOLVE);
    k_seg_scan<<<c->ws.nb, SEGB, 0, c->stream>>>(v, a.xfilt ? 1 : 0);
    k_chain<<<1, TB, 0, c->stream>>>(a.lw, a.w, a.n, n_glob, v, a.step, F, seq);
    k_expand<<<(int)((v.nt + 1 + XB - 2) / (XB - 1)), XB, 0, c->stream>>>(
        v, a.order, a.anc_out, a.out_offset, a.anc_peer, F.world, a.n);
    kt_end(c, KT_RESOLVE);
    c->launches += 6;
    if (a.anc_out) {
        kt_begin(c, KT_ANCESTORS);
        k_ancestors<<<nt, TB, 0, c->stream>>>(a.lw, a.w, a.n, g0, v.rec, &v.info->err);
        k_ancestors_general<<<c->sm_count, TB, 0, c->stream>>>(a.lw, a.w, a.n, g0, v, a.order,
                                                               a.anc_out, a.out_offset, a.anc_peer,
                                                               F.world);
        k_fill_runs<<<c->sm_count * 4, FILL_TB, 0, c->stream>>>(v, a.order, a.anc_out, a.out_offset, a.n,
                                                     a.anc_peer, F.world);
        kt_end(c, KT_ANCESTORS);
        c->launches += 3;
    }
    APOP_CUDA(c, cudaGetLastError());
    return APOP_OK;
}

int resample_fetch_info(apop_ctx* c, ResampleInfo* out) {
    APOP_CUDA(c, cudaMemcpyAsync(c->ws.info_host, c->ws.info, sizeof(ResampleInfo),
                                 cudaMemcpyDeviceToHost, c->stream));
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    *out = *c->ws.info_host;
    return APOP_OK;
}

int info_to_status(apop_ctx* c, const ResampleInfo& I) {
    if (I.err & EB_NAN)
        return set_err(c, APOP_ERR_NAN_WEIGHT,
                       "NaN (or negative) weight: the reference's resample would not terminate");
    if (I.err & EB_INF)
        return set_err(c, APOP_ERR_NAN_WEIGHT,
                       "+inf log-weight in a fused sweep (an observe with infinite density): the "
                       "reference would give those particles weight 1, all others 0 and -inf "
                       "afterwards (smc.cljc:78-82, 98-103); the fused sweep reports it instead "
                       "(the column-level apop_resample follows the reference)");
    if (I.err & EB_ZERO)
        return set_err(c, APOP_ERR_ZERO_WEIGHT, "total weight is not a positive finite number");
    if (I.err & EB_COMM)
        return set_err(c, APOP_ERR_NCCL, "a peer rank never arrived at a fabric exchange (spin timeout; "
                                         "wait %d on rank %d: 1 finalize, 2 sums, 3 fold lap 0, 4 fold "
                                         "lap 1, 5 barrier): results of this sweep are invalid",
                       (I.err >> 16) & 0xff, (I.err >> 24) & 0x7f);
    if (I.err & (EB_CTX | EB_COMB | EB_COUNT))
        return set_err(c, APOP_ERR_INTERNAL, "exact-fold consistency check failed (bits %d)",
                       I.err);
    return APOP_OK;
}

int multinomial_enqueue(apop_ctx* c, const double* lw, const double* w, const double* uni,
                        int64_t n_new, int32_t* anc) {
    APOP_TRY(ensure_resample_ws(c));
    if (!c->ws.edges) APOP_CUDA(c, cudaMalloc(&c->ws.edges, (size_t)c->n * 8));
    ResampleArgs a{};
    a.lw = lw; a.w = w; a.n = c->n; a.n_new = n_new; a.u = 0.0; a.order = APOP_ORDER_EMISSION;
    a.anc_out = nullptr; a.out_offset = 0; a.have_partials = false;
    APOP_TRY(resample_enqueue(c, a));
    WSView v = view_of(c->ws);
    k_edges<<<(int)v.nt, TB, 0, c->stream>>>(lw, w, c->n, v, c->ws.edges);
    int blocks = (int)((n_new + 255) / 256);
    if (blocks > c->sm_count * 16) blocks = c->sm_count * 16;
    k_multinomial_search<<<blocks, 256, 0, c->stream>>>(c->ws.edges, c->n, uni, c->seed, c->step,
                                                        n_new, v.info, anc, c->g0);
    c->launches += 2;
    APOP_CUDA(c, cudaGetLastError());
    return APOP_OK;
}

int gather_enqueue(apop_ctx* c, const int32_t* anc, int64_t n_new, Column* col) {
    const size_t esz = col->dtype == APOP_F64 ? 8 : 4;
    const size_t bytes = (size_t)(c->n + 1) * col->width * esz;
    if (!col->alt) {
        if (cudaMalloc(&col->alt, bytes) != cudaSuccess) {
            col->alt = nullptr;
            return set_err(c, APOP_ERR_OOM, "gather: second buffer of %zu bytes", bytes);
        }
    }
    const size_t row = (size_t)col->width * esz;
    kt_begin(c, KT_GATHER);
    if (col->width == 1) {
        const int blocks = (int)((n_new + 1023) / 1024);
        if (col->dtype == APOP_F64)
            k_gather1<double><<<blocks, 256, 0, c->stream>>>((const double*)col->ptr, (double*)col->alt,
                                                             anc, n_new);
        else
            k_gather1<int32_t><<<blocks, 256, 0, c->stream>>>((const int32_t*)col->ptr,
                                                              (int32_t*)col->alt, anc, n_new);
    } else if (row % 16 == 0) {
        const int chunks = (int)(row / 16);
        int shift = -1;
        if ((chunks & (chunks - 1)) == 0) { shift = 0; while ((1 << shift) < chunks) ++shift; }
        int64_t blocks = (n_new * chunks + 255) / 256;
        if (blocks > (int64_t)c->sm_count * 32) blocks = (int64_t)c->sm_count * 32;
        k_gather_rows16<<<(int)blocks, 256, 0, c->stream>>>((const int4*)col->ptr, (int4*)col->alt, anc,
                                                            n_new, chunks, shift);
    } else {
        int64_t blocks = (n_new * col->width + 255) / 256;
        if (blocks > (int64_t)c->sm_count * 32) blocks = (int64_t)c->sm_count * 32;
        if (col->dtype == APOP_F64)
            k_gather_any<double><<<(int)blocks, 256, 0, c->stream>>>(
                (const double*)col->ptr, (double*)col->alt, anc, n_new, col->width);
        else
            k_gather_any<int32_t><<<(int)blocks, 256, 0, c->stream>>>(
                (const int32_t*)col->ptr, (int32_t*)col->alt, anc, n_new, col->width);
    }
    kt_end(c, KT_GATHER);
    c->launches += 1;
    APOP_CUDA(c, cudaGetLastError());
    // rows beyond n_new (CSMC gathers N-1 and re-inserts the retained particle) keep
    // their old contents in the reference (they are overwritten by the caller): carry them
    if ((size_t)n_new < (size_t)c->n + 1)
        APOP_CUDA(c, cudaMemcpyAsync((char*)col->alt + (size_t)n_new * row,
                                     (char*)col->ptr + (size_t)n_new * row,
                                     ((size_t)c->n + 1 - (size_t)n_new) * row,
                                     cudaMemcpyDeviceToDevice, c->stream));
    void* t = col->ptr; col->ptr = col->alt; col->alt = t;
    return APOP_OK;
}

} // namespace apop
