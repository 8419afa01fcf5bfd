// resample.cu — systematic / multinomial resampling, bit-exact with the
// sequential fp64 wheel of anglican.smc/resample (reference
// src/anglican/smc.cljc:86-137) and particle-weights (smc.cljc:70-84).
//
// Pipeline (all on ctx->stream, no host round trip):
//   K2  k_tile_lse      per-tile (max, sum exp, sum exp^2) of the log-weights
//       k_finalize      global max m, approximate tile prefix, ESS, per-tile
//                       binade classification for both laps of the wheel
//   K3a k_fold_agg      per-tile parity-transducer aggregate of the exact fold
//       k_resolve       chains the aggregates -> exact running sum at every
//                       tile boundary, exact total, pointer-comb table, J
//   K3b/K4 k_ancestors  per-element exact edges, pointer counting, writes the
//                       ancestor index array
//   K5  k_gather        x_new[k] = x[a[k]]
// The arithmetic (why this reproduces a sequential fold exactly) is in
// exact_fold.cuh.
#include <limits.h>
#include <math.h>
#include <stddef.h>

#include "apop_internal.cuh"
#include "rng.cuh"

namespace apop {

struct WSView {
    int64_t nt;
    double *tmax, *tsum, *tsum2, *ain;
    int *cls0, *cls1, *hard0, *hard1;
    Xd *agg0, *agg1;
    double *sin0, *sin1;
    Comb* comb;
    ResampleInfo* info;
};

static WSView view_of(const ResampleWS& w) {
    WSView v;
    v.nt = w.nt; v.tmax = w.tmax; v.tsum = w.tsum; v.tsum2 = w.tsum2; v.ain = w.ain;
    v.cls0 = w.cls0; v.cls1 = w.cls1; v.hard0 = w.hard0; v.hard1 = w.hard1;
    v.agg0 = w.agg0; v.agg1 = w.agg1; v.sin0 = w.sin0; v.sin1 = w.sin1;
    v.comb = w.comb; v.info = w.info;
    return v;
}

struct FoldSmem {
    Xd warp_tot[NWARP];
    double bcast_S;
    int cross;
    double last_edge[TB];
};

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

// Exact sequential fold of one tile (IPT consecutive elements per thread) that
// starts from the exact running sum S_in.  Returns the exact sum after the
// tile; edge[j] receives the exact inclusive running sum after element j.
// Iterates once per binade crossing inside the tile (usually zero).
__device__ double cta_tile_fold(double S_in, const double (&w)[IPT], int nvalid,
                                double (&edge)[IPT], FoldSmem& sm) {
    if (nvalid <= 0) return S_in;
    int pos = 0;
    double S = S_in;
    const int i0 = threadIdx.x * IPT;
    for (;;) {
        const int e = ctx_of(S);
        const i64 s_in = units_of(S);
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
                quantize(w[j], e, f, up, tie);
                qf[j] = f;
                qfl[j] = up | (tie << 1) | 4;
                xd_push(t, f, up, tie);
            }
        }
        if (threadIdx.x == 0) sm.cross = INT_MAX;
        Xd ex = cta_excl_scan_xd(t, sm.warp_tot);      // has a __syncthreads
        i64 s = xd_apply(s_in, ex);
        int mycross = INT_MAX;
        i64 sprev = 0;
#pragma unroll
        for (int j = 0; j < IPT; ++j) {
            if ((qfl[j] & 4) && mycross == INT_MAX) {
                i64 s2 = step_units(s, qf[j], qfl[j] & 1, (qfl[j] >> 1) & 1);
                if (s2 >= kTwo53) { mycross = i0 + j; sprev = s; }
                else { s = s2; edge[j] = from_units(s, e); }
            }
        }
        if (mycross != INT_MAX) atomicMin(&sm.cross, mycross);
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

__device__ __forceinline__ void load_weights(const double* __restrict__ lw,
                                             const double* __restrict__ win, int64_t base,
                                             int64_t n, double m, double (&w)[IPT]) {
    if (win) {
        load_blocked(win, base, n, w, 0.0);
    } else {
        double v[IPT];
        load_blocked(lw, base, n, v, -INFINITY);
        const int64_t i0 = base + (int64_t)threadIdx.x * IPT;
#pragma unroll
        for (int j = 0; j < IPT; ++j) w[j] = (i0 + j < n) ? weight_from_lw(v[j], m) : 0.0;
    }
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
            v.tmax[blockIdx.x] = 0.0; v.tsum[blockIdx.x] = a; v.tsum2[blockIdx.x] = b;
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
        v.tmax[blockIdx.x] = mx; v.tsum[blockIdx.x] = a; v.tsum2[blockIdx.x] = b;
    }
}

// ------------------------------------------------------------------ finalize
constexpr int FB = 1024;

__device__ __forceinline__ double block_excl_scan_f64(double v, double* sm /*[33]*/,
                                                      double& total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    double inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        double o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) sm[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        double ws = (lane < (int)(blockDim.x >> 5)) ? sm[lane] : 0.0;
        double winc = ws;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            double o = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += o;
        }
        sm[lane] = winc - ws;                       // exclusive over warps
        if (lane == 31) sm[32] = winc;
    }
    __syncthreads();
    double r = sm[wid] + (inc - v);
    total = sm[32];
    __syncthreads();
    return r;
}

__device__ __forceinline__ int block_excl_scan_i32(int v, int* sm /*[33]*/, int& total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) sm[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int ws = (lane < (int)(blockDim.x >> 5)) ? sm[lane] : 0;
        int winc = ws;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int o = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += o;
        }
        sm[lane] = winc - ws;
        if (lane == 31) sm[32] = winc;
    }
    __syncthreads();
    int r = sm[wid] + (inc - v);
    total = sm[32];
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(FB) k_finalize(WSView v, int64_t n, int64_t n_new, double u_in,
                                                 uint64_t seed, uint32_t step, int weights_mode) {
    __shared__ double smd[33];
    __shared__ int smi[33];
    __shared__ double sh_m;
    __shared__ int sh_skip;
    const int64_t nt = v.nt;
    const int tid = threadIdx.x;
    const int64_t per = (nt + FB - 1) / FB;
    const int64_t t0 = (int64_t)tid * per;
    const int64_t t1 = (t0 + per < nt) ? t0 + per : nt;

    // ---- global max (smc.cljc:75)
    double m = 0.0;
    if (!weights_mode) {
        double mx = -INFINITY;
        for (int64_t t = t0; t < t1; ++t) mx = fmax(mx, v.tmax[t]);
        mx = warp_max(mx);
        if ((tid & 31) == 0) smd[tid >> 5] = mx;
        __syncthreads();
        if (tid == 0) {
            double a = smd[0];
            for (int i = 1; i < FB / 32; ++i) a = fmax(a, smd[i]);
            sh_m = a;
        }
        __syncthreads();
        m = sh_m;
        __syncthreads();
    }
    auto asum = [&](int64_t t) -> double {
        if (weights_mode) return v.tsum[t];
        double tm = v.tmax[t];
        double fac = (tm == m) ? 1.0 : exp(tm - m);
        return v.tsum[t] * fac;
    };
    auto asq = [&](int64_t t) -> double {
        if (weights_mode) return v.tsum2[t];
        double tm = v.tmax[t];
        double fac = (tm == m) ? 1.0 : exp(2.0 * (tm - m));
        return v.tsum2[t] * fac;
    };
    // ---- approximate exclusive prefix of tile sums
    double loc = 0.0, loc2 = 0.0;
    for (int64_t t = t0; t < t1; ++t) { loc += asum(t); loc2 += asq(t); }
    double atot, asqtot;
    double run = block_excl_scan_f64(loc, smd, atot);
    (void)block_excl_scan_f64(loc2, smd, asqtot);
    for (int64_t t = t0; t < t1; ++t) { v.ain[t] = run; run += asum(t); }
    if (t1 == nt && t0 < nt) v.ain[nt] = run;
    if (nt == 0 && tid == 0) v.ain[0] = 0.0;
    __syncthreads();

    const double u = (u_in != u_in) ? uniform_at(seed, 0, step, STREAM_RESAMPLE, 0) : u_in;
    // x_last < x_0 + total <= total*(u+1): generous upper bound for the skip test
    const double eps_tot = ((double)(2 * n) + 8192.0) * 4.44089209850062616e-16;
    const double xlast_ub = atot * (u + 1.0) * (1.0 + 4.0 * eps_tot);

    // ---- classification of every tile for both laps
    if (tid == 0) sh_skip = (nt < INT_MAX) ? (int)nt : INT_MAX;
    __syncthreads();
    int nh0 = 0, nh1 = 0;
    for (int64_t t = t0; t < t1; ++t) {
        const int64_t hi = ((t + 1) * TILE < n) ? (t + 1) * TILE : n;
        const double a_in = v.ain[t];
        const double ao = a_in + asum(t);
        int c0 = classify_tile(a_in, ao, hi);
        v.cls0[t] = c0;
        nh0 += (c0 == kCtxHard);
        int c1;
        const double eps1 = ((double)(n + hi) + 4096.0) * 4.44089209850062616e-16;
        if ((atot + a_in) * (1.0 - eps1) > xlast_ub) {
            c1 = kCtxSkip;
            atomicMin(&sh_skip, (int)t);
        } else {
            c1 = classify_tile(atot + a_in, atot + ao, n + hi);
        }
        v.cls1[t] = c1;
        nh1 += (c1 == kCtxHard);
    }
    __syncthreads();
    int tot0, tot1;
    int o0 = block_excl_scan_i32(nh0, smi, tot0);
    int o1 = block_excl_scan_i32(nh1, smi, tot1);
    for (int64_t t = t0; t < t1; ++t) {
        if (v.cls0[t] == kCtxHard) v.hard0[o0++] = (int)t;
        if (v.cls1[t] == kCtxHard) v.hard1[o1++] = (int)t;
    }
    if (tid == 0) {
        ResampleInfo* I = v.info;
        I->m = m;
        I->atot = atot;
        I->asq = asqtot;
        I->ess = (asqtot > 0.0) ? atot * atot / asqtot : 0.0;
        I->log_mean_approx = (m > -INFINITY && m < INFINITY) ? log(atot / (double)n) + m
                                                               : -INFINITY;
        I->u = u;
        I->n_new = n_new;
        I->nh0 = tot0;
        I->nh1 = tot1;
        I->first_skip1 = sh_skip;
        I->weights_mode = weights_mode;
        if (!weights_mode && m != m) I->err |= EB_NAN;
    }
}

// ------------------------------------------------------------------ K3a
__global__ void __launch_bounds__(TB) k_fold_agg(const double* __restrict__ lw,
                                                 const double* __restrict__ win, int64_t n,
                                                 WSView v) {
    __shared__ Xd warp_tot[NWARP];
    const int t = blockIdx.x;
    const int c0 = v.cls0[t], c1 = v.cls1[t];
    const bool do0 = (c0 != kCtxHard), do1 = (c1 != kCtxHard && c1 != kCtxSkip);
    if (!do0 && !do1) return;
    const int64_t base = (int64_t)t * TILE;
    double w[IPT];
    load_weights(lw, win, base, n, v.info->m, w);
    Xd a = xd_identity(), b = xd_identity();
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
        i64 f; int up, tie;
        if (do0) { quantize(w[j], c0, f, up, tie); xd_push(a, f, up, tie); }
        if (do1) { quantize(w[j], c1, f, up, tie); xd_push(b, f, up, tie); }
    }
    // (out-of-range elements have w = 0: identity)
    if (do0) {
        (void)cta_excl_scan_xd(a, warp_tot);
        if (threadIdx.x == 0) v.agg0[t] = cta_total_xd(warp_tot);
        __syncthreads();
    }
    if (do1) {
        (void)cta_excl_scan_xd(b, warp_tot);
        if (threadIdx.x == 0) v.agg1[t] = cta_total_xd(warp_tot);
    }
}

// ------------------------------------------------------------------ resolve
constexpr int RI = 4;   // tiles per thread per pass when scanning easy runs

// scans the aggregates of the easy tiles [r0, r1) starting from exact S;
// writes the exact incoming sum of each tile and returns the outgoing sum.
__device__ double resolve_run(int64_t r0, int64_t r1, double S, const Xd* agg, const int* cls,
                              double* sin, FoldSmem& sm, int* err) {
    for (int64_t base = r0; base < r1; base += (int64_t)TB * RI) {
        const int e = ctx_of(S);
        const i64 s_in = units_of(S);
        Xd a[RI];
        Xd loc = xd_identity();
        int bad = 0;
#pragma unroll
        for (int k = 0; k < RI; ++k) {
            const int64_t t = base + (int64_t)threadIdx.x * RI + k;
            a[k] = xd_identity();
            if (t < r1) {
                a[k] = agg[t];
                bad |= (cls[t] != e);
                loc = xd_compose(loc, a[k]);
            }
        }
        Xd ex = cta_excl_scan_xd(loc, sm.warp_tot);
        i64 s = xd_apply(s_in, ex);
#pragma unroll
        for (int k = 0; k < RI; ++k) {
            const int64_t t = base + (int64_t)threadIdx.x * RI + k;
            if (t < r1) {
                bad |= (s >= kTwo53);
                sin[t] = from_units(s, e);
                s = xd_apply(s, a[k]);
            }
        }
        if (bad) atomicOr(err, EB_CTX);
        if (threadIdx.x == 0) {
            i64 st = xd_apply(s_in, cta_total_xd(sm.warp_tot));
            if (st >= kTwo53) atomicOr(err, EB_CTX);
            sm.bcast_S = from_units(st, e);
        }
        __syncthreads();
        S = sm.bcast_S;
        __syncthreads();
    }
    return S;
}

__global__ void __launch_bounds__(TB) k_resolve(const double* __restrict__ lw,
                                                const double* __restrict__ win, int64_t n,
                                                WSView v) {
    __shared__ FoldSmem sm;
    const int64_t nt = v.nt;
    ResampleInfo* I = v.info;
    const double m = I->m;
    double w[IPT], edge[IPT];

    // ---------------- lap 0: exact total (smc.cljc:94)
    double S = 0.0;
    {
        const int nh = I->nh0;
        int64_t r0 = 0;
        for (int h = 0; h <= nh; ++h) {
            const int64_t hend = (h < nh) ? (int64_t)v.hard0[h] : nt;
            if (hend > r0) S = resolve_run(r0, hend, S, v.agg0, v.cls0, v.sin0, sm, &I->err);
            if (h < nh) {
                if (threadIdx.x == 0) v.sin0[hend] = S;
                const int64_t base = hend * TILE;
                const int nvalid = (int)((n - base < TILE) ? (n - base) : TILE);
                load_weights(lw, win, base, n, m, w);
                S = cta_tile_fold(S, w, nvalid, edge, sm);
                __syncthreads();
            }
            r0 = hend + 1;
        }
    }
    const double total = S;
    if (threadIdx.x == 0) {
        v.sin0[nt] = total;
        const i64 n_new = I->n_new;
        int err = 0;
        if (!(total > 0.0) || !(total < INFINITY)) err |= EB_ZERO;
        Comb& c = *v.comb;
        if (!err) {
            comb_build(c, total, I->u, n_new);
            if (c.err) err |= EB_COMB;
        } else {
            c.nseg = 0; c.n_new = n_new; c.x0 = 0; c.xlast = 0; c.step = 0; c.err = 1;
        }
        int hint = 0;
        I->total = total;
        I->step = total / (double)n_new;
        I->J = err ? 0 : comb_count_below(c, total, hint);
        I->new_lw = (m > -INFINITY && m < INFINITY) ? (log(I->step) + m) : -INFINITY;
        if (err) atomicOr(&I->err, err);
    }
    __syncthreads();
    if (I->err & (EB_ZERO | EB_COMB | EB_NAN)) {
        for (int64_t t = threadIdx.x; t <= nt; t += TB) v.sin1[t] = INFINITY;
        return;
    }
    const double xlast = v.comb->xlast;

    // ---------------- lap 1: the wheel wraps with acc NOT reset (smc.cljc:133-136)
    // After the loop S is the exact running sum entering tile `stop`; tiles at or
    // beyond `stop` cannot receive a pointer of the second lap.
    S = total;
    int64_t stop = I->first_skip1;
    if (stop > nt) stop = nt;
    {
        const int nh = I->nh1;
        int h = 0;
        int64_t r0 = 0;
        while (r0 < stop) {
            int64_t hend = (h < nh) ? (int64_t)v.hard1[h] : stop;
            if (hend > stop) hend = stop;
            if (hend > r0) S = resolve_run(r0, hend, S, v.agg1, v.cls1, v.sin1, sm, &I->err);
            if (hend < stop) {
                if (S > xlast) { stop = hend; break; }
                if (threadIdx.x == 0) v.sin1[hend] = S;
                const int64_t base = hend * TILE;
                const int nvalid = (int)((n - base < TILE) ? (n - base) : TILE);
                load_weights(lw, win, base, n, m, w);
                S = cta_tile_fold(S, w, nvalid, edge, sm);
                __syncthreads();
                ++h;
            }
            r0 = hend + 1;
        }
    }
    if (stop < nt && !(S > xlast) && threadIdx.x == 0) atomicOr(&I->err, EB_CTX);
    for (int64_t t = stop + threadIdx.x; t <= nt; t += TB) v.sin1[t] = (t == stop) ? S : INFINITY;
}

// ------------------------------------------------------------------ K3b/K4
__device__ __forceinline__ int64_t slot_of(i64 k, int lap, int order, i64 J, i64 n_new) {
    if (order == APOP_ORDER_EMISSION) return k;
    if (order == APOP_ORDER_REFERENCE) return n_new - 1 - k;      // conj onto nil: smc.cljc:127
    return lap == 0 ? k + (n_new - J) : k - J;                    // sorted physical order
}

__global__ void __launch_bounds__(TB) k_ancestors(const double* __restrict__ lw,
                                                  const double* __restrict__ win, int64_t n,
                                                  int64_t g0, WSView v, int order,
                                                  int32_t* __restrict__ anc, int64_t out_offset) {
    __shared__ FoldSmem sm;
    __shared__ Comb comb;
    const ResampleInfo* I = v.info;
    if (I->err) return;
    const int t = blockIdx.x;
    const double x0 = v.comb->x0, xlast = v.comb->xlast;
    const double si0 = v.sin0[t], so0 = v.sin0[t + 1];
    const double si1 = v.sin1[t];
    const bool need0 = (so0 > x0) && !(si0 > xlast);
    const bool need1 = !(si1 > xlast);
    if (!need0 && !need1) return;
    // copy the comb table (header + used segments)
    {
        const Comb* g = v.comb;
        const int nseg = g->nseg;
        const int hdr_words = (int)(offsetof(Comb, seg) / 8);
        const int words = hdr_words + nseg * (int)(sizeof(CombSeg) / 8);
        const uint64_t* src = reinterpret_cast<const uint64_t*>(g);
        uint64_t* dst = reinterpret_cast<uint64_t*>(&comb);
        for (int i = threadIdx.x; i < words; i += TB) dst[i] = src[i];
    }
    const int64_t base = (int64_t)t * TILE;
    const int nvalid = (int)((n - base < TILE) ? (n - base) : TILE);
    double w[IPT], edge[IPT];
    load_weights(lw, win, base, n, I->m, w);
    __syncthreads();
    const i64 J = I->J, n_new = I->n_new;
    const int lane = threadIdx.x & 31;
    const int i0 = threadIdx.x * IPT;
    for (int lap = 0; lap < 2; ++lap) {
        if (lap == 0 ? !need0 : !need1) continue;
        const double Sin = lap ? si1 : si0;
        (void)cta_tile_fold(Sin, w, nvalid, edge, sm);
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
            const i64 cnt = c - c_prev;
            const int32_t val = (int32_t)(g0 + base + i0 + j);
            const bool heavy = cnt > 8;
            if (!heavy) {
                for (i64 k = c_prev; k < c; ++k)
                    anc[out_offset + slot_of(k, lap, order, J, n_new)] = val;
            }
            unsigned hm = __ballot_sync(0xffffffffu, heavy);
            while (hm) {
                const int src = __ffs(hm) - 1;
                hm &= hm - 1;
                const i64 hs = shfl_i64(c_prev, src);
                const i64 he = shfl_i64(c, src);
                const int32_t hv = __shfl_sync(0xffffffffu, val, src);
                for (i64 k = hs + lane; k < he; k += 32)
                    anc[out_offset + slot_of(k, lap, order, J, n_new)] = hv;
            }
            c_prev = c;
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------ K5
__global__ void k_gather_f64(const double* __restrict__ src, double* __restrict__ dst,
                             const int32_t* __restrict__ anc, int64_t n_new, int width) {
    const int64_t total = n_new * (int64_t)width;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t k = i / width;
        const int c = (int)(i - k * width);
        dst[i] = src[(int64_t)anc[k] * width + c];
    }
}
__global__ void k_gather_i32(const int32_t* __restrict__ src, int32_t* __restrict__ dst,
                             const int32_t* __restrict__ anc, int64_t n_new, int width) {
    const int64_t total = n_new * (int64_t)width;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t k = i / width;
        const int c = (int)(i - k * width);
        dst[i] = src[(int64_t)anc[k] * width + c];
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
    (void)cta_tile_fold(v.sin0[t], w, nvalid, edge, sm);
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

__global__ void k_info_reset(ResampleInfo* I) {
    I->err = 0;
    I->J = 0;
    I->total = 0;
    I->step = 0;
    I->new_lw = 0;
}

// ------------------------------------------------------------------ host side
int ensure_resample_ws(apop_ctx* c) {
    ResampleWS& w = c->ws;
    if (w.info) return APOP_OK;
    const int64_t nt = (c->n + TILE - 1) / TILE;
    w.nt = nt;
    const size_t nd = (size_t)(nt + 1);
    APOP_CUDA(c, cudaMalloc(&w.tmax, nd * 8));
    APOP_CUDA(c, cudaMalloc(&w.tsum, nd * 8));
    APOP_CUDA(c, cudaMalloc(&w.tsum2, nd * 8));
    APOP_CUDA(c, cudaMalloc(&w.ain, nd * 8));
    APOP_CUDA(c, cudaMalloc(&w.cls0, nd * 4));
    APOP_CUDA(c, cudaMalloc(&w.cls1, nd * 4));
    APOP_CUDA(c, cudaMalloc(&w.hard0, nd * 4));
    APOP_CUDA(c, cudaMalloc(&w.hard1, nd * 4));
    APOP_CUDA(c, cudaMalloc(&w.agg0, nd * sizeof(Xd)));
    APOP_CUDA(c, cudaMalloc(&w.agg1, nd * sizeof(Xd)));
    APOP_CUDA(c, cudaMalloc(&w.sin0, nd * 8));
    APOP_CUDA(c, cudaMalloc(&w.sin1, nd * 8));
    APOP_CUDA(c, cudaMalloc(&w.comb, sizeof(Comb)));
    APOP_CUDA(c, cudaMalloc(&w.info, sizeof(ResampleInfo)));
    APOP_CUDA(c, cudaMemsetAsync(w.info, 0, sizeof(ResampleInfo), c->stream));
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
        k_info_reset<<<1, 1, 0, c->stream>>>(v.info);
        kt_begin(c, KT_TILE_LSE);
        k_tile_lse<<<nt, TB, 0, c->stream>>>(a.lw, a.w, a.n, v);
        kt_end(c, KT_TILE_LSE);
        c->launches += 2;
    }
    kt_begin(c, KT_FINALIZE);
    k_finalize<<<1, FB, 0, c->stream>>>(v, a.n, a.n_new, a.u, c->seed, c->step, wm);
    kt_end(c, KT_FINALIZE);
    kt_begin(c, KT_FOLD_AGG);
    k_fold_agg<<<nt, TB, 0, c->stream>>>(a.lw, a.w, a.n, v);
    kt_end(c, KT_FOLD_AGG);
    kt_begin(c, KT_RESOLVE);
    k_resolve<<<1, TB, 0, c->stream>>>(a.lw, a.w, a.n, v);
    kt_end(c, KT_RESOLVE);
    c->launches += 3;
    if (a.anc_out) {
        kt_begin(c, KT_ANCESTORS);
        k_ancestors<<<nt, TB, 0, c->stream>>>(a.lw, a.w, a.n, c->g0, v, a.order, a.anc_out,
                                              a.out_offset);
        kt_end(c, KT_ANCESTORS);
        c->launches += 1;
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
    if (I.err & EB_ZERO)
        return set_err(c, APOP_ERR_ZERO_WEIGHT, "total weight is not a positive finite number");
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
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_multinomial_search<<<blocks, 256, 0, c->stream>>>(c->ws.edges, c->n, uni, c->seed, c->step,
                                                        n_new, v.info, anc, c->g0);
    c->launches += 2;
    APOP_CUDA(c, cudaGetLastError());
    return APOP_OK;
}

int gather_enqueue(apop_ctx* c, const int32_t* anc, int64_t n_new, Column* col) {
    const size_t esz = col->dtype == APOP_F64 ? 8 : 4;
    const size_t bytes = (size_t)c->n * col->width * esz;
    APOP_TRY(ensure_scratch(c, bytes));
    const int64_t total = n_new * col->width;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 148 * 32) blocks = 148 * 32;
    kt_begin(c, KT_GATHER);
    if (col->dtype == APOP_F64)
        k_gather_f64<<<blocks, 256, 0, c->stream>>>((const double*)col->ptr, (double*)c->scratch,
                                                    anc, n_new, col->width);
    else
        k_gather_i32<<<blocks, 256, 0, c->stream>>>((const int32_t*)col->ptr,
                                                    (int32_t*)c->scratch, anc, n_new, col->width);
    kt_end(c, KT_GATHER);
    c->launches += 1;
    APOP_CUDA(c, cudaGetLastError());
    APOP_CUDA(c, cudaMemcpyAsync(col->ptr, c->scratch, (size_t)n_new * col->width * esz,
                                 cudaMemcpyDeviceToDevice, c->stream));
    return APOP_OK;
}

} // namespace apop
