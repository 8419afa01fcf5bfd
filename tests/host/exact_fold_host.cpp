// CPU harness for anglican_b200/csrc/exact_fold.cuh.
//
// Mirrors, sequentially, the structure of the CUDA resampler (classify tiles
// from an approximate prefix -> per-tile parity-transducer aggregates -> chain
// resolve -> per-element replay -> pointer counting) so that the arithmetic of
// the header can be checked on a machine without a GPU against the oracle's
// literal wheel (oracle/anglican_oracle.c: orc_resample_weights).
//
// Built by tests/test_exact_fold_host.py with g++; not part of the product.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../anglican_b200/csrc/exact_fold.cuh"

using namespace apop;

namespace {

const int LANE_ITEMS = 8;

// exact fold of one tile starting from S_in, emulating the device algorithm:
// lanes of LANE_ITEMS elements, lane aggregates, exclusive compose-scan, replay.
// Iterates on binade crossings.  Writes S_i (inclusive edges); returns S_out.
double tile_exact_fold(double S_in, const double* w, int n, double* edges, int* iters) {
    int pos = 0;
    double S = S_in;
    int it = 0;
    std::vector<Xd> agg((n + LANE_ITEMS - 1) / LANE_ITEMS);
    while (pos < n) {
        ++it;
        const int e = ctx_of(S);
        const i64 s_in = units_of(S);
        const int nl = (n + LANE_ITEMS - 1) / LANE_ITEMS;
        for (int l = 0; l < nl; ++l) {
            Xd t = xd_identity();
            for (int j = 0; j < LANE_ITEMS; ++j) {
                int i = l * LANE_ITEMS + j;
                if (i >= n || i < pos) continue;
                i64 f; int up, tie;
                quantize(w[i], e, f, up, tie);
                xd_push(t, f, up, tie);
            }
            agg[l] = t;
        }
        // exclusive scan + replay
        Xd run = xd_identity();
        int cross = -1;
        for (int l = 0; l < nl && cross < 0; ++l) {
            i64 s = xd_apply(s_in, run);
            for (int j = 0; j < LANE_ITEMS; ++j) {
                int i = l * LANE_ITEMS + j;
                if (i >= n || i < pos) continue;
                i64 f; int up, tie;
                quantize(w[i], e, f, up, tie);
                i64 s2 = step_units(s, f, up, tie);
                if (s2 >= kTwo53) { cross = i; break; }
                s = s2;
                edges[i] = from_units(s, e);
            }
            run = xd_compose(run, agg[l]);
        }
        if (cross < 0) { S = (n > pos) ? edges[n - 1] : S; pos = n; break; }
        double Sprev = (cross == pos) ? S : edges[cross - 1];
        S = Sprev + w[cross];                 // the crossing add, hardware fp64
        edges[cross] = S;
        pos = cross + 1;
    }
    if (iters) *iters += it;
    return S;
}

// The FAST iteration of cta_tile_fold (csrc/resample.cu): increments as integer-valued
// doubles (2^52 magic add), prefix sums by ADDITIONS ONLY in the device's tree - 8 elements
// per lane serially, Hillis-Steele over the 32 lanes of a warp, warp totals serially - and
// the crossing decided by "prefix >= 2^53 - s_in".  A prefix below 2^53 is exact, one at or
// above it can only round to >= 2^53, so the decision is exact although the sums are not.
// An element of 2^52 ulps or more counts as 2^53 (it crosses wherever it stands).  Ties add
// floor(w / ulp); those in front of the uncorrected crossing are listed (in an arbitrary
// order, as the device's atomics would) with the parity of their uncorrected sum, ranked by
// element, walked once for the running number of round-ups, and each lane starts from the
// corrections made before its first element.  Returns S_out, or NaN when more than
// `tie_cap` ties are listed (the device then runs the transducer iteration, emulated by
// tile_exact_fold above).  *ties (optional) counts the listed ties over all iterations.
const int TIE_CAP = 32;
double tile_exact_fold_fast(double S_in, const double* w, int n, double* edges, int* iters,
                            int* ties = nullptr) {
    const int LANES = 32;
    int pos = 0;
    double S = S_in;
    const int nl = (n + LANE_ITEMS - 1) / LANE_ITEMS;
    std::vector<double> loc(nl), inc(nl), pre(nl);
    std::vector<double> incD(nl * LANE_ITEMS);
    std::vector<unsigned> tmask(nl);
    while (pos < n) {
        if (iters) *iters += 1;
        const int e = ctx_of(S);
        if (!ctx_fast_ok(e)) return NAN;
        const i64 s_in = units_of(S);
        const double inv_u = pow2(-e), ulp = pow2(e);
        for (int l = 0; l < nl; ++l) {
            double a = 0.0;
            tmask[l] = 0;
            for (int j = 0; j < LANE_ITEMS; ++j) {
                const int i = l * LANE_ITEMS + j;
                double rn = 0.0;
                if (i >= pos && i < n) {
                    const double scaled = w[i] * inv_u;
                    if (!(scaled < 4503599627370496.0)) rn = 9007199254740992.0;
                    else {
                        rn = (scaled + 4503599627370496.0) - 4503599627370496.0;
                        const double diff = scaled - rn;
                        if (diff == 0.5) tmask[l] |= 1u << j;
                        else if (diff == -0.5) { tmask[l] |= 1u << j; rn -= 1.0; }
                    }
                }
                incD[i] = rn;
                a += rn;
            }
            loc[l] = a;
        }
        // inclusive Hillis-Steele scan inside each warp, then serial warp prefixes
        inc = loc;
        for (int d = 1; d < LANES; d <<= 1) {
            std::vector<double> nxt = inc;
            for (int l = 0; l < nl; ++l)
                if ((l % LANES) >= d) nxt[l] = inc[l] + inc[l - d];
            inc = nxt;
        }
        const int nw = (nl + LANES - 1) / LANES;
        std::vector<double> wtot(nw, 0.0);
        for (int wq = 0; wq < nw; ++wq) {
            int last = wq * LANES + LANES - 1;
            if (last >= nl) last = nl - 1;
            wtot[wq] = inc[last];
        }
        for (int l = 0; l < nl; ++l) {
            double p = 0.0;
            for (int wq = 0; wq < l / LANES; ++wq) p += wtot[wq];
            p += (l % LANES) ? inc[l - 1] : 0.0;
            pre[l] = p;
        }
        const double L = (double)(kTwo53 - s_in), sD0 = (double)s_in;
        // tie list: lanes visited from the back so that the list is NOT in element order
        std::vector<int> tie_idx, tie_par;
        for (int l = nl - 1; l >= 0; --l) {
            if (!tmask[l]) continue;
            double b = pre[l];
            for (int j = 0; j < LANE_ITEMS; ++j) {
                b += incD[l * LANE_ITEMS + j];
                if (((tmask[l] >> j) & 1u) && b < L) {
                    tie_idx.push_back(l * LANE_ITEMS + j);
                    tie_par.push_back((int)((s_in + (i64)b) & 1));
                }
            }
        }
        const int K = (int)tie_idx.size();
        if (ties) *ties += K;
        if (K > TIE_CAP) return NAN;
        std::vector<int> ord(K), sidx(K), cum(K);
        for (int q = 0; q < K; ++q) {
            int r = 0;
            for (int k = 0; k < K; ++k) r += tie_idx[k] < tie_idx[q];
            ord[r] = q;
        }
        for (int r = 0, C = 0; r < K; ++r) {
            const int q = ord[r];
            C += (tie_par[q] + C) & 1;
            sidx[r] = tie_idx[q];
            cum[r] = C;
        }
        int cross = -1;
        i64 sprev = 0;
        for (int l = 0; l < nl; ++l) {
            int corr = 0;
            for (int r = 0; r < K; ++r)
                if (sidx[r] < l * LANE_ITEMS) corr = cum[r];
            double sD = pre[l] + (double)corr;
            int mycross = -1;
            i64 mysprev = 0;
            for (int j = 0; j < LANE_ITEMS; ++j) {
                const int i = l * LANE_ITEMS + j;
                if (i < pos || i >= n || mycross >= 0) continue;
                double d = incD[i];
                if ((tmask[l] >> j) & 1u) d += (double)((s_in + (i64)(sD + d)) & 1);
                const double nx = sD + d;
                if (nx >= L) { mycross = i; mysprev = s_in + (i64)sD; }
                else { sD = nx; edges[i] = (sD0 + sD) * ulp; }
            }
            // the block-wide minimum over the lanes' first crossings (atomicMin on the device)
            if (mycross >= 0 && (cross < 0 || mycross < cross)) { cross = mycross; sprev = mysprev; }
        }
        if (cross < 0) { S = edges[n - 1]; pos = n; break; }
        S = from_units(sprev, e) + w[cross];          // the crossing add, hardware fp64
        edges[cross] = S;
        pos = cross + 1;
    }
    return S;
}

} // namespace

extern "C" {

// Full emulation.  Returns 0 on success.  emission[j] = particle of pointer j,
// lap[j] = 0/1.  stats[0]=#hard tiles lap0, stats[1]=#hard tiles lap1,
// stats[2]=crossing iterations, stats[3]=comb segments.
int host_exact_resample(const double* w, int64_t n, int64_t n_new, double u, int tile,
                        int32_t* emission, int32_t* lap_out, double* total_out,
                        int64_t* stats) {
    const int64_t nt = (n + tile - 1) / tile;
    // approximate prefix (pairwise-ish: per tile plain sums, then scan)
    std::vector<double> tsum(nt), ain(nt + 1);
    for (int64_t t = 0; t < nt; ++t) {
        // deliberately a different summation order than the fold (reverse)
        double s = 0;
        int64_t hi = (t + 1) * tile < n ? (t + 1) * tile : n;
        for (int64_t i = hi - 1; i >= t * tile; --i) s += w[i];
        tsum[t] = s;
    }
    ain[0] = 0;
    for (int64_t t = 0; t < nt; ++t) ain[t + 1] = ain[t] + tsum[t];
    const double atot = ain[nt];

    std::vector<double> edges(tile);
    std::vector<double> sin0(nt + 1), sin1(nt + 1);
    int iters = 0;
    int64_t hard0 = 0, hard1 = 0;
    // ---- lap 0 chain
    double S = 0.0;
    for (int64_t t = 0; t < nt; ++t) {
        int64_t lo = t * tile, hi = lo + tile < n ? lo + tile : n;
        int cls = classify_tile(ain[t], ain[t + 1], hi);
        sin0[t] = S;
        if (cls == kCtxHard) {
            ++hard0;
            S = tile_exact_fold(S, w + lo, (int)(hi - lo), edges.data(), &iters);
        } else {
            if (ctx_of(S) != cls) return -11;          // classification must be sound
            Xd a = xd_identity();
            for (int64_t i = lo; i < hi; ++i) {
                i64 f; int up, tie;
                quantize(w[i], cls, f, up, tie);
                xd_push(a, f, up, tie);
            }
            i64 s = xd_apply(units_of(S), a);
            if (s >= kTwo53) return -12;
            S = from_units(s, cls);
        }
    }
    sin0[nt] = S;
    const double total = S;
    if (total_out) *total_out = total;
    Comb comb;
    comb_build(comb, total, u, n_new);
    if (comb.err) return -13;
    // ---- lap 1 chain (continues from total; acc is not reset: smc.cljc:133-136)
    for (int64_t t = 0; t < nt; ++t) {
        int64_t lo = t * tile, hi = lo + tile < n ? lo + tile : n;
        sin1[t] = S;
        if (S > comb.xlast) { sin1[t + 1] = S; for (int64_t q = t + 1; q <= nt; ++q) sin1[q] = S; break; }
        int cls = classify_tile(atot + ain[t], atot + ain[t + 1], n + hi);
        if (cls == kCtxHard) {
            ++hard1;
            S = tile_exact_fold(S, w + lo, (int)(hi - lo), edges.data(), &iters);
        } else {
            if (ctx_of(S) != cls) return -14;
            Xd a = xd_identity();
            for (int64_t i = lo; i < hi; ++i) {
                i64 f; int up, tie;
                quantize(w[i], cls, f, up, tie);
                xd_push(a, f, up, tie);
            }
            i64 s = xd_apply(units_of(S), a);
            if (s >= kTwo53) return -15;
            S = from_units(s, cls);
        }
        sin1[t + 1] = S;
    }
    // ---- replay: per element exact edges -> counts -> emission
    int hint = 0;
    const int64_t J = comb_count_below(comb, total, hint);
    int64_t written = 0;
    for (int lapi = 0; lapi < 2; ++lapi) {
        hint = 0;
        for (int64_t t = 0; t < nt; ++t) {
            int64_t lo = t * tile, hi = lo + tile < n ? lo + tile : n;
            double Sin = lapi ? sin1[t] : sin0[t];
            if (lapi && Sin > comb.xlast) break;
            tile_exact_fold(Sin, w + lo, (int)(hi - lo), edges.data(), nullptr);
            int64_t c_prev = comb_count_below(comb, Sin, hint);
            for (int64_t i = lo; i < hi; ++i) {
                int64_t c = comb_count_below(comb, edges[i - lo], hint);
                if (lapi == 0 && c > J) return -16;
                for (int64_t j = c_prev; j < c; ++j) {
                    if (j >= n_new) return -17;
                    emission[j] = (int32_t)i;
                    lap_out[j] = lapi;
                    ++written;
                }
                c_prev = c;
            }
        }
    }
    if (written != n_new) return -18;
    if (stats) { stats[0] = hard0; stats[1] = hard1; stats[2] = iters; stats[3] = comb.nseg; }
    return 0;
}

// exact fold of w[0..n) from S_in by the fast iteration; returns 0 and the edges, 1 if it
// had to give up (more than TIE_CAP listed ties in one iteration / context outside the fast
// range).  *ties = ties listed over all iterations, *iters = iterations.
int host_fold_fast(double S_in, const double* w, int n, double* edges, double* S_out,
                   int* ties, int* iters) {
    int it = 0, tk = 0;
    const double S = tile_exact_fold_fast(S_in, w, n, edges, &it, &tk);
    *S_out = S;
    if (ties) *ties = tk;
    if (iters) *iters = it;
    return (S != S) ? 1 : 0;
}

// classification of one tile from its approximate prefix (k_finalize)
int host_classify_tile(double a_in, double a_out, int64_t idx_out) {
    return classify_tile(a_in, a_out, idx_out);
}
int host_ctx_of(double S) { return ctx_of(S); }
int host_ctx_hard() { return kCtxHard; }

// ---- sharded fold (k_chain, csrc/resample.cu): a shard that stays in binade `e`
// is one parity transducer; returns 0 and (a0, a1), or 1 if an element saturates
int host_shard_composite(const double* w, int64_t n, int e, int64_t* a0, int64_t* a1) {
    Xd t = xd_identity();
    for (int64_t i = 0; i < n; ++i) {
        i64 f; int up, tie;
        quantize(w[i], e, f, up, tie);
        xd_push(t, f, up, tie);
    }
    *a0 = t.a0; *a1 = t.a1;
    return (t.a0 >= kSat || t.a1 >= kSat) ? 1 : 0;
}
// S' = S + transducer; *ok = 0 if S is not in binade `e` or the sum leaves it
double host_apply_composite(double S, int64_t a0, int64_t a1, int e, int* ok) {
    Xd t; t.a0 = a0; t.a1 = a1;
    *ok = 1;
    if ((a0 | a1) == 0) return S;
    const i64 s = xd_apply(units_of(S), t);
    if (s >= kTwo53 || ctx_of(S) != e) *ok = 0;
    return from_units(s, ctx_of(S));
}

// exact comb values, for direct comparison with repeated addition
void host_comb_values(double total, double u, int64_t n_new, double* out) {
    Comb comb;
    comb_build(comb, total, u, n_new);
    for (int64_t j = 0; j < n_new; ++j) out[j] = comb_value(comb, j);
}

} // extern "C"
