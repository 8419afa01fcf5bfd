// exact_fold.cuh — exact parallel emulation of a sequential fp64 left fold and
// of the systematic-resampling pointer comb.
//
// Replaces the strictly sequential wheel of anglican.smc/resample
// (reference src/anglican/smc.cljc:86-137):
//   total = (reduce + weights)                      left fold, fp64   (:94)
//   acc'  = (+ acc weight)                          running edge      (:120)
//   x'    = (+ x step)                              pointer comb      (:125)
//   (< x next-acc)                                  strict compare    (:122)
//
// Idea.  While the running sum S stays inside one binade [2^k, 2^(k+1)) with
// ulp u = 2^(k-52), the IEEE round-to-nearest-even add  S <- fl(S + w)  (w < 2^k)
// is an INTEGER add in units of u:  s <- s + inc(w, parity(s)),  where
//   w/u = f + r,  inc = f            if r < 1/2
//                 inc = f + 1        if r > 1/2
//                 inc = f + ((s+f)&1) if r == 1/2   (tie -> even)
// so each element is a function {even,odd} -> increment.  These functions
// compose associatively (a "transducer" with one bit of state), hence prefix
// sums of a whole tile can be formed with an ordinary parallel scan and still
// reproduce the sequential fold bit for bit.  Binade crossings (at most ~2100
// in a whole fold, typically ~log2 N) are located exactly and the single
// crossing add is done with a hardware fp64 add.
//
// Everything here is __host__ __device__ so that tests/host/ can exercise the
// arithmetic on the CPU against the oracle's literal loop.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define APOP_HD __host__ __device__ __forceinline__
#else
#define APOP_HD inline
#include <cmath>
#include <cstring>
#endif

namespace apop {

typedef long long i64;
typedef unsigned long long u64;

static constexpr i64 kTwo52 = 1LL << 52;
static constexpr i64 kTwo53 = 1LL << 53;
static constexpr i64 kSat = 1LL << 54;      // saturation value ("crossed")
static constexpr int kCtxSubnormal = -1074; // context of 0 and subnormals

APOP_HD u64 dbits(double x) {
#if defined(__CUDA_ARCH__)
    return (u64)__double_as_longlong(x);
#else
    u64 b; memcpy(&b, &x, 8); return b;
#endif
}
APOP_HD double bits_d(u64 b) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)b);
#else
    double x; memcpy(&x, &b, 8); return x;
#endif
}

// 2^e as a double, e in [-1074, 1023]
APOP_HD double pow2(int e) {
    if (e >= -1022) return bits_d((u64)(e + 1023) << 52);
    return bits_d(1ULL << (e + 1074));
}

// ulp exponent of the binade that holds S (S >= 0 finite): S = s * 2^e with
// s in [2^52, 2^53) for normal S; e = -1074 for zero / subnormal S.
APOP_HD int ctx_of(double S) {
    int be = (int)((dbits(S) >> 52) & 0x7ff);
    return be ? be - 1075 : kCtxSubnormal;
}

// integer significand of S in its own context
APOP_HD i64 units_of(double S) {
    u64 b = dbits(S);
    i64 m = (i64)(b & (u64)(kTwo52 - 1));
    return ((b >> 52) & 0x7ff) ? (m | kTwo52) : m;
}

// exact value s * 2^e (s < 2^53)
APOP_HD double from_units(i64 s, int e) {
    // (double)s is exact; scaling by a power of two is exact in range.
    // Split the scale so that the intermediate never underflows needlessly.
    if (e >= -1022) return (double)s * pow2(e);
    // e < -1022: s*2^e = (s * 2^(e+64)) * 2^-64 would lose bits if subnormal;
    // a direct multiply by the (subnormal) constant 2^e is exact because the
    // result is a multiple of 2^-1074 whenever e >= -1074.
    return (double)s * pow2(e);
}

// Quantise weight w (finite, >= 0) to the grid 2^e.
//   f    = floor(w / 2^e)     (saturated to kSat)
//   up   = 1 if remainder > half
//   tie  = 1 if remainder == half exactly
APOP_HD void quantize(double w, int e, i64& f, int& up, int& tie) {
    u64 b = dbits(w);
    int be = (int)((b >> 52) & 0x7ff);
    u64 m = b & (u64)(kTwo52 - 1);
    int lsb;
    if (be) { m |= (u64)kTwo52; lsb = be - 1075; } else { lsb = -1074; }
    int sh = e - lsb;            // bits of w below the grid
    up = 0; tie = 0;
    if (m == 0) { f = 0; return; }
    if (sh <= 0) {
        // w is a multiple of the grid: exact; may be huge
        int l = -sh;
        if (l >= 11 || (m << l) >= (u64)kSat) f = kSat; else f = (i64)(m << l);
        return;
    }
    if (sh >= 64) { f = 0; return; }       // w < 2^(e+53-64): far below half
    u64 fl = (sh >= 64) ? 0 : (m >> sh);
    u64 rem = m & ((sh >= 64) ? ~0ULL : ((1ULL << sh) - 1ULL));
    u64 half = 1ULL << (sh - 1);
    f = (i64)fl;
    up = rem > half;
    tie = rem == half;
}

// Same result as quantize() for grids 2^e with -1000 <= e <= 0, using fp64
// arithmetic instead of 64-bit shifts: scaled = w * 2^-e is exact (power of
// two, no underflow because 2^-e >= 1); adding 2^52 rounds it to an integer
// (nearest-even) whose value sits in the low mantissa bits.
//   non-tie: f = RN(scaled), up = 0      tie: f = floor(scaled), up = 0
// inv_u = 2^-e.  Must be compiled without FMA contraction.
APOP_HD void quantize_fast(double w, double inv_u, i64& f, int& tie) {
    const double scaled = w * inv_u;
    tie = 0;
    if (!(scaled < 4503599627370496.0)) { f = kSat; return; }     // >= 2^52 (or inf/NaN)
    const double t = scaled + 4503599627370496.0;
    const double rn = t - 4503599627370496.0;
    const double diff = scaled - rn;                               // exact, in [-0.5, 0.5]
    i64 irn = (i64)(dbits(t) - 0x4330000000000000ULL);
    if (diff == 0.5) { tie = 1; }
    else if (diff == -0.5) { tie = 1; irn -= 1; }
    f = irn;
}
APOP_HD bool ctx_fast_ok(int e) { return e >= -1000 && e <= 0; }

// ---- parity transducer: (a0, a1) = total increment for even / odd input ----
struct Xd {
    i64 a0, a1;
};
APOP_HD Xd xd_identity() { Xd r; r.a0 = 0; r.a1 = 0; return r; }
APOP_HD i64 sat(i64 v) { return v > kSat ? kSat : v; }

// append one element (f, up, tie) to a running transducer
APOP_HD void xd_push(Xd& t, i64 f, int up, int tie) {
    if (!tie) {
        i64 inc = f + up;
        t.a0 = sat(t.a0 + inc);
        t.a1 = sat(t.a1 + inc);
    } else {
        t.a0 = sat(t.a0 + f + ((t.a0 + f) & 1));
        t.a1 = sat(t.a1 + f + ((1 + t.a1 + f) & 1));
    }
}
// z = x then y
APOP_HD Xd xd_compose(const Xd& x, const Xd& y) {
    Xd z;
    z.a0 = sat(x.a0 + ((x.a0 & 1) ? y.a1 : y.a0));
    z.a1 = sat(x.a1 + (((1 + x.a1) & 1) ? y.a1 : y.a0));
    return z;
}
// s' = s + t(parity(s))
APOP_HD i64 xd_apply(i64 s, const Xd& t) { return sat(s + ((s & 1) ? t.a1 : t.a0)); }

// one exact step on the integer state (used when replaying elements)
APOP_HD i64 step_units(i64 s, i64 f, int up, int tie) {
    return sat(s + f + (tie ? ((s + f) & 1) : (i64)up));
}

// ---- pointer comb --------------------------------------------------------
// x_0 = fl(total*U), x_{j+1} = fl(x_j + step).  Piecewise closed form:
// for j in [j0, j0+len): x_j = (X0 + (j-j0)*q) * 2^e.
struct CombSeg {
    i64 j0, len, X0, q;
    double xfirst;   // value of x_{j0}
    double xlast;    // value of x_{j0+len-1}
    int e;
    int pad;
};
static constexpr int kMaxCombSegs = 192;

struct Comb {
    int nseg;
    int err;        // 1: table overflow / stuck comb
    i64 n_new;
    double x0, xlast, step;
    CombSeg seg[kMaxCombSegs];
};

APOP_HD void comb_build(Comb& c, double total, double u, i64 n_new) {
    c.nseg = 0; c.err = 0; c.n_new = n_new;
    double step = total / (double)n_new;       // smc.cljc:95
    double x = total * u;                      // (rand total-weight), inference.cljc:124-129
    c.step = step; c.x0 = x; c.xlast = x;
    i64 j = 0;
    while (j < n_new) {
        if (c.nseg >= kMaxCombSegs) { c.err = 1; return; }
        CombSeg& s = c.seg[c.nseg++];
        int e = ctx_of(x);
        i64 X = units_of(x);
        i64 f; int up, tie;
        quantize(step, e, f, up, tie);
        s.j0 = j; s.X0 = X; s.e = e; s.xfirst = x; s.pad = 0;
        bool single = false;
        i64 q = f + up;
        if (f >= kSat) single = true;             // step >= this binade: crosses at once
        else if (tie) {
            if (X & 1) single = true;             // odd state + tie: one parity-dependent step
            else q = f + (f & 1);                 // even state stays even: constant stride
        }
        if (!single && q == 0) {                  // comb cannot advance (n_new > 2^52)
            s.q = 0; s.len = n_new - j; s.xlast = x; c.xlast = x; c.err = 1; return;
        }
        i64 len = 1;
        if (!single) {
            i64 mmax = (kTwo53 - 1 - X) / q;      // X + m*q < 2^53
            len = mmax + 1;
            if (len > n_new - j) len = n_new - j;
        }
        s.q = single ? 0 : q;
        s.len = len;
        double xl = single ? x : from_units(X + (len - 1) * q, e);
        s.xlast = xl;
        c.xlast = xl;
        j += len;
        if (j >= n_new) break;
        x = xl + step;                            // the crossing add, in hardware fp64
    }
}

// exact x_j
APOP_HD double comb_value(const Comb& c, i64 j) {
    int lo = 0, hi = c.nseg - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (c.seg[mid].j0 <= j) lo = mid; else hi = mid - 1;
    }
    const CombSeg& s = c.seg[lo];
    if (s.q == 0) return s.xfirst;
    return from_units(s.X0 + (j - s.j0) * s.q, s.e);
}

// number of pointers of one segment strictly below S, given xfirst < S
APOP_HD i64 seg_count_below(const CombSeg& s, double S) {
    if (S > s.xlast) return s.len;
    // xfirst < S <= xlast < 2^(k+1): S is an exact multiple of 2^e
    i64 Su = (ctx_of(S) == s.e) ? units_of(S) : (units_of(S) << (ctx_of(S) - s.e));
    i64 num = Su - s.X0;                          // > 0
    if (s.q == 0) return s.len;
    // smallest i with i*q >= num  (i = ceil(num/q)), via fp estimate + fix-up
    i64 i = (i64)((double)num / (double)s.q);
    while (i * s.q < num) ++i;
    while (i > 0 && (i - 1) * s.q >= num) --i;
    return i > s.len ? s.len : i;
}

// cnt(S) = #{ j in [0,n_new) : x_j < S }.  `hint` (in/out) is a segment index
// guess; S values that grow monotonically make the search incremental.
APOP_HD i64 comb_count_below(const Comb& c, double S, int& hint) {
    if (!(S > c.x0)) return 0;
    if (S > c.xlast) return c.n_new;
    int g = hint;
    if (g < 0) g = 0;
    if (g >= c.nseg) g = c.nseg - 1;
    // want the last segment with xfirst < S
    if (!(c.seg[g].xfirst < S)) {
        int lo = 0, hi = g;                       // answer in [0, g)
        while (lo < hi) {
            int mid = (lo + hi + 1) >> 1;
            if (c.seg[mid].xfirst < S) lo = mid; else hi = mid - 1;
        }
        g = lo;
    } else {
        while (g + 1 < c.nseg && c.seg[g + 1].xfirst < S) {
            // gallop forward; segments grow geometrically so this is short
            int lo = g + 1, hi = c.nseg - 1;
            while (lo < hi) {
                int mid = (lo + hi + 1) >> 1;
                if (c.seg[mid].xfirst < S) lo = mid; else hi = mid - 1;
            }
            g = lo;
        }
    }
    hint = g;
    return c.seg[g].j0 + seg_count_below(c.seg[g], S);
}

// ---- classification of a tile from an approximate prefix -------------------
// A tile is "easy" in context e when the exact fold provably stays inside one
// normal binade over the whole tile:  2^(e+52) <= S_in  and  S_out < 2^(e+53).
// a_in/a_out: approximate exclusive / inclusive prefix (any summation order),
// idx_out: number of fold terms up to the tile end (bounds the rounding error
// of BOTH the sequential fold and the approximation: (n+slack)*2^-52 relative).
static constexpr int kCtxHard = -100000;
static constexpr int kCtxSkip = -100001;

// rel_eps: extra relative slack when the approximation itself is only accurate
// to rel_eps (fp32 tile partials of the fused propagate kernel).
// e_lo / e_hi (optional): contexts of the lower / upper bound of the running
// sum over the tile, or kCtxHard when they are unusable.  A hard tile with
// e_hi == e_lo + 1 crosses (at most) one binade boundary: its fold can be
// resolved from precomputed prefix tables in e_lo and suffix tables in e_hi.
APOP_HD int classify_tile2(double a_in, double a_out, i64 idx_out, double rel_eps, int* e_lo_out,
                           int* e_hi_out) {
    *e_lo_out = kCtxHard; *e_hi_out = kCtxHard;
    if (!(a_in > 0.0) || !(a_out < 1.0e300)) return kCtxHard;
    double eps = ((double)idx_out + 4096.0) * 4.44089209850062616e-16 + rel_eps; // 2^-51
    double lo = a_in * (1.0 - eps);
    double hi = a_out * (1.0 + eps);
    int e_lo = ctx_of(lo);
    if (e_lo == kCtxSubnormal) return kCtxHard;
    *e_lo_out = e_lo; *e_hi_out = ctx_of(hi);
    if (ctx_of(hi) != e_lo) return kCtxHard;
    return e_lo;
}
APOP_HD int classify_tile(double a_in, double a_out, i64 idx_out, double rel_eps = 0.0) {
    int a, b;
    return classify_tile2(a_in, a_out, idx_out, rel_eps, &a, &b);
}

} // namespace apop
