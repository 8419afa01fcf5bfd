// fastmath.cuh — fp64 exp / log / sincospi / sqrt / reciprocal for the sweep
// kernels.  Same accuracy class as the CUDA math library (<= 2 ulp), but:
//   * every polynomial coefficient is an operand in the constant bank
//     (DFMA R, R, c[3][..], R) instead of a UMOV pair per constant — the
//     library versions spend ~40% of their issue slots on those moves;
//   * no special-case branches on the hot path (callers guarantee the range;
//     out-of-range arguments fall back to the library function).
// The reference evaluates these through java.lang.Math / commons-math FastMath
// (<= 1 ulp, runtime.cljc:403-409 `normal`, smc.cljc:82 `exp`); parity bars are
// 1e-9 relative on log-densities / log-marginals (tests/).
#pragma once
#include <math.h>

#include "fastmath_coef.inc"

namespace apop {
namespace fm {

// [0, NE) exp | [NE, NE+NQ) log Q(z) | sin(pi r)/r | cos(pi r) | misc scalars
enum { OFF_E = 0, OFF_Q = FM_NE, OFF_S = FM_NE + FM_NQ, OFF_C = FM_NE + FM_NQ + FM_NS,
       OFF_M = FM_NE + FM_NQ + FM_NS + FM_NC };
enum { M_L2E = OFF_M, M_MAGIC, M_LN2HI, M_LN2LO, M_ONE, M_HALF, M_TWO, M_NEG2, M_COUNT };

static __constant__ double kC[] = {
    FM_TABLE,
    0x1.71547652b82fep+0,      // log2(e)
    6755399441055744.0,        // 1.5 * 2^52 (round-to-integer magic)
    0x1.62e42fefa39efp-1,      // ln2 hi
    0x1.abc9e3b39803fp-56,     // ln2 lo
    1.0, 0.5, 2.0, -2.0
};

__device__ __forceinline__ double rcp_approx(double d) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
    return y;
}
__device__ __forceinline__ double rsqrt_approx(double d) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
    return y;
}

// a / d for normal, finite d (|d| in [2^-500, 2^500]) and finite a: Newton on
// the MUFU seed, one correction step.  <= 1 ulp.
__device__ __forceinline__ double div_fast(double a, double d) {
    double y = rcp_approx(d);
    double e = fma(-d, y, kC[M_ONE]);
    y = fma(y, e, y);
    e = fma(-d, y, kC[M_ONE]);
    y = fma(y, e, y);
    double q = a * y;
    double r = fma(-d, q, a);
    return fma(r, y, q);
}

// sqrt(a) for normal a > 0 (Goldschmidt on the MUFU seed + one correction)
__device__ __forceinline__ double sqrt_fast(double a) {
    double y = rsqrt_approx(a);
    double g = a * y, h = kC[M_HALF] * y;
    double r = fma(-h, g, kC[M_HALF]);
    g = fma(g, r, g); h = fma(h, r, h);
    r = fma(-h, g, kC[M_HALF]);
    g = fma(g, r, g); h = fma(h, r, h);
    double d = fma(-g, g, a);
    return fma(d, h, g);
}

// exp(x) for |x| < 708 (no overflow / subnormal results)
__device__ __forceinline__ double exp_core(double x) {
    const double t = fma(x, kC[M_L2E], kC[M_MAGIC]);
    const double n = t - kC[M_MAGIC];
    const int ni = __double2loint(t);
    double r = fma(n, -kC[M_LN2HI], x);
    r = fma(n, -kC[M_LN2LO], r);
    double p = kC[OFF_E + FM_NE - 1];
#pragma unroll
    for (int k = FM_NE - 2; k >= 0; --k) p = fma(p, r, kC[OFF_E + k]);
    return __hiloint2double(__double2hiint(p) + (ni << 20), __double2loint(p));
}
// exp(x), any x: library fallback outside the fast range (rare, warp-divergent)
__device__ __forceinline__ double exp_any(double x) {
    if (!(fabs(x) < 708.0)) return exp(x);
    return exp_core(x);
}

// log(u) for normal u > 0
__device__ __forceinline__ double log_core(double u) {
    int hi = __double2hiint(u);
    const int lo = __double2loint(u);
    int e = (hi >> 20) - 1023;
    hi = (hi & 0x000fffff) | 0x3ff00000;
    if (hi > 0x3ff6a09e) { hi -= 0x00100000; e += 1; }          // m in [0.7071, 1.4142)
    const double m = __hiloint2double(hi, lo);
    const double f = m - kC[M_ONE];
    const double s = div_fast(f, kC[M_TWO] + f);
    const double z = s * s;
    double q = kC[OFF_Q + FM_NQ - 1];
#pragma unroll
    for (int k = FM_NQ - 2; k >= 0; --k) q = fma(q, z, kC[OFF_Q + k]);
    const double ed = (double)e;
    const double lm = fma(s * z, q, fma(ed, kC[M_LN2LO], s + s));
    return fma(ed, kC[M_LN2HI], lm);
}

// sin(pi t), cos(pi t) for t in [0, 2]
__device__ __forceinline__ void sincospi_core(double t, double& sn, double& cs) {
    const double kk = fma(t, kC[M_TWO], kC[M_MAGIC]);
    const int qi = __double2loint(kk);
    const double kd = kk - kC[M_MAGIC];
    const double r = fma(kd, -kC[M_HALF], t);                    // exact, |r| <= 1/4
    const double z = r * r;
    double sp = kC[OFF_S + FM_NS - 1], cp = kC[OFF_C + FM_NC - 1];
#pragma unroll
    for (int k = FM_NS - 2; k >= 0; --k) sp = fma(sp, z, kC[OFF_S + k]);
#pragma unroll
    for (int k = FM_NC - 2; k >= 0; --k) cp = fma(cp, z, kC[OFF_C + k]);
    sp = sp * r;
    const double a = (qi & 1) ? cp : sp;
    const double b = (qi & 1) ? sp : cp;
    sn = (qi & 2) ? -a : a;
    cs = ((qi + 1) & 2) ? -b : b;
}

// ---- vectorised forms: the coefficient is fetched once per Horner step and
// applied to N independent values (N-way ILP, 1/N of the constant traffic).
template <int N>
__device__ __forceinline__ void exp_core_v(double (&x)[N]) {     // in place, |x| < 708
    double r[N], p[N];
    int ni[N];
    const double l2e = kC[M_L2E], magic = kC[M_MAGIC], lh = -kC[M_LN2HI], ll = -kC[M_LN2LO];
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const double t = fma(x[i], l2e, magic);
        const double n = t - magic;
        ni[i] = __double2loint(t);
        r[i] = fma(n, ll, fma(n, lh, x[i]));
    }
    const double c0 = kC[OFF_E + FM_NE - 1];
#pragma unroll
    for (int i = 0; i < N; ++i) p[i] = c0;
#pragma unroll
    for (int k = FM_NE - 2; k >= 0; --k) {
        const double c = kC[OFF_E + k];
#pragma unroll
        for (int i = 0; i < N; ++i) p[i] = fma(p[i], r[i], c);
    }
#pragma unroll
    for (int i = 0; i < N; ++i)
        x[i] = __hiloint2double(__double2hiint(p[i]) + (ni[i] << 20), __double2loint(p[i]));
}

// particle-weights (smc.cljc:78-82) for N log-weights against the max m:
//   1.0 if lw == m; 0.0 if lw is -inf or NaN; exp(lw - m) otherwise.
// exp_core(0) is exactly 1, so only d < -708 / NaN needs the explicit rules.
template <int N>
__device__ __forceinline__ void weights_v(const double (&lw)[N], double m, double (&w)[N]) {
    bool slow = false;
#pragma unroll
    for (int i = 0; i < N; ++i) { w[i] = lw[i] - m; slow |= !(w[i] >= -708.0); }
    if (slow) {
#pragma unroll
        for (int i = 0; i < N; ++i)
            w[i] = (lw[i] == m) ? 1.0 : (!(lw[i] > -INFINITY) ? 0.0 : exp(lw[i] - m));
        return;
    }
    exp_core_v<N>(w);
}

// N Box-Muller pairs: (u1, u2) -> (z0, z1) = sqrt(-2 log u1) (cos, sin)(2 pi u2)
template <int N>
__device__ __forceinline__ void box_muller_v(const double (&u1)[N], const double (&u2)[N],
                                             double (&z0)[N], double (&z1)[N]) {
    const double one = kC[M_ONE], half = kC[M_HALF], two = kC[M_TWO], magic = kC[M_MAGIC];
    double s[N], zq[N], ed[N], rad[N];
    // ---- log(u1): atanh series on the reduced mantissa
#pragma unroll
    for (int i = 0; i < N; ++i) {
        int hi = __double2hiint(u1[i]);
        const int lo = __double2loint(u1[i]);
        int e = (hi >> 20) - 1023;
        hi = (hi & 0x000fffff) | 0x3ff00000;
        if (hi > 0x3ff6a09e) { hi -= 0x00100000; e += 1; }
        const double f = __hiloint2double(hi, lo) - one;
        const double d = two + f;
        double y = rcp_approx(d);
        double t = fma(-d, y, one);
        y = fma(y, t, y);
        t = fma(-d, y, one);
        y = fma(y, t, y);
        const double q = f * y;
        s[i] = fma(fma(-d, q, f), y, q);
        zq[i] = s[i] * s[i];
        ed[i] = (double)e;
    }
    double q[N];
    const double cq = kC[OFF_Q + FM_NQ - 1];
#pragma unroll
    for (int i = 0; i < N; ++i) q[i] = cq;
#pragma unroll
    for (int k = FM_NQ - 2; k >= 0; --k) {
        const double c = kC[OFF_Q + k];
#pragma unroll
        for (int i = 0; i < N; ++i) q[i] = fma(q[i], zq[i], c);
    }
    const double lh = kC[M_LN2HI], ll = kC[M_LN2LO], neg2 = kC[M_NEG2];
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const double lm = fma(s[i] * zq[i], q[i], fma(ed[i], ll, s[i] + s[i]));
        const double a = neg2 * fma(ed[i], lh, lm);               // -2 log u1 > 0
        double y = rsqrt_approx(a);
        double g = a * y, h = half * y;
        double r = fma(-h, g, half);
        g = fma(g, r, g); h = fma(h, r, h);
        r = fma(-h, g, half);
        g = fma(g, r, g); h = fma(h, r, h);
        rad[i] = fma(fma(-g, g, a), h, g);
    }
    // ---- sincospi(2 u2)
    double rr[N], sp[N], cp[N];
    int qi[N];
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const double t = u2[i] + u2[i];
        const double kk = fma(t, two, magic);
        qi[i] = __double2loint(kk);
        rr[i] = fma(kk - magic, -half, t);
        zq[i] = rr[i] * rr[i];
    }
    const double cs0 = kC[OFF_S + FM_NS - 1], cc0 = kC[OFF_C + FM_NC - 1];
#pragma unroll
    for (int i = 0; i < N; ++i) { sp[i] = cs0; cp[i] = cc0; }
#pragma unroll
    for (int k = FM_NS - 2; k >= 0; --k) {
        const double c = kC[OFF_S + k];
#pragma unroll
        for (int i = 0; i < N; ++i) sp[i] = fma(sp[i], zq[i], c);
    }
#pragma unroll
    for (int k = FM_NC - 2; k >= 0; --k) {
        const double c = kC[OFF_C + k];
#pragma unroll
        for (int i = 0; i < N; ++i) cp[i] = fma(cp[i], zq[i], c);
    }
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const double sv = sp[i] * rr[i];
        const double a = (qi[i] & 1) ? cp[i] : sv;
        const double b = (qi[i] & 1) ? sv : cp[i];
        const double sn = (qi[i] & 2) ? -a : a;
        const double cs = ((qi[i] + 1) & 2) ? -b : b;
        z0[i] = rad[i] * cs;
        z1[i] = rad[i] * sn;
    }
}


} // namespace fm
} // namespace apop
