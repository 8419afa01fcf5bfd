/*
 * anglican_oracle.c — CPU restatement (the ORACLE) of the particle-population
 * hot path of probprog/anglican.  TEST INFRASTRUCTURE ONLY — see the header.
 *
 * Build: gcc -O2 -std=c11 -ffp-contract=off -fno-fast-math -fPIC -shared
 *        (oracle/build.py).  Strictly sequential fp64; OpenMP (optional) is
 *        used only across particles in the propagate/weight loop of the sweep,
 *        never inside the resampling wheel.
 *
 * Parity pinning: see anglican_oracle.h.  Functions with no reference test or
 * golden vector are marked "parity unpinned" below.
 */
#include "anglican_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif
#ifndef M_E
#define M_E 2.7182818284590452354
#endif

int orc_version(void) { return 1; }

/* ========================================================================
 * Counter-based RNG (engine semantics).  The reference's RNG is an unseeded
 * Well19937c behind a global lock (runtime.cljc:142-147) so sample streams are
 * not reproducible there; only distributional parity is meaningful for draws.
 * Philox4x32-10 (Salmon et al., SC'11), standard constants.
 * ====================================================================== */
static inline void mulhilo32(uint32_t a, uint32_t b, uint32_t* hi, uint32_t* lo) {
    uint64_t p = (uint64_t)a * (uint64_t)b;
    *hi = (uint32_t)(p >> 32);
    *lo = (uint32_t)p;
}

void orc_philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                    uint32_t k0, uint32_t k1, uint32_t out[4]) {
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0, lo0, hi1, lo1;
        mulhilo32(0xD2511F53u, c0, &hi0, &lo0);
        mulhilo32(0xCD9E8D57u, c2, &hi1, &lo1);
        uint32_t n0 = hi1 ^ c1 ^ k0;
        uint32_t n1 = lo1;
        uint32_t n2 = hi0 ^ c3 ^ k1;
        uint32_t n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

double orc_u01_open(uint32_t hi, uint32_t lo) {
    /* 53 bits: 27 from hi, 26 from lo; centred so the value is never 0 or 1 */
    uint64_t bits = ((uint64_t)(hi >> 5) << 26) | (uint64_t)(lo >> 6);
    return ((double)bits + 0.5) * (1.0 / 9007199254740992.0);
}

double orc_u01_half(uint32_t hi, uint32_t lo) {
    uint64_t bits = ((uint64_t)(hi >> 5) << 26) | (uint64_t)(lo >> 6);
    return (double)bits * (1.0 / 9007199254740992.0);
}

/* counter word 3: stream (8 bits) | kind (4 bits: 0 normal pair, 1 uniform) |
 * draw index (20 bits) */
static inline uint32_t ctr3(uint32_t stream, uint32_t kind, uint32_t draw) {
    return (stream << 24) | (kind << 20) | (draw & 0xFFFFFu);
}

/* Gaussian draws (engine semantics; the reference's nextGaussian runs on an unseeded
 * Well19937c, runtime.cljc:142-147, 403-409, so only distributional parity is
 * meaningful): Box-Muller on one Philox block per slot pair (2p, 2p+1), cos -> even
 * slot, sin -> odd slot. */
double orc_normal_at(uint64_t seed, uint64_t slot, uint32_t t, uint32_t stream,
                     uint32_t draw) {
    uint64_t pair = slot >> 1;
    uint32_t o[4];
    orc_philox4x32((uint32_t)pair, (uint32_t)(pair >> 32), t, ctr3(stream, 0, draw),
                   (uint32_t)seed, (uint32_t)(seed >> 32), o);
    double u1 = orc_u01_open(o[0], o[1]);
    double u2 = orc_u01_half(o[2], o[3]);
    double r = sqrt(-2.0 * log(u1));
    double ang = 2.0 * M_PI * u2;
    return (slot & 1) ? r * sin(ang) : r * cos(ang);
}

double orc_uniform_at(uint64_t seed, uint64_t slot, uint32_t t, uint32_t stream,
                      uint32_t draw) {
    uint32_t o[4];
    orc_philox4x32((uint32_t)slot, (uint32_t)(slot >> 32), t, ctr3(stream, 1, draw),
                   (uint32_t)seed, (uint32_t)(seed >> 32), o);
    return orc_u01_half(o[0], o[1]);
}

/* ========================================================================
 * smc.cljc
 * ====================================================================== */

/* particle-weights, smc.cljc:70-84.  `(reduce max ...)` uses clojure.core/max
 * which returns NaN if either argument is NaN.  parity unpinned. */
int orc_particle_weights(const double* lw, int64_t n, double* w, double* max_lw) {
    double m = lw[0];
    for (int64_t i = 1; i < n; ++i) {
        double v = lw[i];
        if (isnan(m) || isnan(v)) m = NAN;
        else if (v > m) m = v;
    }
    *max_lw = m;
    if (isnan(m)) return -1;
    for (int64_t i = 0; i < n; ++i) {
        double v = lw[i];
        if (v == m) w[i] = 1.0;                 /* smc.cljc:80 (all may be -inf) */
        else if (!(v > -INFINITY)) w[i] = 0.0;  /* smc.cljc:81 NaN or -inf */
        else w[i] = exp(v - m);                 /* smc.cljc:82 */
    }
    return 0;
}

/* the wheel of smc.cljc:105-137, literally.  parity unpinned. */
int orc_resample_weights(const double* w, int64_t n, int64_t n_new, double u,
                         int32_t* anc_ref, int32_t* emission, int32_t* lap_of,
                         double* total_out, double* step_out) {
    double total = w[0];                        /* (reduce + weights): left fold */
    for (int64_t i = 1; i < n; ++i) total = total + w[i];
    double step = total / (double)n_new;        /* smc.cljc:95 */
    if (total_out) *total_out = total;
    if (step_out) *step_out = step;
    if (!(total > 0.0) || !(total < INFINITY)) return -2;

    double x = total * u;                       /* (rand total-weight): inference.cljc:124-129 */
    double acc = 0.0;                           /* smc.cljc:112 (integer 0; 0+w == w) */
    int64_t i = 0, cnt = 0;
    int32_t lap = 0;
    while (cnt < n_new) {
        double next_acc = acc + w[i];           /* smc.cljc:120 */
        if (x < next_acc) {                     /* smc.cljc:122, strict */
            if (emission) emission[cnt] = (int32_t)i;
            if (lap_of) lap_of[cnt] = lap;
            if (anc_ref) anc_ref[n_new - 1 - cnt] = (int32_t)i; /* conj onto nil: prepend */
            x = x + step;                       /* smc.cljc:125 */
            ++cnt;
        } else {
            acc = next_acc;                     /* smc.cljc:134 */
            if (++i == n) { i = 0; ++lap; }     /* smc.cljc:135-136: recycle, acc kept */
            if (lap > 4) return -3;
        }
    }
    return 0;
}

int orc_resample(const double* lw, int64_t n, int64_t n_new, double u,
                 int32_t* anc_ref, double* new_lw, double* w_scratch) {
    double m;
    int rc = orc_particle_weights(lw, n, w_scratch, &m);
    if (rc) return rc;
    double total, step;
    rc = orc_resample_weights(w_scratch, n, n_new, u, anc_ref, NULL, NULL, &total, &step);
    /* smc.cljc:98-103 */
    if (new_lw) *new_lw = (m > -INFINITY && m < INFINITY) ? (log(step) + m) : -INFINITY;
    return rc;
}

void orc_emission_to_sorted(const int32_t* emission, const int32_t* lap_of,
                            int64_t n_new, int32_t* anc_sorted, int64_t* n_lap0) {
    int64_t j0 = 0;
    while (j0 < n_new && lap_of[j0] == 0) ++j0;  /* J = emissions of the first lap */
    if (n_lap0) *n_lap0 = j0;
    int64_t k = 0;
    for (int64_t j = j0; j < n_new; ++j) anc_sorted[k++] = emission[j];
    for (int64_t j = 0; j < j0; ++j) anc_sorted[k++] = emission[j];
}

/* n_new independent `discrete` draws: runtime.cljc:277-287 (x = U*total, first
 * k with x < running left-fold sum, strict). */
int orc_multinomial_weights(const double* w, int64_t n, const double* u,
                            int64_t n_new, int32_t* anc) {
    double* edge = (double*)malloc(sizeof(double) * (size_t)n);
    if (!edge) return -9;
    double acc = 0.0;
    for (int64_t i = 0; i < n; ++i) { acc = acc + w[i]; edge[i] = acc; }
    double total = acc;
    for (int64_t k = 0; k < n_new; ++k) {
        double x = u[k] * total;
        int64_t lo = 0, hi = n;          /* first i with x < edge[i] */
        while (lo < hi) {
            int64_t mid = (lo + hi) >> 1;
            if (x < edge[mid]) hi = mid; else lo = mid + 1;
        }
        anc[k] = (int32_t)(lo < n ? lo : n - 1);
    }
    free(edge);
    return 0;
}

/* ========================================================================
 * runtime.cljc:130-138, inference.cljc:181-186, dist_test.cljc:83-94
 * ====================================================================== */
double orc_log_sum_exp(double a, double b) {    /* parity unpinned */
    double mx = (isnan(a) || isnan(b)) ? NAN : (a > b ? a : b);
    if (mx > -INFINITY && mx < INFINITY)
        return mx + log(exp(a - mx) + exp(b - mx));
    return mx;
}

static int cmp_double(const void* a, const void* b) {
    double x = *(const double*)a, y = *(const double*)b;
    return (x > y) - (x < y);
}

double orc_log_marginal(const double* lw, int64_t n) {   /* parity unpinned */
    double* s = (double*)malloc(sizeof(double) * (size_t)n);
    memcpy(s, lw, sizeof(double) * (size_t)n);
    qsort(s, (size_t)n, sizeof(double), cmp_double);     /* (sort ...) ascending */
    double acc = s[0];
    for (int64_t i = 1; i < n; ++i) acc = orc_log_sum_exp(acc, s[i]);
    free(s);
    return acc - log((double)n);
}

double orc_ess(const double* lw, int64_t n) {
    if (n <= 0) return 0.0;
    double m = lw[0];
    for (int64_t i = 1; i < n; ++i) if (lw[i] > m) m = lw[i];
    double s = 0.0, s2 = 0.0;
    for (int64_t i = 0; i < n; ++i) {
        double w = exp(lw[i] - m);
        s = s + w;
        s2 = s2 + w * w;
    }
    return (s * s) / s2;
}

/* ========================================================================
 * Distributions.  normal/gamma/beta/uniform-continuous are commons-math3
 * 3.6.1 objects (project.clj:11) built by `from-apache` (runtime.cljc:207-220);
 * commons-math is NOT under /root/reference, so its published formulas are
 * restated here and cross-checked against scipy in tests/.
 * ====================================================================== */

/* NormalDistribution.logDensity: x1 = (x-mean)/sd;
 * -0.5*x1*x1 - (log(sd) + 0.5*log(2*pi)) */
double orc_normal_logpdf(double x, double mean, double sd) {
    double x0 = x - mean;
    double x1 = x0 / sd;
    return -0.5 * x1 * x1 - (log(sd) + 0.5 * log(2.0 * M_PI));
}

/* Gamma.lanczos (g = 607/128, 15 coefficients) */
static const double LANCZOS_G = 607.0 / 128.0;
static const double LANCZOS[15] = {
    0.99999999999999709182,   57.156235665862923517,   -59.597960355475491248,
    14.136097974741747174,    -0.49191381609762019978,  .33994649984811888699e-4,
    .46523628927048575665e-4, -.98374475304879564677e-4, .15808870322491248884e-3,
    -.21026444172410488319e-3, .21743961811521264320e-3, -.16431810653676389022e-3,
    .84418223983852743293e-4, -.26190838401581408670e-4, .36899182659531622704e-5 };

static double lanczos(double x) {
    double sum = 0.0;
    for (int i = 14; i > 0; --i) sum = sum + LANCZOS[i] / (x + i);
    return sum + LANCZOS[0];
}

/* GammaDistribution(shape, scale = 1/rate).logDensity — Lanczos-prefactor form
 * with the overflow branch (commons-math3 3.6.1 GammaDistribution). */
double orc_gamma_logpdf(double x, double shape, double rate) {
    double scale = 1.0 / rate;                  /* runtime.cljc:327 */
    if (x < 0) return -INFINITY;
    double shifted = shape + LANCZOS_G + 0.5;
    double aux = M_E / (2.0 * M_PI * shifted);
    double lp2 = log(shape) + 0.5 * log(aux) - log(lanczos(shape));
    double lp1 = lp2 - log(scale) - log(shifted) * shape + shape + LANCZOS_G;
    double min_y = shape + LANCZOS_G - log(1.7976931348623157e308);
    double max_log_y = log(1.7976931348623157e308) / (shape - 1.0);
    double y = x / scale;
    if (y <= min_y || log(y) >= max_log_y) {
        double aux1 = (y - shifted) / shifted;
        double aux2 = shape * (log1p(aux1) - aux1);
        double aux3 = -y * (LANCZOS_G + 0.5) / shifted + LANCZOS_G + aux2;
        return lp2 - log(x) + aux3;
    }
    return lp1 - y + log(y) * (shape - 1.0);
}

double orc_gamma_logpdf_direct(double x, double shape, double rate) {
    if (x < 0) return -INFINITY;
    double y = x * rate;
    return -lgamma(shape) + log(rate) + (shape - 1.0) * log(y) - y;
}

/* BetaDistribution.logDensity (z = logGamma(a)+logGamma(b)-logGamma(a+b));
 * commons' Gamma.logGamma agrees with libm lgamma to ~1e-15.  parity unpinned
 * by any reference golden (cross-checked against scipy in tests). */
double orc_beta_logpdf(double x, double alpha, double beta) {
    if (x < 0 || x > 1) return -INFINITY;
    if (x == 0) return (alpha < 1) ? NAN /* NumberIsTooSmallException */ : -INFINITY;
    if (x == 1) return (beta < 1) ? NAN : -INFINITY;
    double z = lgamma(alpha) + lgamma(beta) - lgamma(alpha + beta);
    return (alpha - 1.0) * log(x) + (beta - 1.0) * log1p(-x) - z;
}

/* UniformRealDistribution: log(density), density 1/(hi-lo) on [lo,hi] */
double orc_uniform_logpdf(double x, double lo, double hi) {
    if (x < lo || x > hi) return -INFINITY;
    return log(1.0 / (hi - lo));
}

/* discrete observe*, runtime.cljc:288-290: log(nth(weights, v, 0.0)/total) */
double orc_discrete_logpmf(const double* w, int k, int64_t v) {
    double total = 0.0;
    for (int i = 0; i < k; ++i) total = total + w[i];
    double wv = (v >= 0 && v < k) ? w[v] : 0.0;
    return log(wv / total);
}

int64_t orc_discrete_sample(const double* w, int k, double u) {
    double total = 0.0;
    for (int i = 0; i < k; ++i) total = total + w[i];
    double x = u * total;     /* uniform-continuous 0 total: u*max + (1-u)*min, min = 0 */
    double acc = 0.0;
    for (int i = 0; i < k; ++i) {
        acc = acc + w[i];
        if (x < acc) return i;
    }
    return k - 1;             /* unreachable for u<1 (reference would NPE) */
}

/* dirichlet observe*, runtime.cljc:292-306.  parity unpinned. */
double orc_dirichlet_logpdf(const double* v, const double* alpha, int k) {
    double sl = 0.0, sa = 0.0, acc = 0.0;
    for (int i = 0; i < k; ++i) { sl = sl + lgamma(alpha[i]); sa = sa + alpha[i]; }
    double Z = sl - lgamma(sa);
    for (int i = 0; i < k; ++i) acc = acc + log(v[i]) * (alpha[i] - 1.0);
    return acc - Z;
}

/* mvn, runtime.cljc:349-368.  vectorz's Cholesky is an EJML-derived inner
 * (row) lower decomposition that reads the UPPER triangle of the input
 * (dist_test.cljc:369 passes a non-symmetric matrix, so the triangle matters;
 * which triangle is read is not pinned by any reference test). */
int orc_mvn_prepare(const double* cov, int d, double* L, double* iL, double* Z) {
    memset(L, 0, sizeof(double) * (size_t)d * d);
    for (int i = 0; i < d; ++i) {
        double el_ii = 0.0, div = 0.0;
        for (int j = i; j < d; ++j) {
            double sum = cov[i * d + j];
            for (int k = 0; k < i; ++k) sum = sum - L[i * d + k] * L[j * d + k];
            if (i == j) {
                if (!(sum > 0.0)) return -4;
                el_ii = sqrt(sum);
                L[i * d + i] = el_ii;
                div = 1.0 / el_ii;
            } else {
                L[j * d + i] = sum * div;
            }
        }
    }
    /* iLcov = inverse of lower-triangular L (forward substitution) */
    memset(iL, 0, sizeof(double) * (size_t)d * d);
    for (int c = 0; c < d; ++c) {
        for (int r = c; r < d; ++r) {
            double s = (r == c) ? 1.0 : 0.0;
            for (int k = c; k < r; ++k) s = s - L[r * d + k] * iL[k * d + c];
            iL[r * d + c] = s / L[r * d + r];
        }
    }
    double det = 1.0;                               /* (reduce * (diagonal Lcov)) */
    for (int i = 0; i < d; ++i) det = det * L[i * d + i];
    *Z = 0.5 * d * log(2.0 * M_PI) + log(det);      /* runtime.cljc:355-357 */
    return 0;
}

double orc_mvn_logpdf(const double* x, const double* mean, const double* iL,
                      double Z, int d, double* work) {
    double dot = 0.0;
    for (int i = 0; i < d; ++i) work[i] = x[i] - mean[i];
    for (int r = 0; r < d; ++r) {
        double s = 0.0;
        for (int c = 0; c < d; ++c) s = s + iL[r * d + c] * work[c];   /* (mmul iLcov dx) */
        dot = dot + s * s;
    }
    return -0.5 * dot - Z;                                             /* runtime.cljc:364-366 */
}

/* ---- multivariate-t and wishart (runtime.cljc:371-401, 479-545) ----------
 * The reference forms sigma^-1 with `m/inverse` and the determinant with `m/det`
 * (vectorz: LU with partial pivoting); restated literally.  work: 2 d^2 + d doubles. */
static double lu_det_inverse(const double* a_in, int d, double* lu, double* inv) {
    /* Gauss-Jordan with partial pivoting on [A | I]; returns det(A), inv may be NULL */
    double det = 1.0;
    for (int i = 0; i < d * d; ++i) lu[i] = a_in[i];
    if (inv) for (int i = 0; i < d; ++i) for (int j = 0; j < d; ++j) inv[i * d + j] = (i == j) ? 1.0 : 0.0;
    for (int c = 0; c < d; ++c) {
        int piv = c;
        double best = fabs(lu[c * d + c]);
        for (int r = c + 1; r < d; ++r)
            if (fabs(lu[r * d + c]) > best) { best = fabs(lu[r * d + c]); piv = r; }
        if (piv != c) {
            for (int j = 0; j < d; ++j) {
                double t = lu[c * d + j]; lu[c * d + j] = lu[piv * d + j]; lu[piv * d + j] = t;
                if (inv) { t = inv[c * d + j]; inv[c * d + j] = inv[piv * d + j]; inv[piv * d + j] = t; }
            }
            det = -det;
        }
        const double d0 = lu[c * d + c];
        det = det * d0;
        if (d0 == 0.0) return 0.0;
        for (int r = 0; r < d; ++r) {
            if (r == c) continue;
            if (r < c && !inv) continue;
            const double f = lu[r * d + c] / d0;
            for (int j = 0; j < d; ++j) {
                lu[r * d + j] = lu[r * d + j] - f * lu[c * d + j];
                if (inv) inv[r * d + j] = inv[r * d + j] - f * inv[c * d + j];
            }
        }
    }
    if (inv)
        for (int r = 0; r < d; ++r) {
            const double dd = lu[r * d + r];
            for (int j = 0; j < d; ++j) inv[r * d + j] = inv[r * d + j] / dd;
        }
    return det;
}

/* multivariate-t observe*, runtime.cljc:383-401 */
double orc_mvt_logpdf(const double* y, const double* mu, const double* sigma, double nu, int d,
                      double* work) {
    double* lu = work;
    double* inv = work + (size_t)d * d;
    double* dy = inv + (size_t)d * d;
    const double det = lu_det_inverse(sigma, d, lu, inv);
    for (int i = 0; i < d; ++i) dy[i] = mu[i] - y[i];
    double q = 0.0;                                   /* dy' inv-sigma dy */
    for (int i = 0; i < d; ++i) {
        double s = 0.0;
        for (int j = 0; j < d; ++j) s = s + inv[i * d + j] * dy[j];
        q = q + dy[i] * s;
    }
    return lgamma(0.5 * (nu + d)) -
           (lgamma(0.5 * nu) + 0.5 * d * log(nu) + 0.5 * d * log(M_PI) + 0.5 * log(det) +
            0.5 * (nu + d) * log(1.0 + (1.0 / nu) * q));
}

/* log-mv-gamma-fn, runtime.cljc:479-486 */
double orc_log_mv_gamma(int p, double a) {
    double s = 0.25 * p * (p - 1) * log(M_PI);
    for (int j = 0; j < p; ++j) s = s + lgamma(a - 0.5 * j);
    return s;
}

/* wishart observe*, runtime.cljc:507-509, 536-539.  work: 2 p^2 doubles */
double orc_wishart_logpdf(const double* x, double n, const double* V, int p, double* work) {
    double* lu = work;
    double* inv = work + (size_t)p * p;
    const double detV = lu_det_inverse(V, p, lu, inv);
    const double Z = 0.5 * n * p * log(2.0) + 0.5 * n * log(detV) + orc_log_mv_gamma(p, 0.5 * n);
    double tr = 0.0;                                  /* (trace (mmul inv-V value)) */
    for (int i = 0; i < p; ++i)
        for (int j = 0; j < p; ++j) tr = tr + inv[i * p + j] * x[j * p + i];
    const double detX = lu_det_inverse(x, p, lu, NULL);
    return 0.5 * (n - p - 1.0) * log(detX) - 0.5 * tr - Z;
}

/* ---- samplers (distributional parity only) ----------------------------- */

/* GammaDistribution.sample: Ahrens-Dieter GS (shape<1), Marsaglia-Tsang
 * otherwise; draws indexed by an increasing counter. */
double orc_gamma_sample(double shape, double rate, uint64_t seed, uint64_t slot,
                        uint32_t t, uint32_t stream) {
    double scale = 1.0 / rate;
    uint32_t nu = 0, nn = 0;
    if (shape < 1.0) {
        for (;;) {
            double u = orc_uniform_at(seed, slot, t, stream, nu++);
            double bGS = 1.0 + shape / M_E;
            double p = bGS * u;
            if (p <= 1.0) {
                double x = pow(p, 1.0 / shape);
                double u2 = orc_uniform_at(seed, slot, t, stream, nu++);
                if (u2 > exp(-x)) continue;
                return scale * x;
            } else {
                double x = -log((bGS - p) / shape);
                double u2 = orc_uniform_at(seed, slot, t, stream, nu++);
                if (u2 > pow(x, shape - 1.0)) continue;
                return scale * x;
            }
        }
    }
    double d = shape - 0.333333333333333333;
    double c = 1.0 / (3.0 * sqrt(d));
    for (;;) {
        /* normals here are per-slot (slot*2 so the Box-Muller pair is private) */
        double x = orc_normal_at(seed, slot * 2, t, stream, nn++);
        double v = (1.0 + c * x) * (1.0 + c * x) * (1.0 + c * x);
        if (v <= 0) continue;
        double x2 = x * x;
        double u = orc_uniform_at(seed, slot, t, stream, nu++);
        if (u < 1.0 - 0.0331 * x2 * x2) return scale * d * v;
        if (log(u) < 0.5 * x2 + d * (1.0 - v + log(v))) return scale * d * v;
    }
}

/* BetaDistribution.sample: Cheng (1978) BB / BC */
double orc_beta_sample(double alpha0, double beta0, uint64_t seed, uint64_t slot,
                       uint32_t t, uint32_t stream) {
    uint32_t nu = 0;
    double a = alpha0 < beta0 ? alpha0 : beta0;
    double b = alpha0 < beta0 ? beta0 : alpha0;
    double w;
    if (a > 1.0) {                                   /* BB */
        double alpha = a + b;
        double bet = sqrt((alpha - 2.0) / (2.0 * a * b - alpha));
        double gam = a + 1.0 / bet;
        double r, tt;
        do {
            double u1 = orc_uniform_at(seed, slot, t, stream, nu++);
            double u2 = orc_uniform_at(seed, slot, t, stream, nu++);
            if (u1 <= 0.0) u1 = 1e-300;
            double v = bet * (log(u1) - log1p(-u1));
            w = a * exp(v);
            double z = u1 * u1 * u2;
            r = gam * v - 1.3862944;
            double s = a + r - w;
            if (s + 2.609438 >= 5.0 * z) break;
            tt = log(z);
            if (s >= tt) break;
            if (r + alpha * (log(alpha) - log(b + w)) >= tt) break;
        } while (1);
        if (w > 1.7976931348623157e308) w = 1.7976931348623157e308;
        return (a == alpha0) ? w / (b + w) : b / (b + w);
    }
    /* BC: a = max, b = min */
    double A = b, B = a;
    double alpha = A + B;
    double bet = 1.0 / B;
    double delta = 1.0 + A - B;
    double k1 = delta * (0.0138889 + 0.0416667 * B) / (A * bet - 0.777778);
    double k2 = 0.25 + (0.5 + 0.25 / delta) * B;
    for (;;) {
        double u1 = orc_uniform_at(seed, slot, t, stream, nu++);
        double u2 = orc_uniform_at(seed, slot, t, stream, nu++);
        if (u1 <= 0.0) u1 = 1e-300;
        double y = u1 * u2;
        double z = u1 * y;
        if (u1 < 0.5) {
            if (0.25 * u2 + z - y >= k1) continue;
        } else {
            if (z <= 0.25) {
                double v = bet * (log(u1) - log1p(-u1));
                w = A * exp(v);
                break;
            }
            if (z >= k2) continue;
        }
        double v = bet * (log(u1) - log1p(-u1));
        w = A * exp(v);
        if (alpha * (log(alpha) - log(B + w) + v) - 1.3862944 >= log(z)) break;
    }
    if (w > 1.7976931348623157e308) w = 1.7976931348623157e308;
    return (A == alpha0) ? w / (B + w) : B / (B + w);
}

/* dirichlet sample*, runtime.cljc:298-301: K gamma(alpha_j, 1) draws normalised */
void orc_dirichlet_sample(const double* alpha, int k, uint64_t seed, uint64_t slot,
                          uint32_t t, uint32_t stream, double* out) {
    double tot = 0.0;
    for (int j = 0; j < k; ++j) {
        out[j] = orc_gamma_sample(alpha[j], 1.0, seed, slot * (uint64_t)k + (uint64_t)j, t, stream);
        tot = tot + out[j];
    }
    for (int j = 0; j < k; ++j) out[j] = out[j] / tot;
}

/* ========================================================================
 * lmh.cljc:146-151 accept?  (pimh.cljc:64)
 * ====================================================================== */
int orc_accept(double u_next, double u_prev, double uniform) {
    return (u_prev == -INFINITY) || ((u_next - u_prev) > log(uniform));
}

/* ========================================================================
 * ipmcmc.cljc:85-146
 * ====================================================================== */
void orc_norm_exp(const double* lw, int n, double* ps, double* log_mean) {
    double m = lw[0];
    for (int i = 1; i < n; ++i) if (lw[i] > m) m = lw[i];
    if (m == -INFINITY) {                       /* ipmcmc.cljc:92-95 */
        for (int i = 0; i < n; ++i) ps[i] = 0.0;
        *log_mean = 0.0;
        return;
    }
    double total = 0.0;
    for (int i = 0; i < n; ++i) { ps[i] = exp(lw[i] - m); total = total + ps[i]; }
    for (int i = 0; i < n; ++i) ps[i] = ps[i] / total;
    *log_mean = log(total / (double)n) + m;
}

void orc_ipmcmc_gibbs(const double* log_zs, int m, int p, const double* uniforms,
                      int as_written, int32_t* csmc_idx, double* zeta_sums) {
    int n_smc = m - p;
    int32_t* smc_idx = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n_smc + 1));
    double* lz = (double*)malloc(sizeof(double) * (size_t)(n_smc + 1));
    double* ps = (double*)malloc(sizeof(double) * (size_t)(n_smc + 1));
    for (int i = 0; i < p; ++i) csmc_idx[i] = i;
    for (int i = 0; i < n_smc; ++i) smc_idx[i] = p + i;
    for (int i = 0; i < m; ++i) zeta_sums[i] = 0.0;
    for (int i = 0; i < p; ++i) {
        /* proposal-indices = (conj smc-indices i): NB the slot number i, not
         * (csmc-indices i) — literal, ipmcmc.cljc:121 */
        int np = n_smc + 1;
        for (int j = 0; j < n_smc; ++j) lz[j] = log_zs[smc_idx[j]];
        lz[n_smc] = log_zs[i];
        double lm;
        orc_norm_exp(lz, np, ps, &lm);
        for (int j = 0; j < n_smc; ++j) zeta_sums[smc_idx[j]] += ps[j];
        zeta_sums[i] += ps[n_smc];
        double mx = ps[0];
        for (int j = 1; j < np; ++j) if (ps[j] > mx) mx = ps[j];
        int use_discrete = as_written ? (0.0 > mx) : (mx > 0.0);
        int k;
        if (use_discrete) k = (int)orc_discrete_sample(ps, np, uniforms[i]);
        else {
            k = (int)(uniforms[i] * np);         /* UniformInteger(0, np-1) */
            if (k >= np) k = np - 1;
        }
        if (k != n_smc) {                        /* ipmcmc.cljc:136-146 swap */
            int32_t tmp = smc_idx[k];
            smc_idx[k] = csmc_idx[i];
            csmc_idx[i] = tmp;
        }
    }
    free(smc_idx); free(lz); free(ps);
}

/* ========================================================================
 * State-space sweeps: smc.cljc:35-51 (+ pgibbs.cljc:76-110, ipmcmc.cljc:34-83)
 * specialised to programs of the shape
 *     x0 <- sample; for y_t: x_t <- sample(. | x_{t-1}); observe(. | x_t) y_t
 * with the engine's slot/RNG conventions: physical order sorted,
 * noise keyed by (step, physical slot), resampling uniform keyed by step.
 * ====================================================================== */
static double ssm_obs_logpdf(const orc_ssm* m, const double* x, const double* y,
                             const double* iL, double Z, double* work) {
    switch (m->kind) {
    case ORC_SSM_LGSSM1D: return orc_normal_logpdf(y[0], m->c * x[0] + m->d, m->r);
    case ORC_SSM_SV:      return orc_normal_logpdf(y[0], 0.0, m->beta * exp(x[0] / 2.0));
    case ORC_SSM_HMM: {
        int s = (int)x[0];
        return orc_normal_logpdf(y[0], m->hmm_mu[s], m->hmm_sd[s]);
    }
    case ORC_SSM_LGMVN:   return orc_mvn_logpdf(y, x, iL, Z, m->dim, work);
    }
    return NAN;
}

static void ssm_init(const orc_ssm* m, uint64_t seed, int64_t k, double* x) {
    switch (m->kind) {
    case ORC_SSM_LGSSM1D:
    case ORC_SSM_SV:
        x[0] = m->m0 + m->s0 * orc_normal_at(seed, (uint64_t)k, 0, ORC_STREAM_INIT, 0);
        break;
    case ORC_SSM_HMM:
        x[0] = (double)orc_discrete_sample(m->hmm_init, m->dim,
                    orc_uniform_at(seed, (uint64_t)k, 0, ORC_STREAM_INIT, 0));
        break;
    case ORC_SSM_LGMVN:
        for (int i = 0; i < m->dim; ++i)
            x[i] = m->s0 * orc_normal_at(seed, (uint64_t)k, 0, ORC_STREAM_INIT, (uint32_t)i);
        break;
    }
}

static void ssm_transition(const orc_ssm* m, uint64_t seed, int64_t k, uint32_t t,
                           const double* xp, double* x, double* work) {
    switch (m->kind) {
    case ORC_SSM_LGSSM1D:
    case ORC_SSM_SV:
        x[0] = m->a * xp[0] + m->b
             + m->q * orc_normal_at(seed, (uint64_t)k, t, ORC_STREAM_TRANSITION, 0);
        break;
    case ORC_SSM_HMM: {
        int s = (int)xp[0];
        x[0] = (double)orc_discrete_sample(m->hmm_trans + s * m->dim, m->dim,
                    orc_uniform_at(seed, (uint64_t)k, t, ORC_STREAM_TRANSITION, 0));
        break;
    }
    case ORC_SSM_LGMVN: {
        int d = m->dim;
        for (int i = 0; i < d; ++i)
            work[i] = orc_normal_at(seed, (uint64_t)k, t, ORC_STREAM_TRANSITION, (uint32_t)i);
        for (int r = 0; r < d; ++r) {
            double s = 0.0;
            for (int c = 0; c < d; ++c) s = s + m->A[r * d + c] * xp[c];
            double n = 0.0;
            for (int c = 0; c <= r; ++c) n = n + m->Lq[r * d + c] * work[c];
            x[r] = s + n;                     /* mean + L z, runtime.cljc:360-361 */
        }
        break;
    }
    }
}

/* log f(x_next | x): the `observe*` of the transition `sample` when PGAS rescores
 * the retained value against another particle's state (pgas.cljc:66-82).
 * Scalar-state models. */
static double ssm_transition_logpdf(const orc_ssm* m, const double* x, const double* x_next) {
    switch (m->kind) {
    case ORC_SSM_LGSSM1D:
    case ORC_SSM_SV:
        return orc_normal_logpdf(x_next[0], m->a * x[0] + m->b, m->q);
    case ORC_SSM_HMM: {
        const int s = (int)x[0], c = (int)x_next[0];
        const double* row = m->hmm_trans + s * m->dim;
        double tot = 0.0;
        for (int i = 0; i < m->dim; ++i) tot = tot + row[i];
        return (c >= 0 && c < m->dim) ? log(row[c] / tot) : -INFINITY;   /* runtime.cljc:288-290 */
    }
    default: return NAN;
    }
}

/* Ancestor of the retained particle at a conditional observe (resample-retained,
 * pgas.cljc:107-131): every particle -- the retained one included -- is rescored
 * against the retained random database, i.e. against the retained path's next
 * state; the terms of the later states and observes are the same number for
 * every particle and are left out (the reference adds them before subtracting the
 * max).  Then one `(sample* (discrete weights))`: x = u * total, first k with
 * x < w_0 + ... + w_k (runtime.cljc:281-287). */
static int64_t pgas_ancestor(const orc_ssm* m, const double* x, const double* lw, int64_t n,
                             const double* next_retained, double u, double* v) {
    double mx = -INFINITY;
    for (int64_t k = 0; k < n; ++k) {
        v[k] = lw[k] + (next_retained ? ssm_transition_logpdf(m, x + k, next_retained) : 0.0);
        if (v[k] > mx) mx = v[k];
    }
    double total = 0.0;
    for (int64_t k = 0; k < n; ++k) {
        v[k] = (v[k] == mx) ? 1.0 : (!(v[k] > -INFINITY) ? 0.0 : exp(v[k] - mx));
        total = total + v[k];
    }
    const double xr = u * total;
    double acc = 0.0;
    for (int64_t k = 0; k < n; ++k) {
        acc = acc + v[k];
        if (xr < acc) return k;
    }
    return n - 1;
}

int orc_smc_sweep(const orc_ssm* model, const double* y, const double* retained,
                  const orc_sweep_opts* o, orc_sweep_out* out) {
    const int64_t n = o->n;
    const int T = o->T;
    const int d = (model->kind == ORC_SSM_LGMVN) ? model->dim : 1;
    const int od = d;                       /* observation dimension */
    double* xa = (double*)malloc(sizeof(double) * (size_t)n * d);
    double* xb = (double*)malloc(sizeof(double) * (size_t)n * d);
    double* lw = (double*)malloc(sizeof(double) * (size_t)n);
    double* w = (double*)malloc(sizeof(double) * (size_t)n);
    int32_t* emission = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
    int32_t* lap = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
    int32_t* anc = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
    double *L = NULL, *iL = NULL, Z = 0.0;
    int rc = 0;
    int nthreads = o->threads > 1 ? o->threads : 1;
    (void)nthreads;
    if (model->kind == ORC_SSM_LGMVN) {
        L = (double*)malloc(sizeof(double) * (size_t)d * d);
        iL = (double*)malloc(sizeof(double) * (size_t)d * d);
        rc = orc_mvn_prepare(model->R, d, L, iL, &Z);
        if (rc) goto done;
    }
    /* initial particles: (repeatedly N #(exec ...)), smc.cljc:37-39; the
     * retained particle is first (pgibbs.cljc:78-81) */
    #pragma omp parallel for num_threads(nthreads) schedule(static) if (nthreads > 1)
    for (int64_t k = 0; k < n; ++k) ssm_init(model, o->seed, k, xa + k * d);
    if (o->csmc && retained) memcpy(xa, retained, sizeof(double) * (size_t)d);
    if (o->keep_history && out->x_hist) memcpy(out->x_hist, xa, sizeof(double) * (size_t)n * d);

    double base = 0.0;       /* common accumulated log-weight of all particles */
    double log_z = 0.0;
    int carry = 0;           /* adaptive resampling: the last observe kept the weighted particles */
    for (int t = 1; t <= T; ++t) {
        const double* yt = y + (size_t)(t - 1) * od;
        /* run every particle to the next observe: sample* then observe*
         * (inference.cljc:46-47, smc.cljc:23-27) */
        #pragma omp parallel num_threads(nthreads) if (nthreads > 1)
        {
            double* work = (double*)malloc(sizeof(double) * (size_t)(2 * d + 2));
            #pragma omp for schedule(static)
            for (int64_t k = 0; k < n; ++k) {
                const int64_t par = (t == 1) ? k : anc[k];
                double* xk = xb + k * d;
                if (o->csmc && retained && k == 0)
                    memcpy(xk, retained + (size_t)t * d, sizeof(double) * (size_t)d);
                else
                    ssm_transition(model, o->seed, k, (uint32_t)t, xa + par * d, xk, work);
                lw[k] = (carry ? lw[k] : base) + ssm_obs_logpdf(model, xk, yt, iL, Z, work);  /* add-log-weight */
            }
            free(work);
        }
        if (o->keep_history && out->x_hist)
            memcpy(out->x_hist + (size_t)t * n * d, xb, sizeof(double) * (size_t)n * d);

        double mx;
        rc = orc_particle_weights(lw, n, w, &mx);
        if (rc) goto done;
        double ess_t = 0.0;
        if (out->ess || o->ess_threshold > 0.0) {
            double s = 0.0, s2 = 0.0;
            for (int64_t k = 0; k < n; ++k) { s = s + w[k]; s2 = s2 + w[k] * w[k]; }
            ess_t = s * s / s2;
            if (out->ess) out->ess[t - 1] = ess_t;
        }
        if (out->filt_mean) {
            for (int i = 0; i < d; ++i) {
                double s = 0.0, sx = 0.0;
                for (int64_t k = 0; k < n; ++k) { s = s + w[k]; sx = sx + w[k] * xb[k * d + i]; }
                out->filt_mean[(size_t)(t - 1) * d + i] = sx / s;
            }
        }
        if (o->ess_threshold > 0.0 && ess_t >= o->ess_threshold * (double)n) {
            /* no resample at this observe: every particle is its own ancestor and keeps
             * its weight; the log Z increment is deferred to the next resample */
            for (int64_t k = 0; k < n; ++k) anc[k] = (int32_t)k;
            if (out->log_mean_w) out->log_mean_w[t - 1] = 0.0;
            if (o->keep_history && out->anc_hist)
                memcpy(out->anc_hist + (size_t)(t - 1) * n, anc, sizeof(int32_t) * (size_t)n);
            carry = 1;
            double* tmp0 = xa; xa = xb; xb = tmp0;
            continue;
        }
        carry = 0;
        const int cond = o->csmc && retained && (t <= o->n_cond_steps);
        const int64_t n_new = cond ? n - 1 : n;     /* pgibbs.cljc:92-99 */
        double total, step;
        const double u = orc_uniform_at(o->seed, 0, (uint32_t)t, ORC_STREAM_RESAMPLE, 0);
        rc = orc_resample_weights(w, n, n_new, u, NULL, emission, lap, &total, &step);
        if (rc) goto done;
        const double new_lw = (mx > -INFINITY && mx < INFINITY) ? (log(step) + mx) : -INFINITY;
        if (out->log_mean_w) out->log_mean_w[t - 1] = new_lw - base;
        log_z = log_z + (new_lw - base);
        if (cond) {
            anc[0] = 0;                               /* retained back at position 0 */
            orc_emission_to_sorted(emission, lap, n_new, anc + 1, NULL);
            if (o->ancestor_sampling) {
                if (model->kind == ORC_SSM_LGMVN) { rc = -1; goto done; }
                const double ua = orc_uniform_at(o->seed, 0, (uint32_t)t, ORC_STREAM_AUX, 0);
                anc[0] = (int32_t)pgas_ancestor(model, xb, lw, n,
                                                (t < T) ? retained + (size_t)(t + 1) : NULL, ua, w);
            }
        } else {
            orc_emission_to_sorted(emission, lap, n_new, anc, NULL);
        }
        if (o->keep_history && out->anc_hist)
            memcpy(out->anc_hist + (size_t)(t - 1) * n, anc, sizeof(int32_t) * (size_t)n);
        base = o->reset_weights ? 0.0 : new_lw;       /* pgibbs.cljc:88 */
        double* tmp = xa; xa = xb; xb = tmp;
    }
    out->log_z = o->reset_weights ? log_z : base;
    if (carry) {
        /* the sweep ended on weighted particles: log mean exp of what they carry */
        double mx = -INFINITY, sum = 0.0;
        for (int64_t k = 0; k < n; ++k) if (lw[k] > mx) mx = lw[k];
        for (int64_t k = 0; k < n; ++k) sum = sum + ((lw[k] == mx) ? 1.0 : exp(lw[k] - mx));
        const double L = (mx > -INFINITY && mx < INFINITY) ? log(sum / (double)n) + mx : mx;
        out->log_z = o->reset_weights ? log_z + L : L;
    }
    if (out->lw_final)
        for (int64_t k = 0; k < n; ++k)
            out->lw_final[k] = carry ? lw[k] : (o->reset_weights ? 0.0 : base);
    if (out->x_final) {
        for (int64_t k = 0; k < n; ++k) {
            const int64_t par = (T == 0) ? k : anc[k];
            memcpy(out->x_final + k * d, xa + par * d, sizeof(double) * (size_t)d);
        }
    }
done:
    free(xa); free(xb); free(lw); free(w); free(emission); free(lap); free(anc);
    free(L); free(iL);
    return rc;
}

/* `infer :importance` (importance.cljc:16-22) for a state-space program: every
 * sample runs the whole program on its own; the default `observe` checkpoint
 * adds the log-density to the sample's log-weight and continues
 * (inference.cljc:42-45); nothing is ever resampled.  Sample k uses the same
 * counter-based draws as particle k of a sweep.  Scalar-state models. */
int orc_importance_sweep(const orc_ssm* model, const double* y, int32_t T, int64_t n,
                         uint64_t seed, double* x_final, double* lw_final) {
    if (model->kind == ORC_SSM_LGMVN) return -1;
    double work[4];
    for (int64_t k = 0; k < n; ++k) {
        double x, xn, lw = 0.0;                          /* initial-state: log-weight 0 */
        ssm_init(model, seed, k, &x);
        for (int t = 1; t <= T; ++t) {
            ssm_transition(model, seed, k, (uint32_t)t, &x, &xn, work);
            lw = lw + ssm_obs_logpdf(model, &xn, y + (t - 1), NULL, 0.0, work);
            x = xn;
        }
        x_final[k] = x;
        lw_final[k] = lw;
    }
    return 0;
}

void orc_trace_lineage(const orc_sweep_out* out, int64_t n, int32_t T, int32_t dim,
                       int64_t slot, double* traj) {
    int64_t k = slot;
    for (int t = T; t >= 1; --t) {
        k = out->anc_hist[(size_t)(t - 1) * n + k];         /* slot in x_hist[t] */
        memcpy(traj + (size_t)t * dim, out->x_hist + ((size_t)t * n + k) * dim,
               sizeof(double) * (size_t)dim);
    }
    memcpy(traj, out->x_hist + (size_t)k * dim, sizeof(double) * (size_t)dim);
    /* NB: x_hist[1][k] descends from x_hist[0][k] (identity ancestry at t=1) */
}

/* ========================================================================
 * Exact truths (not from the reference: textbook filters)
 * ====================================================================== */
double orc_kalman1d_logz(const orc_ssm* m, const double* y, int T, double* filt_mean) {
    double mean = m->m0, P = m->s0 * m->s0, logz = 0.0;
    for (int t = 0; t < T; ++t) {
        double mp = m->a * mean + m->b;
        double Pp = m->a * m->a * P + m->q * m->q;
        double v = y[t] - (m->c * mp + m->d);
        double S = m->c * m->c * Pp + m->r * m->r;
        logz += -0.5 * (log(2.0 * M_PI * S) + v * v / S);
        double K = Pp * m->c / S;
        mean = mp + K * v;
        P = (1.0 - K * m->c) * Pp;
        if (filt_mean) filt_mean[t] = mean;
    }
    return logz;
}

/* dense Kalman filter, H = I: x0~N(0,s0^2 I); x' = A x + Lq z; y ~ N(x', R) */
int orc_kalman_mvn_logz(const orc_ssm* m, const double* y, int T, double* logz_out,
                        double* filt_mean) {
    const int d = m->dim;
    const size_t dd = (size_t)d * d;
    double* P = (double*)calloc(dd, sizeof(double));
    double* Pp = (double*)calloc(dd, sizeof(double));
    double* AP = (double*)calloc(dd, sizeof(double));
    double* Q = (double*)calloc(dd, sizeof(double));
    double* S = (double*)calloc(dd, sizeof(double));
    double* Ls = (double*)calloc(dd, sizeof(double));
    double* iLs = (double*)calloc(dd, sizeof(double));
    double* K = (double*)calloc(dd, sizeof(double));
    double* Sinv = (double*)calloc(dd, sizeof(double));
    double* mean = (double*)calloc((size_t)d, sizeof(double));
    double* mp = (double*)calloc((size_t)d, sizeof(double));
    double* v = (double*)calloc((size_t)d, sizeof(double));
    double* work = (double*)calloc((size_t)d * 2, sizeof(double));
    double logz = 0.0;
    int rc = 0;
    for (int i = 0; i < d; ++i) P[i * d + i] = m->s0 * m->s0;
    for (int r = 0; r < d; ++r) for (int c = 0; c < d; ++c) {
        double s = 0.0;
        for (int k = 0; k < d; ++k) s += m->Lq[r * d + k] * m->Lq[c * d + k];
        Q[r * d + c] = s;
    }
    for (int t = 0; t < T; ++t) {
        for (int r = 0; r < d; ++r) {
            double s = 0.0;
            for (int c = 0; c < d; ++c) s += m->A[r * d + c] * mean[c];
            mp[r] = s;
        }
        for (int r = 0; r < d; ++r) for (int c = 0; c < d; ++c) {
            double s = 0.0;
            for (int k = 0; k < d; ++k) s += m->A[r * d + k] * P[k * d + c];
            AP[r * d + c] = s;
        }
        for (int r = 0; r < d; ++r) for (int c = 0; c < d; ++c) {
            double s = 0.0;
            for (int k = 0; k < d; ++k) s += AP[r * d + k] * m->A[c * d + k];
            Pp[r * d + c] = s + Q[r * d + c];
        }
        for (size_t i = 0; i < dd; ++i) S[i] = Pp[i] + m->R[i];
        /* symmetrise to kill round-off drift */
        for (int r = 0; r < d; ++r) for (int c = r + 1; c < d; ++c) {
            double a = 0.5 * (S[r * d + c] + S[c * d + r]);
            S[r * d + c] = a; S[c * d + r] = a;
        }
        double Z;
        rc = orc_mvn_prepare(S, d, Ls, iLs, &Z);
        if (rc) break;
        logz += orc_mvn_logpdf(y + (size_t)t * d, mp, iLs, Z, d, work);
        /* Sinv = iLs^T iLs ; K = Pp Sinv */
        for (int r = 0; r < d; ++r) for (int c = 0; c < d; ++c) {
            double s = 0.0;
            for (int k = 0; k < d; ++k) s += iLs[k * d + r] * iLs[k * d + c];
            Sinv[r * d + c] = s;
        }
        for (int r = 0; r < d; ++r) for (int c = 0; c < d; ++c) {
            double s = 0.0;
            for (int k = 0; k < d; ++k) s += Pp[r * d + k] * Sinv[k * d + c];
            K[r * d + c] = s;
        }
        for (int i = 0; i < d; ++i) v[i] = y[(size_t)t * d + i] - mp[i];
        for (int r = 0; r < d; ++r) {
            double s = 0.0;
            for (int c = 0; c < d; ++c) s += K[r * d + c] * v[c];
            mean[r] = mp[r] + s;
        }
        /* P = Pp - K Pp */
        for (int r = 0; r < d; ++r) for (int c = 0; c < d; ++c) {
            double s = 0.0;
            for (int k = 0; k < d; ++k) s += K[r * d + k] * Pp[k * d + c];
            P[r * d + c] = Pp[r * d + c] - s;
        }
        for (int r = 0; r < d; ++r) for (int c = r + 1; c < d; ++c) {
            double a = 0.5 * (P[r * d + c] + P[c * d + r]);
            P[r * d + c] = a; P[c * d + r] = a;
        }
        if (filt_mean) memcpy(filt_mean + (size_t)t * d, mean, sizeof(double) * (size_t)d);
    }
    *logz_out = logz;
    free(P); free(Pp); free(AP); free(Q); free(S); free(Ls); free(iLs); free(K);
    free(Sinv); free(mean); free(mp); free(v); free(work);
    return rc;
}

static void hmm_rows(const orc_ssm* m, double* init, double* trans) {
    const int K = m->dim;
    double s = 0.0;
    for (int i = 0; i < K; ++i) s += m->hmm_init[i];
    for (int i = 0; i < K; ++i) init[i] = m->hmm_init[i] / s;
    for (int r = 0; r < K; ++r) {
        s = 0.0;
        for (int c = 0; c < K; ++c) s += m->hmm_trans[r * K + c];
        for (int c = 0; c < K; ++c) trans[r * K + c] = m->hmm_trans[r * K + c] / s;
    }
}

double orc_hmm_forward_logz(const orc_ssm* m, const double* y, int T) {
    const int K = m->dim;
    double init[16], trans[256], alpha[16], nxt[16];
    hmm_rows(m, init, trans);
    memcpy(alpha, init, sizeof(double) * (size_t)K);
    double logz = 0.0;
    for (int t = 0; t < T; ++t) {
        double tot = 0.0;
        for (int c = 0; c < K; ++c) {
            double s = 0.0;
            for (int r = 0; r < K; ++r) s += alpha[r] * trans[r * K + c];
            nxt[c] = s * exp(orc_normal_logpdf(y[t], m->hmm_mu[c], m->hmm_sd[c]));
            tot += nxt[c];
        }
        for (int c = 0; c < K; ++c) alpha[c] = nxt[c] / tot;
        logz += log(tot);
    }
    return logz;
}

void orc_hmm_smooth(const orc_ssm* m, const double* y, int T, double* post) {
    const int K = m->dim;
    double init[16], trans[256];
    hmm_rows(m, init, trans);
    double* alpha = (double*)malloc(sizeof(double) * (size_t)(T + 1) * K);
    double* beta = (double*)malloc(sizeof(double) * (size_t)(T + 1) * K);
    memcpy(alpha, init, sizeof(double) * (size_t)K);
    for (int t = 1; t <= T; ++t) {
        double tot = 0.0;
        for (int c = 0; c < K; ++c) {
            double s = 0.0;
            for (int r = 0; r < K; ++r) s += alpha[(t - 1) * K + r] * trans[r * K + c];
            alpha[t * K + c] = s * exp(orc_normal_logpdf(y[t - 1], m->hmm_mu[c], m->hmm_sd[c]));
            tot += alpha[t * K + c];
        }
        for (int c = 0; c < K; ++c) alpha[t * K + c] /= tot;
    }
    for (int c = 0; c < K; ++c) beta[T * K + c] = 1.0;
    for (int t = T - 1; t >= 0; --t) {
        double tot = 0.0;
        for (int r = 0; r < K; ++r) {
            double s = 0.0;
            for (int c = 0; c < K; ++c)
                s += trans[r * K + c] * exp(orc_normal_logpdf(y[t], m->hmm_mu[c], m->hmm_sd[c]))
                     * beta[(t + 1) * K + c];
            beta[t * K + r] = s;
            tot += s;
        }
        for (int r = 0; r < K; ++r) beta[t * K + r] /= tot;
    }
    for (int t = 0; t <= T; ++t) {
        double tot = 0.0;
        for (int c = 0; c < K; ++c) { post[t * K + c] = alpha[t * K + c] * beta[t * K + c]; tot += post[t * K + c]; }
        for (int c = 0; c < K; ++c) post[t * K + c] /= tot;
    }
    free(alpha); free(beta);
}
