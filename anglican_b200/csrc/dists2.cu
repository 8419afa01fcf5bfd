// dists2.cu — the remaining scalar-parameter families of anglican.runtime behind
// ONE pair of entry points (apop_sample_dist / apop_observe_dist, family id +
// up to three parameters), mirroring the `distribution` protocol's dispatch
// (runtime.cljc:119-124):
//   exponential (runtime.cljc:308-313)   laplace (411-416)     poisson (418-423)
//   bernoulli / flip (241-249, 315-323)  student-t (425-444)   student-t-loc-scale (446-448)
//   chi-squared (333-340)                uniform-discrete (464-475)   binomial (258-263)
// Log-densities restate the commons-math 3.6.1 formulas the reference calls
// (`from-apache`, runtime.cljc:207-220); samplers are counter-based (Philox)
// and only distributionally comparable with the reference's Well19937c streams.
#include <math.h>

#include "apop_internal.cuh"
#include "gamma_draw.cuh"
#include "rng.cuh"

namespace apop {

struct DArg2 {
    const double* p;
    double s;
    __device__ __forceinline__ double at(int64_t k) const { return p ? p[k] : s; }
};

constexpr double kLogPi = 1.14472988584940017414;

__device__ __forceinline__ bool is_int(double x) { return x == floor(x) && fabs(x) < 2147483648.0; }

// ---- log-densities ---------------------------------------------------------
template <int FAM>
__device__ __forceinline__ double fam_logp(double x, double a, double b, double c) {
    if (FAM == APOP_DIST_EXPONENTIAL) {
        // ExponentialDistribution(mean = 1/rate).logDensity
        const double mean = 1.0 / a;
        return (x < 0) ? -INFINITY : -x / mean - log(mean);
    } else if (FAM == APOP_DIST_LAPLACE) {
        // LaplaceDistribution(mu, beta): log(density(x)), density = exp(-|x-mu|/beta)/(2 beta)
        return log(exp(-fabs(x - a) / b) / (2.0 * b));
    } else if (FAM == APOP_DIST_POISSON) {
        // PoissonDistribution.logProbability (commons evaluates the same quantity by a
        // saddle-point expansion; direct form, see DESIGN.md "known deviations")
        if (x < 0 || !is_int(x) || x == 2147483647.0) return -INFINITY;
        if (x == 0) return -a;
        return x * log(a) - a - lgamma(x + 1.0);
    } else if (FAM == APOP_DIST_BERNOULLI) {
        // runtime.cljc:245-249 / 319-323: (log (case value 1 p 0 (- 1. p) 0.))
        return (x == 1.0) ? log(a) : (x == 0.0) ? log(1.0 - a) : -INFINITY;
    } else if (FAM == APOP_DIST_STUDENT_T || FAM == APOP_DIST_STUDENT_T_LOC_SCALE) {
        // TDistribution(nu).logDensity; loc-scale: runtime.cljc:446-448
        const double n = a;
        const double v = (FAM == APOP_DIST_STUDENT_T) ? x : (x - b) / c;
        const double h = (n + 1.0) / 2.0;
        const double lp = lgamma(h) - 0.5 * (kLogPi + log(n)) - lgamma(n / 2.0) -
                          h * log(1.0 + v * v / n);
        return (FAM == APOP_DIST_STUDENT_T) ? lp : lp - log(c);
    } else if (FAM == APOP_DIST_CHI_SQUARED) {
        // (gamma (* nu 0.5) 0.5), runtime.cljc:333-340
        if (x < 0) return -INFINITY;
        const double shape = a * 0.5, rate = 0.5, y = x * rate;
        return -lgamma(shape) + log(rate) + (shape - 1.0) * log(y) - y;
    } else if (FAM == APOP_DIST_UNIFORM_DISCRETE) {
        // UniformIntegerDistribution((int min), (dec (int max))): log(1 / (upper - lower + 1))
        const double lo = trunc(a), hi = trunc(b) - 1.0;
        if (!is_int(x) || x < lo || x > hi) return -INFINITY;
        return log(1.0 / (hi - lo + 1.0));
    } else {   // APOP_DIST_BINOMIAL(n, p)
        const double n = trunc(a), p = b;
        if (!is_int(x) || x < 0 || x > n) return -INFINITY;
        if (n == 0) return 0.0;
        double r = lgamma(n + 1.0) - lgamma(x + 1.0) - lgamma(n - x + 1.0);
        if (x > 0) r += x * log(p);
        if (x < n) r += (n - x) * log1p(-p);
        return r;
    }
}

// ---- samplers ---------------------------------------------------------------
__device__ __forceinline__ double u_open(uint64_t seed, uint64_t slot, uint32_t t, uint32_t d) {
    U4 o = philox4x32_10((uint32_t)slot, (uint32_t)(slot >> 32), t, ctr3(STREAM_USER, 1, d),
                         (uint32_t)seed, (uint32_t)(seed >> 32));
    return u01_open(o.x, o.y);                                   // strictly inside (0, 1)
}

__device__ double poisson_draw(double lam, uint64_t seed, uint64_t slot, uint32_t t, uint32_t d0) {
    if (!(lam > 0)) return 0.0;
    if (lam < 30.0) {                                            // Knuth: product of uniforms
        const double L = exp(-lam);
        double p = 1.0;
        int k = 0;
        for (uint32_t i = 0; i < 250; ++i) {
            p *= uniform_at(seed, slot, t, STREAM_USER, d0 + i);
            if (!(p > L)) break;
            ++k;
        }
        return (double)k;
    }
    // PTRS, W. Hormann, "The transformed rejection method for generating Poisson
    // random variables" (1993)
    const double slam = sqrt(lam), loglam = log(lam);
    const double b = 0.931 + 2.53 * slam, a = -0.059 + 0.02483 * b;
    const double invalpha = 1.1239 + 1.1328 / (b - 3.4), vr = 0.9277 - 3.6224 / (b - 2.0);
    for (uint32_t it = 0; it < 120; ++it) {
        const double U = uniform_at(seed, slot, t, STREAM_USER, d0 + 2 * it) - 0.5;
        const double V = u_open(seed, slot, t, d0 + 2 * it + 1);
        const double us = 0.5 - fabs(U);
        const double k = floor((2.0 * a / us + b) * U + lam + 0.43);
        if (us >= 0.07 && V <= vr) return k;
        if (k < 0 || (us < 0.013 && V > us)) continue;
        if (log(V) + log(invalpha) - log(a / (us * us) + b) <= -lam + k * loglam - lgamma(k + 1.0))
            return k;
    }
    return floor(lam);
}

template <int FAM>
__device__ __forceinline__ double fam_draw(double a, double b, double c, uint64_t seed,
                                           uint64_t slot, uint32_t t, uint32_t draw) {
    const uint32_t d0 = draw * 512u;
    if (FAM == APOP_DIST_EXPONENTIAL) {
        return -(1.0 / a) * log(u_open(seed, slot, t, d0));       // inversion
    } else if (FAM == APOP_DIST_LAPLACE) {
        // LaplaceDistribution.inverseCumulativeProbability of a uniform
        const double p = u_open(seed, slot, t, d0);
        const double x = (p > 0.5) ? -log(2.0 - 2.0 * p) : log(2.0 * p);
        return a + b * x;
    } else if (FAM == APOP_DIST_POISSON) {
        return poisson_draw(a, seed, slot, t, d0);
    } else if (FAM == APOP_DIST_BERNOULLI) {
        return (uniform_at(seed, slot, t, STREAM_USER, d0) < a) ? 1.0 : 0.0;   // runtime.cljc:244
    } else if (FAM == APOP_DIST_STUDENT_T || FAM == APOP_DIST_STUDENT_T_LOC_SCALE) {
        // z sqrt(nu / chi2(nu)) (the construction of runtime.cljc:433-437)
        const double z = normal_at(seed, slot * 2, t, STREAM_USER, d0);
        const double y = gamma_draw(a * 0.5, 0.5, seed, slot, t, STREAM_USER, d0 + 16u);
        const double v = z * sqrt(a) / sqrt(y);
        return (FAM == APOP_DIST_STUDENT_T) ? v : v * c + b;
    } else if (FAM == APOP_DIST_CHI_SQUARED) {
        return gamma_draw(a * 0.5, 0.5, seed, slot, t, STREAM_USER, d0);
    } else if (FAM == APOP_DIST_UNIFORM_DISCRETE) {
        const double lo = trunc(a), cnt = trunc(b) - lo;
        double k = floor(uniform_at(seed, slot, t, STREAM_USER, d0) * cnt);
        if (k >= cnt) k = cnt - 1.0;
        return lo + k;
    } else {   // binomial: count of successes (n <= 400 draws per call)
        const int n = (int)a;
        int k = 0;
        for (int i = 0; i < n; ++i) k += (uniform_at(seed, slot, t, STREAM_USER, d0 + i) < b);
        return (double)k;
    }
}

template <int FAM>
__global__ void k_sample_fam(DArg2 a, DArg2 b, DArg2 c, double* __restrict__ out, int64_t n,
                             int64_t g0, uint64_t seed, uint32_t t, uint32_t draw) {
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n;
         k += (int64_t)gridDim.x * blockDim.x)
        out[k] = fam_draw<FAM>(a.at(k), b.at(k), c.at(k), seed, (uint64_t)(g0 + k), t, draw);
}
template <int FAM>
__global__ void k_observe_fam(DArg2 a, DArg2 b, DArg2 c, DArg2 val, double* __restrict__ lw,
                              int64_t n) {
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n;
         k += (int64_t)gridDim.x * blockDim.x)
        lw[k] += fam_logp<FAM>(val.at(k), a.at(k), b.at(k), c.at(k));
}

static int grid2(const apop_ctx* c, int64_t n) {
    int64_t b = (n + 255) / 256;
    const int64_t cap = (int64_t)c->sm_count * 32;
    return (int)(b < cap ? (b > 0 ? b : 1) : cap);
}

static int resolve2(apop_ctx* c, const apop_arg& a, DArg2* out) {
    out->s = a.scalar;
    out->p = nullptr;
    if (a.col >= 0) {
        Column* col = get_col(c, a.col, APOP_F64, 1);
        if (!col) return APOP_ERR_INVALID;
        out->p = (const double*)col->ptr;
    }
    return APOP_OK;
}

} // namespace apop

using namespace apop;

#define FAM_SWITCH(fam, CALL)                                                         \
    switch (fam) {                                                                    \
    case APOP_DIST_EXPONENTIAL: CALL(APOP_DIST_EXPONENTIAL); break;                   \
    case APOP_DIST_LAPLACE: CALL(APOP_DIST_LAPLACE); break;                           \
    case APOP_DIST_POISSON: CALL(APOP_DIST_POISSON); break;                           \
    case APOP_DIST_BERNOULLI: CALL(APOP_DIST_BERNOULLI); break;                       \
    case APOP_DIST_STUDENT_T: CALL(APOP_DIST_STUDENT_T); break;                       \
    case APOP_DIST_STUDENT_T_LOC_SCALE: CALL(APOP_DIST_STUDENT_T_LOC_SCALE); break;   \
    case APOP_DIST_CHI_SQUARED: CALL(APOP_DIST_CHI_SQUARED); break;                   \
    case APOP_DIST_UNIFORM_DISCRETE: CALL(APOP_DIST_UNIFORM_DISCRETE); break;         \
    case APOP_DIST_BINOMIAL: CALL(APOP_DIST_BINOMIAL); break;                         \
    default: return set_err(c, APOP_ERR_UNSUPPORTED, "distribution family %d", fam);  \
    }

extern "C" {

int apop_sample_dist(apop_ctx* c, int32_t family, const apop_arg* params, uint32_t draw,
                     apop_col out) {
    if (!c) return APOP_ERR_INVALID;
    if (!params) return set_err(c, APOP_ERR_INVALID, "sample_dist: params");
    const apop_arg p0 = params[0];
    DArg2 a, b, d;
    APOP_TRY(resolve2(c, params[0], &a)); APOP_TRY(resolve2(c, params[1], &b));
    APOP_TRY(resolve2(c, params[2], &d));
    Column* o = get_col(c, out, APOP_F64, 1);
    if (!o) return APOP_ERR_INVALID;
    if (family == APOP_DIST_BINOMIAL && (p0.col >= 0 || !(p0.scalar >= 0 && p0.scalar <= 400)))
        return set_err(c, APOP_ERR_UNSUPPORTED, "binomial sampler: scalar n <= 400");
    cudaSetDevice(c->device);
#define CALL(F) k_sample_fam<F><<<grid2(c, c->n), 256, 0, c->stream>>>(a, b, d, (double*)o->ptr, c->n, \
                                                                     c->g0, c->seed, c->step, draw)
    FAM_SWITCH(family, CALL)
#undef CALL
    c->launches += 1;
    APOP_CUDA(c, cudaGetLastError());
    return APOP_OK;
}

int apop_observe_dist(apop_ctx* c, int32_t family, const apop_arg* params, apop_arg value,
                      apop_col logw) {
    if (!c) return APOP_ERR_INVALID;
    if (!params) return set_err(c, APOP_ERR_INVALID, "observe_dist: params");
    DArg2 a, b, d, v;
    APOP_TRY(resolve2(c, params[0], &a)); APOP_TRY(resolve2(c, params[1], &b));
    APOP_TRY(resolve2(c, params[2], &d));
    APOP_TRY(resolve2(c, value, &v));
    Column* l = get_col(c, logw, APOP_F64, 1);
    if (!l) return APOP_ERR_INVALID;
    cudaSetDevice(c->device);
#define CALL(F) k_observe_fam<F><<<grid2(c, c->n), 256, 0, c->stream>>>(a, b, d, v, (double*)l->ptr, c->n)
    FAM_SWITCH(family, CALL)
#undef CALL
    c->launches += 1;
    APOP_CUDA(c, cudaGetLastError());
    return APOP_OK;
}

} // extern "C"
