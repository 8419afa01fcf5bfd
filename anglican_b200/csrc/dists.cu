// dists.cu — per-particle `sample*` / `observe*` of the anglican.runtime
// distributions (reference src/anglican/runtime.cljc:119-124 protocol; normal
// 403-409, gamma 325-330, beta 251-256, uniform-continuous 450-453, discrete
// 277-290, categorical 266-274, dirichlet 292-306, mvn 349-368) as streaming
// kernels over device-resident columns.  `observe` adds the log-density into
// the log-weight column (smc.cljc:23-27, state.cljc:27-31).
//
// normal/gamma/beta/uniform are commons-math3 3.6.1 objects in the reference
// (`from-apache`, runtime.cljc:207-220); the log-densities below are the
// closed forms those classes evaluate (lgamma-based; see DESIGN.md for the one
// regime where commons' Lanczos-prefactor branch loses digits).
#include <math.h>

#include "apop_internal.cuh"
#include "rng.cuh"
#include "gamma_draw.cuh"

namespace apop {

struct DArg {
    const double* p;
    double s;
    __device__ __forceinline__ double at(int64_t k) const { return p ? p[k] : s; }
};

constexpr double kHalfLog2Pi = 0.91893853320467274178;

__device__ __forceinline__ double normal_logpdf(double x, double mean, double sd) {
    double x1 = (x - mean) / sd;
    return -0.5 * x1 * x1 - (log(sd) + kHalfLog2Pi);
}
__device__ __forceinline__ double gamma_logpdf(double x, double shape, double rate) {
    if (x < 0) return -INFINITY;
    double y = x * rate;
    return -lgamma(shape) + log(rate) + (shape - 1.0) * log(y) - y;
}
__device__ __forceinline__ double beta_logpdf(double x, double a, double b) {
    if (x < 0 || x > 1) return -INFINITY;
    if (x == 0) return (a < 1) ? NAN : -INFINITY;
    if (x == 1) return (b < 1) ? NAN : -INFINITY;
    double z = lgamma(a) + lgamma(b) - lgamma(a + b);
    return (a - 1.0) * log(x) + (b - 1.0) * log1p(-x) - z;
}
__device__ __forceinline__ double uniform_logpdf(double x, double lo, double hi) {
    if (x < lo || x > hi) return -INFINITY;
    return log(1.0 / (hi - lo));
}

// ------------------------------------------------------------------ kernels
#define GRID_STRIDE(k, n)                                                       \
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < (n);   \
         k += (int64_t)gridDim.x * blockDim.x)

// one thread per slot PAIR: both halves of the Box-Muller pair are used (slots 2p, 2p+1;
// g0 is even for every shard layout the engine accepts)
__global__ void k_sample_normal(DArg mean, DArg sd, double* out, int64_t n, int64_t g0,
                                uint64_t seed, uint32_t t, uint32_t draw) {
    const int64_t np = (n + 1) >> 1;
    GRID_STRIDE(p, np) {
        const int64_t k = 2 * p;
        double z0, z1;
        if ((g0 & 1) == 0) {
            normal_pair_at(seed, (uint64_t)(g0 + k) >> 1, t, STREAM_USER, draw, z0, z1);
        } else {
            z0 = normal_at(seed, (uint64_t)(g0 + k), t, STREAM_USER, draw);
            z1 = normal_at(seed, (uint64_t)(g0 + k + 1), t, STREAM_USER, draw);
        }
        out[k] = sd.at(k) * z0 + mean.at(k);           // NormalDistribution.sample
        if (k + 1 < n) out[k + 1] = sd.at(k + 1) * z1 + mean.at(k + 1);
    }
}
__global__ void k_observe_normal(DArg mean, DArg sd, DArg val, double* lw, int64_t n) {
    GRID_STRIDE(k, n) lw[k] += normal_logpdf(val.at(k), mean.at(k), sd.at(k));
}
__global__ void k_sample_gamma(DArg shape, DArg rate, double* out, int64_t n, int64_t g0,
                               uint64_t seed, uint32_t t, uint32_t draw) {
    GRID_STRIDE(k, n)
        out[k] = gamma_draw(shape.at(k), rate.at(k), seed, (uint64_t)(g0 + k), t, STREAM_USER,
                            draw * 256u);
}
__global__ void k_observe_gamma(DArg shape, DArg rate, DArg val, double* lw, int64_t n) {
    GRID_STRIDE(k, n) lw[k] += gamma_logpdf(val.at(k), shape.at(k), rate.at(k));
}
__global__ void k_sample_beta(DArg a, DArg b, double* out, int64_t n, int64_t g0, uint64_t seed,
                              uint32_t t, uint32_t draw) {
    GRID_STRIDE(k, n) {
        double x = gamma_draw(a.at(k), 1.0, seed, (uint64_t)(g0 + k), t, STREAM_USER, draw * 256u);
        double y = gamma_draw(b.at(k), 1.0, seed, (uint64_t)(g0 + k), t, STREAM_USER,
                              draw * 256u + 128u);
        out[k] = x / (x + y);
    }
}
__global__ void k_observe_beta(DArg a, DArg b, DArg val, double* lw, int64_t n) {
    GRID_STRIDE(k, n) lw[k] += beta_logpdf(val.at(k), a.at(k), b.at(k));
}
__global__ void k_sample_uniform(DArg lo, DArg hi, double* out, int64_t n, int64_t g0,
                                 uint64_t seed, uint32_t t, uint32_t draw) {
    GRID_STRIDE(k, n) {
        double u = uniform_at(seed, (uint64_t)(g0 + k), t, STREAM_USER, draw);
        out[k] = u * hi.at(k) + (1.0 - u) * lo.at(k);   // UniformRealDistribution.sample
    }
}
__global__ void k_observe_uniform(DArg lo, DArg hi, DArg val, double* lw, int64_t n) {
    GRID_STRIDE(k, n) lw[k] += uniform_logpdf(val.at(k), lo.at(k), hi.at(k));
}

// discrete: weights shared (wstride = 0) or per particle (wstride = K)
__global__ void k_sample_discrete(const double* __restrict__ w, int K, int wstride, int32_t* out,
                                  int64_t n, int64_t g0, uint64_t seed, uint32_t t,
                                  uint32_t draw) {
    GRID_STRIDE(k, n) {
        const double* wk = w + (int64_t)wstride * k;
        double total = 0.0;
        for (int i = 0; i < K; ++i) total = total + wk[i];          // (reduce + weights)
        double x = uniform_at(seed, (uint64_t)(g0 + k), t, STREAM_USER, draw) * total;
        double acc = 0.0;
        int v = K - 1;
        for (int i = 0; i < K; ++i) {
            acc = acc + wk[i];
            if (x < acc) { v = i; break; }                           // runtime.cljc:285
        }
        out[k] = v;
    }
}
__global__ void k_observe_discrete(const double* __restrict__ w, int K, int wstride,
                                   const int32_t* __restrict__ val, int32_t vs, double* lw,
                                   int64_t n) {
    GRID_STRIDE(k, n) {
        const double* wk = w + (int64_t)wstride * k;
        double total = 0.0;
        for (int i = 0; i < K; ++i) total = total + wk[i];
        int v = val ? val[k] : vs;
        double wv = (v >= 0 && v < K) ? wk[v] : 0.0;                 // (nth weights value 0.0)
        lw[k] += log(wv / total);
    }
}

__global__ void k_sample_dirichlet(const double* __restrict__ alpha, int K, double* out,
                                   int64_t n, int64_t g0, uint64_t seed, uint32_t t,
                                   uint32_t draw) {
    GRID_STRIDE(k, n) {
        double tot = 0.0;
        for (int j = 0; j < K; ++j) {
            double g = gamma_draw(alpha[j], 1.0, seed, (uint64_t)(g0 + k) * (uint64_t)K + j, t,
                                  STREAM_USER, draw * 256u);
            out[k * K + j] = g;
            tot = tot + g;
        }
        for (int j = 0; j < K; ++j) out[k * K + j] = out[k * K + j] / tot;
    }
}
__global__ void k_observe_dirichlet(const double* __restrict__ alpha, int K, double Z,
                                    const double* __restrict__ val, double* lw, int64_t n) {
    GRID_STRIDE(k, n) {
        double acc = 0.0;
        for (int j = 0; j < K; ++j) acc = acc + log(val[k * K + j]) * (alpha[j] - 1.0);
        lw[k] += acc - Z;
    }
}

// mvn, generic path: one warp per particle.  iL / L are d*d row-major lower
// triangular in global memory (L2/L1 resident).  The tensor-core kernel for the
// benchmark shape lives in mvn.cu.
__global__ void k_observe_mvn(const double* __restrict__ iL, double Z, int d,
                              const double* __restrict__ mean, int mstride,
                              const double* __restrict__ val, int vstride, double* lw,
                              int64_t n) {
    extern __shared__ double sh[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    double* dx = sh + wid * d;
    for (int64_t k = (int64_t)blockIdx.x * nw + wid; k < n; k += (int64_t)gridDim.x * nw) {
        for (int c = lane; c < d; c += 32)
            dx[c] = val[(int64_t)vstride * k + c] - mean[(int64_t)mstride * k + c];
        __syncwarp();
        double acc = 0.0;
        for (int r = lane; r < d; r += 32) {
            double s = 0.0;
            for (int c = 0; c <= r; ++c) s = s + iL[r * d + c] * dx[c];
            acc = acc + s * s;
        }
        acc = warp_sum(acc);
        if (lane == 0) lw[k] += -0.5 * acc - Z;            // runtime.cljc:364-366
        __syncwarp();
    }
}
__global__ void k_sample_mvn(const double* __restrict__ L, int d, const double* __restrict__ mean,
                             int mstride, double* out, int64_t n, int64_t g0, uint64_t seed,
                             uint32_t t, uint32_t draw) {
    extern __shared__ double sh[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    double* z = sh + wid * d;
    for (int64_t k = (int64_t)blockIdx.x * nw + wid; k < n; k += (int64_t)gridDim.x * nw) {
        for (int c = lane; c < d; c += 32)
            z[c] = normal_at(seed, (uint64_t)(g0 + k), t, STREAM_USER, draw * 4096u + (uint32_t)c);
        __syncwarp();
        for (int r = lane; r < d; r += 32) {
            double s = 0.0;
            for (int c = 0; c <= r; ++c) s = s + L[r * d + c] * z[c];
            out[k * d + r] = mean[(int64_t)mstride * k + r] + s;   // mean + L z (360-361)
        }
        __syncwarp();
    }
}

__global__ void k_selftest_math(int fn, const double* __restrict__ in, double* __restrict__ out,
                                int64_t n) {
    GRID_STRIDE(k, n) {
        const double x = in[k];
        double s, c, r = 0.0;
        switch (fn) {
        case 0: r = fm::exp_any(x); break;
        case 1: r = fm::log_core(x); break;
        case 2: r = fm::sqrt_fast(x); break;
        case 3: fm::sincospi_core(x, s, c); r = s; break;
        case 4: fm::sincospi_core(x, s, c); r = c; break;
        default: r = fm::div_fast(1.0, x); break;
        }
        out[k] = r;
    }
}

// ------------------------------------------------------------------ host
static int grid_for(const apop_ctx* c, int64_t n, int tb) {
    int64_t b = (n + tb - 1) / tb;
    const int64_t cap = (int64_t)c->sm_count * 32;
    return (int)(b < cap ? (b > 0 ? b : 1) : cap);
}

int resolve_arg(apop_ctx* c, const apop_arg& a, DArg* out) {
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

#define CHECK_CTX(c) do { if (!(c)) return APOP_ERR_INVALID; } while (0)
#define LAUNCH_DONE(c) do { (c)->launches += 1; APOP_CUDA(c, cudaGetLastError()); } while (0)

extern "C" {

int apop_selftest_math(apop_ctx* c, int fn, apop_col in, apop_col out) {
    CHECK_CTX(c);
    Column* i = get_col(c, in, APOP_F64, 1);
    Column* o = get_col(c, out, APOP_F64, 1);
    if (!i || !o) return APOP_ERR_INVALID;
    k_selftest_math<<<grid_for(c, c->n, 256), 256, 0, c->stream>>>(fn, (const double*)i->ptr,
                                                                 (double*)o->ptr, c->n);
    LAUNCH_DONE(c);
    return APOP_OK;
}

int apop_sample_normal(apop_ctx* c, apop_arg mean, apop_arg sd, uint32_t draw, apop_col out) {
    CHECK_CTX(c);
    DArg m, s;
    APOP_TRY(resolve_arg(c, mean, &m)); APOP_TRY(resolve_arg(c, sd, &s));
    Column* o = get_col(c, out, APOP_F64, 1);
    if (!o) return APOP_ERR_INVALID;
    k_sample_normal<<<grid_for(c, (c->n + 1) / 2, 256), 256, 0, c->stream>>>(m, s, (double*)o->ptr, c->n, c->g0,
                                                                 c->seed, c->step, draw);
    LAUNCH_DONE(c);
    return APOP_OK;
}
int apop_observe_normal(apop_ctx* c, apop_arg mean, apop_arg sd, apop_arg value, apop_col logw) {
    CHECK_CTX(c);
    DArg m, s, v;
    APOP_TRY(resolve_arg(c, mean, &m)); APOP_TRY(resolve_arg(c, sd, &s));
    APOP_TRY(resolve_arg(c, value, &v));
    Column* l = get_col(c, logw, APOP_F64, 1);
    if (!l) return APOP_ERR_INVALID;
    k_observe_normal<<<grid_for(c, c->n, 256), 256, 0, c->stream>>>(m, s, v, (double*)l->ptr, c->n);
    LAUNCH_DONE(c);
    return APOP_OK;
}
int apop_sample_gamma(apop_ctx* c, apop_arg shape, apop_arg rate, uint32_t draw, apop_col out) {
    CHECK_CTX(c);
    DArg a, b;
    APOP_TRY(resolve_arg(c, shape, &a)); APOP_TRY(resolve_arg(c, rate, &b));
    Column* o = get_col(c, out, APOP_F64, 1);
    if (!o) return APOP_ERR_INVALID;
    k_sample_gamma<<<grid_for(c, c->n, 256), 256, 0, c->stream>>>(a, b, (double*)o->ptr, c->n, c->g0,
                                                                c->seed, c->step, draw);
    LAUNCH_DONE(c);
    return APOP_OK;
}
int apop_observe_gamma(apop_ctx* c, apop_arg shape, apop_arg rate, apop_arg value, apop_col logw) {
    CHECK_CTX(c);
    DArg a, b, v;
    APOP_TRY(resolve_arg(c, shape, &a)); APOP_TRY(resolve_arg(c, rate, &b));
    APOP_TRY(resolve_arg(c, value, &v));
    Column* l = get_col(c, logw, APOP_F64, 1);
    if (!l) return APOP_ERR_INVALID;
    k_observe_gamma<<<grid_for(c, c->n, 256), 256, 0, c->stream>>>(a, b, v, (double*)l->ptr, c->n);
    LAUNCH_DONE(c);
    return APOP_OK;
}
int apop_sample_beta(apop_ctx* c, apop_arg alpha, apop_arg beta, uint32_t draw, apop_col out) {
    CHECK_CTX(c);
    DArg a, b;
    APOP_TRY(resolve_arg(c, alpha, &a)); APOP_TRY(resolve_arg(c, beta, &b));
    Column* o = get_col(c, out, APOP_F64, 1);
    if (!o) return APOP_ERR_INVALID;
    k_sample_beta<<<grid_for(c, c->n, 256), 256, 0, c->stream>>>(a, b, (double*)o->ptr, c->n, c->g0,
                                                               c->seed, c->step, draw);
    LAUNCH_DONE(c);
    return APOP_OK;
}
int apop_observe_beta(apop_ctx* c, apop_arg alpha, apop_arg beta, apop_arg value, apop_col logw) {
    CHECK_CTX(c);
    DArg a, b, v;
    APOP_TRY(resolve_arg(c, alpha, &a)); APOP_TRY(resolve_arg(c, beta, &b));
    APOP_TRY(resolve_arg(c, value, &v));
    Column* l = get_col(c, logw, APOP_F64, 1);
    if (!l) return APOP_ERR_INVALID;
    k_observe_beta<<<grid_for(c, c->n, 256), 256, 0, c->stream>>>(a, b, v, (double*)l->ptr, c->n);
    LAUNCH_DONE(c);
    return APOP_OK;
}
int apop_sample_uniform(apop_ctx* c, apop_arg lo, apop_arg hi, uint32_t draw, apop_col out) {
    CHECK_CTX(c);
    DArg a, b;
    APOP_TRY(resolve_arg(c, lo, &a)); APOP_TRY(resolve_arg(c, hi, &b));
    Column* o = get_col(c, out, APOP_F64, 1);
    if (!o) return APOP_ERR_INVALID;
    k_sample_uniform<<<grid_for(c, c->n, 256), 256, 0, c->stream>>>(a, b, (double*)o->ptr, c->n,
                                                                  c->g0, c->seed, c->step, draw);
    LAUNCH_DONE(c);
    return APOP_OK;
}
int apop_observe_uniform(apop_ctx* c, apop_arg lo, apop_arg hi, apop_arg value, apop_col logw) {
    CHECK_CTX(c);
    DArg a, b, v;
    APOP_TRY(resolve_arg(c, lo, &a)); APOP_TRY(resolve_arg(c, hi, &b));
    APOP_TRY(resolve_arg(c, value, &v));
    Column* l = get_col(c, logw, APOP_F64, 1);
    if (!l) return APOP_ERR_INVALID;
    k_observe_uniform<<<grid_for(c, c->n, 256), 256, 0, c->stream>>>(a, b, v, (double*)l->ptr, c->n);
    LAUNCH_DONE(c);
    return APOP_OK;
}

static int discrete_weights(apop_ctx* c, const double* weights, int k, apop_col wcol,
                            const double** dev, int* stride) {
    if (k <= 0) return set_err(c, APOP_ERR_INVALID, "discrete: k must be > 0");
    if (wcol >= 0) {
        Column* w = get_col(c, wcol, APOP_F64, k);
        if (!w) return APOP_ERR_INVALID;
        *dev = (const double*)w->ptr;
        *stride = k;
        return APOP_OK;
    }
    if (!weights) return set_err(c, APOP_ERR_INVALID, "discrete: no weights");
    void* d = nullptr;
    APOP_TRY(param_upload(c, weights, sizeof(double) * (size_t)k, &d));
    *dev = (const double*)d;
    *stride = 0;
    return APOP_OK;
}

int apop_sample_discrete(apop_ctx* c, const double* weights, int k, apop_col wcol, uint32_t draw,
                         apop_col out) {
    CHECK_CTX(c);
    const double* w; int stride;
    APOP_TRY(discrete_weights(c, weights, k, wcol, &w, &stride));
    Column* o = get_col(c, out, APOP_I32, 1);
    if (!o) return APOP_ERR_INVALID;
    k_sample_discrete<<<grid_for(c, c->n, 256), 256, 0, c->stream>>>(w, k, stride, (int32_t*)o->ptr,
                                                                   c->n, c->g0, c->seed, c->step,
                                                                   draw);
    LAUNCH_DONE(c);
    return APOP_OK;
}
int apop_observe_discrete(apop_ctx* c, const double* weights, int k, apop_col wcol,
                          apop_col value, int32_t value_scalar, apop_col logw) {
    CHECK_CTX(c);
    const double* w; int stride;
    APOP_TRY(discrete_weights(c, weights, k, wcol, &w, &stride));
    const int32_t* v = nullptr;
    if (value >= 0) {
        Column* vc = get_col(c, value, APOP_I32, 1);
        if (!vc) return APOP_ERR_INVALID;
        v = (const int32_t*)vc->ptr;
    }
    Column* l = get_col(c, logw, APOP_F64, 1);
    if (!l) return APOP_ERR_INVALID;
    k_observe_discrete<<<grid_for(c, c->n, 256), 256, 0, c->stream>>>(w, k, stride, v, value_scalar,
                                                                    (double*)l->ptr, c->n);
    LAUNCH_DONE(c);
    return APOP_OK;
}

int apop_sample_dirichlet(apop_ctx* c, const double* alpha, int k, uint32_t draw, apop_col out) {
    CHECK_CTX(c);
    if (!alpha || k <= 0) return set_err(c, APOP_ERR_INVALID, "dirichlet: bad alpha");
    Column* o = get_col(c, out, APOP_F64, k);
    if (!o) return APOP_ERR_INVALID;
    void* d = nullptr;
    APOP_TRY(param_upload(c, alpha, sizeof(double) * (size_t)k, &d));
    k_sample_dirichlet<<<grid_for(c, c->n, 128), 128, 0, c->stream>>>((const double*)d, k,
                                                                    (double*)o->ptr, c->n, c->g0,
                                                                    c->seed, c->step, draw);
    LAUNCH_DONE(c);
    return APOP_OK;
}
int apop_observe_dirichlet(apop_ctx* c, const double* alpha, int k, apop_col value,
                           apop_col logw) {
    CHECK_CTX(c);
    if (!alpha || k <= 0) return set_err(c, APOP_ERR_INVALID, "dirichlet: bad alpha");
    Column* v = get_col(c, value, APOP_F64, k);
    Column* l = get_col(c, logw, APOP_F64, 1);
    if (!v || !l) return APOP_ERR_INVALID;
    // Z = sum lgamma(alpha_j) - lgamma(sum alpha_j)   (runtime.cljc:295-296)
    double sl = 0.0, sa = 0.0;
    for (int j = 0; j < k; ++j) { sl = sl + lgamma(alpha[j]); sa = sa + alpha[j]; }
    const double Z = sl - lgamma(sa);
    void* d = nullptr;
    APOP_TRY(param_upload(c, alpha, sizeof(double) * (size_t)k, &d));
    k_observe_dirichlet<<<grid_for(c, c->n, 256), 256, 0, c->stream>>>(
        (const double*)d, k, Z, (const double*)v->ptr, (double*)l->ptr, c->n);
    LAUNCH_DONE(c);
    return APOP_OK;
}

} // extern "C"

namespace apop {

// lower Cholesky factor reading the upper triangle (vectorz / EJML inner
// variant, runtime.cljc:353), its inverse and Z (355-358), all on the host.
int mvn_prepare_host(const double* cov, int d, std::vector<double>& L, std::vector<double>& iL,
                     double* Z) {
    L.assign((size_t)d * d, 0.0);
    iL.assign((size_t)d * d, 0.0);
    for (int i = 0; i < d; ++i) {
        double div = 0.0;
        for (int j = i; j < d; ++j) {
            double sum = cov[i * d + j];
            for (int k = 0; k < i; ++k) sum = sum - L[i * d + k] * L[j * d + k];
            if (i == j) {
                if (!(sum > 0.0)) return -1;
                double el = sqrt(sum);
                L[i * d + i] = el;
                div = 1.0 / el;
            } else {
                L[j * d + i] = sum * div;
            }
        }
    }
    for (int cidx = 0; cidx < d; ++cidx)
        for (int r = cidx; r < d; ++r) {
            double s = (r == cidx) ? 1.0 : 0.0;
            for (int k = cidx; k < r; ++k) s = s - L[r * d + k] * iL[k * d + cidx];
            iL[r * d + cidx] = s / L[r * d + r];
        }
    double det = 1.0;
    for (int i = 0; i < d; ++i) det = det * L[i * d + i];
    *Z = 0.5 * d * log(2.0 * 3.14159265358979323846) + log(det);
    return 0;
}

} // namespace apop

extern "C" {

int apop_sample_mvn(apop_ctx* c, const double* mean, apop_col mean_col, const double* cov, int d,
                    uint32_t draw, apop_col out) {
    CHECK_CTX(c);
    if (!cov || d <= 0 || d > 256) return set_err(c, APOP_ERR_INVALID, "mvn: bad cov/d");
    if (draw >= 256u) return set_err(c, APOP_ERR_INVALID, "mvn: draw_id must be < 256 (each draw of a matrix family takes 4096 counters of a 20-bit field; a larger id would repeat the numbers of draw_id - 256)");
    Column* o = get_col(c, out, APOP_F64, d);
    if (!o) return APOP_ERR_INVALID;
    std::vector<double> L, iL; double Z;
    if (mvn_prepare_host(cov, d, L, iL, &Z)) return set_err(c, APOP_ERR_INVALID, "mvn: cov not PD");
    void *dL = nullptr, *dm = nullptr;
    APOP_TRY(param_upload(c, L.data(), sizeof(double) * (size_t)d * d, &dL));
    const double* mp; int ms;
    if (mean_col >= 0) {
        Column* mc = get_col(c, mean_col, APOP_F64, d);
        if (!mc) return APOP_ERR_INVALID;
        mp = (const double*)mc->ptr; ms = d;
    } else {
        if (!mean) return set_err(c, APOP_ERR_INVALID, "mvn: no mean");
        APOP_TRY(param_upload(c, mean, sizeof(double) * (size_t)d, &dm));
        mp = (const double*)dm; ms = 0;
    }
    const int tb = 128, nw = tb / 32;
    k_sample_mvn<<<grid_for(c, c->n, nw), tb, sizeof(double) * (size_t)d * nw, c->stream>>>(
        (const double*)dL, d, mp, ms, (double*)o->ptr, c->n, c->g0, c->seed, c->step, draw);
    LAUNCH_DONE(c);
    return APOP_OK;
}

int apop_observe_mvn(apop_ctx* c, const double* mean, apop_col mean_col, const double* cov, int d,
                     const double* value, apop_col value_col, apop_col logw) {
    CHECK_CTX(c);
    if (!cov || d <= 0 || d > 256) return set_err(c, APOP_ERR_INVALID, "mvn: bad cov/d");
    Column* l = get_col(c, logw, APOP_F64, 1);
    if (!l) return APOP_ERR_INVALID;
    std::vector<double> L, iL; double Z;
    if (mvn_prepare_host(cov, d, L, iL, &Z)) return set_err(c, APOP_ERR_INVALID, "mvn: cov not PD");
    void *diL = nullptr, *dm = nullptr, *dv = nullptr;
    APOP_TRY(param_upload(c, iL.data(), sizeof(double) * (size_t)d * d, &diL));
    const double *mp, *vp; int ms, vs;
    if (mean_col >= 0) {
        Column* mc = get_col(c, mean_col, APOP_F64, d);
        if (!mc) return APOP_ERR_INVALID;
        mp = (const double*)mc->ptr; ms = d;
    } else {
        if (!mean) return set_err(c, APOP_ERR_INVALID, "mvn: no mean");
        APOP_TRY(param_upload(c, mean, sizeof(double) * (size_t)d, &dm));
        mp = (const double*)dm; ms = 0;
    }
    if (value_col >= 0) {
        Column* vc = get_col(c, value_col, APOP_F64, d);
        if (!vc) return APOP_ERR_INVALID;
        vp = (const double*)vc->ptr; vs = d;
    } else {
        if (!value) return set_err(c, APOP_ERR_INVALID, "mvn: no value");
        APOP_TRY(param_upload(c, value, sizeof(double) * (size_t)d, &dv));
        vp = (const double*)dv; vs = 0;
    }
    if (mvn_dim_supported(d))          // batched whitening on the fp64 tensor pipe (mvn.cu)
        return mvn_observe_dmma_enqueue(c, d, (const double*)diL, Z, mp, ms, vp, vs,
                                        (double*)l->ptr);
    const int tb = 128, nw = tb / 32;
    k_observe_mvn<<<grid_for(c, c->n, nw), tb, sizeof(double) * (size_t)d * nw, c->stream>>>(
        (const double*)diL, Z, d, mp, ms, vp, vs, (double*)l->ptr, c->n);
    LAUNCH_DONE(c);
    return APOP_OK;
}

} // extern "C"
