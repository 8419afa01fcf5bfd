// dists3.cu — the matrix-parameter families that remained of anglican.runtime
//  and the per-particle table lookup that lets a traced
// program turn a discrete state column into a parameter column:
//   multivariate-t  (reference src/anglican/runtime.cljc:371-401)
//   wishart         (runtime.cljc:479-545, log-mv-gamma-fn 479-486)
//   apop_column_take: out[k] = table[index[k]]   (`(nth rows state)` of a program)
// Log-densities restate the reference's formulas; it evaluates the quadratic form
// with an explicit `m/inverse` and `m/det` (vectorz), here the Cholesky factor is
// used for both (same value to fp64 rounding, checked at 1e-9 against the oracle
// and scipy).  Samplers are counter-based and distributionally comparable only
// (runtime.cljc:142-147).
#include <math.h>

#include <vector>

#include "apop_internal.cuh"
#include "gamma_draw.cuh"
#include "rng.cuh"

namespace apop {

constexpr int MAXP = 16;             // wishart: largest matrix order handled per thread

#define GRID_STRIDE3(k, n)                                                      \
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < (n);   \
         k += (int64_t)gridDim.x * blockDim.x)

// ---- multivariate-t ----------------------------------------------------------
// observe (runtime.cljc:383-401):
//   lgamma((nu+d)/2) - [lgamma(nu/2) + d/2 log nu + d/2 log pi + 1/2 log det sigma
//                       + (nu+d)/2 log(1 + (mu-y)' sigma^-1 (mu-y) / nu)]
// one warp per particle; iL = inverse lower Cholesky factor of sigma, so that the
// quadratic form is |iL (mu - y)|^2; C0 = everything that does not depend on y.
__global__ void k_observe_mvt(const double* __restrict__ iL, double C0, double nu, int d,
                              const double* __restrict__ mean, int mstride,
                              const double* __restrict__ val, int vstride, double* lw, int64_t n) {
    extern __shared__ double sh[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    double* dx = sh + wid * d;
    for (int64_t k = (int64_t)blockIdx.x * nw + wid; k < n; k += (int64_t)gridDim.x * nw) {
        for (int c = lane; c < d; c += 32)
            dx[c] = mean[(int64_t)mstride * k + c] - val[(int64_t)vstride * k + c];
        __syncwarp();
        double acc = 0.0;
        for (int r = lane; r < d; r += 32) {
            double s = 0.0;
            for (int c = 0; c <= r; ++c) s = s + iL[r * d + c] * dx[c];
            acc = acc + s * s;
        }
        acc = warp_sum(acc);
        if (lane == 0) lw[k] += C0 - 0.5 * (nu + d) * log(1.0 + (1.0 / nu) * acc);
        __syncwarp();
    }
}
// sample (runtime.cljc:377-382): y ~ mvn(0, sigma), u ~ chi-squared(nu),
// x = mu + y sqrt(nu / u); chi-squared(nu) = gamma(nu/2, rate 1/2) (runtime.cljc:333-340)
__global__ void k_sample_mvt(const double* __restrict__ L, double nu, int d,
                             const double* __restrict__ mean, int mstride, double* out, int64_t n,
                             int64_t g0, uint64_t seed, uint32_t t, uint32_t draw) {
    extern __shared__ double sh[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    double* z = sh + wid * d;
    for (int64_t k = (int64_t)blockIdx.x * nw + wid; k < n; k += (int64_t)gridDim.x * nw) {
        for (int c = lane; c < d; c += 32)
            z[c] = normal_at(seed, (uint64_t)(g0 + k), t, STREAM_USER, draw * 4096u + (uint32_t)c);
        double u = 0.0;
        if (lane == 0)
            u = gamma_draw(0.5 * nu, 0.5, seed, (uint64_t)(g0 + k), t, STREAM_USER,
                           draw * 4096u + 2048u);
        u = __shfl_sync(0xffffffffu, u, 0);
        const double sc = sqrt(nu / u);
        __syncwarp();
        for (int r = lane; r < d; r += 32) {
            double s = 0.0;
            for (int c = 0; c <= r; ++c) s = s + L[r * d + c] * z[c];
            out[k * d + r] = mean[(int64_t)mstride * k + r] + s * sc;
        }
        __syncwarp();
    }
}

// ---- wishart -----------------------------------------------------------------
// observe (runtime.cljc:536-539): 1/2 (n-p-1) log det X - 1/2 tr(V^-1 X) - Z.
// One thread per particle (p <= MAXP): det by LU with partial pivoting (`m/det` is
// signed: a value with det <= 0 gives NaN / -inf exactly as `(log det)` does).
__global__ void k_observe_wishart(const double* __restrict__ invV, double nn, double Z, int p,
                                  const double* __restrict__ val, double* lw, int64_t n) {
    GRID_STRIDE3(k, n) {
        double a[MAXP * MAXP];
        const double* x = val + k * (int64_t)(p * p);
        double tr = 0.0;
        for (int i = 0; i < p; ++i)
            for (int j = 0; j < p; ++j) {
                a[i * p + j] = x[i * p + j];
                tr = tr + invV[i * p + j] * x[j * p + i];
            }
        double det = 1.0;
        for (int c = 0; c < p; ++c) {
            int piv = c;
            double best = fabs(a[c * p + c]);
            for (int r = c + 1; r < p; ++r)
                if (fabs(a[r * p + c]) > best) { best = fabs(a[r * p + c]); piv = r; }
            if (piv != c) {
                for (int j = 0; j < p; ++j) {
                    const double tmp = a[c * p + j]; a[c * p + j] = a[piv * p + j]; a[piv * p + j] = tmp;
                }
                det = -det;
            }
            const double d0 = a[c * p + c];
            det = det * d0;
            if (d0 == 0.0) break;
            for (int r = c + 1; r < p; ++r) {
                const double f = a[r * p + c] / d0;
                for (int j = c + 1; j < p; ++j) a[r * p + j] = a[r * p + j] - f * a[c * p + j];
            }
        }
        lw[k] += 0.5 * (nn - p - 1.0) * log(det) - 0.5 * tr - Z;
    }
}
// sample (runtime.cljc:519-535), Bartlett: A lower triangular, A_ii = sqrt(chi-squared(n - i)),
// A_ij ~ N(0,1) below the diagonal, X = (L A)(L A)'
__global__ void k_sample_wishart(const double* __restrict__ L, double nn, int p, double* out,
                                 int64_t n, int64_t g0, uint64_t seed, uint32_t t, uint32_t draw) {
    GRID_STRIDE3(k, n) {
        double A[MAXP * MAXP], LA[MAXP * MAXP];
        const uint64_t slot = (uint64_t)(g0 + k);
        for (int i = 0; i < p; ++i)
            for (int j = 0; j < p; ++j) {
                double v = 0.0;
                if (i == j)
                    v = sqrt(gamma_draw(0.5 * (nn - i), 0.5, seed, slot * (uint64_t)p + i, t, STREAM_USER,
                                        draw * 4096u + 2048u));
                else if (i > j)
                    v = normal_at(seed, slot, t, STREAM_USER, draw * 4096u + (uint32_t)(i * p + j));
                A[i * p + j] = v;
            }
        for (int i = 0; i < p; ++i)
            for (int j = 0; j < p; ++j) {
                double s = 0.0;
                for (int q = j; q <= i; ++q) s = s + L[i * p + q] * A[q * p + j];
                LA[i * p + j] = s;
            }
        double* x = out + k * (int64_t)(p * p);
        for (int i = 0; i < p; ++i)
            for (int j = 0; j < p; ++j) {
                double s = 0.0;
                const int m = i < j ? i : j;
                for (int q = 0; q <= m; ++q) s = s + LA[i * p + q] * LA[j * p + q];
                x[i * p + j] = s;
            }
    }
}

// ---- per-particle table lookup -------------------------------------------------
// out[k][:] = table[index[k]][:]; an index outside [0, rows) gives NaN rows (the
// reference's `nth` would throw: the error surfaces in whatever consumes the row)
__global__ void k_take_rows(const double* __restrict__ table, int rows, int width,
                            const int32_t* __restrict__ index, double* __restrict__ out, int64_t n) {
    const int64_t total = n * (int64_t)width;
    GRID_STRIDE3(i, total) {
        const int64_t k = i / width;
        const int c = (int)(i - k * width);
        const int r = index[k];
        out[i] = (r >= 0 && r < rows) ? table[(int64_t)r * width + c] : NAN;
    }
}

static int grid_for3(const apop_ctx* c, int64_t n, int tb) {
    int64_t b = (n + tb - 1) / tb;
    const int64_t cap = (int64_t)c->sm_count * 32;
    return (int)(b < cap ? (b > 0 ? b : 1) : cap);
}

// inverse of a symmetric positive definite matrix from its Cholesky factor
static void spd_inverse(const std::vector<double>& iL, int p, std::vector<double>& inv) {
    inv.assign((size_t)p * p, 0.0);
    for (int i = 0; i < p; ++i)
        for (int j = 0; j < p; ++j) {
            double s = 0.0;
            for (int q = (i > j ? i : j); q < p; ++q) s = s + iL[q * p + i] * iL[q * p + j];
            inv[i * p + j] = s;                                   // V^-1 = iL' iL
        }
}

// log-mv-gamma-fn, runtime.cljc:479-486
static double log_mv_gamma(int p, double a) {
    double s = 0.25 * p * (p - 1) * log(3.14159265358979323846);
    for (int j = 0; j < p; ++j) s = s + lgamma(a - 0.5 * j);
    return s;
}

} // namespace apop

using namespace apop;
#define CHECK_CTX(c) do { if (!(c)) return APOP_ERR_INVALID; } while (0)
#define LAUNCH_DONE(c) do { (c)->launches += 1; APOP_CUDA(c, cudaGetLastError()); } while (0)

static int vec_arg(apop_ctx* c, const double* host, apop_col col, int d, const char* what,
                   const double** p, int* stride) {
    if (col >= 0) {
        Column* mc = get_col(c, col, APOP_F64, d);
        if (!mc) return APOP_ERR_INVALID;
        *p = (const double*)mc->ptr; *stride = d;
        return APOP_OK;
    }
    if (!host) return set_err(c, APOP_ERR_INVALID, "%s: neither a host vector nor a column", what);
    void* dv = nullptr;
    APOP_TRY(param_upload(c, host, sizeof(double) * (size_t)d, &dv));
    *p = (const double*)dv; *stride = 0;
    return APOP_OK;
}

extern "C" {

int apop_sample_mvt(apop_ctx* c, double nu, const double* mean, apop_col mean_col,
                    const double* sigma, int d, uint32_t draw, apop_col out) {
    CHECK_CTX(c);
    if (!sigma || d <= 0 || d > 256 || !(nu > 0.0)) return set_err(c, APOP_ERR_INVALID, "multivariate-t: bad nu/sigma/d");
    if (draw >= 256u) return set_err(c, APOP_ERR_INVALID, "multivariate-t: draw_id must be < 256 (each draw of a matrix family takes 4096 counters of a 20-bit field; a larger id would repeat the numbers of draw_id - 256)");
    Column* o = get_col(c, out, APOP_F64, d);
    if (!o) return APOP_ERR_INVALID;
    std::vector<double> L, iL; double Z;
    if (mvn_prepare_host(sigma, d, L, iL, &Z)) return set_err(c, APOP_ERR_INVALID, "multivariate-t: sigma not PD");
    void* dL = nullptr;
    APOP_TRY(param_upload(c, L.data(), sizeof(double) * (size_t)d * d, &dL));
    const double* mp; int ms;
    APOP_TRY(vec_arg(c, mean, mean_col, d, "multivariate-t mean", &mp, &ms));
    const int tb = 128, nw = tb / 32;
    k_sample_mvt<<<grid_for3(c, c->n, nw), tb, sizeof(double) * (size_t)d * nw, c->stream>>>(
        (const double*)dL, nu, d, mp, ms, (double*)o->ptr, c->n, c->g0, c->seed, c->step, draw);
    LAUNCH_DONE(c);
    return APOP_OK;
}

int apop_observe_mvt(apop_ctx* c, double nu, const double* mean, apop_col mean_col,
                     const double* sigma, int d, const double* value, apop_col value_col,
                     apop_col logw) {
    CHECK_CTX(c);
    if (!sigma || d <= 0 || d > 256 || !(nu > 0.0)) return set_err(c, APOP_ERR_INVALID, "multivariate-t: bad nu/sigma/d");
    Column* l = get_col(c, logw, APOP_F64, 1);
    if (!l) return APOP_ERR_INVALID;
    std::vector<double> L, iL; double Z;
    if (mvn_prepare_host(sigma, d, L, iL, &Z)) return set_err(c, APOP_ERR_INVALID, "multivariate-t: sigma not PD");
    double logdet = 0.0;
    for (int i = 0; i < d; ++i) logdet = logdet + 2.0 * log(L[i * d + i]);
    const double C0 = lgamma(0.5 * (nu + d)) -
                      (lgamma(0.5 * nu) + 0.5 * d * log(nu) + 0.5 * d * log(3.14159265358979323846) +
                       0.5 * logdet);
    void* diL = nullptr;
    APOP_TRY(param_upload(c, iL.data(), sizeof(double) * (size_t)d * d, &diL));
    const double *mp, *vp; int ms, vs;
    APOP_TRY(vec_arg(c, mean, mean_col, d, "multivariate-t mean", &mp, &ms));
    APOP_TRY(vec_arg(c, value, value_col, d, "multivariate-t value", &vp, &vs));
    const int tb = 128, nw = tb / 32;
    k_observe_mvt<<<grid_for3(c, c->n, nw), tb, sizeof(double) * (size_t)d * nw, c->stream>>>(
        (const double*)diL, C0, nu, d, mp, ms, vp, vs, (double*)l->ptr, c->n);
    LAUNCH_DONE(c);
    return APOP_OK;
}

int apop_sample_wishart(apop_ctx* c, double n_df, const double* V, int p, uint32_t draw, apop_col out) {
    CHECK_CTX(c);
    if (!V || p <= 0 || p > MAXP || !(n_df > p - 1)) return set_err(c, APOP_ERR_INVALID, "wishart: need 1 <= p <= %d and n > p - 1", MAXP);
    if (draw >= 256u) return set_err(c, APOP_ERR_INVALID, "wishart: draw_id must be < 256 (each draw of a matrix family takes 4096 counters of a 20-bit field; a larger id would repeat the numbers of draw_id - 256)");
    Column* o = get_col(c, out, APOP_F64, p * p);
    if (!o) return APOP_ERR_INVALID;
    std::vector<double> L, iL; double Z;
    if (mvn_prepare_host(V, p, L, iL, &Z)) return set_err(c, APOP_ERR_INVALID, "wishart: V not PD");
    void* dL = nullptr;
    APOP_TRY(param_upload(c, L.data(), sizeof(double) * (size_t)p * p, &dL));
    k_sample_wishart<<<grid_for3(c, c->n, 128), 128, 0, c->stream>>>(
        (const double*)dL, n_df, p, (double*)o->ptr, c->n, c->g0, c->seed, c->step, draw);
    LAUNCH_DONE(c);
    return APOP_OK;
}

int apop_observe_wishart(apop_ctx* c, double n_df, const double* V, int p, apop_col value, apop_col logw) {
    CHECK_CTX(c);
    // the density's normaliser has lgamma((n - j) / 2), j = 0..p-1: at n <= p - 1 that is a pole or
    // a negative argument (log-mv-gamma-fn, runtime.cljc:479-485; commons' logGamma gives NaN there, so the
    // reference's weights turn NaN) - refused here instead of returning a finite but meaningless number
    if (!V || p <= 0 || p > MAXP || !(n_df > p - 1)) return set_err(c, APOP_ERR_INVALID, "wishart: need 1 <= p <= %d and n > p - 1", MAXP);
    Column* l = get_col(c, logw, APOP_F64, 1);
    Column* x = get_col(c, value, APOP_F64, p * p);
    if (!l || !x) return APOP_ERR_INVALID;
    std::vector<double> L, iL, inv; double Zm;
    if (mvn_prepare_host(V, p, L, iL, &Zm)) return set_err(c, APOP_ERR_INVALID, "wishart: V not PD");
    spd_inverse(iL, p, inv);
    double logdet = 0.0;
    for (int i = 0; i < p; ++i) logdet = logdet + 2.0 * log(L[i * p + i]);
    // Z, runtime.cljc:507-509
    const double Z = 0.5 * n_df * p * log(2.0) + 0.5 * n_df * logdet + log_mv_gamma(p, 0.5 * n_df);
    void* dinv = nullptr;
    APOP_TRY(param_upload(c, inv.data(), sizeof(double) * (size_t)p * p, &dinv));
    k_observe_wishart<<<grid_for3(c, c->n, 128), 128, 0, c->stream>>>(
        (const double*)dinv, n_df, Z, p, (const double*)x->ptr, (double*)l->ptr, c->n);
    LAUNCH_DONE(c);
    return APOP_OK;
}

int apop_column_take(apop_ctx* c, const double* table, int rows, int width, apop_col index_i32,
                     apop_col out) {
    CHECK_CTX(c);
    if (!table || rows <= 0 || width <= 0) return set_err(c, APOP_ERR_INVALID, "column_take: bad table");
    Column* ix = get_col(c, index_i32, APOP_I32, 1);
    Column* o = get_col(c, out, APOP_F64, width);
    if (!ix || !o) return APOP_ERR_INVALID;
    void* dt = nullptr;
    APOP_TRY(param_upload(c, table, sizeof(double) * (size_t)rows * width, &dt));
    k_take_rows<<<grid_for3(c, c->n * width, 256), 256, 0, c->stream>>>(
        (const double*)dt, rows, width, (const int32_t*)ix->ptr, (double*)o->ptr, c->n);
    LAUNCH_DONE(c);
    return APOP_OK;
}

} // extern "C"
