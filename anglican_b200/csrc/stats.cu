// stats.cu — result-side reductions of anglican.stat on the device (reference
// src/anglican/stat.cljc): empirical-expectation / -mean / -variance / -std /
// -skew / -kurtosis (56-120), empirical-covariance (84-91), collect-by +
// empirical-distribution for discrete values (7-54).  The reference folds a
// Clojure seq of [x log-w] pairs on the JVM, which would mean downloading every
// particle (1 GiB per column at 2^27); here the particles stay in HBM and only
// the moments cross the boundary.
//
// Weights follow stat.cljc:63-68: w_i = exp(lw_i - max lw).  Sums are plain
// fp64 tree sums in a FIXED order (grid and block shape do not depend on the
// device), so results are reproducible run to run; they differ from the
// reference's left fold only by summation order (tests: 1e-12 relative, the
// tolerance of stat_test.cljc:16).
#include <math.h>

#include "apop_internal.cuh"
#include "fastmath.cuh"

namespace apop {

constexpr int SB = 256;                 // threads per stats CTA
constexpr int SGRID = 592;              // partial-sum slots: 4 resident CTAs on each of 148 SMs (a cap, not a requirement)
constexpr int SCH = 8;                  // columns handled per pass over the particles

__device__ __forceinline__ double stat_weight(const double* __restrict__ lw, int64_t i, double m) {
    if (!lw) return 1.0;
    const double l = lw[i];
    if (l == m) return 1.0;                       // also covers m = +/-inf
    if (!(l > -INFINITY)) return 0.0;
    return fm::exp_any(l - m);
}

__device__ __forceinline__ double block_reduce_sum(double v, double* sm /*[SB/32]*/) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    double r = 0.0;
    if (threadIdx.x < 32) {
        r = (threadIdx.x < SB / 32) ? sm[threadIdx.x] : 0.0;
        r = warp_sum(r);
    }
    return r;                                     // valid in thread 0
}

// partial maxima of the log-weights -> part[blockIdx.x]
__global__ void __launch_bounds__(SB) k_stat_max(const double* __restrict__ lw, int64_t n,
                                                 double* __restrict__ part) {
    __shared__ double sm[SB / 32];
    double mx = -INFINITY;
    for (int64_t i = (int64_t)blockIdx.x * SB + threadIdx.x; i < n; i += (int64_t)gridDim.x * SB) {
        const double l = lw[i];
        mx = (l > mx || l != l) ? l : mx;         // NaN propagates like clojure.core/max
    }
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = sm[0];
        for (int i = 1; i < SB / 32; ++i) a = fmax(a, sm[i]);
        part[blockIdx.x] = a;
    }
}

__device__ __forceinline__ double max_of_parts(const double* __restrict__ part, int g) {
    double a = -INFINITY;
    for (int i = 0; i < g; ++i) a = fmax(a, part[i]);       // every thread: tiny, L1 resident
    return a;
}

// pass 1: sum w and sum w x_c for columns [c0, c0 + nc) -> part[blockIdx.x][1 + SCH]
__global__ void __launch_bounds__(SB) k_stat_m1(const double* __restrict__ x,
                                                const double* __restrict__ lw, int64_t n,
                                                int width, int c0, int nc,
                                                const double* __restrict__ mparts, int gm,
                                                double* __restrict__ part) {
    __shared__ double sm[SB / 32];
    const double m = lw ? max_of_parts(mparts, gm) : 0.0;
    double sw = 0.0, sx[SCH];
#pragma unroll
    for (int k = 0; k < SCH; ++k) sx[k] = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * SB + threadIdx.x; i < n; i += (int64_t)gridDim.x * SB) {
        const double w = stat_weight(lw, i, m);
        sw += w;
        const double* row = x + i * width + c0;
#pragma unroll
        for (int k = 0; k < SCH; ++k)
            if (k < nc) sx[k] += w * row[k];
    }
    double r = block_reduce_sum(sw, sm);
    if (threadIdx.x == 0) part[(size_t)blockIdx.x * (1 + SCH)] = r;
#pragma unroll
    for (int k = 0; k < SCH; ++k) {
        r = block_reduce_sum(sx[k], sm);
        if (threadIdx.x == 0) part[(size_t)blockIdx.x * (1 + SCH) + 1 + k] = r;
    }
}

// pass 2: sum w d^2, w d^3, w d^4 with d = x_c - mean_c -> part[blockIdx.x][3 * SCH]
__global__ void __launch_bounds__(SB) k_stat_m2(const double* __restrict__ x,
                                                const double* __restrict__ lw, int64_t n,
                                                int width, int c0, int nc,
                                                const double* __restrict__ mparts, int gm,
                                                const double* __restrict__ mean,
                                                double* __restrict__ part) {
    __shared__ double sm[SB / 32];
    const double m = lw ? max_of_parts(mparts, gm) : 0.0;
    double s2[SCH], s3[SCH], s4[SCH], mu[SCH];
#pragma unroll
    for (int k = 0; k < SCH; ++k) { s2[k] = s3[k] = s4[k] = 0.0; mu[k] = (k < nc) ? mean[c0 + k] : 0.0; }
    for (int64_t i = (int64_t)blockIdx.x * SB + threadIdx.x; i < n; i += (int64_t)gridDim.x * SB) {
        const double w = stat_weight(lw, i, m);
        const double* row = x + i * width + c0;
#pragma unroll
        for (int k = 0; k < SCH; ++k)
            if (k < nc) {
                const double d = row[k] - mu[k];
                const double d2 = d * d;
                s2[k] += w * d2;
                s3[k] += w * (d2 * d);
                s4[k] += w * (d2 * d2);
            }
    }
#pragma unroll
    for (int k = 0; k < SCH; ++k) {
        double r = block_reduce_sum(s2[k], sm);
        if (threadIdx.x == 0) part[(size_t)blockIdx.x * (3 * SCH) + k] = r;
        r = block_reduce_sum(s3[k], sm);
        if (threadIdx.x == 0) part[(size_t)blockIdx.x * (3 * SCH) + SCH + k] = r;
        r = block_reduce_sum(s4[k], sm);
        if (threadIdx.x == 0) part[(size_t)blockIdx.x * (3 * SCH) + 2 * SCH + k] = r;
    }
}

// fixed-order sum of the per-block partials: out[j] = sum_b part[b][j] (* scale)
__global__ void __launch_bounds__(SB) k_stat_fin(const double* __restrict__ part, int g, int stride,
                                                 int cols, double* __restrict__ out) {
    __shared__ double sm[SB / 32];
    for (int j = blockIdx.x; j < cols; j += gridDim.x) {
        double a = 0.0;
        for (int b = threadIdx.x; b < g; b += SB) a += part[(size_t)b * stride + j];
        const double r = block_reduce_sum(a, sm);
        if (threadIdx.x == 0) out[j] = r;
        __syncthreads();
    }
}

// mean_c = sum w x_c / sum w  (stat.cljc:56-74), written to the device mean vector
__global__ void k_stat_mean(const double* __restrict__ sums /*[1 + SCH]*/, int c0, int nc,
                            double* __restrict__ mean, double* __restrict__ sumw) {
    const int k = threadIdx.x;
    if (k < nc) mean[c0 + k] = sums[1 + k] / sums[0];
    if (k == 0) *sumw = sums[0];
}

// empirical-covariance (stat.cljc:84-91): sum_i w_i (x_i - mu)(x_i - mu)^T.
// A CTA stages CP particles' deviations in shared memory; thread e owns the
// entries e, e + SB, ... of the d x d matrix.  -> part[blockIdx.x][d * d]
constexpr int CP = 32;
constexpr int CMAXD = 64;
__global__ void __launch_bounds__(SB) k_stat_cov(const double* __restrict__ x,
                                                 const double* __restrict__ lw, int64_t n, int d,
                                                 const double* __restrict__ mparts, int gm,
                                                 const double* __restrict__ mean,
                                                 double* __restrict__ part) {
    __shared__ double D[CP][CMAXD + 1];
    __shared__ double W[CP];
    __shared__ double mu[CMAXD];
    const double m = lw ? max_of_parts(mparts, gm) : 0.0;
    const int dd = d * d;
    constexpr int EPT = (CMAXD * CMAXD + SB - 1) / SB;
    double acc[EPT];
#pragma unroll
    for (int k = 0; k < EPT; ++k) acc[k] = 0.0;
    if (threadIdx.x < d) mu[threadIdx.x] = mean[threadIdx.x];
    __syncthreads();
    const int64_t chunks = (n + CP - 1) / CP;
    for (int64_t ch = blockIdx.x; ch < chunks; ch += gridDim.x) {
        const int64_t p0 = ch * CP;
        for (int e = threadIdx.x; e < CP * d; e += SB) {
            const int p = e / d, c = e - p * d;
            D[p][c] = (p0 + p < n) ? x[(p0 + p) * d + c] - mu[c] : 0.0;
        }
        if (threadIdx.x < CP)
            W[threadIdx.x] = (p0 + threadIdx.x < n) ? stat_weight(lw, p0 + threadIdx.x, m) : 0.0;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < EPT; ++k) {
            const int e = threadIdx.x + k * SB;
            if (e < dd) {
                const int a = e / d, b = e - a * d;
                double s = acc[k];
                for (int p = 0; p < CP; ++p) s += W[p] * (D[p][a] * D[p][b]);
                acc[k] = s;
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < EPT; ++k) {
        const int e = threadIdx.x + k * SB;
        if (e < dd) part[(size_t)blockIdx.x * dd + e] = acc[k];
    }
}

// collect-by over discrete values 0..K-1 (stat.cljc:7-24): per-value sum of
// weights.  Shared-memory privatised histogram, one atomicAdd per particle.
// -> part[blockIdx.x][K + 1] (last entry: weight of out-of-range values)
template <typename T>
__global__ void __launch_bounds__(SB) k_stat_hist(const T* __restrict__ v,
                                                  const double* __restrict__ lw, int64_t n, int K,
                                                  const double* __restrict__ mparts, int gm,
                                                  double* __restrict__ part) {
    extern __shared__ double h[];
    const double m = lw ? max_of_parts(mparts, gm) : 0.0;
    for (int k = threadIdx.x; k <= K; k += SB) h[k] = 0.0;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * SB + threadIdx.x; i < n; i += (int64_t)gridDim.x * SB) {
        const double w = stat_weight(lw, i, m);
        const double fv = (double)v[i];
        const int k = (fv >= 0.0 && fv < (double)K && fv == floor(fv)) ? (int)fv : K;
        atomicAdd(&h[k], w);
    }
    __syncthreads();
    for (int k = threadIdx.x; k <= K; k += SB) part[(size_t)blockIdx.x * (K + 1) + k] = h[k];
}

// ------------------------------------------------------------------ host
struct StatBufs {
    double *mparts, *part, *sums, *mean, *sumw;
    int g;
};

static int stat_inputs(apop_ctx* c, apop_col xcol, int want_dtype, int width, apop_col lwcol,
                       const void** x, int* dtype, const double** lw) {
    if (xcol == APOP_COL_SWEEP_FINAL) {
        void* p = nullptr;
        APOP_TRY(apop_sweep_final_device(c, &p));
        *x = p;
        *dtype = APOP_F64;
    } else {
        Column* col = get_col(c, xcol, want_dtype, width);
        if (!col) return APOP_ERR_INVALID;
        *x = col->ptr;
        *dtype = col->dtype;
    }
    *lw = nullptr;
    if (lwcol >= 0) {
        Column* l = get_col(c, lwcol, APOP_F64, 1);
        if (!l) return APOP_ERR_INVALID;
        *lw = (const double*)l->ptr;
    }
    return APOP_OK;
}

static int stat_bufs(apop_ctx* c, size_t part_cols, size_t width, StatBufs* B) {
    const int64_t n = c->n;
    int g = (int)((n + SB - 1) / SB);
    if (g > SGRID) g = SGRID;
    if (g < 1) g = 1;
    const size_t doubles = (size_t)SGRID + (size_t)SGRID * part_cols + part_cols + width + 8;
    APOP_TRY(ensure_scratch(c, doubles * sizeof(double)));
    double* p = (double*)c->scratch;
    B->mparts = p; p += SGRID;
    B->part = p; p += (size_t)SGRID * part_cols;
    B->sums = p; p += part_cols;
    B->mean = p; p += width;
    B->sumw = p;
    B->g = g;
    return APOP_OK;
}

// weighted mean of every column into B.mean (device), sum of weights into B.sumw
static int stat_mean_enqueue(apop_ctx* c, const double* x, const double* lw, int width,
                             const StatBufs& B) {
    const int64_t n = c->n;
    if (lw) { k_stat_max<<<B.g, SB, 0, c->stream>>>(lw, n, B.mparts); c->launches += 1; }
    for (int c0 = 0; c0 < width; c0 += SCH) {
        const int nc = (width - c0 < SCH) ? (width - c0) : SCH;
        k_stat_m1<<<B.g, SB, 0, c->stream>>>(x, lw, n, width, c0, nc, B.mparts, B.g, B.part);
        k_stat_fin<<<1 + SCH, SB, 0, c->stream>>>(B.part, B.g, 1 + SCH, 1 + SCH, B.sums);
        k_stat_mean<<<1, 32, 0, c->stream>>>(B.sums, c0, nc, B.mean, B.sumw);
        c->launches += 3;
    }
    return APOP_OK;
}

__global__ void k_stat_lse_out(const double* __restrict__ mparts, int g,
                               const double* __restrict__ sums, double* __restrict__ out2) {
    out2[0] = max_of_parts(mparts, g);
    out2[1] = sums[0];
}

// {max lw, sum exp(lw - max)} of a log-weight array into out2 (device): the
// pieces of log-sum-exp / log-marginal (runtime.cljc:130-138, inference.cljc:181-186)
// allocations of stat_lse_enqueue, for callers that enqueue it inside a graph capture
int stat_lse_prepare(apop_ctx* c) {
    StatBufs B;
    return stat_bufs(c, 1 + SCH, 1, &B);
}

int stat_lse_enqueue(apop_ctx* c, const double* lw, int64_t n, double* out2) {
    StatBufs B;
    APOP_TRY(stat_bufs(c, 1 + SCH, 1, &B));
    k_stat_max<<<B.g, SB, 0, c->stream>>>(lw, n, B.mparts);
    k_stat_m1<<<B.g, SB, 0, c->stream>>>(lw, lw, n, 1, 0, 0, B.mparts, B.g, B.part);
    k_stat_fin<<<1, SB, 0, c->stream>>>(B.part, B.g, 1 + SCH, 1, B.sums);
    k_stat_lse_out<<<1, 1, 0, c->stream>>>(B.mparts, B.g, B.sums, out2);
    c->launches += 4;
    return APOP_OK;
}

} // namespace apop

using namespace apop;

extern "C" {

int apop_empirical_moments(apop_ctx* c, apop_col xcol, apop_col lwcol, int32_t width, double* mean,
                           double* var, double* skew, double* kurt) {
    if (!c) return APOP_ERR_INVALID;
    if (width < 1 || width > 4096) return set_err(c, APOP_ERR_INVALID, "moments: bad width");
    cudaSetDevice(c->device);
    const void* xv; const double* lw; int dt;
    APOP_TRY(stat_inputs(c, xcol, APOP_F64, width, lwcol, &xv, &dt, &lw));
    const double* x = (const double*)xv;
    StatBufs B;
    APOP_TRY(stat_bufs(c, 3 * SCH + 1, (size_t)width, &B));
    kt_begin(c, KT_STATS);
    APOP_TRY(stat_mean_enqueue(c, x, lw, width, B));
    std::vector<double> hmean(width), hm(3 * (size_t)width);
    double hsumw = 0.0;
    APOP_CUDA(c, cudaMemcpyAsync(hmean.data(), B.mean, sizeof(double) * width,
                                 cudaMemcpyDeviceToHost, c->stream));
    APOP_CUDA(c, cudaMemcpyAsync(&hsumw, B.sumw, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    const bool higher = var || skew || kurt;
    if (higher) {
        for (int c0 = 0; c0 < width; c0 += SCH) {
            const int nc = (width - c0 < SCH) ? (width - c0) : SCH;
            k_stat_m2<<<B.g, SB, 0, c->stream>>>(x, lw, c->n, width, c0, nc, B.mparts, B.g, B.mean,
                                                 B.part);
            k_stat_fin<<<3 * SCH, SB, 0, c->stream>>>(B.part, B.g, 3 * SCH, 3 * SCH, B.sums);
            c->launches += 2;
            for (int q = 0; q < 3; ++q)
                APOP_CUDA(c, cudaMemcpyAsync(hm.data() + (size_t)q * width + c0, B.sums + q * SCH,
                                             sizeof(double) * nc, cudaMemcpyDeviceToHost, c->stream));
        }
    }
    kt_end(c, KT_STATS);
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    if (!(hsumw > 0.0) || hsumw != hsumw)
        return set_err(c, APOP_ERR_ZERO_WEIGHT, "moments: total weight %g", hsumw);
    for (int k = 0; k < width; ++k) {
        if (mean) mean[k] = hmean[k];
        if (!higher) continue;
        const double v = hm[k] / hsumw;                              // stat.cljc:93-98
        const double s = sqrt(v);                                    // :100-104
        if (var) var[k] = v;
        if (skew) skew[k] = (hm[(size_t)width + k] / hsumw) / (s * s * s);          // :106-112
        if (kurt) kurt[k] = (hm[2 * (size_t)width + k] / hsumw) / (v * v);          // :114-120
    }
    return APOP_OK;
}

int apop_empirical_covariance(apop_ctx* c, apop_col xcol, apop_col lwcol, int32_t width,
                              double* mean, double* cov) {
    if (!c) return APOP_ERR_INVALID;
    if (width < 1 || width > CMAXD || !cov)
        return set_err(c, APOP_ERR_INVALID, "covariance: 1 <= width <= %d", CMAXD);
    cudaSetDevice(c->device);
    const void* xv; const double* lw; int dt;
    APOP_TRY(stat_inputs(c, xcol, APOP_F64, width, lwcol, &xv, &dt, &lw));
    const double* x = (const double*)xv;
    const int dd = width * width;
    StatBufs B;
    APOP_TRY(stat_bufs(c, (size_t)((dd > 1 + SCH) ? dd : 1 + SCH), (size_t)width, &B));
    kt_begin(c, KT_STATS);
    APOP_TRY(stat_mean_enqueue(c, x, lw, width, B));
    k_stat_cov<<<B.g, SB, 0, c->stream>>>(x, lw, c->n, width, B.mparts, B.g, B.mean, B.part);
    k_stat_fin<<<(dd < 256) ? dd : 256, SB, 0, c->stream>>>(B.part, B.g, dd, dd, B.sums);
    c->launches += 2;
    kt_end(c, KT_STATS);
    std::vector<double> hmean(width);
    double hsumw = 0.0;
    APOP_CUDA(c, cudaMemcpyAsync(hmean.data(), B.mean, sizeof(double) * width,
                                 cudaMemcpyDeviceToHost, c->stream));
    APOP_CUDA(c, cudaMemcpyAsync(&hsumw, B.sumw, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    APOP_CUDA(c, cudaMemcpyAsync(cov, B.sums, sizeof(double) * dd, cudaMemcpyDeviceToHost,
                                 c->stream));
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    if (!(hsumw > 0.0) || hsumw != hsumw)
        return set_err(c, APOP_ERR_ZERO_WEIGHT, "covariance: total weight %g", hsumw);
    for (int k = 0; k < dd; ++k) cov[k] /= hsumw;
    if (mean) for (int k = 0; k < width; ++k) mean[k] = hmean[k];
    return APOP_OK;
}

int apop_collect_discrete(apop_ctx* c, apop_col xcol, apop_col lwcol, int32_t K,
                          double* log_weight, double* prob) {
    if (!c) return APOP_ERR_INVALID;
    if (K < 1 || K > 4096) return set_err(c, APOP_ERR_INVALID, "collect: 1 <= K <= 4096");
    cudaSetDevice(c->device);
    const void* xv; const double* lw; int dt;
    APOP_TRY(stat_inputs(c, xcol, -1, 1, lwcol, &xv, &dt, &lw));
    StatBufs B;
    APOP_TRY(stat_bufs(c, (size_t)K + 1, 1, &B));
    kt_begin(c, KT_STATS);
    if (lw) { k_stat_max<<<B.g, SB, 0, c->stream>>>(lw, c->n, B.mparts); c->launches += 1; }
    const size_t smem = sizeof(double) * ((size_t)K + 1);
    if (dt == APOP_F64)
        k_stat_hist<double><<<B.g, SB, smem, c->stream>>>((const double*)xv, lw, c->n, K, B.mparts,
                                                          B.g, B.part);
    else
        k_stat_hist<int32_t><<<B.g, SB, smem, c->stream>>>((const int32_t*)xv, lw, c->n, K,
                                                           B.mparts, B.g, B.part);
    k_stat_fin<<<(K + 1 < 256) ? K + 1 : 256, SB, 0, c->stream>>>(B.part, B.g, K + 1, K + 1, B.sums);
    c->launches += 2;
    kt_end(c, KT_STATS);
    std::vector<double> h((size_t)K + 1), mp(B.g);
    APOP_CUDA(c, cudaMemcpyAsync(h.data(), B.sums, sizeof(double) * (K + 1), cudaMemcpyDeviceToHost,
                                 c->stream));
    if (lw)
        APOP_CUDA(c, cudaMemcpyAsync(mp.data(), B.mparts, sizeof(double) * B.g,
                                     cudaMemcpyDeviceToHost, c->stream));
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    double m = 0.0, tot = 0.0;
    if (lw) { m = -INFINITY; for (int i = 0; i < B.g; ++i) m = fmax(m, mp[i]); }
    for (int k = 0; k <= K; ++k) tot += h[k];
    if (!(tot > 0.0) || tot != tot) return set_err(c, APOP_ERR_ZERO_WEIGHT, "collect: total weight %g", tot);
    const double log_norm = log((double)c->n);                      // stat.cljc:14
    for (int k = 0; k < K; ++k) {
        if (log_weight) log_weight[k] = (h[k] > 0.0) ? log(h[k]) + m - log_norm : -INFINITY;
        if (prob) prob[k] = h[k] / tot;                              // stat.cljc:43-54
    }
    return APOP_OK;
}

} // extern "C"

// ===================================================================================
// apop_empirical_moments_noalloc: the moments of apop_empirical_moments for callers
// that must not allocate.
//
// The context-based call above takes its partial-sum workspace from ensure_scratch(),
// i.e. from cudaMalloc, and grows it with cudaFree + cudaMalloc when a wider column
// arrives.  Two kinds of user of this library cannot afford that:
//   * a host that has handed the whole GPU to its own arena (a JVM binding that
//     reserves device memory once at start-up, torch's caching allocator at its cap):
//     any further cudaMalloc fails or fragments what the host thinks it owns;
//   * code that computes result statistics while OTHER populations are sweeping on the
//     same GPU.  cudaMalloc / cudaFree stop the whole device until everything queued
//     has drained; a sharded sweep in flight is spinning on its peers with a bounded
//     wait (fabric.cuh), so such a stall can turn into an APOP_ERR_NCCL timeout on
//     another rank.
// For them the workspace is the fixed array below.  It is part of the library image:
// every GPU that loads libanglican_pop.so carries NOALLOC_DOUBLES * 8 bytes (~123 KB)
// of it from module load until the process exits, used or not.
//
// The kernels are the ones of the context-based call, launched with the same grid and
// in the same order, so the sums - and the results - are the same bit for bit.
// ===================================================================================
#include <stdarg.h>

#include <mutex>

namespace apop {

constexpr int NOALLOC_MAXW = APOP_NOALLOC_MAX_WIDTH;
constexpr int NOALLOC_PARTC = 3 * SCH + 1;                     // widest partial row (k_stat_m2)
// layout (doubles): mparts[SGRID] | part[SGRID][PARTC] | sums[PARTC] | mean[MAXW] | sumw[8]
constexpr size_t NOALLOC_DOUBLES =
    (size_t)SGRID + (size_t)SGRID * NOALLOC_PARTC + NOALLOC_PARTC + NOALLOC_MAXW + 8;
constexpr int NOALLOC_MAXDEV = 64;

static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_DOUBLES];

// In host code `g_noalloc_scratch` is only a handle for the runtime, NOT an address a
// kernel may dereference.  The real address differs per device and is looked up once per
// device, with that device current, and cached here.
static std::mutex g_noalloc_mu;
static double* g_noalloc_dev[NOALLOC_MAXDEV];                  // zero-initialised

static void noalloc_msg(char* buf, int32_t len, const char* fmt, ...) {
    if (!buf || len <= 0) return;
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, (size_t)len, fmt, ap);
    va_end(ap);
}

static int noalloc_scratch_for(int device, double** out, char* err, int32_t errlen) {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_dev[device]) {
        // the caller made `device` current just before
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            noalloc_msg(err, errlen, "moments_noalloc: cudaGetSymbolAddress on device %d failed: %s",
                        device, cudaGetErrorString(e));
            return APOP_ERR_CUDA;
        }
        g_noalloc_dev[device] = (double*)p;
    }
    *out = g_noalloc_dev[device];
    return APOP_OK;
}

// a pointer the kernels can dereference on `device`: device or managed memory
static bool noalloc_device_visible(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// columns [c0, c0 + nc): mean from the scratch copy; variance / skew / kurtosis from the
// three central sums (stat.cljc:93-120) - the arithmetic apop_empirical_moments does on
// the host after its download, done where the numbers already are
__global__ void k_stat_moments_out(const double* __restrict__ sums /*[3*SCH] or null*/,
                                   const double* __restrict__ sumw,
                                   const double* __restrict__ mean, int c0, int nc,
                                   double* __restrict__ o_mean, double* __restrict__ o_var,
                                   double* __restrict__ o_skew, double* __restrict__ o_kurt) {
    const int k = threadIdx.x;
    if (k >= nc) return;
    if (o_mean) o_mean[c0 + k] = mean[c0 + k];
    if (!sums) return;
    const double sw = *sumw;
    const double v = sums[k] / sw;
    const double s = sqrt(v);
    if (o_var) o_var[c0 + k] = v;
    if (o_skew) o_skew[c0 + k] = (sums[SCH + k] / sw) / (s * s * s);
    if (o_kurt) o_kurt[c0 + k] = (sums[2 * SCH + k] / sw) / (v * v);
}

} // namespace apop

extern "C" int apop_empirical_moments_noalloc(int device, void* stream, const double* d_x,
                                              const double* d_logw, int64_t n, int32_t width,
                                              double* d_mean, double* d_var, double* d_skew,
                                              double* d_kurt, char* errbuf, int32_t errbuf_len) {
    if (errbuf && errbuf_len > 0) errbuf[0] = 0;
    // sizes first: a request the fixed workspace cannot serve is refused before CUDA is touched
    if (width < 1 || width > NOALLOC_MAXW) {
        noalloc_msg(errbuf, errbuf_len,
                    "moments_noalloc: width %d is outside [1, %d] (APOP_NOALLOC_MAX_WIDTH): the "
                    "static workspace holds the means of at most %d columns; use "
                    "apop_empirical_moments for wider columns",
                    (int)width, NOALLOC_MAXW, NOALLOC_MAXW);
        return APOP_ERR_INVALID;
    }
    if (n < 1 || n > (int64_t)APOP_NOALLOC_MAX_PARTICLES) {
        noalloc_msg(errbuf, errbuf_len,
                    "moments_noalloc: n = %lld is outside [1, %lld] (APOP_NOALLOC_MAX_PARTICLES)",
                    (long long)n, (long long)APOP_NOALLOC_MAX_PARTICLES);
        return APOP_ERR_INVALID;
    }
    if (!d_x || !(d_mean || d_var || d_skew || d_kurt)) {
        noalloc_msg(errbuf, errbuf_len, "moments_noalloc: d_x and at least one output are required");
        return APOP_ERR_INVALID;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        noalloc_msg(errbuf, errbuf_len, "moments_noalloc: no CUDA device (there is no CPU fallback)");
        return APOP_ERR_CUDA;
    }
    if (device < 0 || device >= ndev || device >= NOALLOC_MAXDEV) {
        noalloc_msg(errbuf, errbuf_len, "moments_noalloc: device %d of %d", device, ndev);
        return APOP_ERR_INVALID;
    }
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) {
        noalloc_msg(errbuf, errbuf_len, "moments_noalloc: cudaSetDevice(%d): %s", device,
                    cudaGetErrorString(e));
        return APOP_ERR_CUDA;
    }
    const void* ptrs[6] = {d_x, d_logw, d_mean, d_var, d_skew, d_kurt};
    static const char* const names[6] = {"d_x", "d_logw", "d_mean", "d_var", "d_skew", "d_kurt"};
    for (int i = 0; i < 6; ++i)
        if (ptrs[i] && !noalloc_device_visible(ptrs[i])) {
            noalloc_msg(errbuf, errbuf_len,
                        "moments_noalloc: %s is not device memory (inputs and outputs stay on the GPU)",
                        names[i]);
            return APOP_ERR_INVALID;
        }
    double* ws = nullptr;
    const int rc = noalloc_scratch_for(device, &ws, errbuf, errbuf_len);
    if (rc != APOP_OK) return rc;

    double* mparts = ws;
    double* part = mparts + SGRID;
    double* sums = part + (size_t)SGRID * NOALLOC_PARTC;
    double* mean = sums + NOALLOC_PARTC;
    double* sumw = mean + NOALLOC_MAXW;
    int g = (int)((n + SB - 1) / SB);                    // the grid rule of stat_bufs()
    if (g > SGRID) g = SGRID;
    cudaStream_t st = (cudaStream_t)stream;

    // Nothing in the workspace has to be zero on entry although the previous call's numbers are
    // still in it: k_stat_max / k_stat_m1 / k_stat_m2 store every slot [0, g) of the rows that
    // k_stat_fin then reads, and mean / sumw / sums are stored before they are read.
    if (d_logw) k_stat_max<<<g, SB, 0, st>>>(d_logw, n, mparts);
    for (int c0 = 0; c0 < width; c0 += SCH) {
        const int nc = (width - c0 < SCH) ? (width - c0) : SCH;
        k_stat_m1<<<g, SB, 0, st>>>(d_x, d_logw, n, width, c0, nc, mparts, g, part);
        k_stat_fin<<<1 + SCH, SB, 0, st>>>(part, g, 1 + SCH, 1 + SCH, sums);
        k_stat_mean<<<1, 32, 0, st>>>(sums, c0, nc, mean, sumw);
    }
    const bool higher = d_var || d_skew || d_kurt;
    for (int c0 = 0; c0 < width; c0 += SCH) {
        const int nc = (width - c0 < SCH) ? (width - c0) : SCH;
        if (higher) {
            k_stat_m2<<<g, SB, 0, st>>>(d_x, d_logw, n, width, c0, nc, mparts, g, mean, part);
            k_stat_fin<<<3 * SCH, SB, 0, st>>>(part, g, 3 * SCH, 3 * SCH, sums);
        }
        k_stat_moments_out<<<1, 32, 0, st>>>(higher ? sums : nullptr, sumw, mean, c0, nc, d_mean,
                                             d_var, d_skew, d_kurt);
    }
    e = cudaGetLastError();
    if (e != cudaSuccess) {
        noalloc_msg(errbuf, errbuf_len, "moments_noalloc: kernel launch failed: %s",
                    cudaGetErrorString(e));
        return APOP_ERR_CUDA;
    }
    return APOP_OK;
}
