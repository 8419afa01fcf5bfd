// ssm.cu — fused sweeps over state-space programs: one call = one `sweep`
// (reference src/anglican/smc.cljc:35-51; conditional variant
// pgibbs.cljc:76-110 / ipmcmc.cljc:34-83) of a program of the shape
//     x0 <- sample;  for y_t:  x_t <- sample(. | x_{t-1});  observe(. | x_t) y_t
//
// Per step t the stream sees
//   K1  k_propagate   x_t[k] = f(x_{t-1}[a_{t-1}[k]], noise(t, slot k))   (sample*,
//                     inference.cljc:46-47), lw[k] = base + log p(y_t | x_t[k])
//                     (observe*, smc.cljc:23-27) and the per-tile
//                     (max, sum exp, sum exp^2, sum exp*x) partials — the
//                     ancestor gather of the previous resample is folded into
//                     the read of x_{t-1} (K5 never materialises).
//   K2..K4            resample_enqueue (resample.cu), ancestors in sorted order
//       k_step_stats  log-mean-weight, ESS, filtering mean, next base weight
#include <math.h>
#include <string.h>

#include "apop_internal.cuh"
#include "fastmath.cuh"
#include "rng.cuh"

namespace apop {

// ------------------------------------------------------------------ state
struct ModelDev {           // model parameters in device memory
    int kind, dim;
    double m0, s0, a, b, q, c, d, r, beta;
    double obs_const;       // log(r) + 0.5 log(2 pi)   |  log(beta) + 0.5 log(2 pi)
    double hmm_init[16], hmm_trans[256], hmm_mu[16], hmm_sd[16], hmm_const[16];
};

} // namespace apop

struct SweepState {
    int64_t n = 0;
    int T = 0, dim = 1, kind = 0;
    bool history = false;
    double *xa = nullptr, *xb = nullptr, *lw = nullptr;
    int32_t* anc = nullptr;
    double* x_hist = nullptr;      // [(T+1) * n * dim]
    int32_t* anc_hist = nullptr;   // [T * n]
    double *y = nullptr, *retained = nullptr;
    double *log_mean_w = nullptr, *ess = nullptr, *filt = nullptr;   // device [T]
    double* tsx = nullptr;         // per-tile sum exp(lw - tmax) * x
    double* x_final = nullptr;     // [n * dim]
    apop::SweepDev* sd = nullptr;
    apop::ModelDev* model = nullptr;
    double* traj = nullptr;
    size_t cap_n = 0, cap_T = 0, cap_hist = 0;
    const double* last_x = nullptr;    // pre-resample states of the last step
    const int32_t* last_anc = nullptr;
};

namespace apop {

void sweep_free(apop_ctx* c) {
    SweepState* s = c->sweep;
    if (!s) return;
    void* ptrs[] = {s->xa, s->xb, s->lw, s->anc, s->x_hist, s->anc_hist, s->y, s->retained,
                    s->log_mean_w, s->ess, s->filt, s->tsx, s->x_final, s->sd, s->model, s->traj};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    delete s;
    c->sweep = nullptr;
}

// ------------------------------------------------------------------ kernels
constexpr double kHalfLog2Pi_ = 0.91893853320467274178;

template <int KIND>
__device__ __forceinline__ double obs_logpdf(const ModelDev& M, double x, double y) {
    if (KIND == APOP_SSM_LGSSM1D) {
        // NormalDistribution.logDensity, runtime.cljc:403-409
        double x1 = (y - (M.c * x + M.d)) / M.r;
        return -0.5 * x1 * x1 - M.obs_const;
    } else if (KIND == APOP_SSM_SV) {
        // (normal 0 (* beta (exp (/ x 2)))):  log(sd) = log(beta) + x/2
        double sd = M.beta * exp(x / 2.0);
        double x1 = y / sd;
        return -0.5 * x1 * x1 - (0.5 * x + M.obs_const);
    } else {
        int s = (int)x;
        double x1 = (y - M.hmm_mu[s]) / M.hmm_sd[s];
        return -0.5 * x1 * x1 - M.hmm_const[s];
    }
}

__device__ __forceinline__ int discrete_draw(const double* w, int K, double u) {
    double total = 0.0;
    for (int i = 0; i < K; ++i) total = total + w[i];      // runtime.cljc:279
    double x = u * total, acc = 0.0;
    for (int i = 0; i < K; ++i) {
        acc = acc + w[i];
        if (x < acc) return i;                              // runtime.cljc:285
    }
    return K - 1;
}

template <int KIND>
__global__ void __launch_bounds__(TB) k_init(const ModelDev* __restrict__ Mp, double* x,
                                             int64_t n, int64_t g0, uint64_t seed,
                                             const double* retained) {
    const ModelDev& M = *Mp;
    const int64_t k0 = ((int64_t)blockIdx.x * TB + threadIdx.x) * 2;
    if (k0 >= n) return;
    double v0, v1;
    if (KIND == APOP_SSM_HMM) {
        v0 = discrete_draw(M.hmm_init, M.dim, uniform_at(seed, (uint64_t)(g0 + k0), 0, STREAM_INIT, 0));
        v1 = discrete_draw(M.hmm_init, M.dim, uniform_at(seed, (uint64_t)(g0 + k0 + 1), 0, STREAM_INIT, 0));
    } else {
        double z0, z1;
        normal_pair_at(seed, (uint64_t)(g0 + k0) >> 1, 0, STREAM_INIT, 0, z0, z1);
        v0 = M.m0 + M.s0 * z0;
        v1 = M.m0 + M.s0 * z1;
    }
    if (retained && g0 + k0 == 0) v0 = retained[0];        // pgibbs.cljc:78-81
    x[k0] = v0;
    if (k0 + 1 < n) x[k0 + 1] = v1;
}

// K1: propagate + reweight + tile partials (scalar-state models).
// The tile partials (max, approximate sum of exp(lw - max)) only steer the
// binade classification of k_finalize, so the exponentials are fp32 MUFU.EX2
// against a per-warp reference; the exact sums (ESS, log-mean-weight) come
// from k_fold_agg.
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <int KIND, bool FULL>
__device__ __forceinline__ void propagate_tile(const ModelDev& M, const double* __restrict__ xp,
                                               const int32_t* __restrict__ anc, int first_step,
                                               double* __restrict__ xn, double* __restrict__ lw,
                                               double y, const double* __restrict__ retained_t,
                                               double wbase, int64_t n, int64_t g0, uint64_t seed,
                                               uint32_t t, double* __restrict__ tmax,
                                               double* __restrict__ tsum, int* errflag) {
    __shared__ double red_mx[NWARP];
    __shared__ float red_ref[NWARP], red_s[NWARP];
    const int64_t base = (int64_t)blockIdx.x * TILE;
    const int64_t i0 = base + (int64_t)threadIdx.x * IPT;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    constexpr int HW = 4;                         // particles per half-step (register budget)
    constexpr float kL2E = 1.4426950408889634f;
    // running per-thread partials: exact max, fp32 (reference, sum of exp) pair
    double mx = -INFINITY;
    float ref = -INFINITY, s = 0.0f;
    bool bad = false;
    // model constants hoisted out of the half loop
    double c_obs = 0.0;
    if (KIND == APOP_SSM_LGSSM1D) c_obs = fm::div_fast(1.0, M.r);
    if (KIND == APOP_SSM_SV) c_obs = fm::div_fast(y, M.beta);
#pragma unroll
    for (int h = 0; h < IPT; h += HW) {
        const int64_t ih = i0 + h;
        const bool inb = FULL || ih + HW <= n;
#define VALID(j) (FULL || ih + (j) < n)
        double x[HW], l[HW];
        // ---- parents (gather folded into the read)
        {
            int32_t par[HW];
            if (first_step) {
#pragma unroll
                for (int j = 0; j < HW; ++j) par[j] = (int32_t)(ih + j);
            } else if (inb) {
                const int4 a = __ldg(reinterpret_cast<const int4*>(anc + ih));
                par[0] = a.x; par[1] = a.y; par[2] = a.z; par[3] = a.w;
            } else {
#pragma unroll
                for (int j = 0; j < HW; ++j) par[j] = (ih + j < n) ? anc[ih + j] : 0;
            }
#pragma unroll
            for (int j = 0; j < HW; ++j) x[j] = VALID(j) ? __ldg(xp + par[j]) : 0.0;
        }
        // ---- sample* (inference.cljc:46-47)
        if (KIND == APOP_SSM_HMM) {
#pragma unroll
            for (int j = 0; j < HW; ++j) {
                double u = uniform_at(seed, (uint64_t)(g0 + ih + j), t, STREAM_TRANSITION, 0);
                x[j] = (double)discrete_draw(M.hmm_trans + (int)x[j] * M.dim, M.dim, u);
            }
        } else {
            double z0[2], z1[2];
            normal_pairs_at<2>(seed, (uint64_t)(g0 + ih) >> 1, t, STREAM_TRANSITION, 0, z0, z1);
#pragma unroll
            for (int j = 0; j < HW; j += 2) {
                x[j] = (M.a * x[j] + M.b) + M.q * z0[j >> 1];              // sd*z + mean
                x[j + 1] = (M.a * x[j + 1] + M.b) + M.q * z1[j >> 1];
            }
        }
        if (h == 0 && retained_t && g0 + i0 == 0) x[0] = retained_t[0];    // conditional SMC: slot 0
        // ---- observe* (smc.cljc:23-27) + add-log-weight (state.cljc:27-31)
        if (KIND == APOP_SSM_LGSSM1D) {
#pragma unroll
            for (int j = 0; j < HW; ++j) {
                const double x1 = (y - (M.c * x[j] + M.d)) * c_obs;
                l[j] = wbase + (-0.5 * x1 * x1 - M.obs_const);
            }
        } else if (KIND == APOP_SSM_SV) {
            // (normal 0 (* beta (exp (/ x 2)))): y/sd = (y/beta) exp(-x/2), log sd = log beta + x/2
            bool slow = false;
#pragma unroll
            for (int j = 0; j < HW; ++j) { l[j] = -0.5 * x[j]; slow |= !(fabs(l[j]) < 708.0); }
            if (slow) {
#pragma unroll
                for (int j = 0; j < HW; ++j) l[j] = exp(l[j]);
            } else {
                fm::exp_core_v<HW>(l);
            }
#pragma unroll
            for (int j = 0; j < HW; ++j) {
                const double x1 = c_obs * l[j];
                l[j] = wbase + (-0.5 * x1 * x1 - (0.5 * x[j] + M.obs_const));
            }
        } else {
#pragma unroll
            for (int j = 0; j < HW; ++j) l[j] = wbase + obs_logpdf<KIND>(M, x[j], y);
        }
        // ---- store
        if (inb) {
            double2* qx = reinterpret_cast<double2*>(xn + ih);
            double2* ql = reinterpret_cast<double2*>(lw + ih);
#pragma unroll
            for (int j = 0; j < HW / 2; ++j) {
                qx[j] = make_double2(x[2 * j], x[2 * j + 1]);
                ql[j] = make_double2(l[2 * j], l[2 * j + 1]);
            }
        } else {
#pragma unroll
            for (int j = 0; j < HW; ++j)
                if (ih + j < n) { xn[ih + j] = x[j]; lw[ih + j] = l[j]; }
        }
        // ---- partials of particle-weights (smc.cljc:70-84): exact max; fp32 online
        //      (reference, sum exp) pair
        float mh = -INFINITY;
#pragma unroll
        for (int j = 0; j < HW; ++j) {
            if (VALID(j)) {
                bad |= !(l[j] < INFINITY);                                 // NaN or +inf
                mx = (l[j] > mx) ? l[j] : mx;
                mh = fmaxf(mh, __double2float_rn(l[j]));
            }
        }
        const float nref = fmaxf(ref, mh);
        if (nref > -INFINITY) {
            s *= ex2_approx((ref - nref) * kL2E);                          // ref = -inf -> 0
            const double refd = (double)nref;
#pragma unroll
            for (int j = 0; j < HW; ++j)
                if (VALID(j)) s += ex2_approx(__double2float_rn(l[j] - refd) * kL2E);
            ref = nref;
        }
#undef VALID
    }
    if (bad) atomicOr(errflag, EB_NAN);
    // ---- warp: common fp32 reference, rescaled sums; exact max
    float wref = ref;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wref = fmaxf(wref, __shfl_xor_sync(0xffffffffu, wref, o));
    if (wref > -INFINITY) s *= ex2_approx((ref - wref) * kL2E); else s = 0.0f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double v = __shfl_xor_sync(0xffffffffu, mx, o);
        mx = (v > mx) ? v : mx;
    }
    if (lane == 0) { red_mx[wid] = mx; red_ref[wid] = (wref > -INFINITY) ? wref : 0.0f; red_s[wid] = s; }
    __syncthreads();
    if (wid == 0) {
        double m = (lane < NWARP) ? red_mx[lane] : -INFINITY;
        m = warp_max(m);
        double a = 0.0;
        if (lane < NWARP && m > -INFINITY) {
            const double d = (double)red_ref[lane] - m;
            a = (double)red_s[lane] * ((d > -700.0) ? fm::exp_core(d) : 0.0);
        }
        a = warp_sum(a);
        // all log-weights -inf: every weight is 1 (smc.cljc:80)
        if (!(m > -INFINITY)) a = (double)((n - base < TILE) ? (n - base) : TILE);
        if (lane == 0) { tmax[blockIdx.x] = m; tsum[blockIdx.x] = a; }
    }
}

template <int KIND>
__global__ void __launch_bounds__(TB, 2) k_propagate(const ModelDev* __restrict__ Mp,
                                                  const double* __restrict__ xp,
                                                  const int32_t* __restrict__ anc, int first_step,
                                                  double* __restrict__ xn, double* __restrict__ lw,
                                                  const double* __restrict__ yt,
                                                  const double* __restrict__ retained_t,
                                                  const SweepDev* __restrict__ sd, int64_t n,
                                                  int64_t g0, uint64_t seed, uint32_t t,
                                                  double* __restrict__ tmax,
                                                  double* __restrict__ tsum, int* errflag) {
    const ModelDev& M = *Mp;
    const double y = yt[0];
    const double wbase = sd->base;
    if ((int64_t)(blockIdx.x + 1) * TILE <= n)
        propagate_tile<KIND, true>(M, xp, anc, first_step, xn, lw, y, retained_t, wbase, n, g0,
                                   seed, t, tmax, tsum, errflag);
    else
        propagate_tile<KIND, false>(M, xp, anc, first_step, xn, lw, y, retained_t, wbase, n, g0,
                                    seed, t, tmax, tsum, errflag);
}

__global__ void k_sweep_reset(SweepDev* sd, ResampleInfo* I) {
    sd->base = 0.0; sd->log_z = 0.0; sd->err = 0;
    I->err = 0;
}

__global__ void k_trace(const int32_t* __restrict__ anc_hist, const double* __restrict__ x_hist,
                        int64_t n, int T, int d, int64_t slot, double* traj) {
    int64_t k = slot;
    for (int t = T; t >= 1; --t) {
        k = anc_hist[(int64_t)(t - 1) * n + k];
        for (int i = threadIdx.x; i < d; i += blockDim.x)
            traj[(int64_t)t * d + i] = x_hist[((int64_t)t * n + k) * d + i];
    }
    for (int i = threadIdx.x; i < d; i += blockDim.x) traj[i] = x_hist[k * d + i];
}

__global__ void k_gather_rows(const double* __restrict__ src, double* __restrict__ dst,
                              const int32_t* __restrict__ anc, int64_t n, int d) {
    const int64_t total = n * (int64_t)d;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t k = i / d;
        dst[i] = src[(int64_t)anc[k] * d + (i - k * d)];
    }
}

// ------------------------------------------------------------------ host
static int dev_alloc(apop_ctx* c, void** p, size_t bytes) {
    if (cudaMalloc(p, bytes) != cudaSuccess) {
        *p = nullptr;
        return set_err(c, APOP_ERR_OOM, "sweep: device allocation of %zu bytes failed", bytes);
    }
    return APOP_OK;
}

static int sweep_prepare(apop_ctx* c, int T, int dim, bool history) {
    SweepState* s = c->sweep;
    if (!s) { s = new SweepState(); c->sweep = s; }
    const size_t n = (size_t)c->n;
    if (s->cap_n < n || s->dim != dim) {
        void* ptrs[] = {s->xa, s->xb, s->lw, s->anc, s->tsx, s->x_final};
        for (void* p : ptrs) if (p) cudaFree(p);
        APOP_TRY(dev_alloc(c, (void**)&s->xa, (n + 8) * dim * 8));
        APOP_TRY(dev_alloc(c, (void**)&s->xb, (n + 8) * dim * 8));
        APOP_TRY(dev_alloc(c, (void**)&s->lw, (n + 8) * 8));
        APOP_TRY(dev_alloc(c, (void**)&s->anc, (n + 8) * 4));
        APOP_TRY(dev_alloc(c, (void**)&s->tsx, ((n + TILE - 1) / TILE + 1) * 8 * (size_t)dim));
        APOP_TRY(dev_alloc(c, (void**)&s->x_final, (n + 8) * dim * 8));
        s->cap_n = n;
    }
    if (!s->sd) {
        APOP_TRY(dev_alloc(c, (void**)&s->sd, sizeof(SweepDev)));
        APOP_TRY(dev_alloc(c, (void**)&s->model, sizeof(ModelDev)));
    }
    if (s->cap_T < (size_t)T || s->dim != dim) {
        void* ptrs[] = {s->y, s->retained, s->log_mean_w, s->ess, s->filt, s->traj};
        for (void* p : ptrs) if (p) cudaFree(p);
        APOP_TRY(dev_alloc(c, (void**)&s->y, (size_t)(T + 1) * dim * 8));
        APOP_TRY(dev_alloc(c, (void**)&s->retained, (size_t)(T + 1) * dim * 8));
        APOP_TRY(dev_alloc(c, (void**)&s->log_mean_w, (size_t)(T + 1) * 8));
        APOP_TRY(dev_alloc(c, (void**)&s->ess, (size_t)(T + 1) * 8));
        APOP_TRY(dev_alloc(c, (void**)&s->filt, (size_t)(T + 1) * dim * 8));
        APOP_TRY(dev_alloc(c, (void**)&s->traj, (size_t)(T + 1) * dim * 8));
        s->cap_T = (size_t)T;
    }
    if (history) {
        const size_t need = (size_t)(T + 1) * n;
        if (s->cap_hist < need || s->dim != dim) {
            if (s->x_hist) cudaFree(s->x_hist);
            if (s->anc_hist) cudaFree(s->anc_hist);
            s->x_hist = nullptr; s->anc_hist = nullptr; s->cap_hist = 0;
            APOP_TRY(dev_alloc(c, (void**)&s->x_hist, need * dim * 8));
            APOP_TRY(dev_alloc(c, (void**)&s->anc_hist, need * 4));
            s->cap_hist = need;
        }
    }
    s->n = c->n; s->T = T; s->dim = dim; s->history = history;
    return APOP_OK;
}

static int upload_model(apop_ctx* c, const apop_ssm* m) {
    ModelDev h;
    memset(&h, 0, sizeof h);
    h.kind = m->kind; h.dim = m->dim;
    h.m0 = m->m0; h.s0 = m->s0; h.a = m->a; h.b = m->b; h.q = m->q;
    h.c = m->c; h.d = m->d; h.r = m->r; h.beta = m->beta;
    if (m->kind == APOP_SSM_LGSSM1D) h.obs_const = log(m->r) + kHalfLog2Pi_;
    if (m->kind == APOP_SSM_SV) h.obs_const = log(m->beta) + kHalfLog2Pi_;
    if (m->kind == APOP_SSM_HMM) {
        const int K = m->dim;
        if (K < 1 || K > 16 || !m->hmm_init || !m->hmm_trans || !m->hmm_mu || !m->hmm_sd)
            return set_err(c, APOP_ERR_INVALID, "hmm: need 1 <= K <= 16 and all tables");
        memcpy(h.hmm_init, m->hmm_init, sizeof(double) * K);
        memcpy(h.hmm_trans, m->hmm_trans, sizeof(double) * K * K);
        memcpy(h.hmm_mu, m->hmm_mu, sizeof(double) * K);
        memcpy(h.hmm_sd, m->hmm_sd, sizeof(double) * K);
        for (int i = 0; i < K; ++i) h.hmm_const[i] = log(m->hmm_sd[i]) + kHalfLog2Pi_;
    }
    void* staged = nullptr;
    APOP_TRY(param_upload(c, &h, sizeof h, &staged));
    APOP_CUDA(c, cudaMemcpyAsync(c->sweep->model, staged, sizeof h, cudaMemcpyDeviceToDevice,
                                 c->stream));
    return APOP_OK;
}

template <int KIND>
static int run_scalar_sweep(apop_ctx* c, const apop_ssm* model, const double* y, int T,
                            const double* retained, const apop_sweep_opts* o) {
    SweepState* s = c->sweep;
    const int64_t n = c->n;
    const int nt = (int)((n + TILE - 1) / TILE);
    APOP_TRY(ensure_resample_ws(c));
    ResampleWS& ws = c->ws;
    const bool csmc = o->csmc && retained;
    void* staged = nullptr;
    APOP_TRY(param_upload(c, y, sizeof(double) * (size_t)T, &staged));
    APOP_CUDA(c, cudaMemcpyAsync(s->y, staged, sizeof(double) * (size_t)T,
                                 cudaMemcpyDeviceToDevice, c->stream));
    if (csmc) {
        APOP_TRY(param_upload(c, retained, sizeof(double) * (size_t)(T + 1), &staged));
        APOP_CUDA(c, cudaMemcpyAsync(s->retained, staged, sizeof(double) * (size_t)(T + 1),
                                     cudaMemcpyDeviceToDevice, c->stream));
    }
    {
        Timer& tm = c->timers[KT_SWEEP];
        if (!tm.open && cudaEventCreate(&tm.a) == cudaSuccess) {
            cudaEventRecord(tm.a, c->stream);
            tm.open = true;
        }
    }
    k_sweep_reset<<<1, 1, 0, c->stream>>>(s->sd, ws.info);
    double* x_prev = s->history ? s->x_hist : s->xa;
    {
        const int64_t pairs = (n + 1) / 2;
        k_init<KIND><<<(int)((pairs + TB - 1) / TB), TB, 0, c->stream>>>(
            s->model, x_prev, n, c->g0, c->seed, csmc ? s->retained : nullptr);
    }
    c->launches += 2;
    int32_t* anc_prev = s->anc;
    for (int t = 1; t <= T; ++t) {
        double* x_new = s->history ? s->x_hist + (size_t)t * n : ((t & 1) ? s->xb : s->xa);
        int32_t* anc_new = s->history ? s->anc_hist + (size_t)(t - 1) * n : s->anc;
        kt_begin(c, KT_PROPAGATE);
        k_propagate<KIND><<<nt, TB, 0, c->stream>>>(
            s->model, x_prev, anc_prev, t == 1, x_new, s->lw, s->y + (t - 1),
            csmc ? s->retained + t : nullptr, s->sd, n, c->g0, c->seed, (uint32_t)t, ws.tmax,
            ws.tsum, &ws.info->err);
        kt_end(c, KT_PROPAGATE);
        c->launches += 1;
        const bool cond = csmc && t <= o->n_cond_steps;
        c->step = (uint32_t)t;
        ResampleArgs a{};
        a.lw = s->lw; a.w = nullptr; a.n = n; a.n_new = cond ? n - 1 : n;
        a.u = NAN; a.order = APOP_ORDER_SORTED;
        a.anc_out = anc_new; a.out_offset = cond ? 1 : 0; a.have_partials = true;
        a.xfilt = o->want_filt ? x_new : nullptr;
        a.rel_eps = 2e-5;            // fp32 tile partials of k_propagate
        a.step.sd = s->sd; a.step.log_mean_w = s->log_mean_w; a.step.ess = s->ess;
        a.step.filt = s->filt; a.step.anc = anc_new; a.step.t = t;
        a.step.reset = o->reset_weights; a.step.cond = cond ? 1 : 0;
        a.step.has_x = o->want_filt ? 1 : 0;
        APOP_TRY(resample_enqueue(c, a));
        x_prev = x_new;
        anc_prev = anc_new;
    }
    s->last_x = x_prev;
    s->last_anc = anc_prev;
    APOP_CUDA(c, cudaGetLastError());
    return APOP_OK;
}

} // namespace apop

using namespace apop;

extern "C" {

int apop_smc_sweep(apop_ctx* c, const apop_ssm* model, const double* y, int32_t T,
                   const double* retained, const apop_sweep_opts* opts,
                   apop_sweep_result* result) {
    if (!c) return APOP_ERR_INVALID;
    if (!model || !y || T < 1 || !opts || !result)
        return set_err(c, APOP_ERR_INVALID, "smc_sweep: bad arguments");
    if (c->world > 1) return set_err(c, APOP_ERR_UNSUPPORTED, "sharded sweep: see apop_smc_sweep_sharded");
    if (opts->csmc && c->n < 2)
        return set_err(c, APOP_ERR_INVALID, "conditional sweep needs >= 2 particles");
    cudaSetDevice(c->device);
    const int dim = (model->kind == APOP_SSM_LGMVN) ? model->dim : 1;
    APOP_TRY(sweep_prepare(c, T, dim, opts->keep_history != 0));
    APOP_TRY(upload_model(c, model));
    switch (model->kind) {
    case APOP_SSM_LGSSM1D:
        APOP_TRY(run_scalar_sweep<APOP_SSM_LGSSM1D>(c, model, y, T, retained, opts)); break;
    case APOP_SSM_SV:
        APOP_TRY(run_scalar_sweep<APOP_SSM_SV>(c, model, y, T, retained, opts)); break;
    case APOP_SSM_HMM:
        APOP_TRY(run_scalar_sweep<APOP_SSM_HMM>(c, model, y, T, retained, opts)); break;
    default:
        return set_err(c, APOP_ERR_UNSUPPORTED, "model kind %d not supported by the fused sweep",
                       model->kind);
    }
    SweepState* s = c->sweep;
    const int64_t n = c->n;
    if (opts->want_final || result->x_final) {
        int blocks = (int)((n * dim + 255) / 256);
        if (blocks > 148 * 32) blocks = 148 * 32;
        kt_begin(c, KT_GATHER);
        k_gather_rows<<<blocks, 256, 0, c->stream>>>(s->last_x, s->x_final, s->last_anc, n, dim);
        kt_end(c, KT_GATHER);
        c->launches += 1;
    }
    {
        Timer& tm = c->timers[KT_SWEEP];
        cudaEvent_t b;
        if (tm.open && cudaEventCreate(&b) == cudaSuccess) {
            cudaEventRecord(b, c->stream);
            tm.pending.push_back({tm.a, b});
            tm.a = nullptr;
            tm.open = false;
        }
    }
    // results: one small D2H per requested array, then a single sync
    SweepDev hsd;
    APOP_CUDA(c, cudaMemcpyAsync(&hsd, s->sd, sizeof hsd, cudaMemcpyDeviceToHost, c->stream));
    if (result->log_mean_w)
        APOP_CUDA(c, cudaMemcpyAsync(result->log_mean_w, s->log_mean_w, sizeof(double) * T,
                                     cudaMemcpyDeviceToHost, c->stream));
    if (result->ess)
        APOP_CUDA(c, cudaMemcpyAsync(result->ess, s->ess, sizeof(double) * T,
                                     cudaMemcpyDeviceToHost, c->stream));
    if (result->filt_mean)
        APOP_CUDA(c, cudaMemcpyAsync(result->filt_mean, s->filt, sizeof(double) * T * dim,
                                     cudaMemcpyDeviceToHost, c->stream));
    if (result->x_final)
        APOP_CUDA(c, cudaMemcpyAsync(result->x_final, s->x_final, sizeof(double) * n * dim,
                                     cudaMemcpyDeviceToHost, c->stream));
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    ResampleInfo I{};
    I.err = hsd.err;
    APOP_TRY(info_to_status(c, I));
    // :smc keeps the accumulated weight on the particles: final weight == log Z-hat
    result->log_z = opts->reset_weights ? hsd.log_z : hsd.base;
    return APOP_OK;
}

int apop_trace_lineage(apop_ctx* c, int64_t slot, double* out_traj) {
    if (!c) return APOP_ERR_INVALID;
    SweepState* s = c->sweep;
    if (!s || !s->history || !out_traj)
        return set_err(c, APOP_ERR_INVALID, "trace_lineage: no keep_history sweep available");
    if (slot < 0 || slot >= s->n) return set_err(c, APOP_ERR_INVALID, "trace_lineage: bad slot");
    k_trace<<<1, 32, 0, c->stream>>>(s->anc_hist, s->x_hist, s->n, s->T, s->dim, slot, s->traj);
    c->launches += 1;
    APOP_CUDA(c, cudaMemcpyAsync(out_traj, s->traj, sizeof(double) * (size_t)(s->T + 1) * s->dim,
                                 cudaMemcpyDeviceToHost, c->stream));
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    return APOP_OK;
}

int apop_sweep_download_history(apop_ctx* c, int32_t* anc, double* x) {
    if (!c) return APOP_ERR_INVALID;
    SweepState* s = c->sweep;
    if (!s || !s->history)
        return set_err(c, APOP_ERR_INVALID, "download_history: no keep_history sweep available");
    const size_t n = (size_t)s->n;
    if (anc)
        APOP_CUDA(c, cudaMemcpyAsync(anc, s->anc_hist, sizeof(int32_t) * n * s->T,
                                     cudaMemcpyDeviceToHost, c->stream));
    if (x)
        APOP_CUDA(c, cudaMemcpyAsync(x, s->x_hist, sizeof(double) * n * (s->T + 1) * s->dim,
                                     cudaMemcpyDeviceToHost, c->stream));
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    return APOP_OK;
}

int apop_sweep_final_device(apop_ctx* c, void** out) {
    if (!c || !out) return APOP_ERR_INVALID;
    if (!c->sweep || !c->sweep->x_final) return set_err(c, APOP_ERR_INVALID, "no sweep yet");
    *out = c->sweep->x_final;
    return APOP_OK;
}

} // extern "C"
