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
#include <stdlib.h>
#include <string.h>

#include <condition_variable>
#include <mutex>

#include "apop_internal.cuh"
#include "fastmath.cuh"
#include "rng.cuh"

// Particles per sub-step / resident CTAs per SM of k_propagate.  HW=2 with 3 CTAs
// is ~3 % faster on a cold GPU but executes 13 % more instructions; in a long
// sweep the B200 sits at its 1000 W power cap and the leaner HW=4 / 2 CTAs
// configuration is ~5 % faster (profiles/r01s_propagate_config.txt).
#ifndef APOP_PROP_HW
#define APOP_PROP_HW 4
#endif
#ifndef APOP_PROP_CTAS
#define APOP_PROP_CTAS 2
#endif

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
    bool heap_backed = false;      // xa/xb/anc live in the fabric heap (world > 1)
    double *xa = nullptr, *xb = nullptr, *lw = nullptr;
    int32_t* anc = nullptr;
    double* x_hist = nullptr;      // [(T+1) * n * dim]
    int32_t* anc_hist = nullptr;   // [T * n]
    // compact lineages (keep_history == 2): two bits per particle and step instead of 12
    // bytes, states recomputed along the one traced lineage from the counter-based RNG
    int hist_mode = 0;             // 0 none, 1 full history, 2 compact
    uint32_t* lin_new = nullptr;   // [T][lin_w] bit j: output slot j >= 1 starts a new ancestor
    uint32_t* lin_alive = nullptr; // [T][lin_w] bit i: particle i has offspring among slots >= 1
    int32_t* lin_anc0 = nullptr;   // [T] ancestor of slot 0 (retained particle / PGAS draw)
    int32_t* lin_idx = nullptr;    // [T+1] particle indices of the traced lineage
    size_t lin_w = 0, cap_lin = 0;
    uint64_t sweep_seed = 0;
    int sweep_csmc = 0;
    // the whole sweep as one CUDA graph, replayed while the shape stays the same
    cudaGraphExec_t gexec = nullptr;
    // everything a captured sweep bakes into its kernel arguments BY VALUE: a sweep that differs
    // in any of these must be captured again.  (Observations, the retained path and the seed are
    // read through device memory and may change between replays.)  Compared field by field:
    // the struct has padding, which `= {}` and assignment need not zero or copy.
    struct GraphKey {
        int kind, T, hist_mode, csmc, n_cond, reset, want_final, want_filt, no_resample;
        int64_t n;
        uint64_t ess_bits;         // ess_threshold bit for bit: k_chain takes it by value, and
                                   // a threshold > 0 adds the final log-sum-exp node
        const void *xa, *xb, *lw, *anc, *y, *ret, *lin, *xf, *ws, *lse, *sd, *model;
        bool same(const GraphKey& o) const {
            return kind == o.kind && T == o.T && hist_mode == o.hist_mode && csmc == o.csmc &&
                   n_cond == o.n_cond && reset == o.reset && want_final == o.want_final &&
                   want_filt == o.want_filt && no_resample == o.no_resample && n == o.n &&
                   ess_bits == o.ess_bits && xa == o.xa && xb == o.xb && lw == o.lw &&
                   anc == o.anc && y == o.y && ret == o.ret && lin == o.lin && xf == o.xf &&
                   ws == o.ws && lse == o.lse && sd == o.sd && model == o.model;
        }
    } gkey = {};
    int64_t glaunches = 0;
    const unsigned long long* seedp = nullptr;   // &sd->seed while a graph is captured / replayed
    double *y = nullptr, *retained = nullptr;
    double *log_mean_w = nullptr, *ess = nullptr, *filt = nullptr;   // device [T]
    double* tsx = nullptr;         // per-tile sum exp(lw - tmax) * x
    double* x_final = nullptr;     // [n * dim]
    apop::SweepDev* sd = nullptr;
    apop::ModelDev* model = nullptr;
    double* traj = nullptr;
    double* lse = nullptr;         // {max, sum exp} of the final log-weights (no_resample sweeps)
    int32_t* as_anc = nullptr;     // PGAS: ancestor drawn for the retained particle (1 index)
    double trans_const = 0.0;      // log(q) + 0.5 log(2 pi): transition log-density constant
    double* mvn_mats = nullptr;    // LGMVN: A | Lq | iL, each [dim*dim]
    double mvn_Z = 0.0, mvn_s0 = 1.0;
    size_t cap_n = 0, cap_T = 0, cap_hist = 0;
    const double* last_x = nullptr;    // pre-resample states of the last step
    const int32_t* last_anc = nullptr;
};

namespace apop {

void sweep_free(apop_ctx* c) {
    SweepState* s = c->sweep;
    if (!s) return;
    if (s->heap_backed) { s->xa = nullptr; s->xb = nullptr; s->anc = nullptr; }
    if (s->gexec) cudaGraphExecDestroy(s->gexec);
    void* ptrs[] = {s->xa, s->xb, s->lw, s->anc, s->x_hist, s->anc_hist, s->y, s->retained,
                    s->log_mean_w, s->ess, s->filt, s->tsx, s->x_final, s->sd, s->model, s->traj,
                    s->mvn_mats, s->lse, s->as_anc, s->lin_new, s->lin_alive, s->lin_anc0, s->lin_idx};
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
                                             const double* retained,
                                             const unsigned long long* __restrict__ seedp) {
    const ModelDev& M = *Mp;
    if (seedp) seed = *seedp;
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
// binade classification of k_fin_b, so the exponentials are fp32 MUFU.EX2
// against a per-warp reference; the exact sums (ESS, log-mean-weight) come
// from k_fold_agg.
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <int KIND, bool FULL, bool SHARD>
__device__ __forceinline__ void propagate_tile(const PeerF64& xpeer,
                                               const ModelDev& M, const double* __restrict__ xp,
                                               const int32_t* __restrict__ anc, int first_step,
                                               double* __restrict__ xn, double* __restrict__ lw,
                                               double y, const double* __restrict__ retained_t,
                                               double wbase, int64_t n, int64_t g0, uint64_t seed,
                                               uint32_t t, double* __restrict__ tmax,
                                               double* __restrict__ tsum, long long* gmax,
                                               int* errflag, int accum, bool aligned) {
    __shared__ double red_mx[NWARP];
    __shared__ float red_ref[NWARP], red_s[NWARP];
    const int64_t base = (int64_t)blockIdx.x * TILE;
    const int64_t i0 = base + (int64_t)threadIdx.x * IPT;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    constexpr int HW = APOP_PROP_HW;              // particles per sub-step (register budget)
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
        const bool inb = FULL || (aligned && ih + HW <= n);     // 128-bit accesses allowed
#define VALID(j) (FULL || ih + (j) < n)
        double x[HW], l[HW];
        // ---- parents (gather folded into the read)
        {
            int32_t par[HW];
            if (first_step) {
#pragma unroll
                for (int j = 0; j < HW; ++j) par[j] = (int32_t)(ih + j);
            } else if (inb) {
                if (HW == 4) {
                    const int4 a = __ldg(reinterpret_cast<const int4*>(anc + ih));
                    par[0] = a.x; par[1] = a.y; par[HW - 2] = a.z; par[HW - 1] = a.w;
                } else {
                    const int2 a = __ldg(reinterpret_cast<const int2*>(anc + ih));
                    par[0] = a.x; par[1] = a.y;
                }
            } else {
#pragma unroll
                for (int j = 0; j < HW; ++j) par[j] = (ih + j < n) ? anc[ih + j] : 0;
            }
            if (!SHARD || first_step) {
#pragma unroll
                for (int j = 0; j < HW; ++j) x[j] = VALID(j) ? __ldg(xp + par[j]) : 0.0;
            } else {
                // parents are global ids: almost all live in this shard; the sliver
                // that migrated is read from its owner's heap over NVLink
#pragma unroll
                for (int j = 0; j < HW; ++j) {
                    x[j] = 0.0;
                    if (VALID(j)) {
                        const int64_t d = (int64_t)par[j] - g0;
                        if ((uint64_t)d < (uint64_t)n) {
                            x[j] = __ldg(xp + d);
                        } else {
                            const int q = (int)((int64_t)par[j] / n);
                            x[j] = __ldcv(xpeer.p[q] + ((int64_t)par[j] - (int64_t)q * n));
                        }
                    }
                }
            }
        }
        // ---- sample* (inference.cljc:46-47)
        if (KIND == APOP_SSM_HMM) {
#pragma unroll
            for (int j = 0; j < HW; ++j) {
                double u = uniform_at(seed, (uint64_t)(g0 + ih + j), t, STREAM_TRANSITION, 0);
                x[j] = (double)discrete_draw(M.hmm_trans + (int)x[j] * M.dim, M.dim, u);
            }
        } else {
            double z0[HW / 2], z1[HW / 2];
            normal_pairs_at<HW / 2>(seed, (uint64_t)(g0 + ih) >> 1, t, STREAM_TRANSITION, 0, z0, z1);
#pragma unroll
            for (int j = 0; j < HW; j += 2) {
                x[j] = (M.a * x[j] + M.b) + M.q * z0[j >> 1];              // sd*z + mean
                x[j + 1] = (M.a * x[j + 1] + M.b) + M.q * z1[j >> 1];
            }
        }
        if (h == 0 && retained_t && g0 + i0 == 0) x[0] = retained_t[0];    // conditional SMC: slot 0
        // accumulated log-weight the observe adds to: common to all particles after
        // a resample (smc.cljc:98-103), the particle's own without one (importance)
        double wb[HW];
#pragma unroll
        for (int j = 0; j < HW; ++j) wb[j] = wbase;
        if (accum) {
#pragma unroll
            for (int j = 0; j < HW; ++j) wb[j] = VALID(j) ? lw[ih + j] : 0.0;
        }
        // ---- observe* (smc.cljc:23-27) + add-log-weight (state.cljc:27-31)
        if (KIND == APOP_SSM_LGSSM1D) {
#pragma unroll
            for (int j = 0; j < HW; ++j) {
                const double x1 = (y - (M.c * x[j] + M.d)) * c_obs;
                l[j] = wb[j] + (-0.5 * x1 * x1 - M.obs_const);
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
                l[j] = wb[j] + (-0.5 * x1 * x1 - (0.5 * x[j] + M.obs_const));
            }
        } else {
#pragma unroll
            for (int j = 0; j < HW; ++j) l[j] = wb[j] + obs_logpdf<KIND>(M, x[j], y);
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
    if (bad) atomicOr(errflag, (mx == INFINITY) ? EB_INF : EB_NAN);
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
        if (lane == 0) {
            tmax[blockIdx.x] = m; tsum[blockIdx.x] = a;
            atomicMax(gmax, dkey(m));                 // global max for k_fin_a (smc.cljc:75)
        }
    }
}

// GSEED: the RNG key is read from SweepDev (sweeps replayed from a CUDA graph); otherwise it
// is a kernel parameter, i.e. a constant-bank operand of the Philox rounds
template <int KIND, bool SHARD, bool GSEED>
__global__ void __launch_bounds__(TB, APOP_PROP_CTAS) k_propagate(const ModelDev* __restrict__ Mp,
                                                  const double* __restrict__ xp,
                                                  const int32_t* __restrict__ anc, int first_step,
                                                  double* __restrict__ xn, double* __restrict__ lw,
                                                  const double* __restrict__ yt,
                                                  const double* __restrict__ retained_t,
                                                  const SweepDev* __restrict__ sd, int64_t n,
                                                  int64_t g0, uint64_t seed, uint32_t t,
                                                  double* __restrict__ tmax,
                                                  double* __restrict__ tsum, long long* gmax,
                                                  int* errflag,
                                                  const __grid_constant__ PeerF64 xpeer,
                                                  int accum) {
    const ModelDev& M = *Mp;
    const double y = yt[0];
    const double wbase = sd->base;
    if (GSEED) seed = sd->seed;
    accum |= sd->carry;          // adaptive resampling: the last observe kept the weighted particles
    // rows of a lineage history start at multiples of n elements: 16-byte aligned
    // only when n is a multiple of 4 (ancestors) / 2 (states)
    const bool aligned = (((uintptr_t)anc | (uintptr_t)xn | (uintptr_t)lw) & 15) == 0;
    if ((int64_t)(blockIdx.x + 1) * TILE <= n && aligned)
        propagate_tile<KIND, true, SHARD>(xpeer, M, xp, anc, first_step, xn, lw, y, retained_t,
                                          wbase, n, g0, seed, t, tmax, tsum, gmax, errflag, accum,
                                          true);
    else
        propagate_tile<KIND, false, SHARD>(xpeer, M, xp, anc, first_step, xn, lw, y, retained_t,
                                           wbase, n, g0, seed, t, tmax, tsum, gmax, errflag, accum,
                                           aligned);
}

__global__ void k_sweep_reset(SweepDev* sd, ResampleInfo* I, long long* gmax, int* ticket) {
    sd->base = 0.0; sd->log_z = 0.0; sd->err = 0; sd->carry = 0;
    gmax[0] = dkey(-INFINITY); ticket[0] = 0; ticket[1] = 0;
    I->err = 0;
}

// ---- PGAS ancestor sampling (resample-retained, pgas.cljc:107-131) ----------
// v_i = logw_i + log f(x*_{t+1} | x_t^i): the rescored transition `sample` of the
// retained path against particle i's state (NormalDistribution.logDensity /
// `discrete` observe*, runtime.cljc:403-409, 288-290).  next == nullptr (last
// observe: nothing left to rescore): v = logw.
template <int KIND>
__global__ void __launch_bounds__(256) k_as_logw(const ModelDev* __restrict__ Mp,
                                                 const double* __restrict__ x,
                                                 const double* __restrict__ lw,
                                                 const double* __restrict__ next, double tconst,
                                                 double* __restrict__ v, int64_t n) {
    const ModelDev& M = *Mp;
    const int64_t k = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (k >= n) return;
    double r = lw[k];
    if (next) {
        const double xn = next[0];
        if (KIND == APOP_SSM_HMM) {
            const int s = (int)x[k], c = (int)xn;
            const double* row = M.hmm_trans + s * M.dim;
            double tot = 0.0;
            for (int i = 0; i < M.dim; ++i) tot = tot + row[i];
            r += (c >= 0 && c < M.dim) ? log(row[c] / tot) : -INFINITY;
        } else {
            const double x1 = (xn - (M.a * x[k] + M.b)) / M.q;
            r += -0.5 * x1 * x1 - tconst;
        }
    }
    v[k] = r;
}
__global__ void k_as_commit(int32_t* anc_step, const int32_t* drawn) { anc_step[0] = drawn[0]; }

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

// ---- compact lineages --------------------------------------------------------
// The ancestors of one step come out of the resampler sorted (slots >= 1; slot 0 is
// the retained particle of a conditional sweep, or a PGAS draw), so two bit-vectors
// hold them: new[j] = slot j starts a new ancestor, alive[i] = particle i is an
// ancestor.  anc[j] = select1(alive, rank1(new, j)).  n / 4 bytes per step instead of
// 12 n (pgibbs.cljc:29-35 only ever follows ONE lineage per sweep).
__global__ void __launch_bounds__(256) k_pack_lineage(const int32_t* __restrict__ anc, int64_t n,
                                                      uint32_t* __restrict__ nw,
                                                      uint32_t* __restrict__ al,
                                                      int32_t* __restrict__ anc0) {
    const int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const int a = (j < n) ? anc[j] : -1;
    int ap = __shfl_up_sync(0xffffffffu, a, 1);
    if (lane == 0) ap = (j > 0 && j < n) ? anc[j - 1] : -1;
    const bool flag = (j >= 1 && j < n) && (j == 1 || a != ap);
    const unsigned act = __ballot_sync(0xffffffffu, flag);
    if (lane == 0 && j < n) nw[j >> 5] = act;
    if (flag) {
        const int word = a >> 5;
        const unsigned peers = __match_any_sync(act, word);
        const unsigned bits = __reduce_or_sync(peers, 1u << (a & 31));
        if (lane == __ffs(peers) - 1) atomicOr(al + word, bits);
    }
    if (j == 0) anc0[0] = a;
}

__device__ __forceinline__ int block_sum_i32(int v, int* sm /*[33]*/) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    int r = 0;
    if (threadIdx.x < 32) {
        r = (threadIdx.x < (blockDim.x >> 5)) ? sm[threadIdx.x] : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
        if (threadIdx.x == 0) sm[32] = r;
    }
    __syncthreads();
    r = sm[32];
    __syncthreads();
    return r;
}

// one state of the traced lineage, recomputed exactly as k_init / k_propagate drew it
template <int KIND>
__device__ double lineage_state(const ModelDev& M, double xprev, uint64_t seed, uint64_t slot,
                                uint32_t t) {
    if (KIND == APOP_SSM_HMM) {
        const double u = uniform_at(seed, slot, t, t ? STREAM_TRANSITION : STREAM_INIT, 0);
        return (double)discrete_draw(t ? M.hmm_trans + (int)xprev * M.dim : M.hmm_init, M.dim, u);
    }
    const double z = normal_at(seed, slot, t, t ? STREAM_TRANSITION : STREAM_INIT, 0);
    return t ? (M.a * xprev + M.b) + M.q * z : M.m0 + M.s0 * z;
}

// trace-back over the bit-vectors (one CTA), then the forward replay of the states
constexpr int TRB = 1024;
template <int KIND>
__global__ void __launch_bounds__(TRB) k_trace_compact(const ModelDev* __restrict__ Mp,
                                                       const uint32_t* __restrict__ nw,
                                                       const uint32_t* __restrict__ al,
                                                       const int32_t* __restrict__ anc0, int W,
                                                       int T, int64_t slot, int64_t g0,
                                                       uint64_t seed,
                                                       const double* __restrict__ retained,
                                                       int32_t* __restrict__ idx,
                                                       double* __restrict__ traj) {
    __shared__ int red[33];
    __shared__ int s_pos, s_ex[TRB / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int k = (int)slot;
    const int chunk = (W + TRB - 1) / TRB;
    for (int t = T; t >= 1; --t) {
        const uint32_t* nwt = nw + (size_t)(t - 1) * W;
        const uint32_t* alt = al + (size_t)(t - 1) * W;
        int i;
        if (k == 0) {
            i = anc0[t - 1];
        } else {
            // r = number of distinct ancestors among slots 1 .. k
            const int lastw = k >> 5;
            int cnt = 0;
            for (int w = tid; w <= lastw; w += TRB) {
                uint32_t v = __ldg(nwt + w);
                if (w == lastw && (k & 31) != 31) v &= (2u << (k & 31)) - 1u;
                cnt += __popc(v);
            }
            const int r = block_sum_i32(cnt, red);
            // position of the r-th set bit of alive
            const int w0 = tid * chunk, w1 = min(w0 + chunk, W);
            int c = 0;
            for (int w = w0; w < w1; ++w) c += __popc(__ldg(alt + w));
            int inc = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int o = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d) inc += o;
            }
            if (lane == 31) s_ex[wid] = inc;
            if (tid == 0) s_pos = -1;
            __syncthreads();
            int pre = 0;
            for (int q = 0; q < wid; ++q) pre += s_ex[q];
            const int ex = pre + inc - c;
            if (ex < r && r <= ex + c) {
                int rem = r - ex;
                for (int w = w0; w < w1; ++w) {
                    uint32_t v = __ldg(alt + w);
                    const int pc = __popc(v);
                    if (rem <= pc) {
                        for (int q = 1; q < rem; ++q) v &= v - 1;
                        s_pos = w * 32 + __ffs(v) - 1;
                        break;
                    }
                    rem -= pc;
                }
            }
            __syncthreads();
            i = s_pos;
            __syncthreads();
            if (i < 0) i = 0;            // inconsistent vectors: never index out of range
        }
        if (tid == 0) idx[t] = i;
        k = i;
    }
    if (tid != 0) return;
    const ModelDev& M = *Mp;
    double x = (retained && g0 + k == 0) ? retained[0]
                                         : lineage_state<KIND>(M, 0.0, seed, (uint64_t)(g0 + k), 0);
    traj[0] = x;
    for (int t = 1; t <= T; ++t) {
        const int i = idx[t];
        x = (retained && g0 + i == 0) ? retained[t]
                                      : lineage_state<KIND>(M, x, seed, (uint64_t)(g0 + i), (uint32_t)t);
        traj[t] = x;
    }
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

// final states of a sharded sweep: parents are global ids
__global__ void k_gather_shard(const __grid_constant__ PeerF64 src, double* __restrict__ dst,
                               const int32_t* __restrict__ anc, int64_t n, int d) {
    const int64_t total = n * (int64_t)d;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t k = i / d;
        const int64_t a = anc[k];
        const int q = (int)(a / n);
        dst[i] = __ldcv(src.p[q] + (a - (int64_t)q * n) * d + (i - k * d));
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

static int sweep_prepare(apop_ctx* c, int T, int dim, int hist_mode) {
    const bool history = hist_mode == 1;
    SweepState* s = c->sweep;
    if (!s) { s = new SweepState(); c->sweep = s; }
    const size_t n = (size_t)c->n;
    if (c->world > 1) {
        // heap columns hold n * dim doubles (the ancestor column is oversized for dim > 1)
        APOP_TRY(fabric_prepare(c, n * (size_t)dim));   // collective: (re)maps the peer heaps
        if (!c->fab.linked)
            return set_err(c, APOP_ERR_INVALID, "sharded sweep: ranks are not linked "
                                                "(apop_link_group / NCCL id)");
    }
    if (s->cap_n < n || s->dim != dim) {
        if (s->heap_backed) { s->xa = nullptr; s->xb = nullptr; s->anc = nullptr; }
        void* ptrs[] = {s->xa, s->xb, s->lw, s->anc, s->tsx, s->x_final};
        for (void* p : ptrs) if (p) cudaFree(p);
        s->xa = s->xb = nullptr; s->anc = nullptr;
        if (c->world == 1) {
            APOP_TRY(dev_alloc(c, (void**)&s->xa, (n + 8) * dim * 8));
            APOP_TRY(dev_alloc(c, (void**)&s->xb, (n + 8) * dim * 8));
            APOP_TRY(dev_alloc(c, (void**)&s->anc, (n + 8) * 4));
        }
        APOP_TRY(dev_alloc(c, (void**)&s->lw, (n + 8) * 8));
        APOP_TRY(dev_alloc(c, (void**)&s->tsx, ((n + TILE - 1) / TILE + 1) * 8 * (size_t)dim));
        APOP_TRY(dev_alloc(c, (void**)&s->x_final, (n + 8) * dim * 8));
        s->cap_n = n;
    }
    if (c->world > 1) {
        char* h = c->fab.heap[c->rank];
        s->xa = (double*)(h + c->fab.off_xa);
        s->xb = (double*)(h + c->fab.off_xb);
        s->anc = (int32_t*)(h + c->fab.off_anc);
        s->heap_backed = true;
    }
    if (!s->sd) {
        APOP_TRY(dev_alloc(c, (void**)&s->sd, sizeof(SweepDev)));
        APOP_TRY(dev_alloc(c, (void**)&s->model, sizeof(ModelDev)));
        APOP_TRY(dev_alloc(c, (void**)&s->lse, 2 * sizeof(double)));
        APOP_TRY(dev_alloc(c, (void**)&s->as_anc, 4 * sizeof(int32_t)));
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
    if (hist_mode == 2) {
        const size_t W = (((size_t)n + 31) / 32 + 3) & ~(size_t)3;
        const size_t need = (size_t)T * W;
        if (s->cap_lin < need) {
            void* ptrs[] = {s->lin_new, s->lin_alive, s->lin_anc0, s->lin_idx};
            for (void* p : ptrs) if (p) cudaFree(p);
            s->lin_new = s->lin_alive = nullptr; s->lin_anc0 = s->lin_idx = nullptr; s->cap_lin = 0;
            APOP_TRY(dev_alloc(c, (void**)&s->lin_new, need * 4));
            APOP_TRY(dev_alloc(c, (void**)&s->lin_alive, need * 4));
            APOP_TRY(dev_alloc(c, (void**)&s->lin_anc0, (size_t)(T + 1) * 4));
            APOP_TRY(dev_alloc(c, (void**)&s->lin_idx, (size_t)(T + 2) * 4));
            s->cap_lin = need;
        }
        s->lin_w = W;
    }
    s->n = c->n; s->T = T; s->dim = dim; s->history = history; s->hist_mode = hist_mode;
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
    c->sweep->trans_const = (m->q > 0.0) ? log(m->q) + kHalfLog2Pi_ : 0.0;
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
    if (m->kind == APOP_SSM_LGMVN) {
        const int d = m->dim;
        if (!mvn_dim_supported(d) || !m->A || !m->Lq || !m->R)
            return set_err(c, APOP_ERR_UNSUPPORTED, "fused mvn sweep: dim must be 8/16/32/64 and "
                                                    "A, Lq, R given (dim %d)", d);
        std::vector<double> L, iL;
        double Z;
        if (mvn_prepare_host(m->R, d, L, iL, &Z))
            return set_err(c, APOP_ERR_INVALID, "mvn: observation covariance not PD");
        SweepState* s = c->sweep;
        const size_t dd = (size_t)d * d;
        if (!s->mvn_mats) APOP_TRY(dev_alloc(c, (void**)&s->mvn_mats, sizeof(double) * 3 * 64 * 64));
        std::vector<double> pack(3 * dd);
        memcpy(pack.data(), m->A, sizeof(double) * dd);
        memcpy(pack.data() + dd, m->Lq, sizeof(double) * dd);
        memcpy(pack.data() + 2 * dd, iL.data(), sizeof(double) * dd);
        void* st = nullptr;
        APOP_TRY(param_upload(c, pack.data(), sizeof(double) * 3 * dd, &st));
        APOP_CUDA(c, cudaMemcpyAsync(s->mvn_mats, st, sizeof(double) * 3 * dd,
                                     cudaMemcpyDeviceToDevice, c->stream));
        s->mvn_Z = Z;
        s->mvn_s0 = m->s0;
    }
    void* staged = nullptr;
    APOP_TRY(param_upload(c, &h, sizeof h, &staged));
    APOP_CUDA(c, cudaMemcpyAsync(c->sweep->model, staged, sizeof h, cudaMemcpyDeviceToDevice,
                                 c->stream));
    return APOP_OK;
}

// peer views of one heap column (identical offsets on every rank)
static PeerF64 peer_f64(const apop_ctx* c, size_t off) {
    PeerF64 P;
    for (int q = 0; q < kMaxWorld; ++q)
        P.p[q] = (q < c->world && c->fab.heap[q]) ? (const double*)(c->fab.heap[q] + off) : nullptr;
    return P;
}
static PeerI32 peer_i32(const apop_ctx* c, size_t off) {
    PeerI32 P;
    for (int q = 0; q < kMaxWorld; ++q)
        P.p[q] = (q < c->world && c->fab.heap[q]) ? (int32_t*)(c->fab.heap[q] + off) : nullptr;
    return P;
}

// ---- one sweep, split in phases so that an in-process group of ranks can be
//      driven in lock step from one host thread (begin / step t / end)
// phase A: allocations and host->device staging.  Everything that may block the
// host (cudaMalloc, pinned allocations, arena wrap-around) happens here, for ALL
// ranks of an in-process group, before any kernel that waits on a peer is
// enqueued — a blocking call issued while a peer's kernel spins would stall it
// until the spin times out.
static int sweep_stage(apop_ctx* c, const double* y, int T, const double* retained,
                       const apop_sweep_opts* o) {
    SweepState* s = c->sweep;
    APOP_TRY(ensure_resample_ws(c));
    if (o->no_resample || o->ess_threshold > 0.0) APOP_TRY(stat_lse_prepare(c));
    const bool csmc = o->csmc && retained;
    void* staged = nullptr;
    const size_t dim = (size_t)s->dim;
    APOP_TRY(param_upload(c, y, sizeof(double) * (size_t)T * dim, &staged));
    APOP_CUDA(c, cudaMemcpyAsync(s->y, staged, sizeof(double) * (size_t)T * dim,
                                 cudaMemcpyDeviceToDevice, c->stream));
    if (csmc) {
        APOP_TRY(param_upload(c, retained, sizeof(double) * (size_t)(T + 1) * dim, &staged));
        APOP_CUDA(c, cudaMemcpyAsync(s->retained, staged, sizeof(double) * (size_t)(T + 1) * dim,
                                     cudaMemcpyDeviceToDevice, c->stream));
    }
    return APOP_OK;
}

// phase B: first kernels of the sweep
template <int KIND>
static int sweep_begin(apop_ctx* c, const double* retained, const apop_sweep_opts* o) {
    SweepState* s = c->sweep;
    const int64_t n = c->n;
    ResampleWS& ws = c->ws;
    const bool csmc = o->csmc && retained;
    k_sweep_reset<<<1, 1, 0, c->stream>>>(s->sd, ws.info, ws.gmax, ws.ticket);
    s->sweep_seed = c->seed;
    s->sweep_csmc = csmc ? 1 : 0;
    if (s->hist_mode == 2)
        APOP_CUDA(c, cudaMemsetAsync(s->lin_alive, 0, (size_t)s->T * s->lin_w * 4, c->stream));
    // sharded: peers may still be reading this rank's buffers (final gather of
    // the previous sweep)
    APOP_TRY(fabric_barrier_enqueue(c));
    double* x0 = s->history ? s->x_hist : s->xa;
    if (KIND == APOP_SSM_LGMVN) {
        APOP_TRY(mvn_init_enqueue(c, x0, s->dim, s->mvn_s0, csmc ? s->retained : nullptr));
    } else {
        const int64_t pairs = (n + 1) / 2;
        k_init<(KIND == APOP_SSM_LGMVN) ? APOP_SSM_LGSSM1D : KIND>
            <<<(int)((pairs + TB - 1) / TB), TB, 0, c->stream>>>(
            s->model, x0, n, c->g0, c->seed, csmc ? s->retained : nullptr, s->seedp);
    }
    c->launches += 2;
    s->last_x = x0;
    s->last_anc = s->anc;
    return APOP_OK;
}

template <int KIND>
static int sweep_step(apop_ctx* c, int t, const apop_sweep_opts* o, bool csmc) {
    SweepState* s = c->sweep;
    const int64_t n = c->n;
    const int nt = (int)((n + TILE - 1) / TILE);
    ResampleWS& ws = c->ws;
    const bool shard = c->world > 1;
    const double* x_prev = s->last_x;
    const int32_t* anc_prev = s->last_anc;
    const size_t dim = (size_t)s->dim;
    double* x_new = s->history ? s->x_hist + (size_t)t * n * dim : ((t & 1) ? s->xb : s->xa);
    int32_t* anc_new = s->history ? s->anc_hist + (size_t)(t - 1) * n : s->anc;
    kt_begin(c, KT_PROPAGATE);
    if (KIND == APOP_SSM_LGMVN) {
        MvnSweepArgs ma;
        const size_t dd = dim * dim;
        ma.A = s->mvn_mats; ma.Lq = s->mvn_mats + dd; ma.iL = s->mvn_mats + 2 * dd;
        ma.Z = s->mvn_Z;
        ma.xp = x_prev; ma.anc = anc_prev; ma.first_step = (t == 1);
        ma.xn = x_new; ma.lw = s->lw; ma.y = s->y + (size_t)(t - 1) * dim;
        ma.retained = csmc ? s->retained + (size_t)t * dim : nullptr;
        ma.sd = s->sd; ma.n = n; ma.g0 = c->g0; ma.seed = c->seed; ma.t = (uint32_t)t;
        ma.world = c->world;
        memset(&ma.xpeer, 0, sizeof ma.xpeer);
        if (shard) ma.xpeer = peer_f64(c, (t & 1) ? c->fab.off_xa : c->fab.off_xb);   // x_{t-1}
        APOP_TRY(mvn_propagate_enqueue(c, s->dim, ma));
        c->launches -= 1;                 // counted below
    } else if (shard) {
        const PeerF64 xpeer = peer_f64(c, (t & 1) ? c->fab.off_xa : c->fab.off_xb);   // x_{t-1}
        k_propagate<(KIND == APOP_SSM_LGMVN) ? APOP_SSM_LGSSM1D : KIND, true, false><<<nt, TB, 0, c->stream>>>(
            s->model, x_prev, anc_prev, t == 1, x_new, s->lw, s->y + (t - 1), nullptr, s->sd, n,
            c->g0, c->seed, (uint32_t)t, ws.tmax, ws.tsum, ws.gmax, &ws.info->err, xpeer, 0);
    } else {
        PeerF64 none;
        memset(&none, 0, sizeof none);
        constexpr int K1 = (KIND == APOP_SSM_LGMVN) ? APOP_SSM_LGSSM1D : KIND;
        auto kern = s->seedp ? k_propagate<K1, false, true> : k_propagate<K1, false, false>;
        kern<<<nt, TB, 0, c->stream>>>(
            s->model, x_prev, anc_prev, t == 1 || o->no_resample, x_new, s->lw, s->y + (t - 1),
            csmc ? s->retained + t : nullptr, s->sd, n, c->g0, c->seed, (uint32_t)t, ws.tmax,
            ws.tsum, ws.gmax, &ws.info->err, none, (o->no_resample && t > 1) ? 1 : 0);
    }
    kt_end(c, KT_PROPAGATE);
    c->launches += 1;
    if (o->no_resample) {          // importance.cljc:16-22: weights accumulate per particle
        c->step = (uint32_t)t;
        s->last_x = x_new;
        return APOP_OK;
    }
    const bool cond = csmc && t <= o->n_cond_steps;
    c->step = (uint32_t)t;
    ResampleArgs a{};
    a.lw = s->lw; a.w = nullptr; a.n = n;
    a.n_new = shard ? c->n_global : (cond ? n - 1 : n);
    a.u = NAN; a.order = APOP_ORDER_SORTED;
    a.anc_out = anc_new; a.out_offset = cond ? 1 : 0;
    a.have_partials = (KIND != APOP_SSM_LGMVN);      // the mvn kernel leaves K2 to k_tile_lse
    a.xfilt = (o->want_filt && dim == 1) ? x_new : nullptr;
    a.rel_eps = (KIND != APOP_SSM_LGMVN) ? 2e-5 : 0.0;   // fp32 tile partials of k_propagate
    a.step.sd = s->sd; a.step.log_mean_w = s->log_mean_w; a.step.ess = s->ess;
    a.step.filt = s->filt; a.step.anc = anc_new; a.step.t = t;
    a.step.reset = o->reset_weights; a.step.cond = cond ? 1 : 0;
    a.step.has_x = (o->want_filt && dim == 1) ? 1 : 0;
    a.step.ess_thr = o->ess_threshold;
    a.sharded = shard;
    a.seedp = s->seedp;
    if (shard) a.anc_peer = peer_i32(c, c->fab.off_anc);
    APOP_TRY(resample_enqueue(c, a));
    if (o->ancestor_sampling && cond && KIND != APOP_SSM_LGMVN) {
        // PGAS: redraw the retained particle's ancestor among the N pre-resample
        // particles.  x_final is free until the sweep ends: it holds v for the draw.
        constexpr int K2 = (KIND == APOP_SSM_LGMVN) ? APOP_SSM_LGSSM1D : KIND;
        double* v = s->x_final;
        k_as_logw<K2><<<(int)((n + 255) / 256), 256, 0, c->stream>>>(
            s->model, x_new, s->lw, (t < s->T) ? s->retained + (t + 1) : nullptr, s->trans_const,
            v, n);
        ResampleArgs b{};
        b.lw = v; b.w = nullptr; b.n = n; b.n_new = 1;
        b.u = uniform_at(c->seed, 0, (uint32_t)t, STREAM_AUX, 0);
        b.order = APOP_ORDER_EMISSION; b.anc_out = s->as_anc; b.out_offset = 0;
        b.have_partials = false; b.keep_err = true; b.rel_eps = 0.0; b.sharded = false;
        APOP_TRY(resample_enqueue(c, b));
        k_as_commit<<<1, 1, 0, c->stream>>>(anc_new, s->as_anc);
        c->launches += 2;
    }
    if (s->hist_mode == 2) {
        k_pack_lineage<<<(int)((n + 255) / 256), 256, 0, c->stream>>>(
            anc_new, n, s->lin_new + (size_t)(t - 1) * s->lin_w,
            s->lin_alive + (size_t)(t - 1) * s->lin_w, s->lin_anc0 + (t - 1));
        c->launches += 1;
    }
    // sharded: offspring indices were stored into peers' arrays and the next
    // propagate reads parents from peers' x_t: all ranks meet here
    if (shard) {
        kt_begin(c, KT_MIGRATE);
        APOP_TRY(fabric_barrier_enqueue(c));
        kt_end(c, KT_MIGRATE);
    }
    s->last_x = x_new;
    s->last_anc = anc_new;
    return APOP_OK;
}

static int sweep_end(apop_ctx* c, int T, const apop_sweep_opts* opts, apop_sweep_result* result) {
    SweepState* s = c->sweep;
    const int64_t n = c->n;
    const int dim = s->dim;
    APOP_CUDA(c, cudaGetLastError());
    if (opts->no_resample) {
        // no resample, no gather: the final states are the last propagated ones
        if (opts->want_final || result->x_final)
            APOP_CUDA(c, cudaMemcpyAsync(s->x_final, s->last_x, sizeof(double) * n * dim,
                                         cudaMemcpyDeviceToDevice, c->stream));
        APOP_TRY(stat_lse_enqueue(c, s->lw, n, s->lse));
    } else if (opts->want_final || result->x_final) {
        if (opts->ess_threshold > 0.0) APOP_TRY(stat_lse_enqueue(c, s->lw, n, s->lse));
        int blocks = (int)((n * dim + 255) / 256);
        if (blocks > c->sm_count * 32) blocks = c->sm_count * 32;
        kt_begin(c, KT_GATHER);
        if (c->world > 1) {
            const PeerF64 src = peer_f64(c, (T & 1) ? c->fab.off_xb : c->fab.off_xa);    // x_T
            k_gather_shard<<<blocks, 256, 0, c->stream>>>(src, s->x_final, s->last_anc, n, dim);
        } else {
            k_gather_rows<<<blocks, 256, 0, c->stream>>>(s->last_x, s->x_final, s->last_anc, n, dim);
        }
        kt_end(c, KT_GATHER);
        c->launches += 1;
    }
    return APOP_OK;
}

// device time of the whole sweep (KT_SWEEP): events around the enqueued work
static void sweep_timer_begin(apop_ctx* c) {
    Timer& tm = c->timers[KT_SWEEP];
    if (!tm.open && cudaEventCreate(&tm.a) == cudaSuccess) {
        cudaEventRecord(tm.a, c->stream);
        tm.open = true;
    }
}
static void sweep_timer_end(apop_ctx* c) {
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
static int sweep_fetch(apop_ctx* c, int T, const apop_sweep_opts* opts, apop_sweep_result* result) {
    SweepState* s = c->sweep;
    const int64_t n = c->n;
    const int dim = s->dim;
    SweepDev hsd;
    int info_err = 0;          // errors raised after the chain kernel (k_ancestors consistency checks)
    APOP_CUDA(c, cudaMemcpyAsync(&hsd, s->sd, sizeof hsd, cudaMemcpyDeviceToHost, c->stream));
    APOP_CUDA(c, cudaMemcpyAsync(&info_err, &c->ws.info->err, sizeof(int), cudaMemcpyDeviceToHost,
                                 c->stream));
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
    double lse[2] = {0.0, 0.0};
    const bool adaptive = opts->ess_threshold > 0.0 && !opts->no_resample;
    if (opts->no_resample || adaptive) {
        APOP_CUDA(c, cudaMemcpyAsync(lse, s->lse, sizeof lse, cudaMemcpyDeviceToHost, c->stream));
        if (result->lw_final)
            APOP_CUDA(c, cudaMemcpyAsync(result->lw_final, s->lw, sizeof(double) * n,
                                         cudaMemcpyDeviceToHost, c->stream));
    }
    APOP_CUDA(c, cudaStreamSynchronize(c->stream));
    ResampleInfo I{};
    I.err = hsd.err | info_err;
    APOP_TRY(info_to_status(c, I));
    // :smc keeps the accumulated weight on the particles: final weight == log Z-hat
    result->log_z = opts->reset_weights ? hsd.log_z : hsd.base;
    if (opts->no_resample || (adaptive && hsd.carry)) {
        // log-marginal of the weighted samples (inference.cljc:181-186): log mean exp
        if (lse[0] != lse[0]) return set_err(c, APOP_ERR_NAN_WEIGHT, "importance: NaN log-weight");
        const double L = (lse[0] > -INFINITY && lse[0] < INFINITY)
                             ? log(lse[1] / (double)n) + lse[0] : lse[0];
        // adaptive sweep that ended without a resample: the particles still carry their
        // weights since the last one (reset mode: on top of the log Z accumulated so far)
        result->log_z = (adaptive && opts->reset_weights) ? hsd.log_z + L : L;
    } else if (adaptive && result->lw_final) {
        // the last observe resampled: all particles share one weight (smc.cljc:98-103)
        const double v = opts->reset_weights ? 0.0 : hsd.base;
        for (int64_t k = 0; k < n; ++k) result->lw_final[k] = v;
    }
    return APOP_OK;
}

__global__ void k_set_sweep_seed(SweepDev* sd, unsigned long long seed) { sd->seed = seed; }

// Small populations are bound by launch latency, not by the kernels (BASELINE config 2:
// ~10 launches per step for 8 MB of state), so the whole sweep - k_sweep_reset, k_init,
// every step's kernels, the final gather - is captured ONCE as a CUDA graph and replayed
// for every later sweep of the same shape (infer :smc / :pimh / :pgibbs / IPMCMC nodes run
// thousands of them).  What changes between sweeps lives in device memory: observations,
// retained path (staged before the launch) and the RNG key (SweepDev::seed).
// opts->use_graph: 0 auto (populations up to 2^22 particles), 1 always, -1 never;
// APOP_GRAPH=0 in the environment disables it.
static bool graph_wanted(const apop_ctx* c, int kind, const apop_sweep_opts* o) {
    if (o->use_graph < 0 || c->world > 1 || c->ktimers || o->ancestor_sampling ||
        kind == APOP_SSM_LGMVN)
        return false;
    // read once; a function-local static initialiser is thread-safe (populations are driven
    // from several host threads)
    static const bool env = [] { const char* e = getenv("APOP_GRAPH"); return !(e && e[0] == '0'); }();
    if (!env) return false;
    return o->use_graph > 0 || c->n <= (1LL << 22);
}

template <int KIND>
static int sweep_body(apop_ctx* c, int T, const double* retained, const apop_sweep_opts* o,
                      apop_sweep_result* result) {
    APOP_TRY(sweep_begin<KIND>(c, retained, o));
    const bool csmc = o->csmc && retained;
    for (int t = 1; t <= T; ++t) APOP_TRY(sweep_step<KIND>(c, t, o, csmc));
    return sweep_end(c, T, o, result);
}

template <int KIND>
static int run_graph(apop_ctx* c, int T, const double* retained, const apop_sweep_opts* o,
                     apop_sweep_result* result) {
    SweepState* s = c->sweep;
    SweepState::GraphKey k = {};
    k.kind = KIND; k.T = T; k.hist_mode = s->hist_mode; k.csmc = (o->csmc && retained) ? 1 : 0;
    k.n_cond = o->n_cond_steps; k.reset = o->reset_weights;
    k.want_final = (o->want_final || result->x_final) ? 1 : 0;
    k.want_filt = o->want_filt; k.no_resample = o->no_resample; k.n = c->n;
    k.xa = s->xa; k.xb = s->xb; k.lw = s->lw; k.anc = s->anc; k.y = s->y; k.ret = s->retained;
    k.lin = s->hist_mode == 2 ? (const void*)s->lin_new : (const void*)s->x_hist;
    k.xf = s->x_final; k.ws = c->ws.slab; k.lse = s->lse; k.sd = s->sd; k.model = s->model;
    memcpy(&k.ess_bits, &o->ess_threshold, sizeof k.ess_bits);
    k_set_sweep_seed<<<1, 1, 0, c->stream>>>(s->sd, c->seed);
    s->sweep_seed = c->seed;
    if (!s->gexec || !k.same(s->gkey)) {
        if (s->gexec) { cudaGraphExecDestroy(s->gexec); s->gexec = nullptr; }
        cudaGraph_t graph = nullptr;
        const int64_t l0 = c->launches;
        APOP_CUDA(c, cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
        s->seedp = &s->sd->seed;
        const int rc = sweep_body<KIND>(c, T, retained, o, result);
        s->seedp = nullptr;
        const cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
        if (rc != APOP_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
        if (e != cudaSuccess || !graph)
            return set_err(c, APOP_ERR_CUDA, "sweep graph capture failed: %s", cudaGetErrorString(e));
        const cudaError_t ei = cudaGraphInstantiate(&s->gexec, graph, 0);
        cudaGraphDestroy(graph);
        if (ei != cudaSuccess) {
            s->gexec = nullptr;
            return set_err(c, APOP_ERR_CUDA, "sweep graph instantiation failed: %s", cudaGetErrorString(ei));
        }
        s->gkey = k;
        s->glaunches = c->launches - l0;
        c->launches = l0;
    }
    sweep_timer_begin(c);
    APOP_CUDA(c, cudaGraphLaunch(s->gexec, c->stream));
    sweep_timer_end(c);
    c->launches += s->glaunches;
    c->step = (uint32_t)T;
    s->sweep_csmc = k.csmc;
    return APOP_OK;
}

template <int KIND>
static int run_group(apop_ctx* const* g, int ng, const double* y, int T, const double* retained,
                     const apop_sweep_opts* o, apop_sweep_result* results) {
    for (int i = 0; i < ng; ++i) {
        cudaSetDevice(g[i]->device);
        APOP_TRY(sweep_stage(g[i], y, T, retained, o));
    }
    if (ng == 1 && graph_wanted(g[0], KIND, o)) {
        APOP_TRY(run_graph<KIND>(g[0], T, retained, o, &results[0]));
        return sweep_fetch(g[0], T, o, &results[0]);
    }
    for (int i = 0; i < ng; ++i) {
        if (ng > 1) cudaSetDevice(g[i]->device);
        sweep_timer_begin(g[i]);
        APOP_TRY(sweep_begin<KIND>(g[i], retained, o));
    }
    const bool csmc = o->csmc && retained;
    for (int t = 1; t <= T; ++t)
        for (int i = 0; i < ng; ++i) {
            if (ng > 1) cudaSetDevice(g[i]->device);
            APOP_TRY(sweep_step<KIND>(g[i], t, o, csmc));
        }
    for (int i = 0; i < ng; ++i) {
        if (ng > 1) cudaSetDevice(g[i]->device);
        APOP_TRY(sweep_end(g[i], T, o, &results[i]));
        sweep_timer_end(g[i]);
    }
    int rc = APOP_OK;
    for (int i = 0; i < ng; ++i) {
        if (ng > 1) cudaSetDevice(g[i]->device);
        const int r = sweep_fetch(g[i], T, o, &results[i]);
        if (r != APOP_OK && rc == APOP_OK) rc = r;
    }
    return rc;
}

// CUDA loads kernels lazily at their first launch, and that load waits for the
// device to go idle: with a peer's kernel spinning on a mailbox it would wait
// for the spin to time out.  Sharded sweeps therefore resolve every kernel they
// will launch before the first one is enqueued.
template <typename K>
static void preload(K kernel) {
    cudaFuncAttributes a;
    cudaFuncGetAttributes(&a, kernel);
}
void preload_resample_kernels();     // resample.cu
void preload_comm_kernels();         // comm.cu
static void preload_sweep_kernels() {
    preload(k_init<APOP_SSM_LGSSM1D>); preload(k_init<APOP_SSM_SV>); preload(k_init<APOP_SSM_HMM>);
    preload(k_propagate<APOP_SSM_LGSSM1D, true, false>); preload(k_propagate<APOP_SSM_SV, true, false>);
    preload(k_propagate<APOP_SSM_HMM, true, false>);
    preload(k_sweep_reset); preload(k_gather_shard);
    preload_resample_kernels();
    preload_comm_kernels();
    cudaGetLastError();
}

// ---- how many SHARDED sweeps this process lets run at once -----------------
// Kernels of a sharded sweep wait for their peers inside the kernel, with a bounded spin.
// Measured on 8 B200: up to 4 sharded populations in flight complete; 6 stall; with the final
// states gathered and downloaded after every sweep even 2 ran into the 3 s bound.  Until one
// communication kernel per rank serves every population, that envelope is enforced here
// instead of being left to the caller to find out through APOP_ERR_NCCL timeouts:
//   at most APOP_SHARDED_MAX_INFLIGHT (default 4) sharded sweeps at a time, and a sweep that
//   returns final states runs alone.
// In-process groups WAIT for their turn: one process drives every rank of the sweep, so an
// admitted sweep always finishes and frees its place.  With one process per GPU waiting is
// not safe - two ranks could admit different populations and wait for each other for ever -
// so the sweep is REFUSED with a message and the caller orders its sweeps (as bench.py does).
// Unsharded populations (IPMCMC nodes) never wait on peers and are not counted.
// APOP_SHARDED_GATE=0 switches the check off (for reproducing the stall).
struct ShardedGate {
    std::mutex mu;
    std::condition_variable cv;
    int inflight = 0, with_final = 0;
};
static ShardedGate g_sharded_gate;
static int sharded_max_inflight() {
    static const int v = [] {
        const char* g = getenv("APOP_SHARDED_GATE");
        if (g && g[0] == '0') return 0;                       // 0 = gate off
        const char* e = getenv("APOP_SHARDED_MAX_INFLIGHT");
        const int k = e ? atoi(e) : 4;
        return k >= 1 ? k : 4;
    }();
    return v;
}
struct ShardedAdmission {
    bool held = false, fin = false;
    // returns false when the sweep may not start now (only when may_wait is false)
    bool enter(bool want_final, bool may_wait) {
        const int cap = sharded_max_inflight();
        if (cap == 0) return true;
        ShardedGate& G = g_sharded_gate;
        std::unique_lock<std::mutex> lk(G.mu);
        auto ok = [&] {
            return G.with_final == 0 && G.inflight < cap && (!want_final || G.inflight == 0);
        };
        if (!ok()) {
            if (!may_wait) return false;
            G.cv.wait(lk, ok);
        }
        G.inflight += 1;
        if (want_final) G.with_final += 1;
        held = true; fin = want_final;
        return true;
    }
    ~ShardedAdmission() {
        if (!held) return;
        ShardedGate& G = g_sharded_gate;
        { std::lock_guard<std::mutex> lk(G.mu); G.inflight -= 1; if (fin) G.with_final -= 1; }
        G.cv.notify_all();
    }
};

static int sweep_group(apop_ctx* const* g, int ng, const apop_ssm* model, const double* y, int T,
                       const double* retained, const apop_sweep_opts* opts,
                       apop_sweep_result* results) {
    apop_ctx* c0 = g[0];
    if (!model || !y || T < 1 || !opts || !results)
        return set_err(c0, APOP_ERR_INVALID, "smc_sweep: bad arguments");
    ShardedAdmission admission;
    if (c0->world > 1) {
        bool fin = opts->want_final != 0;
        for (int i = 0; i < ng; ++i) fin = fin || results[i].x_final;
        if (!admission.enter(fin, /*may_wait=*/c0->fab.in_process))
            return set_err(c0, APOP_ERR_UNSUPPORTED,
                           "sharded sweep refused: another sharded sweep of this process is in "
                           "flight (at most %d at a time, and a sweep that returns final states "
                           "runs alone). Their kernels wait on peer GPUs with a bounded spin and "
                           "stall each other beyond that; run them one after another, or shard "
                           "fewer populations per GPU (unsharded populations are not limited)",
                           sharded_max_inflight());
    }
    const int dim = (model->kind == APOP_SSM_LGMVN) ? model->dim : 1;
    for (int i = 0; i < ng; ++i) {
        apop_ctx* c = g[i];
        if (opts->csmc && c->n < 2)
            return set_err(c, APOP_ERR_INVALID, "conditional sweep needs >= 2 particles");
        if (opts->no_resample && (opts->csmc || opts->keep_history || c->world > 1 ||
                                  model->kind == APOP_SSM_LGMVN))
            return set_err(c, APOP_ERR_UNSUPPORTED,
                           "no_resample (importance) sweeps: unconditional, unsharded, scalar-state, "
                           "no lineages");
        if (opts->ess_threshold < 0.0 || opts->ess_threshold > 1.0 ||
            (opts->ess_threshold > 0.0 && (opts->ancestor_sampling || c->world > 1 ||
                                           model->kind == APOP_SSM_LGMVN || !opts->want_final)))
            return set_err(c, APOP_ERR_UNSUPPORTED,
                           "ess_threshold (adaptive resampling): a fraction of N in [0, 1]; scalar-"
                           "state models on one GPU, final states requested, no ancestor sampling");
        if (opts->ancestor_sampling && (!opts->csmc || c->world > 1 || opts->no_resample ||
                                        model->kind == APOP_SSM_LGMVN))
            return set_err(c, APOP_ERR_UNSUPPORTED,
                           "ancestor sampling (pgas): conditional sweeps of scalar-state models "
                           "on one GPU");
        if (opts->keep_history < 0 || opts->keep_history > 2 ||
            (opts->keep_history == 2 && model->kind == APOP_SSM_LGMVN))
            return set_err(c, APOP_ERR_UNSUPPORTED,
                           "keep_history: 0 none, 1 full, 2 compact (scalar-state models)");
        if (c->world > 1 && (opts->csmc || opts->keep_history))
            return set_err(c, APOP_ERR_UNSUPPORTED,
                           "sharded sweep: conditional sweeps / lineages run one population "
                           "per GPU (IPMCMC nodes), not sharded");
        if (c->world > 1 && (c->n % 8) != 0)
            return set_err(c, APOP_ERR_INVALID, "sharded sweep: particles per rank must be a "
                                                "multiple of 8");
        cudaSetDevice(c->device);
        if (c->world > 1) { preload_sweep_kernels(); if (dim > 1) mvn_preload_kernels(dim); }
        APOP_TRY(sweep_prepare(c, T, dim, opts->keep_history));
        c->sweep->kind = model->kind;
        APOP_TRY(upload_model(c, model));
    }
    switch (model->kind) {
    case APOP_SSM_LGSSM1D:
        return run_group<APOP_SSM_LGSSM1D>(g, ng, y, T, retained, opts, results);
    case APOP_SSM_SV:
        return run_group<APOP_SSM_SV>(g, ng, y, T, retained, opts, results);
    case APOP_SSM_HMM:
        return run_group<APOP_SSM_HMM>(g, ng, y, T, retained, opts, results);
    case APOP_SSM_LGMVN:
        return run_group<APOP_SSM_LGMVN>(g, ng, y, T, retained, opts, results);
    default:
        return set_err(c0, APOP_ERR_UNSUPPORTED, "model kind %d not supported by the fused sweep",
                       model->kind);
    }
}

} // namespace apop

using namespace apop;

extern "C" {

int apop_smc_sweep(apop_ctx* c, const apop_ssm* model, const double* y, int32_t T,
                   const double* retained, const apop_sweep_opts* opts,
                   apop_sweep_result* result) {
    if (!c) return APOP_ERR_INVALID;
    if (c->world > 1 && c->fab.in_process)
        return set_err(c, APOP_ERR_INVALID, "in-process group: call apop_smc_sweep_group");
    apop_ctx* g[1] = {c};
    return sweep_group(g, 1, model, y, T, retained, opts, result);
}

int apop_smc_sweep_group(apop_ctx* const* ctxs, int32_t n_ctx, const apop_ssm* model,
                         const double* y, int32_t T, const apop_sweep_opts* opts,
                         apop_sweep_result* results) {
    if (!ctxs || n_ctx < 1 || n_ctx > kMaxWorld || !ctxs[0]) return APOP_ERR_INVALID;
    for (int i = 0; i < n_ctx; ++i)
        if (!ctxs[i] || ctxs[i]->world != n_ctx || ctxs[i]->rank != i || !ctxs[i]->fab.in_process)
            return set_err(ctxs[0], APOP_ERR_INVALID,
                           "sweep_group: pass all ranks of one apop_link_group, in rank order");
    return sweep_group(ctxs, n_ctx, model, y, T, nullptr, opts, results);
}

int apop_trace_lineage(apop_ctx* c, int64_t slot, double* out_traj) {
    if (!c) return APOP_ERR_INVALID;
    SweepState* s = c->sweep;
    if (!s || !(s->history || s->hist_mode == 2) || !out_traj)
        return set_err(c, APOP_ERR_INVALID, "trace_lineage: no keep_history sweep available");
    if (slot < 0 || slot >= s->n) return set_err(c, APOP_ERR_INVALID, "trace_lineage: bad slot");
    if (s->hist_mode == 2) {
        const double* ret = s->sweep_csmc ? s->retained : nullptr;
#define TRACE(K)                                                                              \
    k_trace_compact<K><<<1, TRB, 0, c->stream>>>(s->model, s->lin_new, s->lin_alive, s->lin_anc0, \
                                                 (int)s->lin_w, s->T, slot, c->g0, s->sweep_seed, \
                                                 ret, s->lin_idx, s->traj)
        if (s->kind == APOP_SSM_HMM) TRACE(APOP_SSM_HMM);
        else if (s->kind == APOP_SSM_SV) TRACE(APOP_SSM_SV);
        else TRACE(APOP_SSM_LGSSM1D);
#undef TRACE
    } else
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
        return set_err(c, APOP_ERR_INVALID, "download_history: no full-history (keep_history = 1) "
                                            "sweep available");
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

int apop_sizeof_sweep_opts(void) { return (int)sizeof(apop_sweep_opts); }

int apop_sweep_final_device(apop_ctx* c, void** out) {
    if (!c || !out) return APOP_ERR_INVALID;
    if (!c->sweep || !c->sweep->x_final) return set_err(c, APOP_ERR_INVALID, "no sweep yet");
    *out = c->sweep->x_final;
    return APOP_OK;
}

} // extern "C"
