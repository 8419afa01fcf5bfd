/*
 * anglican_pop.h — C ABI of libanglican_pop.so, the B200-native particle-
 * population engine that sits behind Anglican's `infer` multimethod
 * (reference src/anglican/inference.cljc:25-28) for :smc / :pimh / :pgibbs /
 * :ipmcmc, and behind the `anglican.runtime` distribution protocol
 * (`sample*` / `observe*`, src/anglican/runtime.cljc:119-124).
 *
 * Conventions
 *  - extern "C", plain pointers and sizes, no C++/torch types.
 *  - every function returns an int status (APOP_OK == 0, errors < 0) and never
 *    throws; apop_last_error(ctx) gives the message of the last failure.
 *  - an apop_ctx owns one CUDA stream, (optionally) one NCCL communicator and
 *    all device memory of one particle population shard.  One ctx per host
 *    thread; ctxs are independent (IPMCMC: one ctx per node).
 *  - particle data live in device-resident columns addressed by integer
 *    handles (SoA; f64 or i32; `width` values per particle, row-major).
 *  - host scalars in/out are caller-owned; an output pointer may be NULL when
 *    the value is not wanted (then the call does not synchronise).
 *  - there is NO CPU fallback: without a CUDA device every compute entry point
 *    fails with APOP_ERR_CUDA.
 *
 * The reference is pure Clojure and has no FFI; INTEGRATION.md shows the
 * JNI/Panama binding a maintainer would add (clj/anglican/gpu.clj sketch).
 */
#ifndef ANGLICAN_POP_H
#define ANGLICAN_POP_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct apop_ctx apop_ctx;
typedef int32_t apop_col;

enum {
    APOP_OK = 0,
    APOP_ERR_INVALID = -1,      /* bad argument / handle */
    APOP_ERR_CUDA = -2,         /* CUDA runtime failure or no device */
    APOP_ERR_NAN_WEIGHT = -3,   /* a log-weight is NaN (reference: resample never terminates) */
    APOP_ERR_ZERO_WEIGHT = -4,  /* total weight not > 0 or not finite */
    APOP_ERR_NCCL = -5,
    APOP_ERR_OOM = -6,
    APOP_ERR_UNSUPPORTED = -7,
    APOP_ERR_INTERNAL = -8     /* internal consistency check failed */
};

enum { APOP_F64 = 0, APOP_I32 = 1 };

/* resampling schemes: systematic = smc.cljc:86-137; multinomial = n_new
 * independent `discrete` draws (runtime.cljc:277-287, inference.cljc:144-147) */
enum { APOP_SYSTEMATIC = 0, APOP_MULTINOMIAL = 1 };

/* output order of ancestor indices.
 * REFERENCE: the list returned by smc.cljc resample (offspring prepended:
 * emission order reversed).  SORTED: the engine's physical order (ancestors
 * monotone non-decreasing: second-lap emissions first, then first-lap). */
enum { APOP_ORDER_REFERENCE = 0, APOP_ORDER_SORTED = 1, APOP_ORDER_EMISSION = 2 };

/* A distribution parameter / observed value: per-particle column or scalar
 * broadcast to all particles (col < 0 => scalar). */
typedef struct {
    apop_col col;
    double scalar;
} apop_arg;

/* ---- lifecycle ---------------------------------------------------------
 * A context owns one particle population: the `particles` sequence that `sweep` builds with
 * `(repeatedly number-of-particles ...)` and threads through its loop (smc.cljc:37-43), as
 * device-resident columns, plus the stream and the RNG key that replace the reference's one
 * global synchronised generator (runtime.cljc:142-147). */
int apop_version(void);
int apop_device_count(void);
/* n_particles_local: this rank's share of `:number-of-particles` (smc.cljc:53-57; >= 1).
 * nccl_id: 128-byte ncclUniqueId shared by all ranks, or NULL when world==1 */
int apop_create(apop_ctx** out, int64_t n_particles_local, uint64_t seed, int device,
                int rank, int world, const void* nccl_id);
/* In-process group: create `world` contexts with ranks 0..world-1 and
 * nccl_id == NULL (same or different devices), then link them; their shards
 * then form ONE population driven through apop_smc_sweep_group.  (The
 * one-process-per-GPU form needs no link call: the NCCL id is enough.) */
int apop_link_group(apop_ctx* const* ctxs, int32_t world);
int apop_destroy(apop_ctx* ctx);
const char* apop_last_error(apop_ctx* ctx);
int apop_get_unique_id(void* out_128_bytes);
int apop_synchronize(apop_ctx* ctx);
int64_t apop_num_particles(apop_ctx* ctx);
/* re-key the counter-based RNG (one key per sweep makes sweeps independent, as successive
 * sweeps of `infer` are: smc.cljc:58-63).  Replaces reseeding runtime.cljc:142-147. */
int apop_set_seed(apop_ctx* ctx, uint64_t seed);
/* Page-locked host buffers for results that the caller reads back every sweep
 * (the final particle states: 8 B x n).  A JVM binding allocates its result
 * segments here once (Panama: MemorySegment.ofAddress; JNI: NewDirectByteBuffer)
 * so that device->host copies run at PCIe speed instead of through the
 * driver's pageable staging.  Replaces nothing in the reference (its states are
 * JVM heap objects, smc.cljc:127-129); free with apop_host_free. */
int apop_host_alloc(apop_ctx* ctx, void** out, uint64_t bytes);
int apop_host_free(apop_ctx* ctx, void* p);
/* number of kernels this ctx has launched so far (bench `gpu_launches`) */
int64_t apop_launch_count(apop_ctx* ctx);
/* raw CUDA stream (cudaStream_t) so callers can record events on it */
void* apop_stream(apop_ctx* ctx);
/* elapsed device ms of the last timed region for tag (see apop_timer_*) */
int apop_timer_begin(apop_ctx* ctx, int tag);
int apop_timer_end(apop_ctx* ctx, int tag);
int apop_timer_elapsed_ms(apop_ctx* ctx, int tag, double* total_ms, int64_t* count);
int apop_timer_reset(apop_ctx* ctx);
/* bracket every kernel class of the sweep / resample pipeline with CUDA events
 * on the ctx stream.  tags: 0 sweep, 1 propagate, 2 tile_lse, 3 finalize,
 * 4 fold_agg, 5 resolve, 6 ancestors, 7 step_stats, 8 gather, 9 migrate */
int apop_enable_kernel_timers(apop_ctx* ctx, int on);

/* ---- columns (state.cljc:9-36: the :log-weight field becomes a column) -- */
int apop_column_alloc(apop_ctx* ctx, int dtype, int width, apop_col* out);
int apop_column_free(apop_ctx* ctx, apop_col col);
int apop_column_upload(apop_ctx* ctx, apop_col col, const void* host, int64_t n_values);
int apop_column_download(apop_ctx* ctx, apop_col col, void* host, int64_t n_values);
int apop_column_fill(apop_ctx* ctx, apop_col col, double value);
int apop_column_device_ptr(apop_ctx* ctx, apop_col col, void** out);

/* ---- runtime distributions (runtime.cljc) ------------------------------
 * sample:  out[k] <- draw            (inference.cljc:46-47)
 * observe: logw[k] += log p(value)   (smc.cljc:23-27, state.cljc:27-31)
 * `draw_id` distinguishes successive sample statements of one program step
 * (counter-based RNG: key = seed, counter = (slot, step, stream, draw)). */
int apop_set_step(apop_ctx* ctx, uint32_t step);
/* normal (runtime.cljc:403-409 -> commons-math NormalDistribution): mean, sd */
int apop_sample_normal(apop_ctx* ctx, apop_arg mean, apop_arg sd, uint32_t draw_id, apop_col out);
int apop_observe_normal(apop_ctx* ctx, apop_arg mean, apop_arg sd, apop_arg value, apop_col logw);
/* gamma (runtime.cljc:325-330 -> GammaDistribution(shape, 1/rate)): shape, RATE */
int apop_sample_gamma(apop_ctx* ctx, apop_arg shape, apop_arg rate, uint32_t draw_id, apop_col out);
int apop_observe_gamma(apop_ctx* ctx, apop_arg shape, apop_arg rate, apop_arg value, apop_col logw);
/* beta (runtime.cljc:251-256 -> BetaDistribution) */
int apop_sample_beta(apop_ctx* ctx, apop_arg alpha, apop_arg beta, uint32_t draw_id, apop_col out);
int apop_observe_beta(apop_ctx* ctx, apop_arg alpha, apop_arg beta, apop_arg value, apop_col logw);
/* uniform-continuous (runtime.cljc:450-453 -> UniformRealDistribution): min, max */
int apop_sample_uniform(apop_ctx* ctx, apop_arg lo, apop_arg hi, uint32_t draw_id, apop_col out);
int apop_observe_uniform(apop_ctx* ctx, apop_arg lo, apop_arg hi, apop_arg value, apop_col logw);
/* discrete / categorical: weights either shared (host array of k doubles,
 * weights_col < 0) or per particle (f64 column of width k).  Values are i32
 * indices; out-of-range / -1 observe gives -inf (runtime.cljc:266-290). */
int apop_sample_discrete(apop_ctx* ctx, const double* weights, int k, apop_col weights_col,
                         uint32_t draw_id, apop_col out_i32);
int apop_observe_discrete(apop_ctx* ctx, const double* weights, int k, apop_col weights_col,
                          apop_col value_i32, int32_t value_scalar, apop_col logw);
/* dirichlet (runtime.cljc:292-306): alpha shared (host, k); values are f64
 * columns of width k */
int apop_sample_dirichlet(apop_ctx* ctx, const double* alpha, int k, uint32_t draw_id, apop_col out);
int apop_observe_dirichlet(apop_ctx* ctx, const double* alpha, int k, apop_col value, apop_col logw);
/* mvn (runtime.cljc:349-368): cov is d*d row-major on the host (upper
 * triangle read); mean/value: shared host vector (col < 0) or f64 column of
 * width d. */
int apop_sample_mvn(apop_ctx* ctx, const double* mean, apop_col mean_col, const double* cov,
                    int d, uint32_t draw_id, apop_col out);
int apop_observe_mvn(apop_ctx* ctx, const double* mean, apop_col mean_col, const double* cov,
                     int d, const double* value, apop_col value_col, apop_col logw);

/* multivariate-t (runtime.cljc:371-401): nu degrees of freedom, mean mu (host vector or f64
 * column of width d), scale matrix sigma (d*d row-major on the host, upper triangle read). */
int apop_sample_mvt(apop_ctx* ctx, double nu, const double* mean, apop_col mean_col,
                    const double* sigma, int d, uint32_t draw_id, apop_col out);
int apop_observe_mvt(apop_ctx* ctx, double nu, const double* mean, apop_col mean_col,
                     const double* sigma, int d, const double* value, apop_col value_col,
                     apop_col logw);
/* wishart (runtime.cljc:479-545): n degrees of freedom, scale V (p*p on the host, p <= 16);
 * values are f64 columns of width p*p (row-major matrices).  Sampling is the Bartlett
 * decomposition the reference uses. */
int apop_sample_wishart(apop_ctx* ctx, double n_df, const double* V, int p, uint32_t draw_id,
                        apop_col out);
int apop_observe_wishart(apop_ctx* ctx, double n_df, const double* V, int p, apop_col value,
                         apop_col logw);
/* per-particle table lookup: out[k][:] = table[index[k]][:] (table: rows*width doubles on the
 * host, index: i32 column, out: f64 column of width `width`).  This is how a traced program
 * turns a discrete state into distribution parameters - `(nth transition-rows state)`. */
int apop_column_take(apop_ctx* ctx, const double* table, int rows, int width, apop_col index_i32,
                     apop_col out);

/* self-test hook for the engine's own fp64 math (csrc/fastmath.cuh), used by
 * tests/test_dists_gpu.py: out[k] = f(in[k]) with fn 0 exp, 1 log, 2 sqrt,
 * 3 sin(pi x), 4 cos(pi x), 5 1/x. */
int apop_selftest_math(apop_ctx* ctx, int fn, apop_col in, apop_col out);

/* diagnostics of the last resample of this ctx: out8 = {ns in lap 0 of the
 * exact chain, ns in sums/exchange/comb table, ns in lap 1, comb segments,
 * hard tiles lap 0, hard tiles lap 1, first skipped tile of lap 1, J} */
int apop_debug_resolve_phases(apop_ctx* ctx, double* out8);
/* per hard-tile event of the last chain: out512[lap*256 + 2h] = SM cycles spent folding
 * hard tile h, out512[lap*256 + 2h + 1] = cycle counter when it began */
int apop_debug_chain_trace(apop_ctx* ctx, double* out512);

/* ---- weights: particle-weights / log-sum-exp / ESS ---------------------
 * smc.cljc:70-84, runtime.cljc:130-138, inference.cljc:181-186,
 * test/anglican/dist_test.cljc:83-94.
 *   max_lw      = max_k logw[k]
 *   log_mean_w  = log(sum_k exp(logw[k]-max)/N) + max   (== log-marginal)
 *   ess         = (sum w)^2 / sum w^2 */
int apop_normalize(apop_ctx* ctx, apop_col logw, double* max_lw, double* log_mean_w, double* ess);

/* ---- resample (smc.cljc:86-137) ----------------------------------------
 * u: the uniform of `(rand total-weight)` (inference.cljc:124-129); NaN =>
 *    drawn from the ctx stream for the current step.
 * n_new: number of offspring (N, or N-1 for conditional SMC, pgibbs.cljc:92-99)
 * anc_out: i32 column, receives n_new indices in `order`.
 * new_logw: log(step)+max, or -inf when max is not finite (smc.cljc:98-103).
 * n_first_lap: #offspring emitted before the wheel wrapped (defines the
 *    closed-form permutation between the three orders).
 * Ancestors are bit-identical to the reference's sequential fp64 wheel on the
 * same weights and uniform. */
int apop_resample(apop_ctx* ctx, apop_col logw, int scheme, double u, int64_t n_new, int order,
                  apop_col anc_out, double* new_logw, int64_t* n_first_lap);
/* same on explicit fp64 weights (already max-normalised or not): the wheel of smc.cljc:105-137
 * on the `weights` that `particle-weights` (smc.cljc:70-84) would have returned */
int apop_resample_weights(apop_ctx* ctx, apop_col weights, int scheme, double u, int64_t n_new,
                          int order, apop_col anc_out, double* total, int64_t* n_first_lap);
/* multinomial with one supplied uniform per offspring (f64 column): n_new independent
 * `(sample* (discrete weights))` draws, runtime.cljc:279-287 (first edge exceeding u * total,
 * strict `<`), the draw pgas.cljc:120-125 uses for its ancestor */
int apop_resample_multinomial_u(apop_ctx* ctx, apop_col weights, apop_col uniforms,
                                int64_t n_new, apop_col anc_out);
/* x_new[k] = x[anc[k]] for each listed column: the offspring copies of smc.cljc:124-129
 * (`(update-in particle [:state] set-log-weight log-weight)` per chosen particle).  The result is written to the column's second
 * buffer and the buffers are swapped: a pointer from apop_column_device_ptr is stale after this. */
int apop_gather(apop_ctx* ctx, apop_col anc, int64_t n_new, const apop_col* cols, int ncols);

/* ---- fused state-space sweeps (the benchmark fast path) -----------------
 * One call = one `sweep` (smc.cljc:35-51 / pgibbs.cljc:76-110 /
 * ipmcmc.cljc:34-83) of a program of the shape
 *   x0 <- sample; for y_t: x_t <- sample(.|x_{t-1}); observe(.|x_t) y_t
 * with every particle at the same global observe (smc.cljc:41-51). */
enum { APOP_SSM_LGSSM1D = 1, APOP_SSM_SV = 2, APOP_SSM_HMM = 3, APOP_SSM_LGMVN = 4 };

typedef struct {
    int32_t kind;
    int32_t dim;               /* state dimension; K for HMM */
    double m0, s0, a, b, q;    /* x0~N(m0,s0); x' = a x + b + q e */
    double c, d, r, beta;      /* LGSSM1D: y~N(c x+d, r);  SV: y~N(0, beta exp(x/2)) */
    const double* hmm_init;    /* [K] */
    const double* hmm_trans;   /* [K*K] row major */
    const double* hmm_mu;      /* [K] */
    const double* hmm_sd;      /* [K] */
    const double* A;           /* [d*d] */
    const double* Lq;          /* [d*d] lower factor of process cov */
    const double* R;           /* [d*d] observation cov */
} apop_ssm;

typedef struct {
    int32_t csmc;              /* conditional sweep: retained particle in slot 0 */
    int32_t n_cond_steps;      /* observes using the conditional resample */
    int32_t reset_weights;     /* pgibbs.cljc:88 */
    int32_t keep_history;      /* lineages for apop_trace_lineage: 0 none; 1 full history (ancestors
                                * + states, 12 B per particle and step, also downloadable);
                                * 2 compact (scalar-state models): two bits per particle and
                                * step, the states of the ONE traced lineage are recomputed
                                * from the counter-based RNG (pgibbs.cljc:29-35, 122-123) */
    int32_t use_graph;         /* the whole sweep as ONE CUDA graph, captured at the first sweep of a
                                * shape and replayed afterwards (small populations are bound by
                                * launch latency): 0 auto (up to 2^22 particles, one GPU, scalar
                                * state), 1 always, -1 never.  Results are bit-identical. */
    int32_t want_final;        /* gather final states into x_final */
    int32_t want_filt;         /* per-step filtering means sum w x / sum w (reads x once more) */
    int32_t no_resample;       /* `infer :importance` (importance.cljc:16-22): every observe only
                                * adds to the particle's own log-weight (inference.cljc:42-45),
                                * nothing is resampled; log_z = log mean exp of the final
                                * log-weights, which lw_final returns */
    int32_t ancestor_sampling; /* `infer :pgas` (pgas.cljc:107-168), conditional sweeps only: at
                                * every conditional observe the retained particle's ancestor is
                                * redrawn, one `(sample* (discrete weights))` with weights
                                * exp(v - max v), v_i = logw_i + log f(retained_{t+1} | x_t^i)
                                * (the rescored transition `sample`; the retained path's later
                                * terms are common to all i and left out).  Scalar-state models,
                                * one GPU.  Only the recorded ancestry changes (lineages). */
    int32_t pad_;
    double ess_threshold;      /* adaptive resampling (the reference resamples at
                                * EVERY observe, smc.cljc:41-43 - this is an extension): 0 = always
                                * resample; in (0, 1]: resample only when ESS < ess_threshold * N,
                                * otherwise the particles keep their own accumulated log-weights
                                * and are their own ancestors.  A sweep that ends without a
                                * resample returns weighted particles (lw_final).  The decision is
                                * taken on the device; scalar-state models, one GPU. */
} apop_sweep_opts;
/* sizeof(apop_sweep_opts) of the loaded library, for bindings that mirror the struct */
int apop_sizeof_sweep_opts(void);

typedef struct {
    double log_z;              /* sum of per-step log-mean-weights */
    double* log_mean_w;        /* [T] or NULL */
    double* ess;               /* [T] or NULL */
    double* filt_mean;         /* [T*dim] or NULL */
    double* x_final;           /* [n*dim] host or NULL */
    double* lw_final;          /* [n] host or NULL: final log-weights (no_resample sweeps) */
} apop_sweep_result;

int apop_smc_sweep(apop_ctx* ctx, const apop_ssm* model, const double* y, int32_t T,
                   const double* retained, const apop_sweep_opts* opts,
                   apop_sweep_result* result);
/* Sharded population: N = world * n_local particles in
 * contiguous shards, global id = rank * n_local + local.  One process per GPU:
 * every rank calls apop_smc_sweep with identical arguments (collective).
 * In-process group: one call drives all ranks in lock step; results[i] belongs
 * to ctxs[i] (x_final = that rank's shard of the final particles).  Results
 * are bit-identical to an unsharded population of N particles. */
int apop_smc_sweep_group(apop_ctx* const* ctxs, int32_t n_ctx, const apop_ssm* model,
                         const double* y, int32_t T, const apop_sweep_opts* opts,
                         apop_sweep_result* results);
/* lineage (x_0..x_T) of final slot `slot` of the last keep_history sweep: the
 * `::past-samples` a retained particle replays (pgibbs.cljc:29-35, 66-72, 122-123) */
int apop_trace_lineage(apop_ctx* ctx, int64_t slot, double* out_traj);
/* whole ancestry of the last keep_history sweep, for host-side result
 * materialisation (the `:result` / `:predicts` of every particle stay on the
 * host side): anc [T*n] i32 physical ancestors, x [(T+1)*n*dim] f64 states
 * before each resample.  Either pointer may be NULL. */
int apop_sweep_download_history(apop_ctx* ctx, int32_t* anc, double* x);
/* device pointer of the final state column of the last sweep (n*dim f64) */
int apop_sweep_final_device(apop_ctx* ctx, void** out);

/* ---- the other scalar-parameter families of anglican.runtime --------------
 * One entry point pair for `sample*` / `observe*` (runtime.cljc:119-124) of
 *   EXPONENTIAL(rate)                 runtime.cljc:308-313
 *   LAPLACE(loc, scale)               :411-416
 *   POISSON(lambda)                   :418-423
 *   BERNOULLI(p) (= flip with 1/0)    :241-249, 315-323
 *   STUDENT_T(nu)                     :425-444
 *   STUDENT_T_LOC_SCALE(nu,loc,scale) :446-448
 *   CHI_SQUARED(nu)                   :333-340
 *   UNIFORM_DISCRETE(min, max)        :464-475  (values min .. max-1)
 *   BINOMIAL(n, p)                    :258-263  (sampler: scalar n <= 400)
 * params points at three apop_arg (unused ones are ignored).  Values and draws are f64 columns (integer
 * valued for the discrete families); observe adds into the log-weight column. */
enum { APOP_DIST_EXPONENTIAL = 1, APOP_DIST_LAPLACE = 2, APOP_DIST_POISSON = 3,
       APOP_DIST_BERNOULLI = 4, APOP_DIST_STUDENT_T = 5, APOP_DIST_STUDENT_T_LOC_SCALE = 6,
       APOP_DIST_CHI_SQUARED = 7, APOP_DIST_UNIFORM_DISCRETE = 8, APOP_DIST_BINOMIAL = 9 };
int apop_sample_dist(apop_ctx* ctx, int32_t family, const apop_arg* params /* [3] */,
                     uint32_t draw_id, apop_col out);
int apop_observe_dist(apop_ctx* ctx, int32_t family, const apop_arg* params /* [3] */,
                      apop_arg value, apop_col logw);

/* ---- result-side reductions on the device (anglican.stat, stat.cljc) -----
 * x: an f64 column of `width` values per particle, or APOP_COL_SWEEP_FINAL for
 * the final states of the last sweep; logw: a width-1 f64 column of
 * log-weights or -1 for equal weights (the states `infer :smc` returns all
 * carry the same weight).  Weights are exp(lw - max lw) (stat.cljc:63-68).
 * Outputs are host arrays, any of which may be NULL.
 *   apop_empirical_moments    empirical-mean / -variance / -skew / -kurtosis
 *                             per column (stat.cljc:76-82, 93-120)
 *   apop_empirical_covariance empirical-mean + empirical-covariance [width^2]
 *                             (stat.cljc:84-91), width <= 64
 *   apop_collect_discrete     collect-by over the values 0..K-1 of an i32 (or
 *                             integer-valued f64) column: log_weight[k] = log
 *                             sum of the weights of value k - log N
 *                             (stat.cljc:7-24), prob[k] = the normalised
 *                             empirical-distribution (stat.cljc:43-54) */
#define APOP_COL_SWEEP_FINAL (-2)
int apop_empirical_moments(apop_ctx* ctx, apop_col x, apop_col logw, int32_t width, double* mean,
                           double* var, double* skew, double* kurt);
int apop_empirical_covariance(apop_ctx* ctx, apop_col x, apop_col logw, int32_t width,
                              double* mean, double* cov);
int apop_collect_discrete(apop_ctx* ctx, apop_col x, apop_col logw, int32_t K, double* log_weight,
                          double* prob);

/* ---- result moments without any allocation -------------------------------
 * NO COUNTERPART IN THE REFERENCE (its statistics fold a JVM seq, stat.cljc:56-120, and
 * the JVM allocates as it pleases); this entry point exists for hosts that must not.
 *
 * Same numbers as apop_empirical_moments - same kernels, same grid, same order of
 * summation, so bit for bit - but no apop_ctx and no cudaMalloc / cudaFree / pinned-host
 * allocation anywhere on the path, not on the first call either.  Use it when
 *   - the process has given all device memory to its own arena (a JVM binding that
 *     reserves the GPU at start-up, torch's caching allocator at its limit), so that a
 *     library-side cudaMalloc would fail or eat into memory the host believes it owns; or
 *   - other populations are sweeping on the same GPU: cudaMalloc and cudaFree hold the
 *     whole device until its queues drain, and a sharded sweep that is waiting on its
 *     peers with a bounded spin can time out (APOP_ERR_NCCL) behind such a stall.
 * The partial sums live in one fixed array inside the library image (about 123 KB of
 * device memory on every GPU that loads the library, held until the process exits).
 *
 * Arguments: d_x [n*width] f64 row-major and d_logw [n] f64 (or NULL for equal weights,
 * otherwise weights are exp(lw - max lw), stat.cljc:63-68) are DEVICE pointers; so are
 * the outputs d_mean / d_var / d_skew / d_kurt, [width] doubles each, any of which may be
 * NULL (not all).  `stream` is the caller's cudaStream_t (NULL: the default stream); the
 * work is queued on it and the call returns without waiting - read the outputs after
 * synchronising that stream.  `device` becomes the calling thread's current device.
 * errbuf (may be NULL) receives the reason of a non-zero return.
 *
 * Limits: 1 <= width <= APOP_NOALLOC_MAX_WIDTH (the array keeps one mean per column);
 * 1 <= n <= APOP_NOALLOC_MAX_PARTICLES (the library's particle indices are 32-bit).
 * Anything larger is refused with APOP_ERR_INVALID and a message; nothing is queued.
 * A population whose total weight is zero gives NaN outputs (the ctx call reports
 * APOP_ERR_ZERO_WEIGHT, which takes a device->host read that this call does not make).
 *
 * Sharing: every call on a GPU uses the SAME array, and it still holds the previous
 * call's partial sums.  Each call overwrites every slot before reading it, so nothing
 * needs clearing, and calls queued one after another on one stream are safe.  Two calls
 * that can run at the same time on one GPU - different streams, or different host
 * threads - overwrite each other's sums: order them yourself (one stream, or events).
 *
 * First call on each device: the library looks up where the array lives on that device
 * (cudaGetSymbolAddress, which may also load the module).  Make that first call during
 * set-up, before any stream capture is begun on the stream you pass. */
#define APOP_NOALLOC_MAX_WIDTH 256
#define APOP_NOALLOC_MAX_PARTICLES 2147483647LL
int apop_empirical_moments_noalloc(int device, void* stream, const double* d_x,
                                   const double* d_logw, int64_t n, int32_t width,
                                   double* d_mean, double* d_var, double* d_skew, double* d_kurt,
                                   char* errbuf, int32_t errbuf_len);

/* ---- PIMH / IPMCMC scalars (host arithmetic, restated on the host side) -
 * accept?: lmh.cljc:146-151; gibbs-update-csmc-indices: ipmcmc.cljc:106-146
 * with the intended sampling rule; exchange: the only data
 * that crosses nodes, ipmcmc.cljc:183-187. */
int apop_pimh_accept(double next_log_z, double prev_log_z, double uniform);
int apop_ipmcmc_gibbs(const double* log_zs, int m, int p, const double* uniforms,
                      int32_t* csmc_idx, double* zeta_sums);
int apop_ipmcmc_exchange_logz(apop_ctx* ctx, const double* local, int n_local, double* out_all);

#ifdef __cplusplus
}
#endif
#endif
