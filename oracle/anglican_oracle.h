/*
 * anglican_oracle.h — CPU restatement (the ORACLE) of the particle-population
 * hot path of probprog/anglican.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, the smoke check
 * in __graft_entry__.py and bench.py's cpu_baseline / --impl reference legs may
 * load it.  The product (anglican_b200 / libanglican_pop.so) never links,
 * imports or calls anything in oracle/.
 *
 * Every function cites the reference file:line it restates (paths relative to
 * the reference tree, commit 583360e).  Arithmetic is fp64, strictly
 * sequential, compiled with -ffp-contract=off so that a+b*c is never fused.
 *
 * Pinning status (see DESIGN.md "Oracle"): densities are pinned against the
 * reference's golden vectors (runtime_test.cljc:41-46, runtime_chi_squared_
 * test.cljc:76-85, runtime_test.cljc:18-24), the IPMCMC index Gibbs step
 * against ipmcmc_test.clj:33-45, the whole path against the HMM posterior
 * table of algorithm_test.cljc:131-147 and exact Kalman/HMM-forward truths.
 * `resample` ordering/ties, `particle-weights` NaN rules, `log-marginal`,
 * `log-sum-exp`, `dirichlet` and `beta` have no reference test or golden
 * vector: for those functions parity is UNPINNED (restatement only).
 */
#ifndef ANGLICAN_ORACLE_H
#define ANGLICAN_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- counter-based RNG shared by oracle and engine (engine semantics, not
 * the reference's unseeded Well19937c: runtime.cljc:142-147) -------------- */
enum { ORC_STREAM_INIT = 0, ORC_STREAM_TRANSITION = 1, ORC_STREAM_RESAMPLE = 2,
       ORC_STREAM_USER = 3, ORC_STREAM_AUX = 4 };

void   orc_philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                      uint32_t k0, uint32_t k1, uint32_t out[4]);
/* uniform in (0,1) from 53 random bits (+0.5 ulp) */
double orc_u01_open(uint32_t hi, uint32_t lo);
/* uniform in [0,1) from 53 random bits */
double orc_u01_half(uint32_t hi, uint32_t lo);
/* j-th standard normal of (slot, t, stream): Box-Muller pair shared by slots
 * 2p and 2p+1 */
double orc_normal_at(uint64_t seed, uint64_t slot, uint32_t t, uint32_t stream,
                     uint32_t draw);
/* j-th uniform [0,1) of (slot, t, stream) */
double orc_uniform_at(uint64_t seed, uint64_t slot, uint32_t t, uint32_t stream,
                      uint32_t draw);

/* ---- smc.cljc ----------------------------------------------------------- */
/* particle-weights, smc.cljc:70-84.  returns 0, or -1 if the max is NaN (the
 * reference would loop forever in resample in that case). */
int orc_particle_weights(const double* lw, int64_t n, double* w, double* max_lw);

/* the literal systematic wheel of smc.cljc:105-137 on given weights.
 * anc_ref[k] (k=0..n_new-1) is the returned list in reference order (offspring
 * are prepended, so emission order reversed).  If lap_of (may be NULL) is given
 * it receives, per emission index j, the lap (0 or 1...) in which pointer j was
 * emitted.  outputs total (left fold) and step.  returns 0 or -2 if total is
 * not > 0 / not finite (reference never terminates). */
int orc_resample_weights(const double* w, int64_t n, int64_t n_new, double u,
                         int32_t* anc_ref, int32_t* emission, int32_t* lap_of,
                         double* total_out, double* step_out);

/* resample, smc.cljc:86-137 from log-weights: anc_ref + the new log-weight
 * (smc.cljc:98-103). */
int orc_resample(const double* lw, int64_t n, int64_t n_new, double u,
                 int32_t* anc_ref, double* new_lw, double* w_scratch);

/* engine physical order: emissions of the second lap first, then the first lap
 * (ancestors monotone non-decreasing).  DESIGN.md "slot permutation". */
void orc_emission_to_sorted(const int32_t* emission, const int32_t* lap_of,
                            int64_t n_new, int32_t* anc_sorted, int64_t* n_lap0);

/* multinomial: n_new independent `discrete` draws (runtime.cljc:277-287) on
 * the same left-fold edges, one supplied uniform each. */
int orc_multinomial_weights(const double* w, int64_t n, const double* u,
                            int64_t n_new, int32_t* anc);

/* ---- runtime.cljc / inference.cljc -------------------------------------- */
double orc_log_sum_exp(double a, double b);                  /* runtime.cljc:130-138 */
double orc_log_marginal(const double* lw, int64_t n);        /* inference.cljc:181-186 */
double orc_ess(const double* lw, int64_t n);                 /* dist_test.cljc:83-94 */

double orc_normal_logpdf(double x, double mean, double sd);          /* runtime.cljc:403-409 */
double orc_gamma_logpdf(double x, double shape, double rate);        /* runtime.cljc:325-330 */
double orc_gamma_logpdf_direct(double x, double shape, double rate); /* textbook lgamma form */
double orc_beta_logpdf(double x, double alpha, double beta);         /* runtime.cljc:251-256 */
double orc_uniform_logpdf(double x, double lo, double hi);           /* runtime.cljc:450-453 */
double orc_discrete_logpmf(const double* w, int k, int64_t v);       /* runtime.cljc:277-290 */
double orc_dirichlet_logpdf(const double* v, const double* alpha, int k); /* 292-306 */
/* mvn, runtime.cljc:349-368.  cov row-major d*d (upper triangle read, EJML-
 * style inner lower Cholesky as used by vectorz).  work: 2*d*d+d doubles. */
int    orc_mvn_prepare(const double* cov, int d, double* L, double* iL, double* Z);
/* multivariate-t / wishart log-densities (runtime.cljc:383-401, 479-486, 536-539); work:
 * 2 d^2 + d doubles.  Pinned on runtime_wishart_test.cljc:161-176 and scipy. */
double orc_mvt_logpdf(const double* y, const double* mu, const double* sigma, double nu, int d,
                      double* work);
double orc_log_mv_gamma(int p, double a);
double orc_wishart_logpdf(const double* x, double n, const double* V, int p, double* work);
double orc_mvn_logpdf(const double* x, const double* mean, const double* iL,
                      double Z, int d, double* work);

/* samplers (distributional parity only: SURVEY §9.15) */
int64_t orc_discrete_sample(const double* w, int k, double u);       /* runtime.cljc:281-287 */
double  orc_gamma_sample(double shape, double rate, uint64_t seed, uint64_t slot,
                         uint32_t t, uint32_t stream);
double  orc_beta_sample(double a, double b, uint64_t seed, uint64_t slot,
                        uint32_t t, uint32_t stream);
void    orc_dirichlet_sample(const double* alpha, int k, uint64_t seed,
                             uint64_t slot, uint32_t t, uint32_t stream, double* out);

/* ---- lmh.cljc:146-151 (used by pimh.cljc:64) ---------------------------- */
int orc_accept(double u_next, double u_prev, double uniform);

/* ---- ipmcmc.cljc -------------------------------------------------------- */
/* norm-exp, ipmcmc.cljc:85-104 */
void orc_norm_exp(const double* lw, int n, double* ps, double* log_mean);
/* gibbs-update-csmc-indices, ipmcmc.cljc:106-146.  uniforms[P] supplies one
 * U[0,1) per conditional slot.  as_written!=0 reproduces the `(> 0.0 max)`
 * test literally (always uniform-discrete); 0 = intended rule (∝ ps when any
 * ps>0) which is what ipmcmc_test.clj:33-45 pins. */
void orc_ipmcmc_gibbs(const double* log_zs, int m, int p, const double* uniforms,
                      int as_written, int32_t* csmc_idx, double* zeta_sums);

/* ---- state-space sweeps (engine semantics over the literal resample) ---- */
enum { ORC_SSM_LGSSM1D = 1, ORC_SSM_SV = 2, ORC_SSM_HMM = 3, ORC_SSM_LGMVN = 4 };
enum { ORC_OBS_NONE = 0 };

typedef struct {
    int32_t kind;
    int32_t dim;        /* state dimension (1 for scalar models), K for HMM */
    /* scalar affine-Gaussian transition: x0 ~ N(m0,s0); x' = a x + b + q e */
    double m0, s0, a, b, q;
    /* observation: LGSSM1D y ~ N(c x + d, r);  SV y ~ N(0, beta exp(x/2)) */
    double c, d, r, beta;
    /* HMM (K = dim <= 16): init weights, row-major transition weights,
     * emission means / sds (algorithm_test.cljc:116-123) */
    const double* hmm_init;
    const double* hmm_trans;
    const double* hmm_mu;
    const double* hmm_sd;
    /* LGMVN: x0 ~ N(0, I s0^2); x' = A x + Lq z;  y ~ mvn(H x, R) with H = I */
    const double* A;     /* d*d row major */
    const double* Lq;    /* d*d lower Cholesky factor of process cov */
    const double* R;     /* d*d observation covariance */
} orc_ssm;

typedef struct {
    int64_t n;              /* particles */
    int32_t T;              /* observations */
    uint64_t seed;
    int32_t csmc;           /* 0 plain SMC; 1 conditional (retained at slot 0) */
    int32_t n_cond_steps;   /* observes at which the conditional resample is used */
    int32_t reset_weights;  /* pgibbs.cljc:88 / ipmcmc.cljc:69-72: lw := 0 after resample */
    int32_t keep_history;   /* store ancestors + states for lineage tracing */
    int32_t threads;        /* >1: OpenMP over particles in propagate/weight */
    int32_t ancestor_sampling; /* pgas.cljc:107-131: at each conditional observe the retained
                             * particle's ancestor is redrawn from discrete(exp(v - max v)),
                             * v_i = lw_i + log f(retained_{t+1} | x_t^i) */
    int32_t pad_;
    double ess_threshold;   /* engine extension (the reference resamples at every
                             * observe): > 0 = resample only when ESS < ess_threshold * N */
} orc_sweep_opts;

typedef struct {
    double log_z;           /* sum_t log-mean-weight (ipmcmc.cljc:67-74) == final lw in :smc */
    double* log_mean_w;     /* [T] per-step log(step)+max  */
    double* ess;            /* [T] */
    double* x_final;        /* [n*dim] resampled final states, sorted physical order */
    int32_t* anc_hist;      /* [T*n] physical ancestors per step (keep_history) */
    double* x_hist;         /* [(T+1)*n*dim] pre-resample states per step (keep_history) */
    double* filt_mean;      /* [T*dim] weighted filtering mean before resampling */
    double* lw_final;       /* [n] or NULL: log-weights of the final particles (adaptive sweeps) */
} orc_sweep_out;

/* y: [T*obs_dim] observations; retained: [(T+1)*dim] trajectory or NULL.
 * returns 0 or a negative error. */
int orc_smc_sweep(const orc_ssm* model, const double* y, const double* retained,
                  const orc_sweep_opts* opts, orc_sweep_out* out);

/* lineage of final slot `slot` from keep_history arrays: out [(T+1)*dim] */
int orc_importance_sweep(const orc_ssm* model, const double* y, int32_t T, int64_t n,
                         uint64_t seed, double* x_final, double* lw_final);
void orc_trace_lineage(const orc_sweep_out* out, int64_t n, int32_t T, int32_t dim,
                       int64_t slot, double* traj);

/* ---- exact truths ------------------------------------------------------- */
double orc_kalman1d_logz(const orc_ssm* m, const double* y, int T, double* filt_mean);
int    orc_kalman_mvn_logz(const orc_ssm* m, const double* y, int T, double* logz,
                           double* filt_mean);
double orc_hmm_forward_logz(const orc_ssm* m, const double* y, int T);
/* smoothing marginals [(T+1)*K] incl. the initial state */
void   orc_hmm_smooth(const orc_ssm* m, const double* y, int T, double* post);

int orc_version(void);

#ifdef __cplusplus
}
#endif
#endif
