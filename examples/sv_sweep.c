/*
 * sv_sweep.c - one `infer :smc` sweep of the stochastic-volatility program of BASELINE
 * config 3 through the C ABI alone (no Python, no torch): what a JNI / Panama binding does
 * per sweep, written as a C99 program.
 *
 *   gcc -std=c99 -Iinclude examples/sv_sweep.c -o sv_sweep \
 *       -Langlican_b200 -langlican_pop -Wl,-rpath,$PWD/anglican_b200 -lm
 *   ./sv_sweep [log2_particles] [T] [observations.f64]
 *
 * observations.f64: T raw little-endian doubles (default: a built-in synthetic series).
 *
 * The Anglican program this replaces (one global observe per time step, smc.cljc:35-51):
 *   (defquery sv [ys]
 *     (let [x0 (sample (normal 0 (/ sigma (sqrt (- 1 (* phi phi))))))]
 *       (reduce (fn [x y] (let [x' (sample (normal (* phi x) sigma))]
 *                           (observe (normal 0 (* beta (exp (/ x' 2)))) y) x'))
 *               x0 ys)))
 * Without a CUDA device apop_create fails with APOP_ERR_CUDA: there is no CPU fallback.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "anglican_pop.h"

#define CHECK(call)                                                                   \
    do {                                                                              \
        int rc_ = (call);                                                             \
        if (rc_ != APOP_OK) {                                                         \
            fprintf(stderr, "%s failed: %d (%s)\n", #call, rc_,                       \
                    ctx ? apop_last_error(ctx) : "no context");                       \
            if (ctx) apop_destroy(ctx);                                               \
            return rc_ == APOP_ERR_CUDA ? 3 : 1;                                      \
        }                                                                             \
    } while (0)

/* synthetic observations: any fixed series will do for a smoke run */
static void make_observations(double* y, int T) {
    uint64_t s = 0x9E3779B97F4A7C15ULL;
    double x = 0.0;
    int t;
    for (t = 0; t < T; ++t) {
        double u1, u2, e1, e2;
        s = s * 6364136223846793005ULL + 1442695040888963407ULL;
        u1 = ((double)(s >> 11) + 0.5) / 9007199254740992.0;
        s = s * 6364136223846793005ULL + 1442695040888963407ULL;
        u2 = ((double)(s >> 11) + 0.5) / 9007199254740992.0;
        e1 = sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
        e2 = sqrt(-2.0 * log(u1)) * sin(6.283185307179586 * u2);
        x = 0.98 * x + 0.16 * e1;
        y[t] = 0.65 * exp(0.5 * x) * e2;
    }
}

int main(int argc, char** argv) {
    const int lg = argc > 1 ? atoi(argv[1]) : 20;
    const int T = argc > 2 ? atoi(argv[2]) : 100;
    const int64_t n = (int64_t)1 << lg;
    apop_ctx* ctx = NULL;
    apop_ssm model;
    apop_sweep_opts opts;
    apop_sweep_result res;
    double *y, *ess, *lmw, *x_final = NULL;
    double mean = 0.0;
    int64_t i;
    int t;

    if (lg < 0 || lg > 30 || T < 1) {
        fprintf(stderr, "usage: sv_sweep [log2_particles] [T] [observations.f64]\n");
        return 2;
    }
    if (apop_sizeof_sweep_opts() != (int)sizeof(apop_sweep_opts)) {
        fprintf(stderr, "header / library mismatch: apop_sweep_opts\n");
        return 2;
    }
    y = (double*)malloc(sizeof(double) * (size_t)T);
    ess = (double*)malloc(sizeof(double) * (size_t)T);
    lmw = (double*)malloc(sizeof(double) * (size_t)T);
    if (!y || !ess || !lmw) return 2;
    if (argc > 3) {
        FILE* f = fopen(argv[3], "rb");
        if (!f || fread(y, sizeof(double), (size_t)T, f) != (size_t)T) {
            fprintf(stderr, "cannot read %d doubles from %s\n", T, argv[3]);
            return 2;
        }
        fclose(f);
    } else {
        make_observations(y, T);
    }

    /* (repeatedly number-of-particles ...) - smc.cljc:37-40 */
    CHECK(apop_create(&ctx, n, 0xA11CEULL, /*device*/ 0, /*rank*/ 0, /*world*/ 1, NULL));
    /* the final states come back into page-locked memory owned by the library */
    CHECK(apop_host_alloc(ctx, (void**)&x_final, sizeof(double) * (uint64_t)n));

    {   /* zero-initialise the plain-C structs field by field (C99, no memset of doubles needed) */
        apop_ssm m0 = {0};
        apop_sweep_opts o0 = {0};
        apop_sweep_result r0 = {0};
        model = m0; opts = o0; res = r0;
    }
    model.kind = APOP_SSM_SV;
    model.dim = 1;
    model.m0 = 0.0;
    model.s0 = 0.16 / sqrt(1.0 - 0.98 * 0.98);
    model.a = 0.98;
    model.b = 0.0;
    model.q = 0.16;
    model.beta = 0.65;
    opts.want_final = 1;
    res.log_mean_w = lmw;
    res.ess = ess;
    res.x_final = x_final;

    CHECK(apop_smc_sweep(ctx, &model, y, T, /*retained*/ NULL, &opts, &res));
    CHECK(apop_synchronize(ctx));

    for (i = 0; i < n; ++i) mean += x_final[i];
    mean /= (double)n;
    printf("particles 2^%d, T %d: log Z = %.17g, ESS at the last observe = %.1f, "
           "mean of the final states = %.6f, kernels launched = %lld\n",
           lg, T, res.log_z, ess[T - 1], mean, (long long)apop_launch_count(ctx));
    for (t = 0; t < T && t < 3; ++t) printf("  step %d: log mean weight %.6f, ESS %.1f\n", t, lmw[t], ess[t]);

    CHECK(apop_host_free(ctx, x_final));
    apop_destroy(ctx);
    free(y); free(ess); free(lmw);
    return 0;
}
