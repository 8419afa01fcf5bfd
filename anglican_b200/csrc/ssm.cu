// ssm.cu — fused state-space sweeps (stub while the resampler is validated)
#include "apop_internal.cuh"
namespace apop { void sweep_free(apop_ctx*) {} }
extern "C" {
int apop_smc_sweep(apop_ctx* c, const apop_ssm*, const double*, int32_t, const double*,
                   const apop_sweep_opts*, apop_sweep_result*) {
    return apop::set_err(c, APOP_ERR_UNSUPPORTED, "sweep not built yet");
}
int apop_trace_lineage(apop_ctx* c, int64_t, double*) { return apop::set_err(c, APOP_ERR_UNSUPPORTED, "n/a"); }
int apop_sweep_final_device(apop_ctx* c, void**) { return apop::set_err(c, APOP_ERR_UNSUPPORTED, "n/a"); }
}
