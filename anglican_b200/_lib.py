"""ctypes binding of libanglican_pop.so (include/anglican_pop.h).

The library is the product: there is no CPU fallback.  Importing this module
loads the shared object (building it with nvcc first if the sources are newer);
compute entry points fail with APOP_ERR_CUDA when no GPU is present.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libanglican_pop.so")

APOP_OK = 0
ERR_NAMES = {0: "APOP_OK", -1: "APOP_ERR_INVALID", -2: "APOP_ERR_CUDA", -3: "APOP_ERR_NAN_WEIGHT",
             -4: "APOP_ERR_ZERO_WEIGHT", -5: "APOP_ERR_NCCL", -6: "APOP_ERR_OOM",
             -7: "APOP_ERR_UNSUPPORTED", -8: "APOP_ERR_INTERNAL"}
F64, I32 = 0, 1
SYSTEMATIC, MULTINOMIAL = 0, 1
ORDER_REFERENCE, ORDER_SORTED, ORDER_EMISSION = 0, 1, 2
SSM_LGSSM1D, SSM_SV, SSM_HMM, SSM_LGMVN = 1, 2, 3, 4

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)
c_int64_p = C.POINTER(C.c_int64)


class ApopArg(C.Structure):
    _fields_ = [("col", C.c_int32), ("scalar", C.c_double)]


class ApopSSM(C.Structure):
    _fields_ = [("kind", C.c_int32), ("dim", C.c_int32),
                ("m0", C.c_double), ("s0", C.c_double), ("a", C.c_double), ("b", C.c_double),
                ("q", C.c_double), ("c", C.c_double), ("d", C.c_double), ("r", C.c_double),
                ("beta", C.c_double),
                ("hmm_init", c_double_p), ("hmm_trans", c_double_p), ("hmm_mu", c_double_p),
                ("hmm_sd", c_double_p), ("A", c_double_p), ("Lq", c_double_p), ("R", c_double_p)]


class ApopSweepOpts(C.Structure):
    _fields_ = [("csmc", C.c_int32), ("n_cond_steps", C.c_int32), ("reset_weights", C.c_int32),
                ("keep_history", C.c_int32), ("use_graph", C.c_int32), ("want_final", C.c_int32),
                ("want_filt", C.c_int32), ("no_resample", C.c_int32),
                ("ancestor_sampling", C.c_int32), ("pad_", C.c_int32),
                ("ess_threshold", C.c_double)]


class ApopSweepResult(C.Structure):
    _fields_ = [("log_z", C.c_double), ("log_mean_w", c_double_p), ("ess", c_double_p),
                ("filt_mean", c_double_p), ("x_final", c_double_p), ("lw_final", c_double_p)]


class ApopError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{ERR_NAMES.get(code, code)}: {msg}")
        self.code = code


# every symbol include/anglican_pop.h declares: (name, restype, argtypes)
_vp = C.c_void_p
_SIGS = [
    ("apop_version", C.c_int, []),
    ("apop_device_count", C.c_int, []),
    ("apop_create", C.c_int, [C.POINTER(_vp), C.c_int64, C.c_uint64, C.c_int, C.c_int, C.c_int, _vp]),
    ("apop_link_group", C.c_int, [C.POINTER(_vp), C.c_int32]),
    ("apop_destroy", C.c_int, [_vp]),
    ("apop_last_error", C.c_char_p, [_vp]),
    ("apop_get_unique_id", C.c_int, [_vp]),
    ("apop_synchronize", C.c_int, [_vp]),
    ("apop_num_particles", C.c_int64, [_vp]),
    ("apop_set_seed", C.c_int, [_vp, C.c_uint64]),
    ("apop_host_alloc", C.c_int, [_vp, C.POINTER(_vp), C.c_uint64]),
    ("apop_host_free", C.c_int, [_vp, _vp]),
    ("apop_launch_count", C.c_int64, [_vp]),
    ("apop_stream", _vp, [_vp]),
    ("apop_timer_begin", C.c_int, [_vp, C.c_int]),
    ("apop_timer_end", C.c_int, [_vp, C.c_int]),
    ("apop_timer_elapsed_ms", C.c_int, [_vp, C.c_int, c_double_p, c_int64_p]),
    ("apop_timer_reset", C.c_int, [_vp]),
    ("apop_enable_kernel_timers", C.c_int, [_vp, C.c_int]),
    ("apop_column_alloc", C.c_int, [_vp, C.c_int, C.c_int, c_int32_p]),
    ("apop_column_free", C.c_int, [_vp, C.c_int32]),
    ("apop_column_upload", C.c_int, [_vp, C.c_int32, _vp, C.c_int64]),
    ("apop_column_download", C.c_int, [_vp, C.c_int32, _vp, C.c_int64]),
    ("apop_column_fill", C.c_int, [_vp, C.c_int32, C.c_double]),
    ("apop_column_device_ptr", C.c_int, [_vp, C.c_int32, C.POINTER(_vp)]),
    ("apop_set_step", C.c_int, [_vp, C.c_uint32]),
    ("apop_sample_normal", C.c_int, [_vp, ApopArg, ApopArg, C.c_uint32, C.c_int32]),
    ("apop_observe_normal", C.c_int, [_vp, ApopArg, ApopArg, ApopArg, C.c_int32]),
    ("apop_sample_gamma", C.c_int, [_vp, ApopArg, ApopArg, C.c_uint32, C.c_int32]),
    ("apop_observe_gamma", C.c_int, [_vp, ApopArg, ApopArg, ApopArg, C.c_int32]),
    ("apop_sample_beta", C.c_int, [_vp, ApopArg, ApopArg, C.c_uint32, C.c_int32]),
    ("apop_observe_beta", C.c_int, [_vp, ApopArg, ApopArg, ApopArg, C.c_int32]),
    ("apop_sample_uniform", C.c_int, [_vp, ApopArg, ApopArg, C.c_uint32, C.c_int32]),
    ("apop_observe_uniform", C.c_int, [_vp, ApopArg, ApopArg, ApopArg, C.c_int32]),
    ("apop_sample_dist", C.c_int, [_vp, C.c_int32, C.POINTER(ApopArg), C.c_uint32, C.c_int32]),
    ("apop_observe_dist", C.c_int, [_vp, C.c_int32, C.POINTER(ApopArg), ApopArg, C.c_int32]),
    ("apop_sample_discrete", C.c_int, [_vp, c_double_p, C.c_int, C.c_int32, C.c_uint32, C.c_int32]),
    ("apop_observe_discrete", C.c_int, [_vp, c_double_p, C.c_int, C.c_int32, C.c_int32, C.c_int32,
                                        C.c_int32]),
    ("apop_sample_dirichlet", C.c_int, [_vp, c_double_p, C.c_int, C.c_uint32, C.c_int32]),
    ("apop_observe_dirichlet", C.c_int, [_vp, c_double_p, C.c_int, C.c_int32, C.c_int32]),
    ("apop_sample_mvn", C.c_int, [_vp, c_double_p, C.c_int32, c_double_p, C.c_int, C.c_uint32,
                                  C.c_int32]),
    ("apop_observe_mvn", C.c_int, [_vp, c_double_p, C.c_int32, c_double_p, C.c_int, c_double_p,
                                   C.c_int32, C.c_int32]),
    ("apop_sample_mvt", C.c_int, [_vp, C.c_double, c_double_p, C.c_int32, c_double_p, C.c_int,
                                  C.c_uint32, C.c_int32]),
    ("apop_observe_mvt", C.c_int, [_vp, C.c_double, c_double_p, C.c_int32, c_double_p, C.c_int,
                                   c_double_p, C.c_int32, C.c_int32]),
    ("apop_sample_wishart", C.c_int, [_vp, C.c_double, c_double_p, C.c_int, C.c_uint32, C.c_int32]),
    ("apop_observe_wishart", C.c_int, [_vp, C.c_double, c_double_p, C.c_int, C.c_int32, C.c_int32]),
    ("apop_column_take", C.c_int, [_vp, c_double_p, C.c_int, C.c_int, C.c_int32, C.c_int32]),
    ("apop_selftest_math", C.c_int, [_vp, C.c_int, C.c_int32, C.c_int32]),
    ("apop_debug_resolve_phases", C.c_int, [_vp, c_double_p]),
    ("apop_debug_chain_trace", C.c_int, [_vp, c_double_p]),
    ("apop_normalize", C.c_int, [_vp, C.c_int32, c_double_p, c_double_p, c_double_p]),
    ("apop_resample", C.c_int, [_vp, C.c_int32, C.c_int, C.c_double, C.c_int64, C.c_int, C.c_int32,
                                c_double_p, c_int64_p]),
    ("apop_resample_weights", C.c_int, [_vp, C.c_int32, C.c_int, C.c_double, C.c_int64, C.c_int,
                                        C.c_int32, c_double_p, c_int64_p]),
    ("apop_resample_multinomial_u", C.c_int, [_vp, C.c_int32, C.c_int32, C.c_int64, C.c_int32]),
    ("apop_gather", C.c_int, [_vp, C.c_int32, C.c_int64, c_int32_p, C.c_int]),
    ("apop_smc_sweep", C.c_int, [_vp, C.POINTER(ApopSSM), c_double_p, C.c_int32, c_double_p,
                                 C.POINTER(ApopSweepOpts), C.POINTER(ApopSweepResult)]),
    ("apop_smc_sweep_group", C.c_int, [C.POINTER(_vp), C.c_int32, C.POINTER(ApopSSM), c_double_p,
                                       C.c_int32, C.POINTER(ApopSweepOpts),
                                       C.POINTER(ApopSweepResult)]),
    ("apop_trace_lineage", C.c_int, [_vp, C.c_int64, c_double_p]),
    ("apop_sweep_download_history", C.c_int, [_vp, c_int32_p, c_double_p]),
    ("apop_sweep_final_device", C.c_int, [_vp, C.POINTER(_vp)]),
    ("apop_sizeof_sweep_opts", C.c_int, []),
    ("apop_empirical_moments", C.c_int, [_vp, C.c_int32, C.c_int32, C.c_int32, c_double_p,
                                         c_double_p, c_double_p, c_double_p]),
    ("apop_empirical_covariance", C.c_int, [_vp, C.c_int32, C.c_int32, C.c_int32, c_double_p,
                                            c_double_p]),
    ("apop_collect_discrete", C.c_int, [_vp, C.c_int32, C.c_int32, C.c_int32, c_double_p,
                                        c_double_p]),
    ("apop_empirical_moments_noalloc", C.c_int, [C.c_int, _vp, _vp, _vp, C.c_int64, C.c_int32,
                                                 _vp, _vp, _vp, _vp, C.c_char_p, C.c_int32]),
    ("apop_pimh_accept", C.c_int, [C.c_double, C.c_double, C.c_double]),
    ("apop_ipmcmc_gibbs", C.c_int, [c_double_p, C.c_int, C.c_int, c_double_p, c_int32_p,
                                    c_double_p]),
    ("apop_ipmcmc_exchange_logz", C.c_int, [_vp, c_double_p, C.c_int, c_double_p]),
]
COL_SWEEP_FINAL = -2
# apop_sample_dist / apop_observe_dist families (include/anglican_pop.h)
DIST_FAMILIES = {"exponential": (1, 1), "laplace": (2, 2), "poisson": (3, 1), "bernoulli": (4, 1),
                 "flip": (4, 1), "student-t": (5, 1), "student-t-loc-scale": (6, 3),
                 "chi-squared": (7, 1), "uniform-discrete": (8, 2), "binomial": (9, 2)}
DECLARED_SYMBOLS = [s[0] for s in _SIGS]

_lib = None


def load(build_if_needed: bool = True):
    """Load libanglican_pop.so; raises if it is missing (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if build_if_needed:
        try:
            from . import build as _b
            if _b.needs_build():
                _b.build()
        except Exception:
            if not os.path.exists(LIB_PATH):
                raise
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python -m anglican_b200.build`; "
                          "anglican_b200 has no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    for name, res, args in _SIGS:
        fn = getattr(lib, name)          # AttributeError if the ABI lost a symbol
        fn.restype = res
        fn.argtypes = args
    if lib.apop_sizeof_sweep_opts() != C.sizeof(ApopSweepOpts):
        raise ImportError(f"{LIB_PATH}: apop_sweep_opts has {lib.apop_sizeof_sweep_opts()} bytes, "
                          f"the binding expects {C.sizeof(ApopSweepOpts)}: rebuild the library")
    _lib = lib
    return lib


def check(ctx, rc):
    if rc != APOP_OK:
        msg = load().apop_last_error(ctx)
        raise ApopError(rc, msg.decode() if msg else "")
    return rc
