"""ctypes wrapper over oracle/_build/liboracle.so.

TEST INFRASTRUCTURE — the CPU restatement of anglican's particle path
(oracle/anglican_oracle.c).  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module; the
product package anglican_b200 must never do so.
"""
import ctypes as C
import math
import os

import numpy as np

from . import build as _build

_lib = None

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)


def lib():
    global _lib
    if _lib is None:
        path = _build.OUT
        if _build.needs_build():
            try:
                path = _build.build()
            except Exception:
                if not os.path.exists(path):
                    raise
        _lib = C.CDLL(path)
        _declare(_lib)
    return _lib


def _dp(a):
    return None if a is None else a.ctypes.data_as(c_double_p)


def _ip(a):
    return None if a is None else a.ctypes.data_as(c_int32_p)


class OrcSSM(C.Structure):
    _fields_ = [("kind", C.c_int32), ("dim", C.c_int32),
                ("m0", C.c_double), ("s0", C.c_double), ("a", C.c_double),
                ("b", C.c_double), ("q", C.c_double),
                ("c", C.c_double), ("d", C.c_double), ("r", C.c_double),
                ("beta", C.c_double),
                ("hmm_init", c_double_p), ("hmm_trans", c_double_p),
                ("hmm_mu", c_double_p), ("hmm_sd", c_double_p),
                ("A", c_double_p), ("Lq", c_double_p), ("R", c_double_p)]


class OrcSweepOpts(C.Structure):
    _fields_ = [("n", C.c_int64), ("T", C.c_int32), ("seed", C.c_uint64),
                ("csmc", C.c_int32), ("n_cond_steps", C.c_int32),
                ("reset_weights", C.c_int32), ("keep_history", C.c_int32),
                ("threads", C.c_int32), ("ancestor_sampling", C.c_int32), ("pad_", C.c_int32),
                ("ess_threshold", C.c_double)]


class OrcSweepOut(C.Structure):
    _fields_ = [("log_z", C.c_double), ("log_mean_w", c_double_p), ("ess", c_double_p),
                ("x_final", c_double_p), ("anc_hist", c_int32_p), ("x_hist", c_double_p),
                ("filt_mean", c_double_p), ("lw_final", c_double_p)]


def _declare(L):
    d, i64, i32, u64, u32 = C.c_double, C.c_int64, C.c_int32, C.c_uint64, C.c_uint32
    L.orc_philox4x32.argtypes = [u32] * 6 + [C.POINTER(u32)]
    L.orc_philox4x32.restype = None
    L.orc_mvt_logpdf.argtypes = [c_double_p, c_double_p, c_double_p, d, C.c_int, c_double_p]
    L.orc_mvt_logpdf.restype = d
    L.orc_wishart_logpdf.argtypes = [c_double_p, d, c_double_p, C.c_int, c_double_p]
    L.orc_wishart_logpdf.restype = d
    L.orc_normal_at.argtypes = [u64, u64, u32, u32, u32]
    L.orc_normal_at.restype = d
    L.orc_uniform_at.argtypes = [u64, u64, u32, u32, u32]
    L.orc_uniform_at.restype = d
    L.orc_particle_weights.argtypes = [c_double_p, i64, c_double_p, c_double_p]
    L.orc_particle_weights.restype = C.c_int
    L.orc_resample_weights.argtypes = [c_double_p, i64, i64, d, c_int32_p, c_int32_p,
                                       c_int32_p, c_double_p, c_double_p]
    L.orc_resample_weights.restype = C.c_int
    L.orc_resample.argtypes = [c_double_p, i64, i64, d, c_int32_p, c_double_p, c_double_p]
    L.orc_resample.restype = C.c_int
    L.orc_emission_to_sorted.argtypes = [c_int32_p, c_int32_p, i64, c_int32_p, C.POINTER(i64)]
    L.orc_emission_to_sorted.restype = None
    L.orc_multinomial_weights.argtypes = [c_double_p, i64, c_double_p, i64, c_int32_p]
    L.orc_multinomial_weights.restype = C.c_int
    L.orc_log_sum_exp.argtypes = [d, d]
    L.orc_log_sum_exp.restype = d
    L.orc_log_marginal.argtypes = [c_double_p, i64]
    L.orc_log_marginal.restype = d
    L.orc_ess.argtypes = [c_double_p, i64]
    L.orc_ess.restype = d
    for name, nargs in (("orc_normal_logpdf", 3), ("orc_gamma_logpdf", 3),
                        ("orc_gamma_logpdf_direct", 3), ("orc_beta_logpdf", 3),
                        ("orc_uniform_logpdf", 3)):
        f = getattr(L, name)
        f.argtypes = [d] * nargs
        f.restype = d
    L.orc_discrete_logpmf.argtypes = [c_double_p, C.c_int, i64]
    L.orc_discrete_logpmf.restype = d
    L.orc_discrete_sample.argtypes = [c_double_p, C.c_int, d]
    L.orc_discrete_sample.restype = i64
    L.orc_dirichlet_logpdf.argtypes = [c_double_p, c_double_p, C.c_int]
    L.orc_dirichlet_logpdf.restype = d
    L.orc_mvn_prepare.argtypes = [c_double_p, C.c_int, c_double_p, c_double_p, c_double_p]
    L.orc_mvn_prepare.restype = C.c_int
    L.orc_mvn_logpdf.argtypes = [c_double_p, c_double_p, c_double_p, d, C.c_int, c_double_p]
    L.orc_mvn_logpdf.restype = d
    L.orc_gamma_sample.argtypes = [d, d, u64, u64, u32, u32]
    L.orc_gamma_sample.restype = d
    L.orc_beta_sample.argtypes = [d, d, u64, u64, u32, u32]
    L.orc_beta_sample.restype = d
    L.orc_dirichlet_sample.argtypes = [c_double_p, C.c_int, u64, u64, u32, u32, c_double_p]
    L.orc_dirichlet_sample.restype = None
    L.orc_accept.argtypes = [d, d, d]
    L.orc_accept.restype = C.c_int
    L.orc_norm_exp.argtypes = [c_double_p, C.c_int, c_double_p, c_double_p]
    L.orc_norm_exp.restype = None
    L.orc_ipmcmc_gibbs.argtypes = [c_double_p, C.c_int, C.c_int, c_double_p, C.c_int,
                                   c_int32_p, c_double_p]
    L.orc_ipmcmc_gibbs.restype = None
    L.orc_smc_sweep.argtypes = [C.POINTER(OrcSSM), c_double_p, c_double_p,
                                C.POINTER(OrcSweepOpts), C.POINTER(OrcSweepOut)]
    L.orc_smc_sweep.restype = C.c_int
    L.orc_trace_lineage.argtypes = [C.POINTER(OrcSweepOut), i64, i32, i32, i64, c_double_p]
    L.orc_trace_lineage.restype = None
    L.orc_importance_sweep.argtypes = [C.POINTER(OrcSSM), c_double_p, i32, i64, C.c_uint64,
                                       c_double_p, c_double_p]
    L.orc_importance_sweep.restype = C.c_int
    L.orc_kalman1d_logz.argtypes = [C.POINTER(OrcSSM), c_double_p, C.c_int, c_double_p]
    L.orc_kalman1d_logz.restype = d
    L.orc_kalman_mvn_logz.argtypes = [C.POINTER(OrcSSM), c_double_p, C.c_int, c_double_p,
                                      c_double_p]
    L.orc_kalman_mvn_logz.restype = C.c_int
    L.orc_hmm_forward_logz.argtypes = [C.POINTER(OrcSSM), c_double_p, C.c_int]
    L.orc_hmm_forward_logz.restype = d
    L.orc_hmm_smooth.argtypes = [C.POINTER(OrcSSM), c_double_p, C.c_int, c_double_p]
    L.orc_hmm_smooth.restype = None


def _f64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


# ---------------------------------------------------------------- RNG
def philox(c, k):
    out = (C.c_uint32 * 4)()
    lib().orc_philox4x32(c[0], c[1], c[2], c[3], k[0], k[1], out)
    return [int(v) for v in out]


def mvt_logpdf(y, mu, sigma, nu):
    """multivariate-t observe* (runtime.cljc:383-401)"""
    y, mu, sigma = _f64(y), _f64(mu), _f64(sigma)
    k = mu.size
    work = np.empty(2 * k * k + k)
    return lib().orc_mvt_logpdf(_dp(y), _dp(mu), _dp(sigma), float(nu), k, _dp(work))


def wishart_logpdf(x, n, V):
    """wishart observe* (runtime.cljc:536-539)"""
    x, V = _f64(x), _f64(V)
    p = int(round(math.sqrt(V.size)))
    work = np.empty(2 * p * p)
    return lib().orc_wishart_logpdf(_dp(x), float(n), _dp(V), p, _dp(work))


def normal_at(seed, slot, t, stream, draw=0):
    return lib().orc_normal_at(seed, slot, t, stream, draw)


def uniform_at(seed, slot, t, stream, draw=0):
    return lib().orc_uniform_at(seed, slot, t, stream, draw)


# ---------------------------------------------------------------- smc.cljc
def particle_weights(lw):
    lw = _f64(lw)
    w = np.empty_like(lw)
    m = C.c_double()
    rc = lib().orc_particle_weights(_dp(lw), lw.size, _dp(w), C.byref(m))
    if rc:
        raise ValueError("oracle: NaN max log-weight (reference would not terminate)")
    return w, m.value


def resample_weights(w, n_new, u):
    """literal wheel on weights -> dict(anc_ref, emission, lap, total, step)"""
    w = _f64(w)
    anc = np.empty(n_new, dtype=np.int32)
    emi = np.empty(n_new, dtype=np.int32)
    lap = np.empty(n_new, dtype=np.int32)
    tot, step = C.c_double(), C.c_double()
    rc = lib().orc_resample_weights(_dp(w), w.size, n_new, float(u), _ip(anc), _ip(emi),
                                    _ip(lap), C.byref(tot), C.byref(step))
    if rc:
        raise ValueError(f"oracle resample failed rc={rc}")
    return dict(anc_ref=anc, emission=emi, lap=lap, total=tot.value, step=step.value)


def resample(lw, n_new, u):
    lw = _f64(lw)
    anc = np.empty(n_new, dtype=np.int32)
    scratch = np.empty_like(lw)
    new_lw = C.c_double()
    rc = lib().orc_resample(_dp(lw), lw.size, n_new, float(u), _ip(anc), C.byref(new_lw),
                            _dp(scratch))
    if rc:
        raise ValueError(f"oracle resample failed rc={rc}")
    return anc, new_lw.value


def emission_to_sorted(emission, lap):
    emission = np.ascontiguousarray(emission, dtype=np.int32)
    lap = np.ascontiguousarray(lap, dtype=np.int32)
    out = np.empty_like(emission)
    j0 = C.c_int64()
    lib().orc_emission_to_sorted(_ip(emission), _ip(lap), emission.size, _ip(out), C.byref(j0))
    return out, j0.value


def multinomial_weights(w, u):
    w, u = _f64(w), _f64(u)
    anc = np.empty(u.size, dtype=np.int32)
    rc = lib().orc_multinomial_weights(_dp(w), w.size, _dp(u), u.size, _ip(anc))
    if rc:
        raise ValueError("oracle multinomial failed")
    return anc


# ---------------------------------------------------------------- runtime
def log_sum_exp(a, b):
    return lib().orc_log_sum_exp(a, b)


def log_marginal(lw):
    lw = _f64(lw)
    return lib().orc_log_marginal(_dp(lw), lw.size)


def ess(lw):
    lw = _f64(lw)
    return lib().orc_ess(_dp(lw), lw.size)


def normal_logpdf(x, mean, sd):
    return lib().orc_normal_logpdf(x, mean, sd)


def gamma_logpdf(x, shape, rate):
    return lib().orc_gamma_logpdf(x, shape, rate)


def gamma_logpdf_direct(x, shape, rate):
    return lib().orc_gamma_logpdf_direct(x, shape, rate)


def beta_logpdf(x, a, b):
    return lib().orc_beta_logpdf(x, a, b)


def uniform_logpdf(x, lo, hi):
    return lib().orc_uniform_logpdf(x, lo, hi)


def discrete_logpmf(w, v):
    w = _f64(w)
    return lib().orc_discrete_logpmf(_dp(w), w.size, int(v))


def discrete_sample(w, u):
    w = _f64(w)
    return lib().orc_discrete_sample(_dp(w), w.size, float(u))


def dirichlet_logpdf(v, alpha):
    v, alpha = _f64(v), _f64(alpha)
    return lib().orc_dirichlet_logpdf(_dp(v), _dp(alpha), alpha.size)


def mvn_prepare(cov):
    cov = _f64(cov)
    d = cov.shape[0]
    L = np.zeros((d, d))
    iL = np.zeros((d, d))
    Z = C.c_double()
    rc = lib().orc_mvn_prepare(_dp(cov), d, _dp(L), _dp(iL), C.byref(Z))
    if rc:
        raise ValueError("covariance not positive definite")
    return L, iL, Z.value


def mvn_logpdf(x, mean, cov):
    x, mean = _f64(x), _f64(mean)
    L, iL, Z = mvn_prepare(cov)
    work = np.empty(2 * x.size)
    return lib().orc_mvn_logpdf(_dp(x), _dp(mean), _dp(iL), Z, x.size, _dp(work))


def gamma_sample(shape, rate, seed, slot, t=0, stream=3):
    return lib().orc_gamma_sample(shape, rate, seed, slot, t, stream)


def beta_sample(a, b, seed, slot, t=0, stream=3):
    return lib().orc_beta_sample(a, b, seed, slot, t, stream)


def dirichlet_sample(alpha, seed, slot, t=0, stream=3):
    alpha = _f64(alpha)
    out = np.empty_like(alpha)
    lib().orc_dirichlet_sample(_dp(alpha), alpha.size, seed, slot, t, stream, _dp(out))
    return out


def accept(u_next, u_prev, uniform):
    return bool(lib().orc_accept(u_next, u_prev, uniform))


def norm_exp(lw):
    lw = _f64(lw)
    ps = np.empty_like(lw)
    lm = C.c_double()
    lib().orc_norm_exp(_dp(lw), lw.size, _dp(ps), C.byref(lm))
    return ps, lm.value


def ipmcmc_gibbs(log_zs, p, uniforms, as_written=False):
    log_zs, uniforms = _f64(log_zs), _f64(uniforms)
    cs = np.empty(p, dtype=np.int32)
    zs = np.empty(log_zs.size)
    lib().orc_ipmcmc_gibbs(_dp(log_zs), log_zs.size, p, _dp(uniforms), int(as_written),
                           _ip(cs), _dp(zs))
    return cs, zs


# ---------------------------------------------------------------- sweeps
SSM_LGSSM1D, SSM_SV, SSM_HMM, SSM_LGMVN = 1, 2, 3, 4


class Model:
    """Holds an OrcSSM plus the numpy arrays it points into."""

    def __init__(self, kind, dim=1, m0=0.0, s0=1.0, a=1.0, b=0.0, q=1.0, c=1.0, d=0.0,
                 r=1.0, beta=1.0, hmm_init=None, hmm_trans=None, hmm_mu=None, hmm_sd=None,
                 A=None, Lq=None, R=None):
        self._keep = {}
        s = OrcSSM()
        s.kind, s.dim = kind, dim
        s.m0, s.s0, s.a, s.b, s.q = m0, s0, a, b, q
        s.c, s.d, s.r, s.beta = c, d, r, beta
        for name, val in (("hmm_init", hmm_init), ("hmm_trans", hmm_trans), ("hmm_mu", hmm_mu),
                          ("hmm_sd", hmm_sd), ("A", A), ("Lq", Lq), ("R", R)):
            if val is not None:
                arr = _f64(val)
                self._keep[name] = arr
                setattr(s, name, _dp(arr))
        self.c = s
        self.kind, self.dim = kind, dim

    @property
    def state_dim(self):
        return self.dim if self.kind == SSM_LGMVN else 1


def smc_sweep(model, y, n, seed, retained=None, csmc=False, n_cond_steps=0,
              reset_weights=False, keep_history=False, threads=1, want_final=True,
              ancestor_sampling=False, ess_threshold=0.0):
    y = _f64(y)
    d = model.state_dim
    T = y.size // d
    o = OrcSweepOpts(n=n, T=T, seed=seed, csmc=int(csmc), n_cond_steps=n_cond_steps,
                     reset_weights=int(reset_weights), keep_history=int(keep_history),
                     threads=threads, ancestor_sampling=int(ancestor_sampling),
                     ess_threshold=float(ess_threshold))
    out = OrcSweepOut()
    res = dict(log_mean_w=np.empty(T), ess=np.empty(T), filt_mean=np.empty((T, d)))
    out.log_mean_w, out.ess = _dp(res["log_mean_w"]), _dp(res["ess"])
    out.filt_mean = _dp(res["filt_mean"])
    if want_final:
        res["x_final"] = np.empty((n, d))
        out.x_final = _dp(res["x_final"])
    if ess_threshold > 0.0:
        res["lw_final"] = np.empty(n)
        out.lw_final = _dp(res["lw_final"])
    if keep_history:
        res["anc_hist"] = np.empty((T, n), dtype=np.int32)
        res["x_hist"] = np.empty((T + 1, n, d))
        out.anc_hist, out.x_hist = _ip(res["anc_hist"]), _dp(res["x_hist"])
    ret = _f64(retained) if retained is not None else None
    rc = lib().orc_smc_sweep(C.byref(model.c), _dp(y), _dp(ret), C.byref(o), C.byref(out))
    if rc:
        raise ValueError(f"oracle sweep failed rc={rc}")
    res["log_z"] = out.log_z
    res["_out"] = out
    res["_n"], res["_T"], res["_d"] = n, T, d
    return res


def importance_sweep(model, y, n, seed):
    """`infer :importance` over a state-space program (importance.cljc:16-22)"""
    y = _f64(y)
    x, lw = np.empty(n), np.empty(n)
    rc = lib().orc_importance_sweep(C.byref(model.c), _dp(y), y.size, n, seed, _dp(x), _dp(lw))
    if rc:
        raise ValueError(f"oracle importance sweep failed rc={rc}")
    return {"x_final": x, "lw_final": lw, "log_z": log_marginal(lw)}


def trace_lineage(res, slot):
    T, d = res["_T"], res["_d"]
    traj = np.empty((T + 1, d))
    lib().orc_trace_lineage(C.byref(res["_out"]), res["_n"], T, d, int(slot), _dp(traj))
    return traj


def kalman1d(model, y):
    y = _f64(y)
    fm = np.empty(y.size)
    lz = lib().orc_kalman1d_logz(C.byref(model.c), _dp(y), y.size, _dp(fm))
    return lz, fm


def kalman_mvn(model, y):
    y = _f64(y)
    d = model.dim
    T = y.size // d
    fm = np.empty((T, d))
    lz = C.c_double()
    rc = lib().orc_kalman_mvn_logz(C.byref(model.c), _dp(y), T, C.byref(lz), _dp(fm))
    if rc:
        raise ValueError("kalman failed")
    return lz.value, fm


def hmm_forward_logz(model, y):
    y = _f64(y)
    return lib().orc_hmm_forward_logz(C.byref(model.c), _dp(y), y.size)


def hmm_smooth(model, y):
    y = _f64(y)
    post = np.empty((y.size + 1, model.dim))
    lib().orc_hmm_smooth(C.byref(model.c), _dp(y), y.size, _dp(post))
    return post


# ---------------------------------------------------------------- anglican.stat
# Literal restatements of src/anglican/stat.cljc (test infrastructure; loops in
# the reference's own left-fold order, so only for small cases).
def empirical_expectation(f, weighted):
    """stat.cljc:56-74: weights exp(lw - max lw), left fold of (v*w + sum) and w"""
    xs = [np.asarray(f(x), dtype=np.float64) for x, _ in weighted]
    lws = [float(lw) for _, lw in weighted]
    max_lw = lws[0]
    for l in lws[1:]:
        max_lw = l if l > max_lw else max_lw
    sum_wv, sum_w = 0.0, 0.0
    for v, lw in zip(xs, lws):
        w = math.exp(lw - max_lw)
        sum_wv = v * w + sum_wv
        sum_w = sum_w + w
    return sum_wv / sum_w


def empirical_mean(weighted):
    return empirical_expectation(lambda x: x, weighted)          # stat.cljc:76-82


def empirical_variance(weighted):
    mu = empirical_mean(weighted)                                # stat.cljc:93-98
    return empirical_expectation(lambda x: (np.asarray(x, float) - mu) ** 2, weighted)


def empirical_std(weighted):
    return np.sqrt(empirical_variance(weighted))                 # stat.cljc:100-104


def empirical_skew(weighted):
    mu, s = empirical_mean(weighted), empirical_std(weighted)    # stat.cljc:106-112
    return empirical_expectation(lambda x: ((np.asarray(x, float) - mu) / s) ** 3, weighted)


def empirical_kurtosis(weighted):
    mu, s = empirical_mean(weighted), empirical_std(weighted)    # stat.cljc:114-120
    return empirical_expectation(lambda x: ((np.asarray(x, float) - mu) / s) ** 4, weighted)


def empirical_covariance(weighted):
    mu = empirical_mean(weighted)                                # stat.cljc:84-91
    dx = [(np.asarray(x, float) - mu, lw) for x, lw in weighted]
    return empirical_expectation(lambda d: np.outer(d, d), dx)


def collect_by(values, lws):
    """stat.cljc:7-24: {v: log sum exp(lw - log N)} in order of first appearance"""
    log_norm = math.log(len(values))
    out = {}
    for v, lw in zip(values, lws):
        l = lw - log_norm
        out[v] = log_sum_exp(out[v], l) if v in out else l
    return out


def empirical_distribution(weighted):
    """stat.cljc:43-54: weighted = iterable of (x, log-w)"""
    weighted = list(weighted)
    log_z = None
    for lw in sorted(lw for _, lw in weighted):
        log_z = lw if log_z is None else log_sum_exp(log_z, lw)
    dist = {}
    for x, lw in weighted:
        dist[x] = dist.get(x, 0.0) + math.exp(lw - log_z)
    return dist


# ---------------------------------------------------------------- more runtime distributions
# commons-math 3.6.1 log-densities behind `from-apache` (runtime.cljc:207-220)
# and the defdist bodies of runtime.cljc:241-475.  Scalar restatements
# (test infrastructure); checked against scipy in tests/test_oracle_golden.py.
def exponential_logpdf(x, rate):
    mean = 1.0 / rate                                   # (Exponential (/ 1. rate)), runtime.cljc:310-311
    return -math.inf if x < 0 else -x / mean - math.log(mean)


def laplace_logpdf(x, loc, scale):
    d = math.exp(-abs(x - loc) / scale) / (2.0 * scale)  # LaplaceDistribution.density
    return math.log(d) if d > 0 else -math.inf


def poisson_logpmf(x, lam):
    if x < 0 or x != math.floor(x):
        return -math.inf
    if x == 0:
        return -lam
    return x * math.log(lam) - lam - math.lgamma(x + 1.0)


def bernoulli_logpmf(x, p):
    v = p if x == 1 else (1.0 - p) if x == 0 else 0.0    # runtime.cljc:245-249
    return math.log(v) if v > 0 else -math.inf


def student_t_logpdf(x, nu, loc=0.0, scale=1.0, loc_scale=False):
    v = (x - loc) / scale if loc_scale else x
    h = (nu + 1.0) / 2.0
    lp = (math.lgamma(h) - 0.5 * (math.log(math.pi) + math.log(nu)) - math.lgamma(nu / 2.0)
          - h * math.log(1.0 + v * v / nu))
    return lp - math.log(scale) if loc_scale else lp     # runtime.cljc:446-448


def chi_squared_logpdf(x, nu):
    return gamma_logpdf_direct(x, nu * 0.5, 0.5)         # runtime.cljc:333-340


def uniform_discrete_logpmf(x, lo, hi):
    lo, up = int(lo), int(hi) - 1                        # runtime.cljc:466-467
    if x != math.floor(x) or x < lo or x > up:
        return -math.inf
    return math.log(1.0 / (up - lo + 1))


def binomial_logpmf(x, n, p):
    n = int(n)
    if x != math.floor(x) or x < 0 or x > n:
        return -math.inf
    r = math.lgamma(n + 1.0) - math.lgamma(x + 1.0) - math.lgamma(n - x + 1.0)
    if x > 0:
        r += x * math.log(p)
    if x < n:
        r += (n - x) * math.log1p(-p)
    return r


DIST_LOGP = {
    "exponential": exponential_logpdf, "laplace": laplace_logpdf, "poisson": poisson_logpmf,
    "bernoulli": bernoulli_logpmf, "student-t": student_t_logpdf,
    "student-t-loc-scale": lambda x, nu, loc, scale: student_t_logpdf(x, nu, loc, scale, True),
    "chi-squared": chi_squared_logpdf, "uniform-discrete": uniform_discrete_logpmf,
    "binomial": binomial_logpmf,
}
