"""Device-resident particle population: the host-side handle over one apop_ctx.

Mirrors the data plane of the reference's particle list: the `:log-weight`
field of every particle's state map (src/anglican/state.cljc:22-36) and any
per-particle program values become SoA columns in HBM.
"""
import ctypes as C
import math

import numpy as np

from . import _lib as L


class Column:
    """A device column (f64 or i32, `width` values per particle)."""

    def __init__(self, pop, handle, dtype, width):
        self.pop, self.handle, self.dtype, self.width = pop, handle, dtype, width

    @property
    def np_dtype(self):
        return np.float64 if self.dtype == L.F64 else np.int32

    def upload(self, data):
        a = np.ascontiguousarray(data, dtype=self.np_dtype)
        L.check(self.pop.ctx, self.pop.lib.apop_column_upload(
            self.pop.ctx, self.handle, a.ctypes.data_as(C.c_void_p), a.size))
        return self

    def download(self, n=None):
        n = self.pop.n if n is None else n
        out = np.empty(n * self.width, dtype=self.np_dtype)
        L.check(self.pop.ctx, self.pop.lib.apop_column_download(
            self.pop.ctx, self.handle, out.ctypes.data_as(C.c_void_p), out.size))
        return out.reshape(n, self.width) if self.width > 1 else out

    def fill(self, v):
        L.check(self.pop.ctx, self.pop.lib.apop_column_fill(self.pop.ctx, self.handle, float(v)))
        return self

    def free(self):
        if self.handle is not None:
            L.check(self.pop.ctx, self.pop.lib.apop_column_free(self.pop.ctx, self.handle))
            self.handle = None


def _arg(v):
    """scalar or Column -> ApopArg"""
    if isinstance(v, Column):
        return L.ApopArg(v.handle, 0.0)
    return L.ApopArg(-1, float(v))


def _dp(a):
    return None if a is None else a.ctypes.data_as(L.c_double_p)


class Population:
    def __init__(self, n, seed=0xA11CE, device=0, rank=0, world=1, nccl_id=None):
        self.lib = L.load()
        self.ctx = C.c_void_p()
        idbuf = None
        if nccl_id is not None:
            idbuf = C.create_string_buffer(bytes(nccl_id), 128)
        rc = self.lib.apop_create(C.byref(self.ctx), int(n), int(seed), int(device), int(rank),
                                  int(world), idbuf)
        if rc != L.APOP_OK:
            raise L.ApopError(rc, "apop_create failed (no CUDA device? there is no CPU fallback)")
        self.n = int(n)
        self.seed = int(seed)
        self.rank, self.world = rank, world
        self._pin = None

    # ------------------------------------------------------------ anglican.stat on the device
    @staticmethod
    def _colh(col):
        if col is None:
            return -1
        if isinstance(col, str):
            assert col == "final"
            return L.COL_SWEEP_FINAL
        return col.handle

    def empirical_moments(self, x, logw=None, width=None):
        """empirical-mean / -variance / -skew / -kurtosis of stat.cljc:76-120 per
        column of `x` (a Column, or "final" for the last sweep's final states),
        weighted by exp(logw - max) (logw None: equal weights)."""
        w = width if width is not None else (x.width if not isinstance(x, str) else 1)
        out = [np.empty(w) for _ in range(4)]
        L.check(self.ctx, self.lib.apop_empirical_moments(self.ctx, self._colh(x), self._colh(logw),
                                                          int(w), *[_dp(o) for o in out]))
        return {"mean": out[0], "variance": out[1], "std": np.sqrt(out[1]), "skew": out[2],
                "kurtosis": out[3]}

    def empirical_covariance(self, x, logw=None, width=None):
        """empirical-mean and empirical-covariance (stat.cljc:84-91)"""
        w = width if width is not None else (x.width if not isinstance(x, str) else 1)
        mean, cov = np.empty(w), np.empty((w, w))
        L.check(self.ctx, self.lib.apop_empirical_covariance(self.ctx, self._colh(x),
                                                             self._colh(logw), int(w), _dp(mean),
                                                             _dp(cov)))
        return mean, cov

    def collect_discrete(self, x, K, logw=None):
        """collect-by + empirical-distribution over the values 0..K-1
        (stat.cljc:7-54): (log weight per value, normalised probabilities)"""
        lw, pr = np.empty(K), np.empty(K)
        L.check(self.ctx, self.lib.apop_collect_discrete(self.ctx, self._colh(x), self._colh(logw),
                                                         int(K), _dp(lw), _dp(pr)))
        return lw, pr

    # ------------------------------------------------------------ lifecycle
    def close(self):
        if self.ctx:
            if getattr(self, "_pin", None) is not None:
                self.lib.apop_host_free(self.ctx, self._pin[0])
                self._pin = None
            self.lib.apop_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def synchronize(self):
        L.check(self.ctx, self.lib.apop_synchronize(self.ctx))

    @property
    def launches(self):
        return int(self.lib.apop_launch_count(self.ctx))

    def set_step(self, t):
        L.check(self.ctx, self.lib.apop_set_step(self.ctx, int(t)))

    # timers on the ctx stream
    def timer_begin(self, tag):
        L.check(self.ctx, self.lib.apop_timer_begin(self.ctx, tag))

    def timer_end(self, tag):
        L.check(self.ctx, self.lib.apop_timer_end(self.ctx, tag))

    def timer_ms(self, tag):
        ms, cnt = C.c_double(), C.c_int64()
        L.check(self.ctx, self.lib.apop_timer_elapsed_ms(self.ctx, tag, C.byref(ms), C.byref(cnt)))
        return ms.value, cnt.value

    def enable_kernel_timers(self, on=True):
        L.check(self.ctx, self.lib.apop_enable_kernel_timers(self.ctx, int(on)))

    def timer_reset(self):
        L.check(self.ctx, self.lib.apop_timer_reset(self.ctx))

    # ------------------------------------------------------------ columns
    def column(self, dtype=L.F64, width=1, data=None, fill=None):
        h = C.c_int32()
        L.check(self.ctx, self.lib.apop_column_alloc(self.ctx, dtype, width, C.byref(h)))
        col = Column(self, h.value, dtype, width)
        if data is not None:
            col.upload(data)
        elif fill is not None:
            col.fill(fill)
        return col

    # ------------------------------------------------------------ weights
    def normalize(self, logw):
        m, lm, ess = C.c_double(), C.c_double(), C.c_double()
        L.check(self.ctx, self.lib.apop_normalize(self.ctx, logw.handle, C.byref(m), C.byref(lm),
                                                  C.byref(ess)))
        return m.value, lm.value, ess.value

    def resample(self, logw, u=math.nan, n_new=None, scheme=L.SYSTEMATIC, order=L.ORDER_REFERENCE,
                 anc=None):
        n_new = self.n if n_new is None else n_new
        anc = anc or self.column(L.I32)
        new_lw, j = C.c_double(), C.c_int64()
        L.check(self.ctx, self.lib.apop_resample(self.ctx, logw.handle, scheme, float(u), n_new,
                                                 order, anc.handle, C.byref(new_lw), C.byref(j)))
        return anc, new_lw.value, j.value

    def resample_weights(self, w, u=math.nan, n_new=None, scheme=L.SYSTEMATIC,
                         order=L.ORDER_REFERENCE, anc=None):
        n_new = self.n if n_new is None else n_new
        anc = anc or self.column(L.I32)
        total, j = C.c_double(), C.c_int64()
        L.check(self.ctx, self.lib.apop_resample_weights(self.ctx, w.handle, scheme, float(u),
                                                         n_new, order, anc.handle,
                                                         C.byref(total), C.byref(j)))
        return anc, total.value, j.value

    def resample_multinomial_u(self, w, uniforms, n_new=None, anc=None):
        n_new = self.n if n_new is None else n_new
        anc = anc or self.column(L.I32)
        L.check(self.ctx, self.lib.apop_resample_multinomial_u(self.ctx, w.handle, uniforms.handle,
                                                               n_new, anc.handle))
        return anc

    def gather(self, anc, cols, n_new=None):
        n_new = self.n if n_new is None else n_new
        arr = (C.c_int32 * len(cols))(*[c.handle for c in cols])
        L.check(self.ctx, self.lib.apop_gather(self.ctx, anc.handle, n_new, arr, len(cols)))

    # ------------------------------------------------------------ distributions
    def sample_normal(self, mean, sd, out, draw=0):
        L.check(self.ctx, self.lib.apop_sample_normal(self.ctx, _arg(mean), _arg(sd), draw, out.handle))

    def observe_normal(self, mean, sd, value, logw):
        L.check(self.ctx, self.lib.apop_observe_normal(self.ctx, _arg(mean), _arg(sd), _arg(value),
                                                       logw.handle))

    def sample_gamma(self, shape, rate, out, draw=0):
        L.check(self.ctx, self.lib.apop_sample_gamma(self.ctx, _arg(shape), _arg(rate), draw, out.handle))

    def observe_gamma(self, shape, rate, value, logw):
        L.check(self.ctx, self.lib.apop_observe_gamma(self.ctx, _arg(shape), _arg(rate), _arg(value),
                                                      logw.handle))

    def sample_beta(self, a, b, out, draw=0):
        L.check(self.ctx, self.lib.apop_sample_beta(self.ctx, _arg(a), _arg(b), draw, out.handle))

    def observe_beta(self, a, b, value, logw):
        L.check(self.ctx, self.lib.apop_observe_beta(self.ctx, _arg(a), _arg(b), _arg(value),
                                                     logw.handle))

    def sample_uniform(self, lo, hi, out, draw=0):
        L.check(self.ctx, self.lib.apop_sample_uniform(self.ctx, _arg(lo), _arg(hi), draw, out.handle))

    def observe_uniform(self, lo, hi, value, logw):
        L.check(self.ctx, self.lib.apop_observe_uniform(self.ctx, _arg(lo), _arg(hi), _arg(value),
                                                        logw.handle))

    def _family(self, name, params):
        fam, npar = L.DIST_FAMILIES[str(name).lstrip(":").replace("_", "-")]
        if len(params) != npar:
            raise TypeError(f"{name} takes {npar} parameter(s)")           # defdist arity
        a = (L.ApopArg * 3)(*([_arg(p) for p in params] + [_arg(0.0)] * (3 - npar)))
        return fam, a

    def sample_dist(self, name, params, out, draw=0):
        """sample* of exponential / laplace / poisson / bernoulli (flip) / student-t /
        student-t-loc-scale / chi-squared / uniform-discrete / binomial
        (runtime.cljc:241-475); params are scalars or Columns"""
        fam, a = self._family(name, params)
        L.check(self.ctx, self.lib.apop_sample_dist(self.ctx, fam, a, int(draw), out.handle))
        return out

    def observe_dist(self, name, params, value, logw):
        """observe* of the same families: logw += log p(value)"""
        fam, a = self._family(name, params)
        L.check(self.ctx, self.lib.apop_observe_dist(self.ctx, fam, a, _arg(value), logw.handle))
        return logw

    def sample_discrete(self, weights, out, draw=0):
        if isinstance(weights, Column):
            L.check(self.ctx, self.lib.apop_sample_discrete(self.ctx, None, weights.width,
                                                            weights.handle, draw, out.handle))
        else:
            w = np.ascontiguousarray(weights, dtype=np.float64)
            L.check(self.ctx, self.lib.apop_sample_discrete(self.ctx, _dp(w), w.size, -1, draw,
                                                            out.handle))

    def observe_discrete(self, weights, value, logw):
        vcol, vs = (value.handle, 0) if isinstance(value, Column) else (-1, int(value))
        if isinstance(weights, Column):
            L.check(self.ctx, self.lib.apop_observe_discrete(self.ctx, None, weights.width,
                                                             weights.handle, vcol, vs, logw.handle))
        else:
            w = np.ascontiguousarray(weights, dtype=np.float64)
            L.check(self.ctx, self.lib.apop_observe_discrete(self.ctx, _dp(w), w.size, -1, vcol, vs,
                                                             logw.handle))

    def sample_dirichlet(self, alpha, out, draw=0):
        a = np.ascontiguousarray(alpha, dtype=np.float64)
        L.check(self.ctx, self.lib.apop_sample_dirichlet(self.ctx, _dp(a), a.size, draw, out.handle))

    def observe_dirichlet(self, alpha, value, logw):
        a = np.ascontiguousarray(alpha, dtype=np.float64)
        L.check(self.ctx, self.lib.apop_observe_dirichlet(self.ctx, _dp(a), a.size, value.handle,
                                                          logw.handle))

    def sample_mvn(self, mean, cov, out, draw=0):
        cov = np.ascontiguousarray(cov, dtype=np.float64)
        d = cov.shape[0]
        if isinstance(mean, Column):
            L.check(self.ctx, self.lib.apop_sample_mvn(self.ctx, None, mean.handle, _dp(cov), d,
                                                       draw, out.handle))
        else:
            m = np.ascontiguousarray(mean, dtype=np.float64)
            L.check(self.ctx, self.lib.apop_sample_mvn(self.ctx, _dp(m), -1, _dp(cov), d, draw,
                                                       out.handle))

    def observe_mvn(self, mean, cov, value, logw):
        cov = np.ascontiguousarray(cov, dtype=np.float64)
        d = cov.shape[0]
        mh, mp, vh, vp = -1, None, -1, None
        keep = []
        if isinstance(mean, Column):
            mh = mean.handle
        else:
            keep.append(np.ascontiguousarray(mean, dtype=np.float64))
            mp = _dp(keep[-1])
        if isinstance(value, Column):
            vh = value.handle
        else:
            keep.append(np.ascontiguousarray(value, dtype=np.float64))
            vp = _dp(keep[-1])
        L.check(self.ctx, self.lib.apop_observe_mvn(self.ctx, mp, mh, _dp(cov), d, vp, vh,
                                                    logw.handle))

    # ------------------------------------------------------------ fused sweeps
    @staticmethod
    def _vec_or_col(v, d):
        """(host pointer, column handle) of a shared vector or a per-particle column"""
        if isinstance(v, Column):
            return None, v.handle, None
        a = np.ascontiguousarray(v, dtype=np.float64).reshape(-1)
        assert a.size == d
        return _dp(a), -1, a

    def sample_mvt(self, nu, mean, sigma, out, draw=0):
        """`(sample* (multivariate-t nu mean sigma))`, runtime.cljc:371-382"""
        sg = np.ascontiguousarray(sigma, dtype=np.float64)
        d = sg.shape[0]
        mp, mc, _keep = self._vec_or_col(mean, d)
        L.check(self.ctx, self.lib.apop_sample_mvt(self.ctx, float(nu), mp, mc, _dp(sg), d, draw,
                                                   out.handle))

    def observe_mvt(self, nu, mean, sigma, value, logw):
        """`(observe* (multivariate-t nu mean sigma) value)`, runtime.cljc:383-401"""
        sg = np.ascontiguousarray(sigma, dtype=np.float64)
        d = sg.shape[0]
        mp, mc, _k1 = self._vec_or_col(mean, d)
        vp, vc, _k2 = self._vec_or_col(value, d)
        L.check(self.ctx, self.lib.apop_observe_mvt(self.ctx, float(nu), mp, mc, _dp(sg), d, vp, vc,
                                                    logw.handle))

    def sample_wishart(self, n_df, V, out, draw=0):
        """`(sample* (wishart n V))`: Bartlett decomposition, runtime.cljc:519-535"""
        v = np.ascontiguousarray(V, dtype=np.float64)
        L.check(self.ctx, self.lib.apop_sample_wishart(self.ctx, float(n_df), _dp(v), v.shape[0], draw,
                                                       out.handle))

    def observe_wishart(self, n_df, V, value, logw):
        """`(observe* (wishart n V) value)`, runtime.cljc:536-539; value: column of width p*p"""
        v = np.ascontiguousarray(V, dtype=np.float64)
        L.check(self.ctx, self.lib.apop_observe_wishart(self.ctx, float(n_df), _dp(v), v.shape[0],
                                                        value.handle, logw.handle))

    def take(self, table, index, out):
        """out[k] = table[index[k]]: per-particle parameters from a discrete state column"""
        t = np.ascontiguousarray(table, dtype=np.float64)
        t2 = t.reshape(t.shape[0], -1)
        L.check(self.ctx, self.lib.apop_column_take(self.ctx, _dp(t2), t2.shape[0], t2.shape[1],
                                                    index.handle, out.handle))

    def set_seed(self, seed):
        L.check(self.ctx, self.lib.apop_set_seed(self.ctx, int(seed)))
        self.seed = int(seed)

    def _pinned_f64(self, count):
        """page-locked result buffer owned by this population (reused by every
        sweep that asks for it; freed on close)"""
        if self._pin is None or self._pin[1] < count:
            if self._pin is not None:
                L.check(self.ctx, self.lib.apop_host_free(self.ctx, self._pin[0]))
                self._pin = None
            p = C.c_void_p()
            L.check(self.ctx, self.lib.apop_host_alloc(self.ctx, C.byref(p), int(count) * 8))
            self._pin = (p, int(count))
        buf = (C.c_double * count).from_address(self._pin[0].value)
        return np.frombuffer(buf, dtype=np.float64, count=count)

    def sweep(self, model, y, retained=None, csmc=False, n_cond_steps=0, reset_weights=False,
              keep_history=False, want_final=True, want_stats=True, want_filt=None,
              pinned_final=False, no_resample=False, ancestor_sampling=False, use_graph=0,
              ess_threshold=0.0):
        """One `sweep` (smc.cljc:35-51; conditional: pgibbs.cljc:76-110,
        ipmcmc.cljc:34-83).  Returns a SweepOutput.  pinned_final: x_final is a
        view of a page-locked buffer owned by the population, overwritten by its
        next sweep (copy it to keep it).  no_resample: `infer :importance`
        (importance.cljc:16-22) — weights accumulate per particle, the output
        carries lw_final and log_z = log mean exp(lw_final).  keep_history: False,
        True / 1 (ancestors and states of every step stay in HBM: trace, history,
        trajectories) or "compact" / 2 (two bits per particle and step, trace only;
        scalar-state models).  ancestor_sampling:
        `infer :pgas` (pgas.cljc:107-131) — in a conditional sweep the retained
        particle's ancestor is redrawn at every conditional observe.  ess_threshold
        (extension; the reference resamples at every observe): in (0, 1], resample only
        when ESS < ess_threshold * N; the output then carries lw_final, the log-weights of
        the final particles (all equal if the last observe resampled)."""
        y = np.ascontiguousarray(y, dtype=np.float64)
        d = model.state_dim
        T = y.size // d
        if no_resample:                      # no resample, no per-step weight statistics
            want_stats, want_filt = False, False
        if want_filt is None:
            want_filt = want_stats and d == 1
        if keep_history == "compact":
            keep_history = 2 if d == 1 else 1
        opts = L.ApopSweepOpts(int(csmc), int(n_cond_steps), int(reset_weights),
                               int(keep_history), int(use_graph), int(want_final), int(bool(want_filt)),
                               int(bool(no_resample)), int(bool(ancestor_sampling)), 0,
                               float(ess_threshold))
        res = L.ApopSweepResult()
        out = SweepOutput(self, T, d)
        if want_stats:
            out.log_mean_w = np.empty(T)
            out.ess = np.empty(T)
            res.log_mean_w, res.ess = _dp(out.log_mean_w), _dp(out.ess)
        if want_filt:
            out.filt_mean = np.empty(T * d)
            res.filt_mean = _dp(out.filt_mean)
        if want_final:
            if pinned_final:
                flat = self._pinned_f64(self.n * d)
                out.x_final = flat.reshape(self.n, d) if d > 1 else flat
            else:
                out.x_final = np.empty((self.n, d)) if d > 1 else np.empty(self.n)
            res.x_final = _dp(out.x_final)
        if no_resample or ess_threshold > 0.0:
            out.lw_final = np.empty(self.n)
            res.lw_final = _dp(out.lw_final)
        ret = None
        if retained is not None:
            ret = np.ascontiguousarray(retained, dtype=np.float64)
        L.check(self.ctx, self.lib.apop_smc_sweep(self.ctx, C.byref(model.c), _dp(y), T, _dp(ret),
                                                  C.byref(opts), C.byref(res)))
        out.log_z = res.log_z
        out.has_history = bool(keep_history)
        return out


class SweepOutput:
    """Result of one sweep; ancestry stays on the device until asked for."""

    def __init__(self, pop, T, d):
        self.pop, self.T, self.d = pop, T, d
        self.log_z = None
        self.log_mean_w = self.ess = self.filt_mean = self.x_final = self.lw_final = None
        self.has_history = False
        self._hist = None

    def trace(self, slot):
        """x_0..x_T of the final particle in `slot` (its `::past-samples`,
        pgibbs.cljc:31-35)."""
        out = np.empty((self.T + 1) * self.d)
        L.check(self.pop.ctx, self.pop.lib.apop_trace_lineage(self.pop.ctx, int(slot), _dp(out)))
        return out.reshape(self.T + 1, self.d) if self.d > 1 else out

    def history(self):
        if self._hist is None:
            n = self.pop.n
            anc = np.empty((self.T, n), dtype=np.int32)
            x = np.empty((self.T + 1, n, self.d))
            L.check(self.pop.ctx, self.pop.lib.apop_sweep_download_history(
                self.pop.ctx, anc.ctypes.data_as(L.c_int32_p), _dp(x)))
            self._hist = (anc, x)
        return self._hist

    def trajectories(self):
        """[n, T+1, d] state sequences of all final particles (host-side
        materialisation of every particle's `:result`)."""
        anc, x = self.history()
        n = self.pop.n
        out = np.empty((n, self.T + 1, self.d))
        k = np.arange(n)
        for t in range(self.T, 0, -1):
            k = anc[t - 1][k]
            out[:, t, :] = x[t][k]
        out[:, 0, :] = x[0][k]
        return out


class PopulationGroup:
    """`world` shards of ONE population inside this process.

    `devices[r]` is the GPU of rank r (all ranks may share one device: the
    shards then exchange through the same peer-memory protocol, which is how the
    sharded path is exercised on a single-GPU box).  One process per GPU needs
    no group: every rank builds `Population(n_local, rank=r, world=R,
    nccl_id=...)` and calls `sweep` collectively."""

    def __init__(self, n_local, world, seed=0xA11CE, devices=None):
        devices = list(devices) if devices is not None else [0] * world
        assert len(devices) == world and world >= 2
        self.lib = L.load()
        self.world, self.n_local, self.n = world, int(n_local), int(n_local) * world
        self.ranks = [Population(n_local, seed=seed, device=devices[r], rank=r, world=world)
                      for r in range(world)]
        self._arr = (C.c_void_p * world)(*[p.ctx for p in self.ranks])
        rc = self.lib.apop_link_group(self._arr, world)
        if rc != L.APOP_OK:
            L.check(self.ranks[0].ctx, rc)

    def close(self):
        for p in self.ranks:
            p.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_seed(self, seed):
        for p in self.ranks:
            p.set_seed(seed)

    def sweep(self, model, y, want_final=True, want_stats=True, want_filt=False):
        y = np.ascontiguousarray(y, dtype=np.float64)
        d = model.state_dim
        T = y.size // d
        want_filt = bool(want_filt) and d == 1
        opts = L.ApopSweepOpts(0, 0, 0, 0, 0, int(want_final), int(want_filt), 0)
        res = (L.ApopSweepResult * self.world)()
        outs = []
        for r, p in enumerate(self.ranks):
            out = SweepOutput(p, T, d)
            if want_stats:
                out.log_mean_w, out.ess = np.empty(T), np.empty(T)
                res[r].log_mean_w, res[r].ess = _dp(out.log_mean_w), _dp(out.ess)
            if want_filt:
                out.filt_mean = np.empty(T)
                res[r].filt_mean = _dp(out.filt_mean)
            if want_final:
                out.x_final = np.empty((self.n_local, d)) if d > 1 else np.empty(self.n_local)
                res[r].x_final = _dp(out.x_final)
            outs.append(out)
        rc = self.lib.apop_smc_sweep_group(self._arr, self.world, C.byref(model.c), _dp(y), T,
                                           C.byref(opts), res)
        if rc != L.APOP_OK:
            for p in self.ranks:
                msg = self.lib.apop_last_error(p.ctx)
                if msg:
                    raise L.ApopError(rc, msg.decode())
            raise L.ApopError(rc, "")
        for r, out in enumerate(outs):
            out.log_z = res[r].log_z
        return outs
