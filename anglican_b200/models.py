"""State-space program shapes served by the fused sweep (apop_smc_sweep).

Each class describes an Anglican program of the form
    (let [x0 (sample init)]
      (reduce (fn [x y] (let [x' (sample (trans x))] (observe (obs x') y) x')) x0 ys))
i.e. every particle reaches the same global `observe` at every step — the
condition smc.cljc:41-51 enforces.  Parameter conventions follow
anglican.runtime: `normal` is (mean, sd) (runtime.cljc:403-409), `discrete`
takes unnormalised weights (runtime.cljc:277-290), `mvn` is (mean, cov).
"""
import ctypes as C

import numpy as np

from . import _lib as L


def _f64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


class StateSpaceModel:
    kind = 0
    state_dim = 1

    def __init__(self):
        self._keep = {}
        self.c = L.ApopSSM()
        self.c.kind = self.kind
        self.c.dim = 1

    def _ptr(self, name, arr):
        arr = _f64(arr)
        self._keep[name] = arr
        setattr(self.c, name, arr.ctypes.data_as(L.c_double_p))
        return arr


class LinearGaussian1D(StateSpaceModel):
    """x0 ~ normal(m0, s0); x' ~ normal(a x + b, q); y ~ normal(c x' + d, r)
    (BASELINE config 2: a=0.9, q=1, c=1, r=1, m0=0, s0=1)."""
    kind = L.SSM_LGSSM1D

    def __init__(self, m0=0.0, s0=1.0, a=0.9, b=0.0, q=1.0, c=1.0, d=0.0, r=1.0):
        super().__init__()
        s = self.c
        s.m0, s.s0, s.a, s.b, s.q, s.c, s.d, s.r = m0, s0, a, b, q, c, d, r
        self.params = dict(m0=m0, s0=s0, a=a, b=b, q=q, c=c, d=d, r=r)


class StochasticVolatility(StateSpaceModel):
    """x0 ~ normal(0, sigma/sqrt(1-phi^2)); x' ~ normal(phi x, sigma);
    y ~ normal(0, beta exp(x'/2))   (BASELINE config 3: phi=.98 sigma=.16 beta=.65)."""
    kind = L.SSM_SV

    def __init__(self, phi=0.98, sigma=0.16, beta=0.65):
        super().__init__()
        s = self.c
        s.m0, s.s0 = 0.0, sigma / np.sqrt(1.0 - phi * phi)
        s.a, s.b, s.q, s.beta = phi, 0.0, sigma, beta
        self.params = dict(phi=phi, sigma=sigma, beta=beta)


class GaussianHMM(StateSpaceModel):
    """x0 ~ discrete(init); x' ~ discrete(trans[x]); y ~ normal(mu[x'], sd[x'])
    (BASELINE config 1 = algorithm_test.cljc:102-123)."""
    kind = L.SSM_HMM

    def __init__(self, init, trans, mu, sd):
        super().__init__()
        init = self._ptr("hmm_init", init)
        self._ptr("hmm_trans", trans)
        self._ptr("hmm_mu", mu)
        self._ptr("hmm_sd", sd)
        self.c.dim = init.size
        self.K = init.size


class LinearGaussianMVN(StateSpaceModel):
    """x0 ~ mvn(0, s0^2 I); x' ~ mvn(A x, Q); y ~ mvn(x', R)  (BASELINE config 5)."""
    kind = L.SSM_LGMVN

    def __init__(self, A, Q, R, s0=1.0):
        super().__init__()
        A = self._ptr("A", A)
        d = A.shape[0]
        Lq = np.linalg.cholesky(_f64(Q))
        self._ptr("Lq", Lq)
        self._ptr("R", R)
        self.c.dim = d
        self.c.s0 = s0
        self.state_dim = d


def as_apop(model):
    return C.byref(model.c)
