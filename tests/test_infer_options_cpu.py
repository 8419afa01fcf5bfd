"""Option handling of the `infer` mirror (anglican_b200/inference.py): the
reference's keywords, defaults and assertions, checked without a GPU (every
assertion fires before a population is created; a valid call must then fail
LOUDLY for want of a device -- there is no CPU path)."""
import math

import numpy as np
import pytest

from anglican_b200 import _lib as L
from anglican_b200.inference import State, doquery, infer, log_marginal
from anglican_b200.models import LinearGaussian1D

Y = np.array([0.1, -0.2, 0.3])


def _first(gen):
    try:
        return next(gen)
    finally:
        gen.close()


@pytest.mark.parametrize("alg,opts,msg", [
    ("smc", dict(number_of_particles=0), ":number-of-particles must be at least 1"),      # smc.cljc:56-57
    ("pimh", dict(number_of_particles=0), ":number-of-particles must be at least 1"),     # pimh.cljc:44-45
    ("pgibbs", dict(number_of_particles=1), ":number-of-particles must be at least 2"),   # pgibbs.cljc:115-116
    ("pgas", dict(number_of_particles=1), ":number-of-particles must be at least 2"),     # pgas.cljc:158-159
    ("ipmcmc", dict(number_of_particles=1), ":number-of-particles must be larger than 1"),  # ipmcmc.cljc:157-160
    ("ipmcmc", dict(number_of_nodes=1), ":number-of-nodes must be larger 1"),
    ("ipmcmc", dict(number_of_nodes=4, number_of_csmc_nodes=4),
     ":number-of-csmc-nodes must be smaller than :number-of-nodes"),                      # ipmcmc.cljc:166-171
    (":ipmcmc", dict(number_of_nodes=4, number_of_csmc_nodes=7),
     ":number-of-csmc-nodes must be smaller than :number-of-nodes"),
])
def test_reference_assertions(alg, opts, msg):
    with pytest.raises(AssertionError) as e:
        _first(infer(alg, LinearGaussian1D(), Y, **opts))
    assert msg in str(e.value)


def test_unknown_algorithm():
    with pytest.raises(ValueError):
        infer("lmh", LinearGaussian1D(), Y)            # core.cljc:15-26: no such namespace here


@pytest.mark.parametrize("alg", ["smc", ":smc", "pimh", "pgibbs", "pgas", "ipmcmc", "importance",
                                 "pfilter"])
def test_valid_call_without_device_fails_loudly(alg):
    if L.load().apop_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(L.ApopError):
        _first(infer(alg, LinearGaussian1D(), Y))      # defaults of the reference: N = 1 / 2, M = 32


def test_doquery_drops_invalid_and_log_marginal():
    states = [State(-math.inf, 0.0), State(math.log(0.25), 1.0), State(math.log(0.75), 2.0)]
    # inference.cljc:181-186: log mean weight over ALL states
    assert abs(log_marginal(states) - math.log(1.0 / 3.0)) < 1e-12
    import anglican_b200.inference as inf
    real = inf._ALGORITHMS["smc"]
    inf._ALGORITHMS["smc"] = lambda model, y, **kw: iter(states)
    try:
        kept = list(doquery("smc", None, Y))           # inference.cljc:151-157
        assert [s.result for s in kept] == [1.0, 2.0]
        assert len(list(doquery("smc", None, Y, drop_invalid=False))) == 3
    finally:
        inf._ALGORITHMS["smc"] = real
