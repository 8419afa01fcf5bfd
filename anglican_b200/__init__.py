"""anglican_b200 — B200-native particle-population engine behind Anglican's
`infer` (:smc / :pimh / :pgibbs / :ipmcmc) and `anglican.runtime`
distribution protocol.  All particle arithmetic runs in hand-written CUDA
(sm_100a) inside libanglican_pop.so; this package is the host-side mirror of
the reference interface.  There is no CPU fallback."""
from . import _lib
from .population import Column, Population, PopulationGroup

__all__ = ["_lib", "Column", "Population", "PopulationGroup"]
