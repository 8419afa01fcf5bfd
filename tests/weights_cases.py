"""Shared weight families for the resampler parity tests (seeded)."""
import numpy as np


def families(n, rng):
    fam = {}
    lw = rng.normal(0, 3, n)
    fam["lognormal3"] = np.exp(lw - lw.max())
    fam["ones"] = np.ones(n)
    fam["halves"] = rng.integers(0, 5, n) * 0.5 + (np.arange(n) == 0)
    fam["zeros_mixed"] = np.where(rng.random(n) < 0.7, 0.0, rng.random(n)) + (np.arange(n) == n // 2)
    k = n // 4
    fam["leading_zeros"] = np.concatenate([np.zeros(k), rng.random(n - k) + 1e-3])
    t = np.array([1e-300, 1e-200, 1e-100, 1e-50])[: min(4, n)]
    fam["tiny_first"] = np.concatenate([t, rng.random(n - t.size) + 1e-3]) if n > 4 else np.ones(n)
    fam["ascending_range"] = np.exp(np.linspace(-700, 0, n))
    fam["descending_range"] = np.exp(np.linspace(0, -700, n))
    oh = rng.random(n) * 1e-12
    oh[n // 3] = 1.0
    fam["one_heavy"] = oh
    fam["pow2"] = 2.0 ** rng.integers(-60, 1, n)
    fam["dyadic_ties"] = rng.integers(1, 2 ** 20, n) * 2.0 ** -30
    return fam


U_VALUES = [0.0, 0.3, 0.77, 1 - 2 ** -53, 1e-12]
