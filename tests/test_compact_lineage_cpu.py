"""Compact lineages (csrc/ssm.cu k_pack_lineage / k_trace_compact): the sorted
ancestors of one step are stored as two bit-vectors - new[j]: output slot j >= 1 starts a new
ancestor, alive[i]: particle i is an ancestor of a slot >= 1 - plus the ancestor of slot 0
(the retained particle of a conditional sweep, pgibbs.cljc:92-99).  This checks the algebra
the kernels rely on, anc[j] = select1(alive, rank1(new, j)), on ancestors produced by the
oracle's literal wheel (smc.cljc:86-137), in numpy: no GPU, no engine code."""
import numpy as np
import pytest

from oracle import pyoracle as orc


def encode(anc):
    """what k_pack_lineage writes for one step"""
    n = anc.size
    new = np.zeros(n, bool)
    if n > 1:
        new[1] = True
        new[2:] = anc[2:] != anc[1:-1]
    alive = np.zeros(n, bool)
    alive[anc[new]] = True
    return new, alive, int(anc[0])


def decode_slot(new, alive, anc0, j):
    """what k_trace_compact computes for one step of the trace-back"""
    if j == 0:
        return anc0
    r = int(np.count_nonzero(new[1:j + 1]))                  # rank1(new, j), 1-based
    return int(np.flatnonzero(alive)[r - 1])                 # select1(alive, r)


def sorted_ancestors(lw, n_new, u):
    w, _ = orc.particle_weights(lw)
    r = orc.resample_weights(w, n_new, u)
    s, _ = orc.emission_to_sorted(r["emission"], r["lap"])
    assert np.all(np.diff(s) >= 0)                           # the engine's physical order
    return s


def weight_cases(rng, n):
    yield "gaussian", rng.normal(0, 3, n)
    yield "equal", np.zeros(n)
    dom = rng.normal(0, 1, n); dom[n // 3] = 60.0
    yield "one dominant", dom
    z = rng.normal(0, 2, n); z[rng.random(n) < 0.4] = -np.inf
    yield "zeros", z
    yield "two survivors", np.where(np.arange(n) % (n // 2 + 1) == 0, 0.0, -np.inf)


@pytest.mark.parametrize("n", [2, 33, 64, 1000, 4097])
def test_two_bit_vectors_reproduce_the_sorted_ancestors(n):
    rng = np.random.default_rng(n)
    for name, lw in weight_cases(rng, n):
        for u in (0.0, 0.37, 0.999999):
            for conditional in (False, True):
                if conditional:                              # N - 1 offspring behind the retained slot
                    anc = np.concatenate([[0], sorted_ancestors(lw, n - 1, u)]).astype(np.int32)
                else:
                    anc = sorted_ancestors(lw, n, u)
                    anc = np.concatenate([[int(rng.integers(0, n))], anc[1:]]).astype(np.int32)
                new, alive, anc0 = encode(anc)
                assert np.count_nonzero(new) == np.count_nonzero(alive), (name, n, u)
                got = np.array([decode_slot(new, alive, anc0, j) for j in range(n)])
                assert np.array_equal(got, anc), (name, n, u, conditional)


def test_trace_back_through_many_steps():
    rng = np.random.default_rng(3)
    n, T = 257, 40
    steps = []
    for t in range(T):
        anc = sorted_ancestors(rng.normal(0, 2.5, n), n, float(rng.random()))
        steps.append((anc, encode(anc)))
    for slot in (0, 1, 100, n - 1):
        a = b = slot
        for anc, (new, alive, anc0) in reversed(steps):
            a = int(anc[a])
            b = decode_slot(new, alive, anc0, b)
            assert a == b
