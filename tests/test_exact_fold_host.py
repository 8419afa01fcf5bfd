"""CPU check of anglican_b200/csrc/exact_fold.cuh (the arithmetic the CUDA
resampler is built on) against the oracle's literal wheel, via the sequential
emulation in tests/host/exact_fold_host.cpp."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import pyoracle as orc
from tests.weights_cases import U_VALUES, families

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host", "exact_fold_host.cpp")
OUT = os.path.join(HERE, "host", "_build", "libexact_fold_host.so")
HDR = os.path.join(HERE, "..", "anglican_b200", "csrc", "exact_fold.cuh")
dp, ip, lp = C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_int64)


@pytest.fixture(scope="module")
def host():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    if (not os.path.exists(OUT) or os.path.getmtime(OUT) < max(os.path.getmtime(SRC),
                                                               os.path.getmtime(HDR))):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared",
                               "-x", "c++", SRC, "-o", OUT])
    lib = C.CDLL(OUT)
    lib.host_exact_resample.argtypes = [dp, C.c_int64, C.c_int64, C.c_double, C.c_int, ip, ip, dp, lp]
    lib.host_comb_values.argtypes = [C.c_double, C.c_double, C.c_int64, dp]
    lib.host_fold_fast.argtypes = [C.c_double, dp, C.c_int, dp, dp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    return lib


def _run(lib, w, n_new, u, tile):
    w = np.ascontiguousarray(w, dtype=np.float64)
    em = np.full(n_new, -1, np.int32)
    lap = np.full(n_new, -1, np.int32)
    tot = C.c_double()
    st = (C.c_int64 * 4)()
    rc = lib.host_exact_resample(w.ctypes.data_as(dp), w.size, n_new, u, tile,
                                 em.ctypes.data_as(ip), lap.ctypes.data_as(ip), C.byref(tot), st)
    return rc, em, lap, tot.value


@pytest.mark.parametrize("n", [1, 2, 3, 17, 256, 257, 4097, 30011])
def test_exact_fold_emulation_bit_exact(host, n):
    rng = np.random.default_rng(n)
    for name, w in families(n, rng).items():
        for u in U_VALUES:
            for n_new in sorted({n, max(1, n - 1), 7}):
                for tile in (256, 4096):
                    ref = orc.resample_weights(w, n_new, u)
                    rc, em, lap, tot = _run(host, w, n_new, u, tile)
                    assert rc == 0, (name, u, n_new, rc)
                    assert tot == ref["total"]
                    assert np.array_equal(em, ref["emission"]), (name, u, n_new, tile)
                    assert np.array_equal(lap, ref["lap"])


def test_comb_closed_form_equals_repeated_addition(host):
    for tot, u, n in [(1234.567, 0.3, 100000), (1.0, 0.0, 1000), (7.3e5, 1 - 2 ** -53, 300000),
                      (3.0, 1e-300, 1000), (2.0 ** 20, 0.5, 2 ** 20), (0.1, 0.999, 77777)]:
        out = np.empty(n)
        host.host_comb_values(tot, u, n, out.ctypes.data_as(dp))
        step, x = tot / float(n), tot * u
        ref = np.empty(n)
        for j in range(n):
            ref[j] = x
            x = x + step
        assert np.array_equal(out, ref), (tot, u, n)


@pytest.mark.parametrize("n,small,pos,giant", [
    (600000, 1e-30, 360000, 1.0), (600000, 1e-30, 570000, 1.0), (1 << 20, 1e-30, 777777, 1.5),
    (1 << 20, 1e-20, 777777, 1.0), (1 << 20, 1e-9, 777777, 1e6), (3 * 4096, 1e-30, 3 * 4096 - 1, 1.0),
    (100000, 1e-300, 5, 1e300), (100000, 5e-324, 99999, 1.0),
])
def test_dominant_weight_shapes(host, n, small, pos, giant):
    """one particle outweighing the rest by more than 2^53 (every other increment
    vanishes against the running sum); cf. tests/test_zz_dominant_weight_gpu.py"""
    w = np.full(n, small)
    w[pos] = giant
    for u in (0.37, 0.999999, 0.0):
        ref = orc.resample_weights(w, n, u)
        rc, em, lap, tot = _run(host, w, n, u, 4096)
        assert rc == 0, (u, rc)
        assert tot == ref["total"]
        assert np.array_equal(em, ref["emission"]) and np.array_equal(lap, ref["lap"])


def test_single_pointer_is_a_discrete_draw(host):
    """n_new = 1: the wheel is `(sample* (discrete weights))` (runtime.cljc:281-287),
    the draw `infer :pgas` uses for the retained particle's ancestor"""
    rng = np.random.default_rng(12)
    for n in (5, 4096, 30011):
        for name, w in families(n, rng).items():
            for u in (0.0, 0.25, 0.6180339887, 1 - 2 ** -53):
                ref = orc.resample_weights(w, 1, u)
                rc, em, lap, tot = _run(host, w, 1, u, 4096)
                assert rc == 0 and tot == ref["total"], (name, u, rc)
                assert em[0] == ref["emission"][0] and lap[0] == 0
                # first k with u*total < w_0 + ... + w_k (left fold, strict)
                edges = np.cumsum(w)
                assert em[0] == int(np.searchsorted(edges, u * edges[-1], side="right"))


def _fold_fast(host, s_in, w):
    w = np.ascontiguousarray(w, dtype=np.float64)
    edges = np.empty(w.size)
    s_out, ties, iters = C.c_double(), C.c_int(), C.c_int()
    rc = host.host_fold_fast(float(s_in), w.ctypes.data_as(dp), w.size, edges.ctypes.data_as(dp),
                             C.byref(s_out), C.byref(ties), C.byref(iters))
    return rc, edges, s_out.value, ties.value, iters.value


def test_fast_iteration_is_exact_across_binade_crossings(host):
    """cta_tile_fold's fast iteration (integer-valued doubles, additions only) against the
    literal left fold `(+ acc weight)` of smc.cljc:120: every edge bit for bit, for tiles
    that cross one or several binades, start just below a power of two, start from a sum
    smaller than their elements (elements saturate the binade of the running sum), or hold
    one weight that dwarfs the rest (partial sums far above 2^53 units, i.e. inexact, beyond
    the crossing)."""
    rng = np.random.default_rng(5)
    cases = []
    for n in (1, 7, 8, 255, 256, 1000, 4096):
        w = np.exp(rng.normal(0, 2.0, n))
        for s_in in (0.75, 1.0, 1.9999999, 3.0e5, 1e-7 * w.sum(), w.sum() * 0.99, 2.0 ** 30 - 0.5,
                     1e-30, 1e-280):
            cases.append((s_in, w))
        big = w.copy()
        big[n // 2] = w.sum() * 1000.0                      # one element jumps several binades
        cases.append((1.0, big))
        cases.append((w.sum(), np.full(n, w.mean())))        # equal weights
    crossed = 0
    for s_in, w in cases:
        rc, edges, s_out, ties, iters = _fold_fast(host, s_in, w)
        ref = np.cumsum(np.concatenate([[s_in], w]))[1:]      # sequential fp64 fold
        assert rc == 0, (s_in, w.size, ties)
        assert np.array_equal(edges, ref), (s_in, w.size)
        assert s_out == ref[-1]
        crossed += int(s_in > 0 and np.floor(np.log2(ref[-1])) > np.floor(np.log2(s_in)))
    assert crossed >= 20, (crossed, len(cases))


def test_fast_iteration_resolves_ties_by_parity(host):
    """Round half to even inside the fast iteration: weights that are odd multiples of half
    an ulp of the running sum tie, and whether a tie rounds up depends on the parity of the
    exact sum before it, i.e. on every earlier tie.  Sparse ties (a few per iteration, as at
    the head of a real fold) must come out bit for bit; a tile with more ties in front of one
    crossing than the device lists (32) must report that, not a wrong sum."""
    rng = np.random.default_rng(11)
    half = 2.0 ** -53                                     # half an ulp of sums in [1, 2)
    seen_ties = seen_runs = 0
    for n in (8, 64, 1000, 4096):
        for p_tie in (0.0005, 0.002, 0.006):
            for rep in range(6):
                # multiples of 64 half-ulps: exact increments up to sums of 32; adding 1, 2 or 4
                # half-ulps makes a tie in the binade [1, 2), [2, 4) or [4, 8) respectively
                m = rng.integers(1, 2 ** 38, n) * 64
                odd = rng.random(n) < p_tie
                m = np.where(odd, m + 2 ** rng.integers(0, 3, n), m)
                if rep % 2:                                # runs of consecutive ties
                    k = int(rng.integers(0, n))
                    m[k:k + 3] |= 1
                    seen_runs += 1
                w = m.astype(np.float64) * half
                s_in = 1.0 + float(rng.integers(0, 2 ** 30)) * 2.0 ** -52
                rc, edges, s_out, ties, iters = _fold_fast(host, s_in, w)
                ref = np.cumsum(np.concatenate([[s_in], w]))[1:]
                assert rc == 0, (n, p_tie, rep, ties)
                assert np.array_equal(edges, ref), (n, p_tie, rep, ties, iters)
                seen_ties += ties
    assert seen_ties >= 100 and seen_runs >= 10, (seen_ties, seen_runs)
    # every element ties: more than the list holds in front of the first crossing
    m = rng.integers(1, 2 ** 30, 4096) * 2 + 1
    rc, _, _, ties, _ = _fold_fast(host, 1.0, m.astype(np.float64) * half)
    assert rc == 1 and ties > 32
