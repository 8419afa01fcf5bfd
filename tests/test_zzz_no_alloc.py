"""apop_empirical_moments_noalloc: result moments for callers that may not allocate.

Who needs it: a host process that owns all device memory itself (a JVM binding with an
up-front arena, torch's caching allocator at its cap) cannot let this library call
cudaMalloc behind its back; and a thread that summarises one population while other,
sharded populations are sweeping on the same GPU must not trigger cudaMalloc / cudaFree,
because those hold the whole device and the sweeps' bounded peer waits can expire.  The
context-based apop_empirical_moments does allocate (its partial-sum scratch); this entry
point keeps that scratch in a fixed array inside the library instead.

What is checked, through the C ABI only: its outputs equal the context-based entry
point's (the tolerance of tests/test_stat_gpu.py, 1e-12 - in fact the two run the same
kernels in the same order) on two calls in a row that share the array, with different
shapes so that stale sums of the first call would show in the second; and that sizes
beyond the stated maximum are refused with a message, before any CUDA call.

This file is named to be collected after every other test file: `pytest -x` stops at the
first failure, and a failure here must not hide the parity tests.  It passes on its own in
a fresh process (where its first call is also the per-device set-up) and after the rest.
"""
import ctypes as C

import numpy as np
import pytest

from anglican_b200 import _lib as L

TOL = 1e-12                                            # tests/test_stat_gpu.py, stat_test.cljc:16
MAX_WIDTH = 256                                        # APOP_NOALLOC_MAX_WIDTH


def _rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.abs(b))))


def _dev_ptr(pop, col):
    p = C.c_void_p()
    L.check(pop.ctx, pop.lib.apop_column_device_ptr(pop.ctx, col.handle, C.byref(p)))
    return p


def _noalloc(lib, stream, x_ptr, lw_ptr, n, width, outs):
    err = C.create_string_buffer(512)
    rc = lib.apop_empirical_moments_noalloc(0, stream, x_ptr, lw_ptr, n, width, *outs, err, 512)
    return rc, err.value.decode()


def test_sizes_beyond_the_maximum_are_refused_with_a_reason():
    """No GPU needed: the size checks come before the first CUDA call, and nothing is queued."""
    lib = L.load()
    fake = C.c_void_p(256)                             # never dereferenced on this path
    rc, msg = _noalloc(lib, None, fake, None, 1000, MAX_WIDTH + 1, [fake, None, None, None])
    assert rc == -1 and str(MAX_WIDTH) in msg and "width" in msg          # APOP_ERR_INVALID
    rc, msg = _noalloc(lib, None, fake, None, 0, 4, [fake, None, None, None])
    assert rc == -1 and "n = 0" in msg
    rc, msg = _noalloc(lib, None, fake, None, 2 ** 31, 1, [fake, None, None, None])
    assert rc == -1 and "APOP_NOALLOC_MAX_PARTICLES" in msg
    rc, msg = _noalloc(lib, None, fake, None, 10, 1, [None, None, None, None])
    assert rc == -1 and "output" in msg
    # a NULL message buffer is allowed
    assert lib.apop_empirical_moments_noalloc(0, None, fake, None, 10, MAX_WIDTH + 1, fake, None,
                                              None, None, None, 0) == -1


@pytest.mark.gpu
def test_host_pointers_are_refused():
    """Inputs and outputs stay on the GPU; a host pointer is reported, not dereferenced."""
    lib = L.load()
    host = np.zeros(16)
    hp = C.c_void_p(host.ctypes.data)
    rc, msg = _noalloc(lib, None, hp, None, 16, 1, [hp, None, None, None])
    assert rc == -1 and "d_x is not device memory" in msg


@pytest.mark.gpu
@pytest.mark.parametrize("first,second", [
    ((200_000, 11, True), (4097, 3, True)),            # the second call has fewer CTAs and columns
    ((1, 1, False), (300_001, MAX_WIDTH, True)),       # a single particle, then the widest column
    ((70_000, 8, True), (70_000, 8, False)),           # same shape, with and without weights
])
def test_two_calls_in_a_row_match_the_context_entry_point(first, second):
    from anglican_b200 import Population
    lib = L.load()
    rng = np.random.default_rng(first[0] + second[1])
    pops, cases = [], []
    try:
        for n, width, weighted in (first, second):
            xs = rng.normal(1.5, 2.0, (n, width)) * np.arange(1, width + 1)
            lws = rng.normal(0, 2.0, n)
            if n > 1:
                lws[rng.integers(0, n)] = -np.inf      # a dead particle weighs nothing
            pop = Population(n)
            pops.append(pop)
            x = pop.column(L.F64, width=width, data=xs)
            lw = pop.column(L.F64, data=lws) if weighted else None
            want = pop.empirical_moments(x, lw)        # the existing entry point (allocates)
            # caller-owned outputs: four [width] device vectors, poisoned so that a slot the
            # new entry point failed to write cannot pass
            out_pop = Population(1)
            pops.append(out_pop)
            outs = [out_pop.column(L.F64, width=width, fill=-7.0) for _ in range(4)]
            cases.append((pop, out_pop, x, lw, n, width, outs, want))
        # everything is resident; now the two calls back to back on ONE stream, no
        # synchronisation in between: they share the static array, stream order keeps them apart
        stream = C.c_void_p(lib.apop_stream(cases[0][0].ctx))
        for pop, out_pop, x, lw, n, width, outs, _ in cases:
            pop.synchronize()
            out_pop.synchronize()
        for pop, out_pop, x, lw, n, width, outs, _ in cases:
            rc, msg = _noalloc(lib, stream, _dev_ptr(pop, x),
                               _dev_ptr(pop, lw) if lw is not None else None, n, width,
                               [_dev_ptr(out_pop, o) for o in outs])
            assert rc == 0, msg
        cases[0][0].synchronize()                      # the stream both calls went to
        for pop, out_pop, x, lw, n, width, outs, want in cases:
            got = [o.download(1).reshape(-1) for o in outs]
            assert _rel(got[0], want["mean"]) < TOL
            if n > 1:
                assert _rel(got[1], want["variance"]) < TOL
                assert _rel(got[2], want["skew"]) < TOL
                assert _rel(got[3], want["kurtosis"]) < TOL
                # same kernels, same grid, same order: not merely close
                assert np.array_equal(got[0], want["mean"])
                assert np.array_equal(got[1], want["variance"])
    finally:
        for p in pops:
            p.close()


@pytest.mark.gpu
def test_mean_only_leaves_the_other_outputs_alone():
    from anglican_b200 import Population
    lib = L.load()
    rng = np.random.default_rng(3)
    xs = rng.normal(size=(5000, 5))
    with Population(5000) as pop, Population(1) as out_pop:
        x = pop.column(L.F64, width=5, data=xs)
        want = pop.empirical_moments(x)
        mean = out_pop.column(L.F64, width=5, fill=-7.0)
        pop.synchronize()
        out_pop.synchronize()
        rc, msg = _noalloc(lib, C.c_void_p(lib.apop_stream(pop.ctx)), _dev_ptr(pop, x), None, 5000,
                           5, [_dev_ptr(out_pop, mean), None, None, None])
        assert rc == 0, msg
        pop.synchronize()
        assert _rel(mean.download(1).reshape(-1), want["mean"]) < TOL
