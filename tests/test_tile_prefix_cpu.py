"""Soundness of the tile classification of the CUDA resampler (k_finalize in
anglican_b200/csrc/resample.cu) next to a dominant weight, on the CPU.

A tile may be classified "easy in binade e" only if the exact sequential fold
(smc.cljc:94, 120) enters and leaves it inside that binade.  classify_tile
(exact_fold.cuh) guarantees this as long as the approximate prefix it is given
has a small RELATIVE error.  The block scan that forms the prefix is restated
here in numpy in the kernel's order (1024 threads x 4 tiles, warp scans, scan
of the warp totals) in two variants:
  * exclusive = neighbour's inclusive value      (what the kernel does)
  * exclusive = inclusive - own value            (what it did before the fix:
    one dominant tile sum cancels the whole prefix -> wrong binade)
The second variant must be caught by this checker on the shapes that failed on
the GPU (tools/dbg_giant.py), the first must be sound on all of them."""
import ctypes as C

import numpy as np
import pytest

from tests.test_exact_fold_host import host  # noqa: F401  (fixture: builds the shim)

TILE, FB, FCH = 4096, 1024, 4


def block_scan_prefix(tsum, by_subtraction):
    nt = len(tsum)
    ain = np.zeros(nt + 1)
    carry = 0.0
    for base in range(0, nt, FB * FCH):
        a = np.zeros(FB * FCH)
        m = min(FB * FCH, nt - base)
        a[:m] = tsum[base:base + m]
        a = a.reshape(FB, FCH)
        loc = a.sum(axis=1)
        inc = np.concatenate([np.cumsum(loc[w * 32:(w + 1) * 32]) for w in range(FB // 32)])
        wt = inc[31::32]
        winc = np.cumsum(wt)
        if by_subtraction:
            wex, ex = winc - wt, inc - loc
        else:
            wex = np.concatenate([[0.0], winc[:-1]])
            ex = np.concatenate([[0.0], inc[:-1]])
            ex[0::32] = 0.0
        for t in range(min(FB, (m + FCH - 1) // FCH)):
            run = carry + wex[t // 32] + ex[t]
            for k in range(FCH):
                idx = base + t * FCH + k
                if idx < nt:
                    ain[idx] = run
                run += a[t, k]
        carry += winc[-1]
    ain[nt] = carry
    return ain


def unsound_tiles(lib, w, by_subtraction):
    n = w.size
    nt = (n + TILE - 1) // TILE
    pad = np.zeros(nt * TILE)
    pad[:n] = w
    tsum = pad.reshape(nt, TILE).sum(axis=1)
    ain = block_scan_prefix(tsum, by_subtraction)
    edges = np.cumsum(w)                      # sequential left fold, like the reference
    hard = lib.host_ctx_hard()
    bad = []
    for lap, off in ((0, 0.0), (1, float(ain[nt]))):
        S_in = 0.0 if lap == 0 else float(edges[-1])
        for t in range(nt):
            hi = min((t + 1) * TILE, n)
            if lap == 0:
                S_out = float(edges[hi - 1])
            else:                              # the wheel wraps with acc not reset
                S_out = float(np.cumsum(np.concatenate([[S_in], w[t * TILE:hi]]))[-1])
            cls = lib.host_classify_tile(off + ain[t], off + ain[t + 1], lap * n + hi)
            if cls != hard and S_out != S_in:
                if lib.host_ctx_of(S_in) != cls or lib.host_ctx_of(S_out) != cls:
                    bad.append((lap, t))
            S_in = S_out
    return bad


SHAPES = [  # n, background, position of the heavy particle, its weight, failed on the GPU before
    (450000, 1e-30, 270000, 1.0, False),
    (600000, 1e-30, 360000, 1.0, True),      # tile 87: 4th tile of its scan thread
    (600000, 1e-30, 450000, 1.0, False),
    (600000, 1e-30, 570000, 1.0, True),      # tile 139: second warp of the scan
    (800000, 1e-30, 600000, 1.0, True),
    (1 << 20, 1e-30, 777777, 1.5, True),
    (1 << 20, 1e-9, 777777, 1e6, False),
]


@pytest.mark.parametrize("n,small,pos,giant,failed_before", SHAPES)
def test_prefix_next_to_a_dominant_weight(host, n, small, pos, giant, failed_before):  # noqa: F811
    host.host_classify_tile.argtypes = [C.c_double, C.c_double, C.c_int64]
    host.host_ctx_of.argtypes = [C.c_double]
    w = np.full(n, small)
    w[pos] = giant
    assert unsound_tiles(host, w, by_subtraction=False) == []
    if failed_before:
        assert unsound_tiles(host, w, by_subtraction=True) != []


def test_prefix_on_ordinary_weights(host):  # noqa: F811
    host.host_classify_tile.argtypes = [C.c_double, C.c_double, C.c_int64]
    host.host_ctx_of.argtypes = [C.c_double]
    rng = np.random.default_rng(3)
    for w in (np.exp(rng.normal(0, 3, 700001)), np.ones(1 << 20), rng.random(333333) * 1e-12):
        assert unsound_tiles(host, w, by_subtraction=False) == []
