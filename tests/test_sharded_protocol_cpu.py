"""World-size-2/4 `gloo` test of the sharded resampling PROTOCOL,
restated on the host and checked against the oracle's literal wheel
(smc.cljc:86-137) on the unsharded population:

  1. allgather of (local max, local sum) -> global max m, approximate weight of
     the lower ranks (k_fin_a)
  2. every rank classifies its WHOLE shard from the approximate prefix
     (exact_fold.cuh classify_tile): a shard that provably stays inside one
     binade publishes its parity-transducer composite at once (CK_EASY); a shard
     that owns a binade crossing waits for its exact input - the composites and
     exact outputs of the lower ranks applied in rank order - folds its
     particles literally and publishes the exact sum leaving it (CK_EXACT).
     Every rank applies all parts itself and so obtains the same exact total
     (k_chain, csrc/resample.cu).  Lap 1 repeats this from `total`
     (acc NOT reset, smc.cljc:133-136).
  3. every rank builds the same pointer comb (total, U) and counts the
     pointers below its own edges -> offspring of ITS particles and their
     global output slots (sorted physical order), stored into the owner's
     ancestor array (here: gathered on rank 0 and compared).

Only host logic runs here (the transducer arithmetic is the header the kernels
use, compiled for the CPU: tests/host/exact_fold_host.cpp); the CUDA kernels
implement the same steps with mailboxes over peer memory (csrc/fabric.cuh) and
are covered by tests/test_sharded_gpu.py."""
import ctypes as C
import os
import socket
import subprocess

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host", "exact_fold_host.cpp")
OUT = os.path.join(HERE, "host", "_build", "libexact_fold_host.so")
HDR = os.path.join(HERE, "..", "anglican_b200", "csrc", "exact_fold.cuh")
dp, lp = C.POINTER(C.c_double), C.POINTER(C.c_int64)
CK_EASY, CK_EXACT = 1, 2


def _host_lib():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    if (not os.path.exists(OUT) or os.path.getmtime(OUT) < max(os.path.getmtime(SRC),
                                                               os.path.getmtime(HDR))):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared",
                               "-x", "c++", SRC, "-o", OUT + ".tmp%d" % os.getpid()])
        os.replace(OUT + ".tmp%d" % os.getpid(), OUT)
    lib = C.CDLL(OUT)
    lib.host_shard_composite.argtypes = [dp, C.c_int64, C.c_int, lp, lp]
    lib.host_apply_composite.argtypes = [C.c_double, C.c_int64, C.c_int64, C.c_int, C.POINTER(C.c_int)]
    lib.host_apply_composite.restype = C.c_double
    lib.host_classify_tile.argtypes = [C.c_double, C.c_double, C.c_int64]
    return lib


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _seq_fold(s_in, w):
    """exact left fold: edges[i] = ((s_in + w0) + w1) + ...  (np.cumsum on 1-D
    float64 is sequential)"""
    return np.cumsum(np.concatenate([[s_in], w]))[1:]


def _lap(lib, rank, world, w, a_in, a_out, idx_out, s_start, hard_code):
    """one lap of the sharded fold: returns (exact sum entering this rank, exact
    sum after the last rank, kinds of all ranks)"""
    cls = lib.host_classify_tile(a_in, a_out, idx_out)
    part = None
    if cls != hard_code:
        a0, a1 = C.c_int64(), C.c_int64()
        if lib.host_shard_composite(w.ctypes.data_as(dp), w.size, cls, C.byref(a0), C.byref(a1)) == 0:
            part = (CK_EASY, a0.value, a1.value, cls)
    easy = [None] * world
    dist.all_gather_object(easy, part)                 # composites are published up front

    def apply(S, q, exact):
        if easy[q] is not None:
            ok = C.c_int()
            S = lib.host_apply_composite(S, easy[q][1], easy[q][2], easy[q][3], C.byref(ok))
            assert ok.value == 1, "composite applied outside its binade"
            return S
        return exact[q]

    exact = {}
    S = s_start
    s_in = None
    for q in range(world):                             # hard shards resolve in rank order
        if q == rank:
            s_in = S
        if easy[q] is None:
            t = torch.zeros(1, dtype=torch.float64)
            if q == rank:
                t[0] = _seq_fold(S, w)[-1]             # literal fold of this shard (smc.cljc:120)
            dist.broadcast(t, src=q)
            exact[q] = float(t)
        S = apply(S, q, exact)
    return s_in, S, [CK_EASY if e is not None else CK_EXACT for e in easy]


def _worker(rank, world, port, n_local, u, seed, scale, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from anglican_b200.distributed import shard_range
    lib = _host_lib()
    hard_code = lib.host_ctx_hard()
    n = n_local * world
    lw_all = np.random.default_rng(seed).normal(0, scale, n)
    lo, hi = shard_range(n, rank, world)
    lw = lw_all[lo:hi]
    # 1. global max, approximate rank offsets
    mx = torch.tensor([lw.max()], dtype=torch.float64)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    m = float(mx)
    w = np.where(lw == m, 1.0, np.exp(lw - m))                 # smc.cljc:78-82
    sums = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(sums, torch.tensor([w[::-1].sum()], dtype=torch.float64))   # any summation order
    sums = [float(s) for s in sums]
    off, atot = sum(sums[:rank]), sum(sums)
    # 2. the two laps
    s_in, total, kinds0 = _lap(lib, rank, world, w, off, off + sums[rank], hi, 0.0, hard_code)
    s1, _, kinds1 = _lap(lib, rank, world, w, atot + off, atot + off + sums[rank], n + hi, total,
                         hard_code)
    e0 = _seq_fold(s_in, w)
    e1 = _seq_fold(s1, w)
    # 3. the comb, identical on every rank: x_0 = total*U, x_{k+1} = x_k + step
    step = total / float(n)
    x = np.empty(n)
    x[0] = total * u
    for k in range(1, n):
        x[k] = x[k - 1] + step
    # emission index k selects particle i of lap L iff edge_{i-1} <= x_k < edge_i
    c0 = np.searchsorted(x, e0, side="left")                   # pointers strictly below each edge
    c1 = np.searchsorted(x, e1, side="left")
    p0 = np.searchsorted(x, s_in, side="left")
    p1 = np.searchsorted(x, s1, side="left")
    J = int(np.searchsorted(x, total, side="left"))            # emissions of the first lap
    cnt0 = np.diff(np.concatenate([[p0], c0]))
    cnt1 = np.diff(np.concatenate([[p1], c1]))
    ids = np.arange(lo, hi, dtype=np.int64)
    # sorted physical order: second-lap emissions first (slots k - J), then first lap
    slots1 = np.arange(p1, c1[-1]) - J if c1[-1] > p1 else np.empty(0, np.int64)
    slots0 = np.arange(p0, c0[-1]) + (n - J) if c0[-1] > p0 else np.empty(0, np.int64)
    vals1, vals0 = np.repeat(ids, cnt1), np.repeat(ids, cnt0)
    slots = np.concatenate([slots1, slots0]).astype(np.int64)
    vals = np.concatenate([vals1, vals0]).astype(np.int64)
    gathered = [None] * world
    dist.all_gather_object(gathered, (slots, vals))
    if rank == 0:
        anc = np.full(n, -1, dtype=np.int64)
        for sl, va in gathered:
            anc[sl] = va
        q.put((anc, total, lw_all, kinds0, kinds1))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,u,scale", [(2, 0.0, 3.0), (2, 0.37, 3.0), (2, 0.999, 3.0),
                                           (4, 0.37, 0.1), (4, 0.81, 3.0)])
def test_sharded_protocol_matches_oracle_wheel(world, u, scale):
    from oracle import pyoracle as orc
    n_local = 512
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    _host_lib()                                        # build once, before the ranks race for it
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_local, u, 11, scale, q))
             for r in range(world)]
    for p in procs:
        p.start()
    anc, total, lw_all, kinds0, kinds1 = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    w, m = orc.particle_weights(lw_all)
    ref = orc.resample_weights(w, w.size, u)
    srt, j0 = orc.emission_to_sorted(ref["emission"], ref["lap"])
    assert total == ref["total"]
    assert np.array_equal(anc, srt)
    # rank 0 starts the fold at zero (a crossing at every doubling): always exact
    assert kinds0[0] == CK_EXACT
    if world == 4 and scale < 1.0:
        # near-equal weights: the total lies in [2^10, 2^11), crossings at 1/2, 1/4, ... of it
        # fall into ranks 0..2 at most - the protocol must have used composites somewhere
        assert CK_EASY in kinds0 + kinds1


def test_shard_range():
    from anglican_b200.distributed import shard_range
    assert shard_range(16, 0, 2) == (0, 8) and shard_range(16, 1, 2) == (8, 16)
    with pytest.raises(ValueError):
        shard_range(10, 0, 4)
