"""World-size-2 `gloo` test of the sharded resampling PROTOCOL,
restated in numpy and checked against the oracle's literal wheel
(smc.cljc:86-137) on the unsharded population:

  1. allgather of the local max log-weight -> global max m
  2. the exact sequential fold is chained rank to rank: rank r starts its
     left fold from the exact running sum S_in sent by rank r-1 (lap 0), the
     last rank's result is the total; lap 1 chains again from `total`
  3. every rank builds the same pointer comb (total, U) and counts the
     pointers below its own edges -> offspring of ITS particles and their
     global output slots (sorted physical order), stored into the owner's
     ancestor array (here: gathered on rank 0 and compared).

Only host logic runs here; the CUDA kernels implement the same steps with
mailboxes over peer memory (csrc/fabric.cuh) and are covered by
tests/test_sharded_gpu.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


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


def _worker(rank, world, port, n_local, u, seed, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from anglican_b200.distributed import shard_range
    n = n_local * world
    lw_all = np.random.default_rng(seed).normal(0, 3, n)
    lo, hi = shard_range(n, rank, world)
    lw = lw_all[lo:hi]
    # 1. global max
    mx = torch.tensor([lw.max()], dtype=torch.float64)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    m = float(mx)
    w = np.where(lw == m, 1.0, np.exp(lw - m))                 # smc.cljc:78-82
    # 2. lap 0 chain
    s_in = torch.zeros(1, dtype=torch.float64)
    if rank > 0:
        dist.recv(s_in, src=rank - 1)
    e0 = _seq_fold(float(s_in), w)
    if rank < world - 1:
        dist.send(torch.tensor([e0[-1]], dtype=torch.float64), dst=rank + 1)
    tot = torch.tensor([e0[-1]], dtype=torch.float64)
    dist.broadcast(tot, src=world - 1)
    total = float(tot)
    #    lap 1 chain (acc NOT reset, smc.cljc:133-136)
    s1 = torch.tensor([total], dtype=torch.float64)
    if rank > 0:
        dist.recv(s1, src=rank - 1)
    e1 = _seq_fold(float(s1), w)
    if rank < world - 1:
        dist.send(torch.tensor([e1[-1]], dtype=torch.float64), dst=rank + 1)
    # 3. the comb, identical on every rank: x_0 = total*U, x_{k+1} = x_k + step
    step = total / float(n)
    x = np.empty(n)
    x[0] = total * u
    for k in range(1, n):
        x[k] = x[k - 1] + step
    # emission index k selects particle i of lap L iff edge_{i-1} <= x_k < edge_i
    c0 = np.searchsorted(x, e0, side="left")                   # pointers strictly below each edge
    c1 = np.searchsorted(x, e1, side="left")
    p0 = np.searchsorted(x, float(s_in), side="left")
    p1 = np.searchsorted(x, float(s1), side="left")
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
        q.put((anc, total, lw_all))
    dist.destroy_process_group()


@pytest.mark.parametrize("u", [0.0, 0.37, 0.999])
def test_sharded_protocol_matches_oracle_wheel(u):
    from oracle import pyoracle as orc
    world, n_local = 2, 512
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_local, u, 11, q))
             for r in range(world)]
    for p in procs:
        p.start()
    anc, total, lw_all = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    w, m = orc.particle_weights(lw_all)
    ref = orc.resample_weights(w, w.size, u)
    srt, j0 = orc.emission_to_sorted(ref["emission"], ref["lap"])
    assert total == ref["total"]
    assert np.array_equal(anc, srt)


def test_shard_range():
    from anglican_b200.distributed import shard_range
    assert shard_range(16, 0, 2) == (0, 8) and shard_range(16, 1, 2) == (8, 16)
    with pytest.raises(ValueError):
        shard_range(10, 0, 4)
