#!/usr/bin/env python
"""Secondary measurements on the other BASELINE configs (bench.py measures
config 3).  One JSON line per config; copied into profiles/ by hand.

    python tools/bench_configs.py c2 | c4 | c5 [--particles N] [--T T]
"""
import argparse
import json
import math
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from anglican_b200 import Population                                   # noqa: E402
from anglican_b200.models import LinearGaussian1D, LinearGaussianMVN   # noqa: E402

KT = {1: "k_propagate", 2: "k_tile_lse", 3: "k_fin_a+k_fin_b", 4: "k_fold_agg", 5: "k_seg_scan+k_chain+k_expand",
      6: "k_ancestors", 8: "k_gather"}


def timed_sweeps(pop, model, y, steps, warmup):
    for i in range(warmup):
        pop.set_seed(10 + i)
        pop.sweep(model, y, want_final=False, want_stats=False)
    pop.timer_reset()
    pop.enable_kernel_timers(True)
    zs = []
    for i in range(steps):
        pop.set_seed(100 + i)
        zs.append(pop.sweep(model, y, want_final=False, want_stats=False).log_z)
    pop.synchronize()
    ms, _ = pop.timer_ms(0)
    kt = {}
    for tag, name in KT.items():
        m, c = pop.timer_ms(tag)
        if c:
            kt[name] = 1e3 * m / c
    pop.enable_kernel_timers(False)
    return ms / 1e3, zs, kt


def c2(args):
    """config 2: 1-D linear-Gaussian SSM, T=1000, 2^20 particles, log Z vs exact Kalman"""
    n, T = args.particles or (1 << 20), args.T or 1000
    rng = np.random.default_rng(20240601)
    x, ys = rng.normal(), []
    for _ in range(T):
        x = 0.9 * x + rng.normal()
        ys.append(x + rng.normal())
    y = np.array(ys)
    m = LinearGaussian1D()
    # exact Kalman log-marginal (textbook filter)
    mean, P, lz = 0.0, 1.0, 0.0
    for v in y:
        mp, Pp = 0.9 * mean, 0.81 * P + 1.0
        S = Pp + 1.0
        lz += -0.5 * (math.log(2 * math.pi * S) + (v - mp) ** 2 / S)
        K = Pp / S
        mean, P = mp + K * (v - mp), (1 - K) * Pp
    with Population(n, seed=7) as pop:
        dev_s, zs, kt = timed_sweeps(pop, m, y, args.steps, args.warmup)
    return {"config": "configs[1]: linear-Gaussian SSM 1-D, T=%d, %d particles :smc, 1 B200" % (T, n),
            "particle_steps_per_s": n * T * args.steps / dev_s, "ms_per_sweep": 1e3 * dev_s / args.steps,
            "log_z": zs, "kalman_log_z": lz, "abs_err": [abs(z - lz) for z in zs],
            "kernel_avg_us": kt}


def c5(args):
    """config 5 on one GPU: 64-dim linear-Gaussian SSM, mvn observe on the fp64 tensor pipe"""
    d, T = 64, args.T or 100
    n = args.particles or (1 << 24)
    rng = np.random.default_rng(20240603)
    Q, _ = np.linalg.qr(rng.normal(size=(d, d)))
    A = 0.9 * Q
    Qc = 0.1 * np.eye(d)
    R = 0.5 * np.eye(d) + 0.5 * np.ones((d, d)) / d
    x = rng.normal(size=d)
    Lr = np.linalg.cholesky(R)
    ys = []
    for _ in range(T):
        x = A @ x + math.sqrt(0.1) * rng.normal(size=d)
        ys.append(x + Lr @ rng.normal(size=d))
    y = np.array(ys).ravel()
    m = LinearGaussianMVN(A, Qc, R)
    with Population(n, seed=9) as pop:
        dev_s, zs, kt = timed_sweeps(pop, m, y, args.steps, args.warmup)
    flop = 2.0 * 2.0 * n * d * d * T * args.steps          # A x and the whitening product
    byts = (4 + 8 * d + 8 * d + 8) * float(n) * T * args.steps
    return {"config": "configs[4] on 1 B200: %d-dim linear-Gaussian SSM, mvn observe, T=%d, %d particles"
                      % (d, T, n),
            "particle_steps_per_s": n * T * args.steps / dev_s, "ms_per_sweep": 1e3 * dev_s / args.steps,
            "dmma_tflops_whole_step": flop / dev_s / 1e12,
            "compulsory_gbs_whole_step": byts / dev_s / 1e9, "log_z": zs, "kernel_avg_us": kt}


def c4(args):
    """config 4 on one GPU: IPMCMC on the SV model; the nodes of this GPU run one after
    the other on the device (conditional and unconditional sweeps with lineages kept),
    only the node log-marginals and the retained paths leave it (ipmcmc.cljc:34-208)."""
    import ctypes as C
    from anglican_b200 import _lib as L
    from anglican_b200.models import StochasticVolatility
    from bench import sv_observations
    M = args.nodes
    P = M // 2
    n, T = args.particles or (1 << 22), args.T or 200
    y = sv_observations(2000)[:T]
    m = StochasticVolatility()
    lib = L.load()
    rng = np.random.default_rng(1)
    from concurrent.futures import ThreadPoolExecutor
    # one context (own stream) per node, driven by a pool like the reference's
    # cp/pmap over nodes (ipmcmc.cljc:176-182): the single-CTA phases of one node's
    # resampler overlap with the streaming kernels of the others
    W = M if args.concurrent else 1
    pops = [Population(n, seed=77 + w) for w in range(W)]
    pool = ThreadPoolExecutor(max_workers=W)
    try:
        retained = [None] * P
        log_hist = []
        t_all = []
        for it in range(args.warmup + args.steps):
            for p_ in pops:
                p_.synchronize()
            t0 = time.perf_counter()
            picks = [int(rng.random() * n) for _ in range(M)]

            def node(j):
                pop = pops[j % W]
                ret = retained[j] if j < P else None
                pop.set_seed(1000 * it + j + 1)
                out = pop.sweep(m, y, retained=ret, csmc=ret is not None, n_cond_steps=T,
                                reset_weights=True, keep_history=True, want_final=False,
                                want_stats=False)
                return out.log_z, out.trace(picks[j])               # rand-nth, ipmcmc.cljc:189

            res = list(pool.map(node, range(M)))
            log_zs = np.array([r[0] for r in res])
            cand = [r[1] for r in res]
            u = rng.random(P)
            cs, zs = np.empty(P, dtype=np.int32), np.empty(M)
            rc = lib.apop_ipmcmc_gibbs(log_zs.ctypes.data_as(L.c_double_p), M, P,
                                       u.ctypes.data_as(L.c_double_p),
                                       cs.ctypes.data_as(L.c_int32_p), zs.ctypes.data_as(L.c_double_p))
            assert rc == 0
            retained = [cand[int(c)] for c in cs]
            dt = time.perf_counter() - t0
            if it >= args.warmup:
                t_all.append(dt)
                log_hist.append(log_zs.tolist())
    finally:
        pool.shutdown()
        for p_ in pops:
            p_.close()
    tot = sum(t_all)
    return {"config": "configs[3] on 1 B200: IPMCMC on the SV model, %d nodes (%d conditional) x %d "
                      "particles, T=%d, lineages kept, %s" % (M, P, n, T, "nodes concurrent (one ctx + stream each)" if args.concurrent else "nodes one after the other"),
            "particle_steps_per_s": M * n * T * len(t_all) / tot, "s_per_mcmc_iteration": tot / len(t_all),
            "node_log_z": log_hist, "timing": "wall clock around the iteration, device synchronised"}


def c4dist(args):
    """config 4 over the GPUs of the box (one process per GPU under torchrun):
    M nodes split over the ranks, `--width` node sweeps in flight per GPU,
    log-marginals all-reduced, retained paths moved point to point
    (anglican_b200/ipmcmc.py).  Prints on rank 0.
        python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 \
            tools/bench_configs.py c4dist --nodes 32 [--particles N] [--T T]"""
    import os
    import torch
    import torch.distributed as dist
    from anglican_b200.ipmcmc import ipmcmc_iterations
    from anglican_b200.models import StochasticVolatility
    from bench import sv_observations
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    M = args.nodes
    n, T = args.particles or (1 << 22), args.T or 200
    y = sv_observations(2000)[:T]
    gen = ipmcmc_iterations(StochasticVolatility(), y, n, M, M // 2, seed=77, device=local,
                            width=args.width, abi_exchange=True)
    times, zs, moved = [], [], 0
    for it in range(args.warmup + args.steps):
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        st = next(gen)
        torch.cuda.synchronize()
        dist.barrier()
        dt = time.perf_counter() - t0
        if it >= args.warmup:
            times.append(dt)
            zs.append(st["log_zs"].tolist())
            moved += st["paths_moved"]
    gen.close()
    t = torch.tensor([sum(times)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tot = float(t)
    dist.destroy_process_group()
    if rank != 0:
        return None
    return {"config": "configs[3] on %d B200: IPMCMC on the SV model, %d nodes (%d conditional) x %d "
                      "particles, T=%d, %d node sweeps in flight per GPU" % (world, M, M // 2, n, T,
                                                                            args.width),
            "particle_steps_per_s": M * n * T * len(times) / tot,
            "s_per_mcmc_iteration": tot / len(times), "node_log_z": zs,
            "paths_moved_rank0": moved,
            "log_z_exchange": "apop_ipmcmc_exchange_logz (C ABI, NCCL allgather of M/world doubles)",
            "lineages": "compact (2 bits per particle and step)",
            "timing": "wall clock around the iteration, devices synchronised, max over ranks"}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("config", choices=["c2", "c4", "c4dist", "c5"])
    ap.add_argument("--nodes", type=int, default=4)
    ap.add_argument("--concurrent", action="store_true")
    ap.add_argument("--width", type=int, default=4, help="c4dist: node sweeps in flight per GPU")
    ap.add_argument("--particles", type=int, default=0)
    ap.add_argument("--T", type=int, default=0)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=2)
    a = ap.parse_args()
    r = {"c2": c2, "c4": c4, "c4dist": c4dist, "c5": c5}[a.config](a)
    if r is not None:
        print(json.dumps(r), flush=True)
