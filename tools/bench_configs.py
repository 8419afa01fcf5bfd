#!/usr/bin/env python
"""Secondary measurements on the other BASELINE configs (bench.py measures
config 3).  One JSON line per config; copied into profiles/ by hand.

    python tools/bench_configs.py c2 | c5 [--particles N] [--T T]
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

KT = {1: "k_propagate", 2: "k_tile_lse", 3: "k_finalize", 4: "k_fold_agg", 5: "k_resolve",
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


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("config", choices=["c2", "c5"])
    ap.add_argument("--particles", type=int, default=0)
    ap.add_argument("--T", type=int, default=0)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=2)
    a = ap.parse_args()
    print(json.dumps({"c2": c2, "c5": c5}[a.config](a)), flush=True)
