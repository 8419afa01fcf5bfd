#!/usr/bin/env python
"""bench.py — SMC particle-steps/s of the particle-population hot path.

    python bench.py --gpus N --steps K --warmup W          (our arm)
    python bench.py --impl reference --gpus N --steps K --warmup W

One "step" = `--inflight` (default 3) independent full `sweep`s (smc.cljc:35-51;
`infer :smc` is an endless sequence of independent sweeps, smc.cljc:58-63) of
BASELINE config 3: the stochastic-volatility SSM, T=2000 observations, 2^27
particles in total per sweep, sharded contiguously over the N ranks (strong
scaling, north_star), each sweep on its own population (context + stream) so
that the serial phases of one overlap the streaming kernels of the other.
`--workload lgmvn64` measures BASELINE config 5 instead.  `value` is
whole-job particle-steps/s timed on the device (CUDA events on the engine's
stream, max over ranks) with inputs resident in HBM; `e2e` is the same metric
through the public call (host observations in, log-marginal + per-step stats +
final particle states out, host<->device copies inside the timed region).
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "smc_particle_steps_per_sec"
UNIT = "particle-steps/s"
# timer tags of apop_enable_kernel_timers -> the kernels each tag brackets today
KT_NAMES = {1: "k_propagate", 2: "k_tile_lse", 3: "k_fin_a+k_fin_b", 4: "k_fold_agg",
            5: "k_seg_scan+k_chain+k_expand", 6: "k_ancestors", 7: "k_stat_lse", 8: "k_gather",
            9: "migrate"}
# algorithmic bytes per particle per launch (d = 1); the three streaming
# kernels of a step sum to the 64 B of the whole particle-step, each stage counted ONCE:
#   k_propagate K1 propagate+reweight (16d+8 = 24) + K5 ancestor gather (16d+4 = 20) = 44
#   k_fold_agg  K2 normalise: the pass that re-reads the log-weights (its fp32 tile
#               partials ride along in k_propagate at no HBM cost)                    =  8
#   k_ancestors K3+K4 scan + ancestor search (R 8 + W 4)                              = 12
# (column-level API only: k_tile_lse is K2 there, k_gather is K5)
ALGO_BYTES = {"k_propagate": 44.0, "k_fold_agg": 8.0, "k_ancestors": 12.0, "k_tile_lse": 8.0,
              "k_gather": 20.0}
STREAMING = ("k_propagate", "k_fold_agg", "k_ancestors")
# `roofline.achieved` follows SURVEY §8(d): ALGORITHMIC bytes of the stages the kernel
# covers / its launch time.  The fused k_propagate never materialises the gathered
# state, so the bytes it MUST move (compulsory) are R anc 4 + R x 8 + W x 8 + W lw 8 =
# 28 of its 44: `compulsory_frac` and, from the tracked ncu capture, `dram_frac` are
# reported beside it.  The figure that answers the north star ("resample+reweight at
# >= 70 % of HBM") is `path_roofline` (the whole particle-step, 64 B).
MOVED_BYTES = {"k_propagate": 28.0, "k_fold_agg": 8.0, "k_ancestors": 12.0, "k_tile_lse": 8.0,
               "k_gather": 20.0}
# dram__bytes_read.sum + dram__bytes_write.sum per launch from the tracked
# `ncu --set full` capture of the shipped kernels (tools/ncu_summary.py writes the file)
NCU_TRAFFIC_FILE = os.path.join(ROOT, "profiles", "ncu_traffic.json")
STEP_BYTES = 64.0          # whole particle-step: 32d + 32
RESAMPLE_REWEIGHT_BYTES = 48.0   # BASELINE's second metric: logw W 8 + K2 8 + K3/4 12 + K5 20 (16d + 32)


def ncu_traffic_per_particle():
    try:
        d = json.load(open(NCU_TRAFFIC_FILE))
        return {k: v["dram_bytes_per_particle"] for k, v in d["kernels"].items()}, d.get("source")
    except Exception:
        return {}, None


def sv_observations(T, seed=20240602, phi=0.98, sigma=0.16, beta=0.65):
    rng = np.random.default_rng(seed)
    x = rng.normal() * sigma / math.sqrt(1 - phi * phi)
    ys = np.empty(T)
    for t in range(T):
        x = phi * x + sigma * rng.normal()
        ys[t] = beta * math.exp(x / 2) * rng.normal()
    return ys


def lgssm_observations(T, seed=20240601):
    """BASELINE config 2: x0 ~ N(0,1), x_t = 0.9 x_{t-1} + N(0,1), y_t = x_t + N(0,1)"""
    rng = np.random.default_rng(seed)
    x = rng.normal()
    ys = np.empty(T)
    for t in range(T):
        x = 0.9 * x + rng.normal()
        ys[t] = x + rng.normal()
    return ys


def kalman_logz(y, a=0.9, q=1.0, c=1.0, r=1.0, m0=0.0, s0=1.0):
    """exact log-marginal of the 1-D linear-Gaussian model (the check BASELINE config 2 names)"""
    m, P, lz = m0, s0 * s0, 0.0
    for yt in y:
        m, P = a * m, a * a * P + q * q
        S = c * c * P + r * r
        lz += -0.5 * math.log(2 * math.pi * S) - 0.5 * (yt - c * m) ** 2 / S
        K = P * c / S
        m, P = m + K * (yt - c * m), (1 - K * c) * P
    return lz


def lgmvn_workload(T, d=64):
    """BASELINE config 5: A = 0.9 Q, process cov 0.1 I, obs cov 0.5 I + 0.5 11^T/d"""
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
    return A, Qc, R, np.array(ys).ravel()


def measured_peak():
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(pk["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                pass
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                    "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        busy = sorted(sm)[len(sm) // 2:]
        return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": max(mx),
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def oracle_sample(args, T_cap):
    """the bounded sample the CPU arms run: (oracle model, observations, particles, steps, label)"""
    from oracle import pyoracle as orc
    if args.workload == "lgmvn64":
        n_s, T_s = 1 << 13, min(args.T, 20)
        A, Qc, R, y = lgmvn_workload(args.T, 64)
        m = orc.Model(orc.SSM_LGMVN, dim=64, s0=1.0, A=A, Lq=np.linalg.cholesky(Qc), R=R)
        return m, y[:T_s * 64], n_s, T_s, "64-dim linear-Gaussian sweep"
    # 2^20 particles: 24 MB of state per array, beyond the host's L2
    n_s, T_s = 1 << 20, min(args.T, T_cap)
    if args.workload == "lgssm1d":
        m = orc.Model(orc.SSM_LGSSM1D, m0=0.0, s0=1.0, a=0.9, b=0.0, q=1.0, c=1.0, d=0.0, r=1.0)
        return m, lgssm_observations(args.T)[:T_s], n_s, T_s, "1-D linear-Gaussian sweep"
    phi, sigma, beta = 0.98, 0.16, 0.65
    m = orc.Model(orc.SSM_SV, m0=0.0, s0=sigma / math.sqrt(1 - phi * phi), a=phi, b=0.0, q=sigma,
                  beta=beta)
    return m, sv_observations(args.T)[:T_s], n_s, T_s, "SV sweep"


def resample_microbench(device, sizes, peak):
    """SURVEY §8(d) resample-only microbench: lw ~ N(0, 3^2) from default_rng(7), one
    reweight (`observe* (normal 0 1)` added into the log-weights), systematic resample
    and ancestor gather of one state column through the column-level C ABI.
    Algorithmic bytes: 48 per particle (16d + 32, d = 1).  Device time by CUDA events
    on the engine stream."""
    from anglican_b200 import Population, _lib as L
    out = []
    for n in sizes:
        lw0 = np.random.default_rng(7).normal(0.0, 3.0, n)
        with Population(n, seed=7, device=device) as pop:
            lw = pop.column(L.F64, data=lw0)
            x = pop.column(L.F64, data=lw0)
            anc = pop.column(L.I32)
            best = None
            for it in range(6):
                lw.upload(lw0)
                pop.synchronize()
                pop.timer_reset()
                pop.timer_begin(15)
                pop.observe_normal(0.0, 1.0, x, lw)
                pop.resample(lw, u=0.37, order=L.ORDER_SORTED, anc=anc)
                pop.gather(anc, [x])
                pop.timer_end(15)
                pop.synchronize()
                ms, _ = pop.timer_ms(15)
                if it >= 2:
                    best = ms if best is None else min(best, ms)
            gbs = RESAMPLE_REWEIGHT_BYTES * n / (best * 1e-3) / 1e9
            out.append({"particles": n, "ms": best, "achieved_gbs": gbs, "frac": gbs / peak})
    return out


def dgemm_peak_tflops(torch, n=8192, reps=5):
    """fp64 tensor-pipe yard-stick for the mvn kernels: cuBLAS DGEMM n^3 on this GPU"""
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    torch.matmul(a, b)
    torch.cuda.synchronize()
    best = None
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
    del a, b
    torch.cuda.empty_cache()
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12


# --------------------------------------------------------------------- reference arm
def run_reference(args):
    """The reference's CPU path for the same workload.  The Clojure reference
    cannot run here (no JVM, SURVEY header), so this times the oracle port of
    smc.cljc on the host cores — the one place besides cpu_baseline where bench
    executes oracle/."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    from oracle import pyoracle as orc
    T = args.T
    cores = os.cpu_count() or 1
    m, y, n_s, T_s, label = oracle_sample(args, 500)       # a few seconds per step on all threads
    times = []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        orc.smc_sweep(m, y, n_s, seed=100 + i, threads=cores, want_final=True)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            times.append(dt)
    total = sum(times)
    v = n_s * T_s * len(times) / total
    sample = (f"{label}, {n_s} particles x {T_s} of {T} steps per bench step: a bounded sample of "
              f"the 2^{int(math.log2(args.particles))}-particle workload, the rate is a LINEAR "
              "EXTRAPOLATION in particles and steps (the sequential wheel is O(N) per step)")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": args.workload_name, "particles_total": args.particles,
                       "T": T, "note": "reference JVM path unavailable on this box (no JVM); "
                                       "oracle C port of smc.cljc timed instead"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------- our arm
def run_ours(args):
    rank, world, local = dist_env()
    import torch
    import torch.distributed as dist
    from anglican_b200 import Population
    from anglican_b200.models import StochasticVolatility

    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    nccl_id = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        from anglican_b200.distributed import nccl_unique_id
        nccl_id = nccl_unique_id(rank)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    n_total, T = args.particles, args.T
    assert n_total % world == 0
    n_local = n_total // world
    mvn = args.workload == "lgmvn64"
    dim = 64 if mvn else 1
    if mvn:
        from anglican_b200.models import LinearGaussianMVN
        A, Qc, R, y = lgmvn_workload(T, dim)
        model = LinearGaussianMVN(A, Qc, R)
    elif args.workload == "lgssm1d":
        from anglican_b200.models import LinearGaussian1D
        y = lgssm_observations(T)
        model = LinearGaussian1D()
    else:
        y = sv_observations(T)
        model = StochasticVolatility()
    # `infer :smc` is an endless sequence of INDEPENDENT sweeps (sample-seq,
    # smc.cljc:58-63).  F of them are kept in flight, each on its own population
    # (context + stream): the single-CTA phases of one sweep's exact resampler
    # (k_fin_a/b tails, k_chain) overlap with the streaming kernels of the others.
    from concurrent.futures import ThreadPoolExecutor
    F = max(1, min(args.inflight, 8))
    pops = []
    for f in range(F):
        nid = nccl_id
        if world > 1 and f > 0:
            from anglican_b200.distributed import nccl_unique_id
            nid = nccl_unique_id(rank)            # one communicator per population
        pops.append(Population(n_local, seed=0xA11CE + f, device=local, rank=rank, world=world,
                               nccl_id=nid))
    pop = pops[0]
    pool = ThreadPoolExecutor(max_workers=F)

    e2e_seen = {}          # "f<f>_i<i>" -> (log Z, checksum of this rank's final states)

    def sweeps(f, count, seed0, **kw):
        zs = []
        for i in range(count):
            pops[f].set_seed(seed0 + 100 * f + i)
            o = pops[f].sweep(model, y, **kw)
            if kw.get("want_final"):
                # the host reads the result: every 4096th final state, summed as 64-bit
                # patterns modulo 2^64 - exact, and independent of how the population is
                # sharded (shards are multiples of 4096 values; the ranks' sums add up)
                bits = np.ascontiguousarray(o.x_final.reshape(-1)[::4096]).view(np.uint64)
                e2e_seen["f%d_i%d" % (f, i)] = (o.log_z, int(bits.sum(dtype=np.uint64)))
            zs.append(o.log_z)
        return zs

    # ---- warm-up (>= 3 sweeps per population): the first sweep of each population
    # runs alone (allocations, and for N > 1 the NCCL / IPC set-up, are collective)
    for f in range(F):
        sweeps(f, 1, 900, want_final=False, want_stats=False)
        barrier()
    if args.warmup > 1:
        list(pool.map(lambda f: sweeps(f, args.warmup - 1, 1000, want_final=False, want_stats=False),
                      range(F)))
    # ---- per-kernel timers (roofline, kernel shares): one population alone when
    # several are in flight, so that launches of another stream do not overlap
    kt = {}
    if F > 1:
        barrier()
        pop.timer_reset()
        pop.enable_kernel_timers(True)
        sweeps(0, 1, 1500, want_final=False, want_stats=False)
        pop.synchronize()
        for tag, name in KT_NAMES.items():
            ms, c = pop.timer_ms(tag)
            if c:
                kt[name] = (ms, c)
        pop.enable_kernel_timers(False)
    for p_ in pops:
        p_.timer_reset()
    if F == 1:
        pop.enable_kernel_timers(True)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = sum(p_.launches for p_ in pops)
    barrier()
    wall0 = time.perf_counter()
    zs = list(pool.map(lambda f: sweeps(f, args.steps, 2000, want_final=False, want_stats=False),
                       range(F)))
    barrier()
    wall = time.perf_counter() - wall0
    log_zs = [z for l in zs for z in l]
    launches = sum(p_.launches for p_ in pops) - launches0
    # device time of the timed sweeps: CUDA events on each population's stream
    # (sum over its back-to-back sweeps); populations run concurrently -> max
    dev_ms = max(p_.timer_ms(0)[0] for p_ in pops)
    if F == 1:
        for tag, name in KT_NAMES.items():
            ms, c = pop.timer_ms(tag)
            if c:
                kt[name] = (ms, c)
        pop.enable_kernel_timers(False)
    # ---- the same with ONE sweep in flight (population 0 alone), reported beside the
    # headline so that the gain from overlapping independent sweeps is visible
    dev1_ms = None
    if F > 1:
        barrier()
        pop.timer_reset()
        z1 = sweeps(0, args.steps, 2000, want_final=False, want_stats=False)
        barrier()
        dev1_ms = pop.timer_ms(0)[0]
        if z1 != zs[0]:
            raise SystemExit("log Z of a sweep depends on what else is in flight: %r vs %r" % (z1, zs[0]))
    clocks = sampler.stop() if rank == 0 else None

    # ---- e2e: public call, host buffers in/out, copies inside the timed region
    x_host_bytes = n_local * 8 * dim
    # untimed: allocates the populations' page-locked result buffers
    for f in range(F):
        pops[f].sweep(model, y[:4 * dim], want_final=True, want_stats=True, pinned_final=True)
        barrier()
    # Sharded populations (N > 1) run the e2e leg ONE population at a time: with final
    # states requested, two sharded populations in flight have stalled each other across
    # ranks on an 8-GPU box (a barrier wait beyond the 3 s spin bound, gpurun_out/r03s, r03u;
    # the device-timed legs above do not show it up to 4 in flight) - DESIGN.md section 5.
    Fe = F if world == 1 else 1
    e0 = time.perf_counter()
    e2e_error = None
    try:
        if Fe == F:
            list(pool.map(lambda f: sweeps(f, args.steps, 3000, want_final=True, want_stats=True,
                                           pinned_final=True), range(F)))
        else:
            for f in range(F):
                sweeps(f, args.steps, 3000, want_final=True, want_stats=True, pinned_final=True)
        barrier()
    except Exception as ex:                       # reported in the line, never silently dropped
        e2e_error = "%s: %s" % (type(ex).__name__, ex)
    e2e_s = time.perf_counter() - e0
    pool.shutdown()

    t_dev = torch.tensor([dev_ms / 1e3, e2e_s, wall, (dev1_ms or 0.0) / 1e3], dtype=torch.float64,
                         device="cuda")
    if world > 1 and e2e_error is None:
        dist.all_reduce(t_dev, op=dist.ReduceOp.MAX)
    dev_s, e2e_s, wall, dev1_s = [float(v) for v in t_dev.cpu()]
    work = float(n_total) * T * args.steps * F
    # log Z of a sweep is a function of (seed, T, particles) only - not of how the
    # population is sharded nor of what else is in flight.  Checked against the values
    # recorded by a 1-GPU run (profiles/logz_reference.json); a mismatch fails the run.
    logz_note = check_logz(args, rank, world, F, zs)
    # the e2e leg returns final states: its log Z and a checksum of the states of EVERY timed
    # sweep are compared with the values a 1-GPU run recorded, so that a sharded (IPC / NVLink)
    # run is checked on the particles it returns, not on log Z alone
    x_final_check = None
    if e2e_error is None and e2e_seen:
        keys = sorted(e2e_seen)
        sums = np.array([e2e_seen[k][1] for k in keys], dtype=np.uint64).view(np.int64)
        if world > 1:
            if (n_local * dim) % 4096 == 0:
                tsum = torch.from_numpy(sums.copy()).cuda()
                dist.all_reduce(tsum)                       # int64 addition wraps: sum modulo 2^64
                sums = tsum.cpu().numpy()
            else:
                sums = None
        if sums is not None:
            usums = sums.view(np.uint64)
            x_final_check = check_recorded(
                args, rank, world, "/e2e",
                dict([("logz_" + k, e2e_seen[k][0]) for k in keys] +
                     [("xsum_" + k, int(usums[j])) for j, k in enumerate(keys)]))

    if rank == 0:
        peak, peak_kind = measured_peak()
        dom = max(((n, v) for n, v in kt.items() if n in ALGO_BYTES), key=lambda kv: kv[1][0],
                  default=None)
        roof = None
        step_bytes = 32.0 * dim + 32.0                 # 32 d + 32
        if mvn and "k_propagate" in kt:
            # K6: the two d x d products per particle on the fp64 tensor pipe (DMMA)
            ms, c = kt["k_propagate"]
            avg_s = ms / 1e3 / c
            flop = 4.0 * n_local * dim * dim
            pk64 = dgemm_peak_tflops(torch)
            roof = {"kernel": "k_propagate_mvn<64>", "bound": "tensor", "achieved": flop / avg_s / 1e12,
                    "peak": pk64, "unit": "TFLOP/s", "frac": flop / avg_s / 1e12 / pk64,
                    "traffic": None, "peak_source": "measured in this run: cuBLAS DGEMM 8192^3, best of 5 "
                    "(fp64 tensor pipe; MEASURED_PEAKS.json has no fp64 figure)",
                    "avg_launch_us": avg_s * 1e6, "launches_timed": c,
                    "algorithmic_flop_per_particle": 4.0 * dim * dim}
            dom = None
        trs, tr_src = ncu_traffic_per_particle()

        def kernel_roof(name, ms, c):
            avg_s = ms / 1e3 / c
            achieved = ALGO_BYTES[name] * n_local / avg_s / 1e9
            moved = MOVED_BYTES[name] * n_local / avg_s / 1e9
            tr = trs.get(name)
            return {"kernel": name, "bound": "hbm", "achieved": achieved, "peak": peak,
                    "unit": "GB/s", "frac": achieved / peak, "algorithmic_frac": achieved / peak,
                    "traffic": tr * n_local if tr else None, "traffic_source": tr_src,
                    "dram_frac": (tr * n_local / avg_s / 1e9 / peak) if tr else None,
                    "algorithmic_bytes_per_particle": ALGO_BYTES[name],
                    "compulsory_bytes_per_particle": MOVED_BYTES[name],
                    "compulsory_gbs": moved, "compulsory_frac": moved / peak,
                    "avg_launch_us": avg_s * 1e6, "launches_timed": c}

        roof_all = None
        if dom:
            name, (ms, c) = dom
            roof = kernel_roof(name, ms, c)
            roof["note"] = ("achieved/frac: SURVEY 8(d) algorithmic bytes of the stages this kernel "
                            "covers, every stage counted once over the step's three streaming kernels "
                            "(k_propagate K1+K5 = 44, k_fold_agg K2 = 8, k_ancestors K3+K4 = 12; sum 64). "
                            "compulsory_*: the bytes the fused kernel must move (the gathered state "
                            "never touches HBM). dram_frac: DRAM bytes per launch of the tracked "
                            "`ncu --set full` capture (traffic_source; not re-measured in this run) over "
                            "this run's launch time. The whole-step figure is path_roofline.frac")
            roof["peak_source"] = f"{peak_kind} (MEASURED_PEAKS.json hbm_gbs)"
            roof_all = [kernel_roof(k, *kt[k]) for k in STREAMING if k in kt]
        ksum_ms = sum(v[0] for v in kt.values())
        path = {"bytes_per_particle_step": step_bytes,
                "achieved_gbs": step_bytes * n_local * T * args.steps * F / (dev_s) / 1e9,
                "frac": step_bytes * n_local * T * args.steps * F / dev_s / 1e9 / peak,
                "kernel_share": {n: v[0] / ksum_ms for n, v in kt.items()} if ksum_ms else {},
                "kernel_avg_us": {n: 1e3 * v[0] / v[1] for n, v in kt.items()}}
        # BASELINE's second metric inside the sweep: everything but the state traffic of
        # K1 (16 B of its 24): log-weight write 8 + K2 8 + K3/4 12 + K5 20 = 48 of the 64 B
        if not mvn and ksum_ms:
            path["resample_reweight_share_of_step_bytes"] = RESAMPLE_REWEIGHT_BYTES / STEP_BYTES
        line = {"metric": METRIC, "value": work / dev_s, "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": 1e3 * dev_s / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": args.workload_name, "particles_total": n_total,
                           "particles_per_gpu": n_local, "T": T,
                           "model": ("64-dim linear-Gaussian SSM, mvn observe (DMMA)" if mvn else
                                     "1-D linear-Gaussian a=0.9 q=1 r=1" if args.workload == "lgssm1d" else
                                     "stochastic volatility phi=0.98 sigma=0.16 beta=0.65"),
                           "resampling": "systematic, every step",
                           "sweeps_in_flight": F,
                           "step": "one step = %d independent sweep(s) of the workload in flight, "
                                   "one population (context + stream) each; `value` times the sweeps "
                                   "up to log Z (no final-state gather / download), `e2e` includes "
                                   "the gather, the download of all final states and the per-step "
                                   "statistics" % F,
                           "l2": "inputs exceed L2 (columns of %d MiB per GPU)" % (n_local * 8 >> 20),
                           "timing": "CUDA events on the engine stream, max over ranks"},
                "e2e": {"value": (work / e2e_s) if e2e_error is None else None, "unit": UNIT,
                        "error": e2e_error, "sweeps_in_flight": Fe,
                        "h2d_bytes_per_step": int(T * 8 * dim + 4096) * F,
                        "d2h_bytes_per_step": int(x_host_bytes + 2 * T * 8 + 64) * F},
                "gpu_launches": int(launches), "clocks": clocks, "roofline": roof,
                "roofline_streaming_kernels": roof_all,
                "path_roofline": path, "wall_ms_per_step": 1e3 * wall / args.steps,
                "log_z_mean": float(np.mean(log_zs)), "log_z_check": logz_note,
                "x_final_check": x_final_check,
                "log_zs": {"f%d_i%d" % (f, i): z for f in range(len(zs)) for i, z in enumerate(zs[f])}}
        if args.workload == "lgssm1d":
            kz = kalman_logz(y)
            line["kalman_check"] = {"exact_log_z": kz, "max_abs_error_of_timed_sweeps":
                                    float(max(abs(z - kz) for z in log_zs))}
        if dev1_s > 0.0:
            v1 = float(n_total) * T * args.steps / dev1_s
            line["one_sweep_in_flight"] = {"value": v1, "unit": UNIT,
                                           "path_roofline_frac": step_bytes * v1 / 1e9 / peak / world}
        if world == 1 and not mvn and not args.no_microbench:
            rr = resample_microbench(local, [1 << 20, 1 << 24, 1 << 27], peak)
            line["resample_reweight"] = {
                "metric": "resample+reweight HBM GB/s vs peak (BASELINE.json metric, second part)",
                "bytes_per_particle": RESAMPLE_REWEIGHT_BYTES, "peak": peak, "unit": "GB/s",
                "microbench": rr,
                "what": "column-level C ABI: apop_observe_normal + apop_resample (systematic, sorted "
                        "order) + apop_gather of one f64 column; lw ~ N(0, 3^2), default_rng(7)"}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args)
            line["vs_cpu_baseline"] = {"e2e_over_cpu": (work / e2e_s / line["cpu_baseline"]["value"])
                                       if e2e_error is None else None,
                                       "note": "reported baseline (1 core, C port), not a target; "
                                               "vs_baseline stays null: BASELINE.md holds no published number"}
        print(json.dumps(line), flush=True)
    for p_ in pops:
        p_.close()
    if world > 1:
        dist.destroy_process_group()


LOGZ_FILE = os.path.join(ROOT, "profiles", "logz_reference.json")


def check_logz(args, rank, world, F, zs):
    """zs[f][i] = log Z of timed sweep i of population f (seed 0xA11CE.. + 2000 + 100 f + i)"""
    return check_recorded(args, rank, world, "",
                          {"f%d_i%d" % (f, i): z for f in range(len(zs)) for i, z in enumerate(zs[f])})


def check_recorded(args, rank, world, suffix, mine):
    """Compare this run's per-sweep values (log Z; for the e2e leg also the checksum of the
    final states) with those a 1-GPU run recorded in profiles/logz_reference.json.  They are
    functions of (seed, T, particles) only, so every N must reproduce them exactly; a mismatch
    fails the run.  Values with no recorded counterpart are counted, not hidden."""
    if rank != 0:
        return None
    key = "%s/n%d/T%d%s" % (args.workload, args.particles, args.T, suffix)
    try:
        ref = json.load(open(LOGZ_FILE))
    except Exception:
        ref = {}
    have = ref.get(key, {})
    bad = {k: (v, have[k]) for k, v in mine.items() if k in have and have[k] != v}
    if bad:
        raise SystemExit("values differ from the recorded 1-GPU ones (%s, %s): %r" % (LOGZ_FILE, key, bad))
    checked = sum(1 for k in mine if k in have)
    if world == 1 and args.record_logz:
        have.update(mine)
        ref[key] = have
        json.dump(ref, open(LOGZ_FILE, "w"), indent=1, sort_keys=True)
    return {"compared_with_recorded_1gpu_values": checked, "of": len(mine),
            "not_on_file": len(mine) - checked, "file": "profiles/logz_reference.json", "key": key}


def cpu_baseline(args):
    """oracle port of smc.cljc, 1 core (the reference's `infer :smc` is single
    threaded), on a bounded sample of the same workload."""
    from oracle import pyoracle as orc
    m, y, n_s, T_s, label = oracle_sample(args, 200)       # ~15 s of one core
    t0 = time.perf_counter()
    orc.smc_sweep(m, y, n_s, seed=1, threads=1, want_final=True)
    dt = time.perf_counter() - t0
    return {"value": n_s * T_s / dt, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"{label} of {n_s} particles x {T_s} steps ({dt:.1f} s), oracle C port of "
                      "smc.cljc, extrapolated linearly to the full workload; reference JVM path "
                      "unavailable on this box (no JVM)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="sv", choices=["sv", "lgmvn64", "lgssm1d"],
                    help="sv: BASELINE configs[2] (the headline); lgmvn64: configs[4]; "
                         "lgssm1d: configs[1] (2^20 particles, T=1000, Kalman check)")
    ap.add_argument("--particles", type=int, default=0)
    ap.add_argument("--T", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-microbench", action="store_true",
                    help="skip the resample-only microbench (N = 1, sv workload)")
    ap.add_argument("--record-logz", action="store_true",
                    help="N = 1: store the log Z of every timed sweep in profiles/logz_reference.json; "
                         "runs at any N compare against it")
    ap.add_argument("--inflight", type=int, default=3,
                    help="independent sweeps kept in flight, one population each (round 2, "
                         "G particle-steps/s: 1 GPU 65.6 (1) -> 69.3 (2) = (3); 8 GPUs 253 (1) -> 356 (2) "
                         "-> 425 (3); 4 complete, 6 stall on peer waits at 8 GPUs - DESIGN.md section 5)")
    ap.add_argument("--max-connections", type=int, default=0,
                    help="experiment: CUDA_DEVICE_MAX_CONNECTIONS for this process (0 = leave the "
                         "driver default); to try with --inflight > 3 on several GPUs")
    args = ap.parse_args()
    # one hardware queue per engine stream (see apop_create); must be set before torch
    # creates the CUDA context
    os.environ["CUDA_DEVICE_MAX_CONNECTIONS"] = str(args.max_connections if args.max_connections > 0 else 32)
    if args.workload == "lgmvn64":
        args.particles = args.particles or (1 << 24)
        args.T = args.T or 100
        args.workload_name = ("configs[4]: 64-dim linear-Gaussian SSM, mvn observe, T=%d, %d particles "
                              ":smc sharded over %d B200" % (args.T, args.particles, args.gpus))
    elif args.workload == "lgssm1d":
        args.particles = args.particles or (1 << 20)
        args.T = args.T or 1000
        args.workload_name = ("configs[1]: linear-Gaussian SSM, 1-D latent, T=%d, %d particles :smc on "
                              "%d B200 (log-marginal checked vs the exact Kalman filter)"
                              % (args.T, args.particles, args.gpus))
    else:
        args.particles = args.particles or (1 << 27)
        args.T = args.T or 2000
        args.workload_name = ("configs[2]: stochastic-volatility SSM, T=%d, %d particles :smc sharded "
                              "over %d B200" % (args.T, args.particles, args.gpus))
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
