"""Host-side mirror of `anglican.inference/infer` for the particle algorithms.

    infer(:smc | :pimh | :pgibbs | :pgas | :ipmcmc | :importance | :pfilter, model, observations, **options)

keeps the reference's keywords, defaults and assertions
(smc.cljc:53-57, pimh.cljc:41-45, pgibbs.cljc:112-116, ipmcmc.cljc:148-171,
importance.cljc:16-22, pfilter.cljc:22-62) and
its return contract: a lazy (generator) sequence of states carrying
`log_weight`, `result`, `predicts` (stripdown, inference.cljc:114-117).  The
control flow below (MH accept, retained-trajectory bookkeeping, node index
Gibbs) is the part that stays on the JVM in the real integration; every sweep
over the particle population runs in libanglican_pop.so.
"""
import math
from dataclasses import dataclass, field

import numpy as np

from . import _lib as L
from .population import Population


@dataclass
class State:
    """the stripped-down state map {:log-weight :result :predicts}"""
    log_weight: float
    result: object
    predicts: list = field(default_factory=list)


class _SeedSeq:
    """one RNG key per sweep; host-side draws (accept, rand-nth, index Gibbs)
    come from a numpy generator (the reference uses its unseeded runtime RNG:
    only distributional parity applies, runtime.cljc:142-147)."""

    def __init__(self, seed):
        self.base = int(seed)
        self.k = 0
        self.rng = np.random.default_rng(self.base ^ 0x5EED)

    def next_key(self):
        self.k += 1
        return (self.base + 0x9E3779B97F4A7C15 * self.k) & 0xFFFFFFFFFFFFFFFF


def _result_of(out, k, trajectories):
    if trajectories is not None:
        r = trajectories[k]
        return r[:, 0].copy() if r.shape[1] == 1 else r.copy()
    return out.x_final[k].copy() if np.ndim(out.x_final[k]) else float(out.x_final[k])


def _states(out, log_weight, with_paths):
    traj = out.trajectories() if with_paths else None
    for k in range(out.pop.n):
        yield State(log_weight, _result_of(out, k, traj))


# ---------------------------------------------------------------- :smc
def _infer_smc(model, y, number_of_particles=1, seed=0xA11CE, device=0, paths=False, pop=None,
               ess_threshold=0.0):
    """ess_threshold (engine extension; the reference resamples at every
    observe): resample only when ESS < ess_threshold * N; the states of a sweep that ends
    without a resample carry their own log-weights."""
    assert number_of_particles >= 1, ":number-of-particles must be at least 1"   # smc.cljc:56-57
    own = pop is None
    pop = pop or Population(number_of_particles, seed=seed, device=device)
    seeds = _SeedSeq(seed)
    try:
        while True:                                   # (sample-seq), smc.cljc:58-63
            pop.set_seed(seeds.next_key())
            out = pop.sweep(model, y, keep_history=paths, ess_threshold=ess_threshold)
            if ess_threshold > 0.0:
                traj = out.trajectories() if paths else None
                for k in range(pop.n):
                    yield State(float(out.lw_final[k]), _result_of(out, k, traj))
            else:
                yield from _states(out, out.log_z, paths)
    finally:
        if own:
            pop.close()


# ---------------------------------------------------------------- :pimh
def _infer_pimh(model, y, number_of_particles=1, seed=0xA11CE, device=0, paths=False):
    assert number_of_particles >= 1, ":number-of-particles must be at least 1"   # pimh.cljc:44-45
    n = number_of_particles
    pop = Population(n, seed=seed, device=device)
    seeds = _SeedSeq(seed)
    lib = L.load()

    def run():
        pop.set_seed(seeds.next_key())
        out = pop.sweep(model, y, keep_history=paths)
        traj = out.trajectories() if paths else None
        res = [_result_of(out, k, traj) for k in range(n)]
        # get-log-Z, pimh.cljc:48-50: all final weights equal => log(N) + log Z-hat
        log_z = math.log(float(n)) + out.log_z if math.isfinite(out.log_z) else out.log_z
        return res, log_z

    try:
        cur, log_z = run()
        while True:
            nxt, nxt_log_z = run()
            u = seeds.rng.random()
            if lib.apop_pimh_accept(nxt_log_z, log_z, u):           # lmh.cljc:146-151
                cur, log_z = nxt, nxt_log_z
            for r in cur:                                           # pimh.cljc:53-55
                yield State(0.0, r)
    finally:
        pop.close()


# ---------------------------------------------------------------- :pgibbs
def _infer_pgibbs(model, y, number_of_particles=2, seed=0xA11CE, device=0, paths=False):
    assert number_of_particles >= 2, ":number-of-particles must be at least 2"   # pgibbs.cljc:115-116
    n = number_of_particles
    pop = Population(n, seed=seed, device=device)
    seeds = _SeedSeq(seed)
    T = len(y) // model.state_dim
    retained = None
    try:
        while True:
            pop.set_seed(seeds.next_key())
            # the retained particle runs out of stored samples before the last
            # observe, so that resample is unconditional (pgibbs.cljc:89)
            out = pop.sweep(model, y, retained=retained, csmc=retained is not None,
                            n_cond_steps=T - 1, reset_weights=True,
                            keep_history=True if paths else "compact")
            b = int(seeds.rng.random() * n)                         # rand-nth, pgibbs.cljc:122-123
            retained = out.trace(b)
            yield from _states(out, 0.0 if math.isfinite(out.log_z) else -math.inf, paths)
    finally:
        pop.close()


# ---------------------------------------------------------------- :pgas
def _infer_pgas(model, y, number_of_particles=2, seed=0xA11CE, device=0, paths=False):
    """Particle Gibbs with ancestor sampling (pgas.cljc:107-168) for state-space
    programs.  Every observe is a conditional resample (N-1 of N, pgas.cljc:146-151)
    and the retained particle's ancestor is redrawn among all N particles rescored
    against the retained path (resample-retained, 107-131).  The first sweep has an
    empty retained database (pgas.cljc:168): an unconditional sweep here."""
    assert number_of_particles >= 2, ":number-of-particles must be at least 2"   # pgas.cljc:158-159
    n = number_of_particles
    pop = Population(n, seed=seed, device=device)
    seeds = _SeedSeq(seed)
    T = len(y) // model.state_dim
    retained = None
    try:
        while True:
            pop.set_seed(seeds.next_key())
            cond = retained is not None
            out = pop.sweep(model, y, retained=retained, csmc=cond, n_cond_steps=T,
                            reset_weights=True, keep_history=True if paths else "compact",
                            ancestor_sampling=cond)
            b = int(seeds.rng.random() * n)                         # rand-nth, pgas.cljc:166-167
            retained = out.trace(b)
            yield from _states(out, 0.0 if math.isfinite(out.log_z) else -math.inf, paths)
    finally:
        pop.close()


# ---------------------------------------------------------------- :ipmcmc
def _dist_world():
    try:
        import torch.distributed as dist
    except ImportError:
        return 1
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def _infer_ipmcmc(model, y, number_of_particles=2, number_of_nodes=32,
                  number_of_csmc_nodes=None, all_particles=True, pool=None, seed=0xA11CE,
                  device=0, paths=False, exchange=None, group=None, width=4):
    """one process: the nodes run one after the other on one population.  Under
    `torch.distributed` with more than one rank (or an explicit `group`): the
    nodes are partitioned over the ranks' GPUs (anglican_b200/ipmcmc.py) and this
    rank yields the states of its own nodes."""
    if group is not None or _dist_world() > 1:
        from .ipmcmc import infer_ipmcmc
        yield from infer_ipmcmc(model, y, number_of_particles, number_of_nodes,
                                number_of_csmc_nodes, all_particles, seed, None, group, paths,
                                width)
        return
    assert number_of_particles > 1, ":number-of-particles must be larger than 1"
    assert number_of_nodes > 1, ":number-of-nodes must be larger 1"
    M, n = number_of_nodes, number_of_particles
    P = number_of_csmc_nodes or int(M / 2)                          # ipmcmc.cljc:162-165
    assert P < M, ":number-of-csmc-nodes must be smaller than :number-of-nodes"
    pop = Population(n, seed=seed, device=device)
    seeds = _SeedSeq(seed)
    lib = L.load()
    T = len(y) // model.state_dim
    retained = [None] * P
    try:
        while True:
            results, cand, picks, log_zs = [], [], [], np.empty(M)
            for j in range(M):                                      # cp/pmap over nodes, 176-182
                ret = retained[j] if j < P else None
                pop.set_seed(seeds.next_key())
                out = pop.sweep(model, y, retained=ret, csmc=ret is not None, n_cond_steps=T,
                                reset_weights=True, keep_history=True if paths else "compact")
                log_zs[j] = out.log_z
                traj = out.trajectories() if paths else None
                results.append([_result_of(out, k, traj) for k in range(n)])
                # candidate next retained path of this node, drawn now so the
                # ancestry need not outlive the sweep (rand-nth, ipmcmc.cljc:189)
                picks.append(int(seeds.rng.random() * n))
                cand.append(out.trace(picks[-1]))
            if exchange is not None:
                log_zs = exchange(log_zs)                           # only log-Zs cross nodes, 183-187
            u = seeds.rng.random(P)
            cs = np.empty(P, dtype=np.int32)
            zs = np.empty(M)
            rc = lib.apop_ipmcmc_gibbs(log_zs.ctypes.data_as(L.c_double_p), M, P,
                                       u.ctypes.data_as(L.c_double_p),
                                       cs.ctypes.data_as(L.c_int32_p),
                                       zs.ctypes.data_as(L.c_double_p))
            assert rc == 0
            retained = [cand[int(c)] for c in cs]
            for j in range(M):                                      # ipmcmc.cljc:193-203
                with np.errstate(divide="ignore"):
                    lw = float(np.log(zs[j] / n))
                if all_particles:
                    for r in results[j]:
                        yield State(lw, r)
                else:
                    # one particle per node, the very one whose path is the node's candidate
                    # retained path (ipmcmc.cljc:76-78, 189)
                    yield State(lw, results[j][picks[j]])
    finally:
        pop.close()


# ---------------------------------------------------------------- :importance
def _infer_importance(model, y, number_of_particles=1024, seed=0xA11CE, device=0):
    """importance.cljc:16-22: an endless sequence of independent samples, each
    carrying its own log-weight (the default observe checkpoint,
    inference.cljc:42-45).  The reference runs one sample at a time and takes
    no options; `number_of_particles` here is only the device batch size."""
    assert number_of_particles >= 1
    pop = Population(number_of_particles, seed=seed, device=device)
    seeds = _SeedSeq(seed)
    try:
        while True:
            pop.set_seed(seeds.next_key())
            out = pop.sweep(model, y, no_resample=True)
            for k in range(pop.n):
                yield State(float(out.lw_final[k]), _result_of(out, k, None))
    finally:
        pop.close()


# ---------------------------------------------------------------- :pfilter
def _infer_pfilter(model, y, number_of_particles=1, seed=0xA11CE, device=0):
    """pfilter.cljc:22-62: every `observe` emits the batch of particle states
    as they stand after the observe and BEFORE the resample, with the weight
    reset to 0 ("so that all predicts have the same weight", pfilter.cljc:41-44);
    when the program ends the particles restart.  A state-space program
    predicts its latent state at every step, so each batch is x_t of all
    particles; `predicts` carries (t, x_t)."""
    assert number_of_particles >= 1, ":number-of-particles must be at least 1"   # pfilter.cljc:25-26
    pop = Population(number_of_particles, seed=seed, device=device)
    seeds = _SeedSeq(seed)
    d = model.state_dim
    try:
        while True:                                   # "Restart the particles", pfilter.cljc:54-56
            pop.set_seed(seeds.next_key())
            out = pop.sweep(model, y, keep_history=True)
            _, x = out.history()                      # x[t]: states before the t-th resample
            for t in range(1, out.T + 1):
                for k in range(pop.n):
                    v = x[t][k, 0] if d == 1 else x[t][k].copy()
                    yield State(0.0, None, [(t, float(v) if d == 1 else v)])
    finally:
        pop.close()


_ALGORITHMS = {"smc": _infer_smc, "pimh": _infer_pimh, "pgibbs": _infer_pgibbs, "pgas": _infer_pgas,
               "ipmcmc": _infer_ipmcmc, "importance": _infer_importance,
               "pfilter": _infer_pfilter}


def infer(algorithm, model, observations, **options):
    """(infer algorithm prog value & options) — inference.cljc:25-28.

    `algorithm`: "smc" / ":smc" / "pimh" / "pgibbs" / "ipmcmc" / "importance" / "pfilter".  Options use the
    reference's names with `-` -> `_` (number_of_particles, number_of_nodes,
    number_of_csmc_nodes, all_particles, pool)."""
    key = str(algorithm).lstrip(":")
    if key not in _ALGORITHMS:
        raise ValueError(f"no particle engine behind algorithm :{key}")   # core.cljc:15-26
    y = np.ascontiguousarray(observations, dtype=np.float64).ravel()
    return _ALGORITHMS[key](model, y, **options)


def doquery(algorithm, model, observations, drop_invalid=True, **options):
    """doquery (core.cljc:223-252): infer, then drop states with log-weight
    -inf unless :drop-invalid false (inference.cljc:151-157)."""
    for s in infer(algorithm, model, observations, **options):
        if drop_invalid and not (s.log_weight > -math.inf):
            continue
        yield s


def log_marginal(states):
    """inference.cljc:181-186 on a finite collection of states"""
    lw = np.sort(np.array([s.log_weight for s in states], dtype=np.float64))
    acc = lw[0]
    for v in lw[1:]:
        mx = max(acc, v)
        acc = mx + math.log(math.exp(acc - mx) + math.exp(v - mx)) if math.isfinite(mx) else mx
    return acc - math.log(len(lw))
