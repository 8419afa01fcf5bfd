"""`infer :ipmcmc` over the GPUs of a box (ipmcmc.cljc:148-208).

One process per GPU (`torch.distributed`, any backend).  The M nodes are
split into contiguous blocks, one block per rank; each node is an independent
(conditional) SMC sweep on that rank's GPU, several nodes of a rank in flight
at once (one population = context + stream each, like the reference's
`cp/pmap` over nodes, ipmcmc.cljc:176-182).  Per MCMC iteration the ranks
exchange

  * the M node log-marginals (one all-reduce of M doubles, ipmcmc.cljc:183-187),
  * a retained path ((T+1)*d doubles), point to point, only when the index
    Gibbs update hands conditional slot j a path from a node of another rank;

no particle ever crosses a node.  The index Gibbs update
(`gibbs-update-csmc-indices`, ipmcmc.cljc:106-146) runs redundantly on every
rank from a stream derived from (seed, iteration), so all ranks agree on it
without a broadcast.  All host-side draws are keyed by (seed, iteration[, node])
and every node sweep by (seed, iteration, node): the emitted states do not
depend on the number of ranks.
"""
import math
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass

import numpy as np

from . import _lib as L


# ------------------------------------------------------------ partition, keys
def node_range(number_of_nodes, rank, world):
    """contiguous block [lo, hi) of nodes owned by `rank` (balanced: the first
    M mod world ranks own one node more)"""
    q, r = divmod(number_of_nodes, world)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def owner_of(node, number_of_nodes, world):
    q, r = divmod(number_of_nodes, world)
    edge = r * (q + 1)
    return node // (q + 1) if node < edge else r + (node - edge) // q


_M64 = 0xFFFFFFFFFFFFFFFF


def _mix(x):
    """splitmix64 finaliser"""
    x = (x + 0x9E3779B97F4A7C15) & _M64
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & _M64
    return x ^ (x >> 31)


def node_key(seed, iteration, node):
    """RNG key of the sweep of `node` in `iteration` (same on any rank layout)"""
    return _mix(_mix(_mix(int(seed) & _M64) ^ (iteration + 1)) ^ ((node + 1) << 20))


def shared_draws(seed, iteration, number_of_nodes, number_of_csmc_nodes):
    """host-side uniforms of one iteration, identical on every rank:
    picks[j]  -> rand-nth over the particles of node j   (ipmcmc.cljc:189); with
                 all-particles? false also the one particle the node emits (76-78)
    u[i]      -> index Gibbs draw of conditional slot i   (ipmcmc.cljc:125-133)
    (third array: spare)"""
    g = np.random.default_rng([int(seed) & _M64, int(iteration), 0x1B3C])
    return g.random(number_of_nodes), g.random(number_of_csmc_nodes), g.random(number_of_nodes)


# ------------------------------------------------------------------ node runs
@dataclass
class NodeJob:
    node: int
    key: int
    retained: object          # (T+1)*d doubles or None (unconditional sweep)
    pick: float               # uniform for the candidate retained path


@dataclass
class NodeRun:
    log_z: float
    candidate: np.ndarray     # (T+1)*d doubles: path of one final particle
    results: np.ndarray       # final states [n] / [n, d], or paths [n, T+1, d]


class PopulationNodes:
    """Runs node sweeps on one GPU through libanglican_pop.so: `width`
    populations (context + engine stream each), one host thread per population.
    There is no CPU path: constructing it without the CUDA library fails."""

    def __init__(self, model, y, number_of_particles, device=0, width=4, paths=False):
        from .population import Population
        self.model, self.y, self.n, self.paths = model, np.asarray(y, dtype=np.float64), \
            number_of_particles, paths
        self.T = self.y.size // model.state_dim
        self.pops = [Population(number_of_particles, seed=1 + i, device=device)
                     for i in range(max(1, width))]
        self.pool = ThreadPoolExecutor(max_workers=len(self.pops))

    def _one(self, pop, job):
        pop.set_seed(job.key)
        ret = job.retained
        out = pop.sweep(self.model, self.y, retained=ret, csmc=ret is not None,
                        n_cond_steps=self.T, reset_weights=True,
                        keep_history=True if self.paths else "compact")
        cand = np.asarray(out.trace(int(job.pick * self.n)), dtype=np.float64).reshape(-1)
        res = out.trajectories() if self.paths else np.array(out.x_final, copy=True)
        return NodeRun(float(out.log_z), cand, res)

    def run(self, jobs):
        W = len(self.pops)
        lanes = [jobs[i::W] for i in range(W)]          # a population runs one sweep at a time

        def lane(i):
            return [self._one(self.pops[i], j) for j in lanes[i]]
        done = list(self.pool.map(lane, range(W)))
        out = [None] * len(jobs)
        for i in range(W):
            out[i::W] = done[i]
        return out

    def close(self):
        self.pool.shutdown()
        for p in self.pops:
            p.close()


# ------------------------------------------------------------------ exchange
class _Group:
    """the three things that cross ranks, on `torch.distributed` (gloo: host
    tensors; nccl: device tensors) or nothing at all for a single process"""

    def __init__(self, group=None, enabled=True, abi_exchange=False, device=0):
        self.dist = None
        self.abi = None
        self.rank, self.world = 0, 1
        if enabled:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                import torch
                self.dist, self.torch, self.group = dist, torch, group
                self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
                self.dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
                if abi_exchange and self.world > 1:
                    # the call a JVM binding makes: apop_ipmcmc_exchange_logz on a
                    # context that carries this rank's NCCL communicator (a few particles only:
                    # the context exists for the exchange)
                    from .distributed import nccl_unique_id
                    from .population import Population
                    self.abi = Population(8, device=device, rank=self.rank, world=self.world,
                                          nccl_id=nccl_unique_id(self.rank, group=group))

    def close(self):
        if self.abi is not None:
            self.abi.close()
            self.abi = None

    def gather_log_zs(self, local, lo, number_of_nodes):
        """every rank gets all M log-marginals: all-reduce(sum) of a vector that
        is zero outside the own block (x + 0.0 == x, also for -inf)"""
        full = np.zeros(number_of_nodes)
        full[lo:lo + len(local)] = local
        if self.dist is None:
            return full
        if self.abi is not None and number_of_nodes % self.world == 0:
            # equal blocks in rank order: one NCCL allgather of M / world doubles per rank
            loc = np.ascontiguousarray(local, dtype=np.float64)
            out = np.empty(number_of_nodes)
            L.check(self.abi.ctx, self.abi.lib.apop_ipmcmc_exchange_logz(
                self.abi.ctx, loc.ctypes.data_as(L.c_double_p), loc.size,
                out.ctypes.data_as(L.c_double_p)))
            return out
        t = self.torch.from_numpy(full).to(self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        return t.cpu().numpy()

    def move_paths(self, moves, have, length):
        """moves: [(slot, src_rank, dst_rank, src_node)] with src_rank != dst_rank, in
        the same order on every rank; have(node) -> path held by this rank.
        Returns {slot: path} for the slots this rank receives."""
        if not moves or self.dist is None:
            return {}
        ops, bufs = [], {}
        for slot, src, dst, node in moves:
            if src == self.rank:
                t = self.torch.from_numpy(np.ascontiguousarray(have(node))).to(self.dev)
                ops.append(self.dist.P2POp(self.dist.isend, t, self._global(dst), self.group, slot))
            elif dst == self.rank:
                t = self.torch.empty(length, dtype=self.torch.float64, device=self.dev)
                bufs[slot] = t
                ops.append(self.dist.P2POp(self.dist.irecv, t, self._global(src), self.group, slot))
        if ops:
            for w in self.dist.batch_isend_irecv(ops):
                w.wait()
        return {s: t.cpu().numpy() for s, t in bufs.items()}

    def _global(self, r):
        return r if self.group is None else self.dist.get_global_rank(self.group, r)


# -------------------------------------------------------------------- driver
def ipmcmc_iterations(model, y, number_of_particles=2, number_of_nodes=32,
                      number_of_csmc_nodes=None, seed=0xA11CE, device=None, group=None,
                      paths=False, width=4, runner=None, distributed=True, abi_exchange=False):
    """Generator over MCMC iterations.  Yields, on every rank, a dict with
    `log_zs` [M] (all nodes), `zeta_sums` [M], `csmc_indices` [P], `nodes`
    (the range this rank owns) and `runs` (NodeRun per owned node).
    `runner`: object with run(jobs) -> [NodeRun] and close(); default
    PopulationNodes on this rank's GPU."""
    assert number_of_particles > 1, ":number-of-particles must be larger than 1"   # ipmcmc.cljc:157-160
    assert number_of_nodes > 1, ":number-of-nodes must be larger 1"
    M, n = number_of_nodes, number_of_particles
    P = number_of_csmc_nodes or int(M / 2)                          # ipmcmc.cljc:162-165
    assert P < M, ":number-of-csmc-nodes must be smaller than :number-of-nodes"
    if device is None and distributed:
        try:
            import torch.distributed as _d
            device = _d.get_rank(group) if _d.is_initialized() else 0
        except Exception:
            device = 0
    g = _Group(group, distributed, abi_exchange, device or 0)
    lo, hi = node_range(M, g.rank, g.world)
    d = model.state_dim
    T = np.asarray(y).size // d
    own = runner is None
    if own:
        if device is None:
            device = g.rank
        runner = PopulationNodes(model, y, n, device=device, width=min(width, max(1, hi - lo)),
                                 paths=paths)
    lib = L.load()
    retained = {}                        # conditional slot (node index < P) -> path, owned slots only
    it = 0
    try:
        while True:
            picks, u, _ = shared_draws(seed, it, M, P)
            jobs = [NodeJob(j, node_key(seed, it, j), retained.get(j), float(picks[j]))
                    for j in range(lo, hi)]
            runs = runner.run(jobs)
            log_zs = g.gather_log_zs([r.log_z for r in runs], lo, M)
            cs = np.empty(P, dtype=np.int32)
            zs = np.empty(M)
            rc = lib.apop_ipmcmc_gibbs(np.ascontiguousarray(log_zs).ctypes.data_as(L.c_double_p), M, P,
                                       np.ascontiguousarray(u).ctypes.data_as(L.c_double_p),
                                       cs.ctypes.data_as(L.c_int32_p), zs.ctypes.data_as(L.c_double_p))
            if rc != L.APOP_OK:
                raise L.ApopError(rc, "apop_ipmcmc_gibbs")
            # next retained path of conditional slot j: the candidate of node cs[j]
            # (rand-nth over that node's particles, ipmcmc.cljc:189-190)
            moves, nxt = [], {}
            for j in range(P):
                src_node = int(cs[j])
                src, dst = owner_of(src_node, M, g.world), owner_of(j, M, g.world)
                if src == dst:
                    if dst == g.rank:
                        nxt[j] = runs[src_node - lo].candidate
                else:
                    moves.append((j, src, dst, src_node))
            nxt.update(g.move_paths(moves, lambda node: runs[node - lo].candidate, (T + 1) * d))
            retained = nxt
            yield {"iteration": it, "log_zs": log_zs, "zeta_sums": zs, "csmc_indices": cs,
                   "nodes": range(lo, hi), "runs": runs,
                   "paths_moved": sum(1 for m_ in moves if g.rank in (m_[1], m_[2]))}
            it += 1
    finally:
        g.close()
        if own:
            runner.close()


def infer_ipmcmc(model, y, number_of_particles=2, number_of_nodes=32, number_of_csmc_nodes=None,
                 all_particles=True, seed=0xA11CE, device=None, group=None, paths=False, width=4,
                 runner=None, distributed=True):
    """Lazy sequence of states of the nodes THIS rank owns, weighted
    log(zeta_node / N) (ipmcmc.cljc:193-203).  Concatenating the ranks' sequences
    per iteration in rank order gives the reference's sequence."""
    from .inference import State
    n = number_of_particles
    for step in ipmcmc_iterations(model, y, n, number_of_nodes, number_of_csmc_nodes, seed, device,
                                  group, paths, width, runner, distributed):
        # all-particles? false: the reference keeps ONE particle per node (ipmcmc.cljc:76-78) and
        # the next retained state is rand-nth of that one-element list - the emitted particle
        # and the candidate retained path are the same particle (picks), not two draws
        picks, _, _ = shared_draws(seed, step["iteration"], number_of_nodes, len(step["csmc_indices"]))
        for j, run in zip(step["nodes"], step["runs"]):
            z = step["zeta_sums"][j]
            lw = math.log(z / n) if z > 0.0 else -math.inf
            res = run.results
            if all_particles:
                for k in range(n):
                    yield State(lw, _value(res[k]))
            else:
                yield State(lw, _value(res[int(picks[j] * n)]))


def _value(r):
    r = np.asarray(r)
    if r.ndim == 0:
        return float(r)
    if r.ndim == 2 and r.shape[1] == 1:
        return r[:, 0].copy()
    return r.copy()
