"""IPMCMC over several ranks (anglican_b200/ipmcmc.py): world-size-2
and -3 `gloo` runs of the PROTOCOL on the CPU -- node partition, all-reduce of the
node log-marginals, redundant index Gibbs (libanglican_pop.so host entry point),
point-to-point moves of retained paths -- with the node sweeps supplied by the
oracle (test seam `runner=`; the product default runs them on the GPU and has no
CPU path).  The emitted sequence must not depend on the number of ranks:
every rank layout reproduces the single-process run bit for bit."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

M_NODES, P_COND, N_PART, T_OBS, ITERS, SEED = 6, 3, 64, 12, 6, 4242


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _problem():
    from anglican_b200.models import LinearGaussian1D
    from oracle import pyoracle as orc
    rng = np.random.default_rng(11)
    x, ys = 0.0, []
    for _ in range(T_OBS):
        x = 0.9 * x + rng.normal()
        ys.append(x + rng.normal())
    model = LinearGaussian1D(m0=0.0, s0=1.0, a=0.9, b=0.0, q=1.0, c=1.0, d=0.0, r=1.0)
    om = orc.Model(orc.SSM_LGSSM1D, m0=0.0, s0=1.0, a=0.9, b=0.0, q=1.0, c=1.0, d=0.0, r=1.0)
    return model, om, np.array(ys)


class OracleNodes:
    """node sweeps by the CPU oracle (conditional SMC of pgibbs.cljc:76-110 /
    ipmcmc.cljc:34-83): test infrastructure only"""

    def __init__(self, om, y, n):
        self.om, self.y, self.n = om, y, n

    def run(self, jobs):
        from anglican_b200.ipmcmc import NodeRun
        from oracle import pyoracle as orc
        out = []
        for j in jobs:
            r = orc.smc_sweep(self.om, self.y, self.n, seed=j.key, retained=j.retained,
                              csmc=j.retained is not None, n_cond_steps=self.y.size,
                              reset_weights=True, keep_history=True)
            cand = orc.trace_lineage(r, int(j.pick * self.n)).reshape(-1)
            out.append(NodeRun(float(r["log_z"]), cand, r["x_final"][:, 0].copy()))
        return out

    def close(self):
        pass


def _collect(distributed, m_nodes=M_NODES, p_cond=P_COND):
    from anglican_b200.ipmcmc import ipmcmc_iterations
    model, om, y = _problem()
    steps = []
    gen = ipmcmc_iterations(model, y, N_PART, m_nodes, p_cond, seed=SEED,
                            runner=OracleNodes(om, y, N_PART), distributed=distributed)
    for _, st in zip(range(ITERS), gen):
        steps.append({"log_zs": st["log_zs"].copy(), "zs": st["zeta_sums"].copy(),
                      "cs": st["csmc_indices"].copy(), "nodes": list(st["nodes"]),
                      "moved": st["paths_moved"],
                      "res": {j: r.results.copy() for j, r in zip(st["nodes"], st["runs"])},
                      "cand": {j: r.candidate.copy() for j, r in zip(st["nodes"], st["runs"])}})
    gen.close()
    return steps


def _worker(rank, world, port, m_nodes, p_cond, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        q.put((rank, _collect(True, m_nodes, p_cond)))
    finally:
        dist.barrier()
        dist.destroy_process_group()


@pytest.mark.parametrize("world,m_nodes,p_cond", [(2, 6, 3), (2, 5, 2), (3, 7, 3)])
def test_rank_layout_does_not_change_the_chain(world, m_nodes, p_cond):
    ref = _collect(False, m_nodes, p_cond)
    assert all(len(s["nodes"]) == m_nodes for s in ref)
    # the conditional sweeps actually condition: their log-marginals differ from
    # what an unconditional sweep with the same key would give only through the
    # retained path, so at least the index update must have moved something
    assert any(list(s["cs"]) != list(range(p_cond)) for s in ref)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, m_nodes, p_cond, q))
             for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    moved = 0
    for it in range(ITERS):
        seen = []
        for r in range(world):
            s = got[r][it]
            assert np.array_equal(s["log_zs"], ref[it]["log_zs"]), (it, r)
            assert np.array_equal(s["cs"], ref[it]["cs"])
            assert np.array_equal(s["zs"], ref[it]["zs"])
            for j in s["nodes"]:
                assert np.array_equal(s["res"][j], ref[it]["res"][j]), (it, r, j)
                assert np.array_equal(s["cand"][j], ref[it]["cand"][j])
            seen += s["nodes"]
            moved += s["moved"]
        assert seen == list(range(m_nodes))                 # contiguous blocks in rank order
        assert abs(ref[it]["zs"].sum() - p_cond) < 1e-12     # one unit of probability per slot
    assert moved > 0                                        # retained paths did cross ranks


def test_partition_helpers():
    from anglican_b200.ipmcmc import node_key, node_range, owner_of
    for m in range(2, 40):
        for world in range(1, 9):
            owners = [owner_of(j, m, world) for j in range(m)]
            for r in range(world):
                lo, hi = node_range(m, r, world)
                assert owners[lo:hi] == [r] * (hi - lo)
            assert node_range(m, world - 1, world)[1] == m
    keys = {node_key(7, it, j) for it in range(50) for j in range(64)}
    assert len(keys) == 50 * 64


def test_states_and_weights_single_process():
    from anglican_b200.ipmcmc import infer_ipmcmc
    model, om, y = _problem()
    gen = infer_ipmcmc(model, y, N_PART, M_NODES, P_COND, seed=SEED,
                       runner=OracleNodes(om, y, N_PART), distributed=False)
    states = [s for _, s in zip(range(2 * M_NODES * N_PART), gen)]
    gen.close()
    for it in range(2):
        blk = states[it * M_NODES * N_PART:(it + 1) * M_NODES * N_PART]
        w = np.exp([s.log_weight for s in blk])
        assert abs(w.sum() - P_COND) < 1e-9                 # sum_j N * zeta_j / N
        assert all(isinstance(s.result, float) for s in blk)
