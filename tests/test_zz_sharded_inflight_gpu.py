"""Several SHARDED populations driven from independent host threads, with the final states
gathered and downloaded after every sweep.  Kernels of a sharded sweep wait on their peers
inside the kernel; on 8 B200 two such populations in flight with final-state download ran into
the spin bound.  The library now admits sharded sweeps itself (csrc/ssm.cu,
ShardedAdmission): at most 4 at a time, a sweep that returns final states alone; in-process
groups wait for their turn.  So any number of threads may call `sweep` concurrently and every
sweep completes and equals the unsharded population bit for bit (smc.cljc:35-51: a sweep does
not depend on what else runs)."""
import threading

import numpy as np
import pytest

from tests.test_sharded_gpu import _sv_data

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world,n_groups", [(2, 4), (4, 6), (8, 4)])
@pytest.mark.parametrize("want_final", [True, False])
def test_many_sharded_populations_from_threads(world, n_groups, want_final):
    from anglican_b200 import Population, PopulationGroup
    from anglican_b200.models import StochasticVolatility
    n_local, T, rounds = 8 * 2048, 30, 3
    y = _sv_data(T)
    m = StochasticVolatility()
    refs = {}
    for gi in range(n_groups):
        for r in range(rounds):
            seed = 1000 + 17 * gi + r
            with Population(n_local * world, seed=seed) as pop:
                o = pop.sweep(m, y)
                refs[(gi, r)] = (o.log_z, o.x_final.copy())
    groups = [PopulationGroup(n_local, world, seed=1) for _ in range(n_groups)]
    errors, got = [], {}
    start = threading.Barrier(n_groups)

    def drive(gi):
        try:
            start.wait()
            for r in range(rounds):
                groups[gi].set_seed(1000 + 17 * gi + r)
                outs = groups[gi].sweep(m, y, want_final=want_final)
                x = np.concatenate([o.x_final for o in outs]) if want_final else None
                got[(gi, r)] = (outs[0].log_z, x)
        except Exception as ex:                       # surfaced below, with the thread's number
            errors.append((gi, repr(ex)))

    threads = [threading.Thread(target=drive, args=(gi,)) for gi in range(n_groups)]
    try:
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=300)
        assert not any(t.is_alive() for t in threads), "a sharded sweep did not return"
        assert not errors, errors
        for key, (lz, x) in refs.items():
            assert got[key][0] == lz, key
            if want_final:
                assert np.array_equal(got[key][1], x), key
    finally:
        for g in groups:
            g.close()
