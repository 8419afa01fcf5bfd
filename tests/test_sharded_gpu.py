"""Sharded population: `world` contiguous shards exchanging
through the peer-memory fabric must reproduce the unsharded sweep BIT FOR BIT
(the sequential fp64 fold of smc.cljc:94,118-137 is chained exactly from rank
to rank; the counter-based noise depends on the global slot only).

All ranks live in this process on cuda:0, so the N>1 path is covered on a
single-GPU box; the protocol (mailboxes, flags, peer loads/stores) is the one
used across GPUs."""
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _sv_data(T, seed=20240602, phi=0.98, sigma=0.16, beta=0.65):
    rng = np.random.default_rng(seed)
    x = rng.normal() * sigma / math.sqrt(1 - phi * phi)
    ys = []
    for _ in range(T):
        x = phi * x + sigma * rng.normal()
        ys.append(beta * math.exp(x / 2) * rng.normal())
    return np.array(ys)


@pytest.mark.parametrize("world,n_local", [(2, 4096), (2, 50000 * 8), (4, 8 * 1025), (8, 8192)])
def test_sharded_sv_equals_unsharded(world, n_local):
    from anglican_b200 import Population, PopulationGroup
    from anglican_b200.models import StochasticVolatility
    y = _sv_data(25)
    m = StochasticVolatility()
    with Population(n_local * world, seed=31337) as pop:
        ref = pop.sweep(m, y, want_filt=True)
    with PopulationGroup(n_local, world, seed=31337) as grp:
        outs = grp.sweep(m, y, want_filt=True)
    x = np.concatenate([o.x_final for o in outs])
    for o in outs:
        assert o.log_z == ref.log_z
        assert np.array_equal(o.log_mean_w, ref.log_mean_w)
        assert np.allclose(o.ess, ref.ess, rtol=1e-12)
        assert np.allclose(o.filt_mean, ref.filt_mean, rtol=1e-11, atol=1e-13)
    assert np.array_equal(x, ref.x_final)


def test_sharded_lgssm_vs_oracle_and_kalman():
    from anglican_b200 import PopulationGroup
    from anglican_b200.models import LinearGaussian1D
    from oracle import pyoracle as orc
    rng = np.random.default_rng(20240601)
    xs, ys = rng.normal(), []
    for _ in range(30):
        xs = 0.9 * xs + rng.normal()
        ys.append(xs + rng.normal())
    y = np.array(ys)
    m = LinearGaussian1D()
    om = orc.Model(orc.SSM_LGSSM1D, **m.params)
    world, n_local = 2, 32768
    with PopulationGroup(n_local, world, seed=7) as grp:
        outs = grp.sweep(m, y)
    ref = orc.smc_sweep(om, y, n_local * world, seed=7)
    assert abs(outs[0].log_z - ref["log_z"]) <= 1e-9 * abs(ref["log_z"])
    x = np.concatenate([o.x_final for o in outs])
    assert np.max(np.abs(x - ref["x_final"][:, 0])) < 1e-9
    exact, _ = orc.kalman1d(om, y)                 # exact Kalman filter log-marginal
    assert abs(outs[0].log_z - exact) < 0.05


def test_sharded_repeat_sweeps_and_seed():
    """two sweeps on one group (mailbox sequence numbers keep advancing)"""
    from anglican_b200 import Population, PopulationGroup
    from anglican_b200.models import StochasticVolatility
    y = _sv_data(10)
    m = StochasticVolatility()
    with PopulationGroup(8192, 2, seed=1) as grp, Population(16384, seed=1) as pop:
        for seed in (11, 12, 13):
            grp.set_seed(seed)
            pop.set_seed(seed)
            outs = grp.sweep(m, y)
            ref = pop.sweep(m, y, want_filt=False)
            assert outs[0].log_z == ref.log_z == outs[1].log_z
            assert np.array_equal(np.concatenate([o.x_final for o in outs]), ref.x_final)


def test_sharded_with_noise_prefetch(monkeypatch):
    """sharded sweep with the side-stream noise prefetch forced on for every
    full tile == unsharded sweep with it off"""
    from anglican_b200 import Population, PopulationGroup
    from anglican_b200.models import StochasticVolatility
    y = _sv_data(12)
    m = StochasticVolatility()
    monkeypatch.setenv("APOP_NOISE_PREFETCH_TILES", "0")
    with Population(4 * 3 * 4096, seed=5) as pop:
        ref = pop.sweep(m, y)
    monkeypatch.setenv("APOP_NOISE_PREFETCH_TILES", "1000")
    with PopulationGroup(3 * 4096, 4, seed=5) as grp:
        outs = grp.sweep(m, y)
    for o in outs:
        assert o.log_z == ref.log_z
    assert np.array_equal(np.concatenate([o.x_final for o in outs]), ref.x_final)


@pytest.mark.parametrize("world,n_local,d", [(2, 4096, 8), (4, 8 * 700, 16), (2, 8 * 1111, 64)])
def test_sharded_lgmvn_equals_unsharded(world, n_local, d):
    """BASELINE config 5 shape: vector states shard too — parent rows that
    migrated are read from the owner's heap; results are bit-identical to the
    unsharded population (same global Philox slots, exact chained fold)."""
    from anglican_b200 import Population, PopulationGroup
    from anglican_b200.models import LinearGaussianMVN
    rng = np.random.default_rng(d)
    Q, _ = np.linalg.qr(rng.normal(size=(d, d)))
    A = 0.9 * Q
    Qc = 0.1 * np.eye(d)
    R = 0.5 * np.eye(d) + 0.5 * np.ones((d, d)) / d
    T = 6
    y = rng.normal(size=(T, d)).ravel()
    m = LinearGaussianMVN(A, Qc, R)
    with Population(n_local * world, seed=99) as pop:
        ref = pop.sweep(m, y)
    with PopulationGroup(n_local, world, seed=99) as grp:
        outs = grp.sweep(m, y)
    for o in outs:
        assert o.log_z == ref.log_z
        assert np.array_equal(o.log_mean_w, ref.log_mean_w)
    x = np.concatenate([o.x_final.reshape(-1, d) for o in outs])
    assert np.array_equal(x, ref.x_final.reshape(-1, d))
