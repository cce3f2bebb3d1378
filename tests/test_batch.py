"""
Replica ensembles (BASELINE.json configs[3]): replica k = frame
(k mod 80) of the reference's data/ka_N150_rho1.200.xyz (L = 5 = 2 r_c,AA: all
pairs, strict cut-off) + uniform jitter +-0.05 with seed 100 + k; KA cut & shifted.
CPU part: the batch entry points of the oracle against Oracle-A.  GPU part: the
one-block-per-replica kernel against the oracle.
"""
import numpy
import pytest

from conftest import golden
from oracle import oracle_np as onp
from atooms_b200 import capi
from atooms_b200.batch import ReplicaEnsemble


def replicas(count):
    d = golden('ka_N150_frames')
    pos = numpy.empty((count, 150, 3))
    for k in range(count):
        rng = numpy.random.default_rng(100 + k)
        pos[k] = d['position'][k % 80] + rng.uniform(-0.05, 0.05, size=(150, 3))
    return pos, d['species'], d['side']


def test_oracle_batch_matches_numpy_fire(oracle_lib):
    pos, spe, side = replicas(2)
    ens = ReplicaEnsemble(pos, spe, side, 'kob_andersen', library=oracle_lib)
    res = ens.minimize('fire', tolerance=1e-30, steps=150)
    for k in range(2):
        _, it, conv, e, fn2 = onp.fire(pos[k], numpy.ones(150), spe, side, onp.kob_andersen(), 0.002, 150, 1e-30)
        assert res['steps'][k] == it == 150 and not res['converged'][k]
        assert abs(res['energy'][k] - e) < 1e-9 * abs(e)
        assert abs(res['fn2'][k] - fn2) < 1e-6 * fn2


def test_round_robin_split(oracle_lib):
    pos, spe, side = replicas(5)
    parts = [ReplicaEnsemble(pos, spe, side, 'kob_andersen', library=oracle_lib, rank=r, world=2) for r in range(2)]
    assert [list(p.index) for p in parts] == [[0, 2, 4], [1, 3]]
    assert parts[0].nrep == 3 and parts[1].nrep == 2


@pytest.mark.gpu
@pytest.mark.parametrize('method,steps', [('fire', 200), ('sd', 200)])
@pytest.mark.parametrize('precision', [64, 32])
def test_replica_kernel_follows_the_oracle(cuda_lib, oracle_lib, method, steps, precision):
    """A fixed number of iterations: the two implementations stay on the same path."""
    pos, spe, side = replicas(16)
    got = ReplicaEnsemble(pos, spe, side, 'kob_andersen', precision=precision).minimize(method, 1e-30, steps)
    ref = ReplicaEnsemble(pos, spe, side, 'kob_andersen', library=oracle_lib).minimize(method, 1e-30, steps)
    assert numpy.all(got['steps'] == steps) and not got['converged'].any()
    tol = 1e-8 if precision == 64 else 2e-4
    assert numpy.abs(got['energy'] - ref['energy']).max() < tol * numpy.abs(ref['energy']).max()
    assert numpy.abs(got['virial'] - ref['virial']).max() < 100 * tol * numpy.abs(ref['virial']).max()


@pytest.mark.gpu
def test_replicas_converge_to_the_oracle_minima(cuda_lib, oracle_lib):
    """FIRE to sum F^2 / N < 1e-10 (dryrun.py:140) on 64 replicas."""
    pos, spe, side = replicas(64)
    ens = ReplicaEnsemble(pos, spe, side, 'kob_andersen', precision=64)
    got = ens.minimize('fire', tolerance=1e-10, steps=100000)
    ref = ReplicaEnsemble(pos, spe, side, 'kob_andersen', library=oracle_lib).minimize('fire', 1e-10, 100000)
    assert got['converged'].all() and ref['converged'].all()
    assert (got['fn2'] < 1e-10).all()
    same = numpy.abs(got['energy'] - ref['energy']) < 1e-6 * numpy.abs(ref['energy'])
    assert same.mean() >= 0.9                      # round-off may send a few descents into a neighbouring basin
    assert numpy.abs(got['steps'][same] - ref['steps'][same]).max() <= 0.05 * ref['steps'][same].max() + 5
    # the minimised configurations really are minima of the oracle's energy surface
    final = ens.positions()
    e = onp.compute(final[0], spe, side, onp.kob_andersen())
    assert abs(e['energy'] - got['energy'][0]) < 1e-9 * abs(e['energy'])
    assert numpy.sum(e['forces'] ** 2) / 150 < 1e-9


@pytest.mark.gpu
def test_replica_edge_cases(cuda_lib):
    pos, spe, side = replicas(3)
    ens = ReplicaEnsemble(pos, spe, side, 'kob_andersen')
    res = ens.minimize('fire', tolerance=1e30, steps=10)        # already "converged": zero iterations
    assert res['converged'].all() and (res['steps'] == 0).all()
    assert numpy.abs(ens.positions() - (pos - numpy.rint(pos / side) * side)).max() < 1e-12
    with pytest.raises(ValueError):
        ens.minimize('cg')
    with pytest.raises(ValueError):
        ReplicaEnsemble(numpy.zeros((1, 2000, 3)), numpy.zeros(2000, dtype=int), side, 'kob_andersen')
