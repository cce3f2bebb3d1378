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


# ---- NVE of small systems inside one launch (BASELINE.json configs[0]) ----
def lj256():
    d = golden('lj_N256')
    vel = onp.maxwell_boltzmann(256, 2.0, 1.0, seed=1)
    return d['position'][None], d['species'], d['side'], vel[None]


def test_oracle_batch_nve_matches_engine(oracle_lib):
    from conftest import make_engine
    pos, spe, side, vel = lj256()
    ens = ReplicaEnsemble(pos, spe, side, 'lennard_jones', library=oracle_lib)
    ens.set_velocities(vel)
    r = ens.run_nve(50, 0.002)
    eng = make_engine(oracle_lib, {'position': pos[0], 'velocity': vel[0], 'species': spe, 'side': side,
                                   'center': numpy.zeros(3)}, onp.lennard_jones(2.5, True))
    eng.run_nve(50, 0.002)
    t = eng.thermo()
    assert abs(r['energy'][0] - t['potential']) < 1e-9 * abs(t['potential'])
    assert abs(r['kinetic'][0] - t['kinetic']) < 1e-9 * t['kinetic']
    assert numpy.abs(ens.positions()[0] - eng.download()[0]).max() < 1e-10


@pytest.mark.gpu
@pytest.mark.parametrize('precision', [64, 32])
def test_batch_nve_follows_the_oracle_and_conserves_energy(cuda_lib, oracle_lib, precision):
    """C1: LJ N=256 rho=1.0 (data/lj_N256_rho1.0.xyz), T=2.0 velocities seed 1, cut & shifted 2.5,
    dt=0.002: 50 steps step-for-step against the oracle, then 1e4 steps for the drift."""
    pos, spe, side, vel = lj256()
    got = ReplicaEnsemble(pos, spe, side, 'lennard_jones', precision=precision)
    ref = ReplicaEnsemble(pos, spe, side, 'lennard_jones', library=oracle_lib)
    for e in (got, ref):
        e.set_velocities(vel)
    a, b = got.run_nve(50, 0.002), ref.run_nve(50, 0.002)
    tol = 1e-9 if precision == 64 else 2e-4
    assert numpy.abs(got.positions() - ref.positions()).max() < tol
    assert numpy.abs(got.velocities() - ref.velocities()).max() < 10 * tol
    assert abs(a['energy'][0] - b['energy'][0]) < (1e-9 if precision == 64 else 1e-4) * abs(b['energy'][0])
    e0 = (a['energy'][0] + a['kinetic'][0]) / 256
    long = got.run_nve(10000, 0.002)
    e1 = (long['energy'][0] + long['kinetic'][0]) / 256
    assert abs(e1 - e0) < (5e-4 if precision == 64 else 2e-3)
    # momentum stays zero, images are tracked
    v = got.velocities()[0]
    assert numpy.abs(v.sum(axis=0)).max() < (1e-9 if precision == 64 else 1e-2)
    assert numpy.abs(got.batch.state()[1]).max() >= 1


@pytest.mark.gpu
def test_batch_nve_many_replicas_are_independent(cuda_lib):
    pos, spe, side = replicas(8)
    vel = numpy.stack([onp.maxwell_boltzmann(150, 0.5, 1.0, seed=50 + k) for k in range(8)])
    ens = ReplicaEnsemble(pos, spe, side, 'kob_andersen')
    ens.set_velocities(vel)
    r8 = ens.run_nve(200, 0.002)
    one = ReplicaEnsemble(pos[5:6], spe, side, 'kob_andersen')
    one.set_velocities(vel[5:6])
    r1 = one.run_nve(200, 0.002)
    assert r8['energy'][5] == r1['energy'][0] and r8['kinetic'][5] == r1['kinetic'][0]
