"""
Parity at the shapes bench.py times, through the C ABI, against the
C oracle (oracle/oracle_c.c), which evaluates a million particles in seconds.

  C3  Kob-Andersen N = 1 200 000 (bench.ka_workload(): BASELINE.json configs[2]) — code paths
      the small fixtures never reach: three radix passes, a 293-block histogram scan,
      2^18 codes of cell blocks in the list build, list capacity growth.
  C5  LJ FCC rho = 0.85, dt = 0.005, jitter 0.05 a (BASELINE.json configs[4]) at 60^3 cells,
      N = 864 000: the per-GPU share of the 8M system at 8 GPUs is 1M; the full 8 001 504
      would take the CPU side minutes per evaluation.

Tolerances as in test_gpu_parity.py (north_star): fp64 forces 1e-10, totals 1e-12; fp32 forces
1e-5, totals 1e-6 (against the oracle on the fixed-point positions the fp32 engine holds);
neighbour lists bit-exact, every row of every particle.
"""
import os
import sys

import numpy
import pytest

from conftest import make_engine
from oracle import oracle_np as onp
from test_gpu_parity import assert_observables, rounded

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu


def c3_workload():
    import bench
    pos, spe, side, vel = bench.ka_workload()
    return {'position': pos, 'species': spe, 'side': side, 'center': numpy.zeros(3), 'velocity': vel,
            'mass': numpy.ones(len(spe))}, onp.kob_andersen(), 0.002


def c5_workload(cells=60):
    a = (4.0 / 0.85) ** (1.0 / 3.0)
    basis = numpy.array([[0, 0, 0], [0.5, 0.5, 0], [0.5, 0, 0.5], [0, 0.5, 0.5]])
    ix = numpy.arange(cells)
    grid = numpy.stack(numpy.meshgrid(ix, ix, ix, indexing='ij'), -1).reshape(-1, 3)
    pos = ((grid[:, None, :] + basis[None, :, :]) * a).reshape(-1, 3) - 0.5 * cells * a
    pos += numpy.random.default_rng(5).uniform(-0.05 * a, 0.05 * a, size=pos.shape)
    n = len(pos)
    return {'position': pos, 'species': numpy.zeros(n, dtype=numpy.int32), 'side': numpy.full(3, cells * a),
            'center': numpy.zeros(3), 'velocity': onp.maxwell_boltzmann(n, 1.0, 1.0, seed=6),
            'mass': numpy.ones(n)}, onp.lennard_jones(2.5, True), 0.005


WORKLOADS = {'C3_ka_1200000': c3_workload, 'C5_ljfcc_864000': c5_workload}


@pytest.fixture(scope='module', params=sorted(WORKLOADS))
def workload(request):
    return (request.param,) + WORKLOADS[request.param]()


@pytest.mark.parametrize('precision', [64, 32])
def test_single_point_and_list_at_bench_shape(cuda_lib, oracle_lib, workload, precision):
    name, data, table, dt = workload
    d = rounded(data, precision)
    n = len(d['species'])
    eng = make_engine(cuda_lib, d, table, precision)
    ora = make_engine(oracle_lib, d, table, 64)
    got, ref = eng.compute(), ora.compute()
    assert_observables(got, ref, precision)
    ora.close()
    # the whole Verlet list, bit for bit, in the engine's own precision of the predicate
    cnt, nbr = eng.neighbors_matrix()
    orl = make_engine(oracle_lib, d, table, precision)
    rcnt, rnbr = orl.neighbors_matrix()
    orl.close()
    assert cnt.shape == (n,) and numpy.array_equal(cnt, rcnt)
    assert nbr.shape == rnbr.shape and numpy.array_equal(nbr, rnbr)
    st = eng.stats()
    # fp32: the windowed list (b2_win.cuh), fp64: the global-index list
    assert st['list_mode'] == (2 if precision == 32 else 0) and st['max_neighbors'] == int(cnt.max())
    eng.close()


@pytest.mark.parametrize('precision', [64, 32])
def test_nve_at_bench_shape(cuda_lib, oracle_lib, workload, precision):
    """200 steps with several list refreshes at the benchmarked size; the configuration the engine
    then holds is re-evaluated from scratch by the oracle, and the first 20 steps are compared
    step for step (before chaos amplifies round-off)."""
    name, data, table, dt = workload
    d = rounded(data, precision)
    n = len(d['species'])
    eng = make_engine(cuda_lib, d, table, precision)
    if precision == 64:
        ora = make_engine(oracle_lib, d, table, 64)
        eng.run_nve(20, dt)
        ora.run_nve(20, dt)
        (p, v, _), (po, vo, _) = eng.download(), ora.download()
        dp = p - po
        dp -= numpy.rint(dp / d['side']) * d['side']
        assert numpy.abs(dp).max() < 1e-9 and numpy.abs(v - vo).max() < 1e-8
        ora.close()
        done = 20
    else:
        done = 0
    t0 = eng.thermo()
    eng.run_nve(200 - done, dt)
    t1 = eng.thermo()
    st = eng.stats()
    assert st['steps'] == 200 and st['rebuilds'] >= 5
    e0, e1 = (t0['kinetic'] + t0['potential']) / n, (t1['kinetic'] + t1['potential']) / n
    assert abs(e1 - e0) < (2e-4 if precision == 64 else 1e-3)
    pos, vel, _ = eng.download()
    got = eng.compute()
    ref = make_engine(oracle_lib, dict(d, position=pos, velocity=vel), table, 64).compute()
    assert_observables(got, ref, precision)
    assert abs(t1['kinetic'] - onp.kinetic_energy(vel, d['mass'])) < 1e-6 * t1['kinetic']
    eng.close()
