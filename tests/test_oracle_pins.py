"""
Pin the oracles to every known-answer value the reference's tests and fixtures
hold for the hot path (values + file:line in golden/pins.json).
Both oracles are checked: Oracle-A (NumPy O(N^2)) and Oracle-B (C, Verlet list,
behind the product's C ABI).
"""
import numpy
import pytest

from conftest import golden, make_engine
from oracle import oracle_np as onp


def _places(a, b, places):
    assert round(abs(a - b), places) == 0, (a, b)


@pytest.fixture(scope='module')
def lj_unshifted():
    return onp.lennard_jones(2.5, shifted=False)


@pytest.fixture(scope='module')
def lj_shifted():
    return onp.lennard_jones(2.5, shifted=True)


def _both(oracle_lib, data, table):
    a = onp.compute(data['position'], data['species'], data['side'], table, per_particle=True)
    b = make_engine(oracle_lib, data, table).compute(per_particle=True, stress=True)
    return a, b


def test_lj_N1000_energy(oracle_lib, pins, lj_unshifted):
    pin = pins['lj_N1000_unshifted_U_per_particle']
    for r in _both(oracle_lib, golden('lj_N1000'), lj_unshifted):
        # the reference asserts equality with an 8-significant-digit literal
        assert float('%.7f' % (r['energy'] / 1000)) == pin['value']


def test_lj_fcc_energy_pressure_force(oracle_lib, pins, lj_unshifted):
    d = golden('lj_fcc_N108')
    volume = numpy.prod(d['side'])
    for r in _both(oracle_lib, d, lj_unshifted):
        _places(r['energy'] / 108, pins['lj_fcc_N108_unshifted_U_per_particle']['value'], 7)
        # T = 0: P = (N T + W/d) / V   (atooms/system/system.py:348)
        _places(r['virial'] / 3 / volume, pins['lj_fcc_N108_unshifted_pressure']['value'], 7)
        fn = numpy.sum(r['forces'] ** 2) ** 0.5 / 108
        _places(fn, pins['lj_fcc_N108_force_norm_per_particle']['value'], 2)


def test_pair_energy_docs(pins, lj_unshifted):
    r = onp.compute(numpy.array([[0., 0., 0.], [1.122, 0., 0.]]), [0, 0], [10., 10., 10.], lj_unshifted)
    _places(r['energy'], pins['pair_r1.122_unshifted_U']['value'], 7)


def test_three_particles(oracle_lib, pins, lj_shifted):
    d = golden('lj_N3_inline')
    for r in _both(oracle_lib, d, lj_shifted):
        _places(r['energy'], pins['lj_N3_shifted_U']['value'], 4)
    _places(onp.temperature(d['velocity'], d['mass']), pins['lj_N3_temperature']['value'], 7)
    th = make_engine(oracle_lib, d, lj_shifted).thermo()
    _places(2 * th['kinetic'] / ((3 - 1) * 3), 1.0, 7)


def test_temperature_with_masses(oracle_lib, pins, lj_shifted):
    d = golden('lj_N4_inline')
    table = onp.PairTable(numpy.ones((2, 2)), numpy.ones((2, 2)), 2.5 * numpy.ones((2, 2)))
    _places(onp.temperature(d['velocity'], d['mass']), pins['lj_N4_temperature_mass']['value'], 7)
    th = make_engine(oracle_lib, d, table).thermo()
    _places(2 * th['kinetic'] / ((4 - 1) * 3), pins['lj_N4_temperature_mass']['value'], 7)


def test_ka_N256_energy_and_rumd_columns(oracle_lib, pins):
    d = golden('ka_N256_rumd')
    ka = onp.kob_andersen()
    fref = numpy.c_[d['fx'], d['fy'], d['fz']]
    for r in _both(oracle_lib, d, ka):
        _places(r['energy'] / 256, pins['ka_N256_shifted_U_per_particle']['value'], 3)
        # RUMD is fp32: 1e-5 relative to the largest force
        assert numpy.abs(r['forces'] - fref).max() / numpy.abs(fref).max() < 1e-5
        assert numpy.abs(r['pe'] - d['pe']).max() < 5e-5
        assert numpy.abs(r['vir'] - d['vir']).max() / numpy.abs(d['vir']).max() < 1e-5
        # file columns are per-particle halves: pe_i = 1/2 sum u, vir_i = 1/2 sum r.F
        assert abs(r['pe'].sum() - r['energy']) < 1e-9
        assert abs(r['vir'].sum() - r['virial']) < 1e-8


def test_nve_trajectory_pin(oracle_lib, pins, lj_unshifted):
    """x of particle 0 after 10 and 20 velocity-Verlet steps (tests/test_lammps.py:33-40)."""
    pin = pins['lj_N1000_nve_x0']
    d = golden('lj_N1000')
    _places(d['position'][0, 0], pin['value'][0], 5)
    eng = make_engine(oracle_lib, d, lj_unshifted)
    for target in pin['value'][1:]:
        eng.run_nve(10, pin['dt'])
        pos, _, _ = eng.download()
        _places(pos[0, 0], target, 5)
    # Oracle-A, one shot of 20 steps
    pos, vel, _ = onp.nve(d['position'], d['velocity'], d['mass'], d['species'], d['side'],
                          lj_unshifted, pin['dt'], 20)
    _places(pos[0, 0], pin['value'][2], 5)


def test_run_zero_is_noop(oracle_lib, lj_unshifted):
    """run(10); run(0) leaves x unchanged (tests/test_lammps.py:123-142)."""
    eng = make_engine(oracle_lib, golden('lj_N1000'), lj_unshifted)
    eng.run_nve(10, 0.005)
    x1 = eng.download()[0]
    eng.run_nve(0, 0.005)
    x2 = eng.download()[0]
    assert numpy.abs(x2 - x1).max() < 1e-12


def test_minimisation_inequalities(oracle_lib, pins, lj_shifted):
    """tests/test_optimization.py:20-35: LAMMPS `minimize 0.0 1e-4` stops on
    |F|_2 < 1e-4, i.e. sum F^2 / N < 1e-8 / N."""
    pin = pins['lj_N256_minimised']
    d = golden('lj_N256')
    from atooms_b200 import capi
    eng = make_engine(oracle_lib, d, lj_shifted, velocities=False)
    steps, converged = eng.minimize(capi.MIN_FIRE, 100000, 0.002, 1e-8 / 256)
    assert converged
    th = eng.thermo()
    assert th['potential'] / 256 < pin['U_per_particle_below']
    assert th['fnorm2'] / 256 < pin['fn2_per_particle_below']


def test_oracles_agree(oracle_lib):
    """Two independent formulations (O(N^2) NumPy vs cell/Verlet-list C)."""
    for name, table in [('ka_N1200', onp.kob_andersen()),
                        ('ka_N1212_rumd', onp.kob_andersen(nsp=3)),
                        ('lj_N256', onp.lennard_jones())]:
        d = golden(name)
        a, b = _both(oracle_lib, d, table)
        assert abs(a['energy'] - b['energy']) < 1e-12 * abs(a['energy'])
        assert abs(a['virial'] - b['virial']) < 1e-12 * abs(a['virial'])
        assert numpy.abs(a['forces'] - b['forces']).max() < 1e-10 * numpy.abs(a['forces']).max()
        assert numpy.abs(a['stress'] - b['stress']).max() < 1e-10 * numpy.abs(a['stress']).max()
        assert abs(numpy.trace(b['stress']) - b['virial']) < 1e-9 * abs(b['virial'])


@pytest.mark.parametrize('precision,dtype', [(64, numpy.float64), (32, numpy.float32)])
def test_neighbor_lists_agree(oracle_lib, precision, dtype):
    ka = onp.kob_andersen()
    for name in ['ka_N256_rumd', 'ka_N1200']:
        d = golden(name)
        eng = make_engine(oracle_lib, d, ka, precision=precision)
        ref = onp.neighbor_list(d['position'], d['species'], d['side'], ka, 0.3, dtype=dtype)
        got = eng.neighbors()
        assert all(numpy.array_equal(a, b) for a, b in zip(got, ref))


def test_minimisers_agree(oracle_lib, lj_shifted):
    from atooms_b200 import capi
    d = golden('lj_N256')
    eng = make_engine(oracle_lib, d, lj_shifted, velocities=False)
    steps, conv = eng.minimize(capi.MIN_FIRE, 300, 0.002, 1e-4)
    pos, it, conv_np, e_np, fn2 = onp.fire(d['position'], d['mass'], d['species'], d['side'],
                                           lj_shifted, 0.002, 300, 1e-4)
    assert steps == it and conv == conv_np
    assert abs(eng.thermo()['potential'] - e_np) < 1e-9 * abs(e_np)


@pytest.mark.parametrize('kind', [1, 0], ids=['nose-hoover', 'berendsen'])
def test_nvt_reaches_the_target_temperature(oracle_lib, pins, lj_unshifted, kind):
    """tests/test_lammps.py:42-98: start at T = 1.5, thermostat at 2.0 with coupling time 0.2,
    dt = 0.002; the reference averages T sampled every 500 steps after 1500 steps (1 place)."""
    pin = pins['lj_N1000_nvt_temperature']
    d = dict(golden('lj_N1000').items())
    d['velocity'] = onp.maxwell_boltzmann(1000, 1.5, 1.0, seed=1)
    eng = make_engine(oracle_lib, d, lj_unshifted)
    T = []
    for _ in range(40):
        eng.run_nvt(100, 0.002, 2.0, 0.2, kind)
        T.append(2 * eng.thermo()['kinetic'] / (3 * 999))
    assert abs(numpy.mean(T[15:]) - pin['value']) < 0.05
    # the thermostat really acts: NVE from the same start stays near its own temperature
    nve = make_engine(oracle_lib, d, lj_unshifted)
    nve.run_nve(1000, 0.002)
    assert abs(2 * nve.thermo()['kinetic'] / (3 * 999) - 2.0) > 0.1
