"""
Host-side logic of the drop-in (Interaction / System / B200 / EnergyMinimization /
Trajectory) on CPU.  The numerical engine behind the C ABI is the CPU oracle here
(injected with `library=`; the product default is the CUDA library), so what is
tested is the Python layer: protocol, synchronisation contract, checkpointing and,
when /root/reference is present, the reference's OWN drivers
(atooms.simulation.Simulation, atooms.optimization.Optimization) running our
backend unchanged.  GPU runs of the same classes are in test_gpu_backend.py.
"""
import copy
import os
import sys

import numpy
import pytest

from conftest import golden
import atooms_b200
from atooms_b200 import B200, EnergyMinimization, Interaction, System, Cell, Trajectory

# The reference's own drivers: from the read-only tree when it is there (development container),
# else from the unmodified install under baseline/_ref (recipe: tools/install_reference.py; it
# travels to the GPU box).  B2_NO_REFERENCE_TREE=1 forces the second route, for testing it here.
REF = '/root/reference'
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_TREE = os.path.isdir(os.path.join(REF, 'atooms')) and not os.environ.get('B2_NO_REFERENCE_TREE')
_INSTALLED = os.path.join(ROOT, 'baseline', '_ref')
REF_IMPORT = REF if _TREE else (_INSTALLED if os.path.isdir(os.path.join(_INSTALLED, 'atooms')) else None)
HAVE_REF = REF_IMPORT is not None
if HAVE_REF:
    if REF_IMPORT not in sys.path:
        sys.path.insert(0, REF_IMPORT)
    if not hasattr(numpy, 'set_string_function'):     # numpy >= 2 shim for the reference's xyz reader
        numpy.set_string_function = lambda *a, **k: None
needs_ref = pytest.mark.skipif(not HAVE_REF, reason='reference neither at /root/reference nor installed in baseline/_ref')

# data files of the reference the driver tests start from -> the golden fixture made from each
_DATA = {'lj_N1000_rho1.0.xyz': 'lj_N1000', 'lj_N256_rho1.0.xyz': 'lj_N256',
         'ka_N256_rho1.185_rumd.xyz.gz': 'ka_N256_rumd'}


def ref_data(name, tmp_path):
    """Path of a reference data file: the file itself when the tree is present, else an xyz file
    written from the golden fixture that tests/golden/make_golden.py made from it (same numbers)."""
    if _TREE:
        return os.path.join(REF, 'data', name)
    d = golden(_DATA[name])
    path = str(tmp_path / name.replace('.gz', ''))
    names = d['species_names'][d['species']]
    cols = [names.astype(str), d['position']] + ([d['velocity']] if 'velocity' in d else [])
    with open(path, 'w') as fh:
        fh.write('%d\nstep:0 columns:species,position%s dt:0.005 cell:%s\n'
                 % (len(names), ',velocity' if 'velocity' in d else '', ','.join(repr(float(x)) for x in d['side'])))
        for row in zip(*cols):
            fh.write(' '.join([row[0]] + [repr(float(x)) for block in row[1:] for x in block]) + '\n')
    return path


def make_system(name, lib, model, precision=64, velocities=True):
    d = golden(name)
    names = d['species_names'][d['species']]
    return System(d['position'], names, Cell(d['side'], d['center']),
                  velocity=d['velocity'] if velocities else None, mass=d['mass'],
                  model=model, precision=precision, library=lib)


def test_backend_surface(oracle_lib):
    """Members Simulation probes (atooms/backends/dryrun.py:13-40; simulation/core.py:89,225,263,277,443)."""
    bck = B200(make_system('lj_N1000', oracle_lib, 'lennard_jones'), 'lennard_jones', dt=0.005, library=oracle_lib)
    for name in ('system', 'run', 'rmsd', 'timestep', 'version', 'trajectory_class',
                 'write_checkpoint', 'read_checkpoint'):
        assert hasattr(bck, name)
    assert 'B200' in str(bck)
    assert bck.rmsd == 0.0
    s = bck.system
    assert len(s.particle) == 1000 and s.number_of_dimensions == 3
    assert s.distinct_species == ['A'] and s.composition == {'A': 1000}
    assert abs(s.density - 1.0) < 1e-12
    assert 'N=1000' in str(s)


def test_nve_pin_through_the_backend(oracle_lib, pins):
    """tests/test_lammps.py:20-40 with B200 in place of LAMMPS (unshifted LJ 2.5, dt 0.005)."""
    pin = pins['lj_N1000_nve_x0']
    model = atooms_b200.lennard_jones(2.5, shifted=False)
    bck = B200(make_system('lj_N1000', oracle_lib, model), model, dt=pin['dt'], library=oracle_lib)
    assert round(abs(bck.system.particle[0].position[0] - pin['value'][0]), 5) == 0
    for target in pin['value'][1:]:
        bck.run(10)
        assert round(abs(bck.system.particle[0].position[0] - target), 5) == 0
    x1 = bck.system.particle[3].position[0]
    bck.run(0)                                   # tests/test_lammps.py:123-142
    assert abs(bck.system.particle[3].position[0] - x1) < 1e-12
    assert bck.rmsd > 0.0


def test_model_registry_and_reference_schema(oracle_lib):
    """named models, and the reference's own description of a pair interaction
    (func / params / species / cutoff.scheme / cutoff.radius, trajectory/hdf5.py:21-43) in and out."""
    from atooms_b200 import models
    ka = models.get('kob_andersen')
    phis = ka.potentials()
    assert [p.species for p in phis] == [[1, 1], [1, 2], [2, 2]]
    assert phis[1].func == 'lennard_jones' and phis[1].params == {'epsilon': 1.5, 'sigma': 0.8}
    assert phis[1].cutoff.scheme == 'cs' and abs(phis[1].cutoff.radius - 2.0) < 1e-12
    back = models.from_potentials(phis)
    assert numpy.array_equal(back.eps, ka.eps) and numpy.array_equal(back.rcut, ka.rcut) and back.shifted
    as_dicts = [{'func': 'lennard_jones', 'params': {'epsilon': 1.0, 'sigma': 1.0}, 'species': [1, 1],
                 'cutoff': {'scheme': 'c', 'radius': 2.5}}]
    lj = models.from_potentials(as_dicts)
    assert lj.nsp == 1 and not lj.shifted and lj.rcut[0, 0] == 2.5
    assert models.from_potentials([dict(as_dicts[0], func='gaussian')]).tabulated     # other functions: lookup tables
    with pytest.raises(ValueError):
        models.get('no_such_model')
    inter = Interaction('kob_andersen', library=oracle_lib)
    assert inter.name == 'atomic' and len(inter.potential) == 3
    # WCA: the repulsive LJ core, u(sigma) = epsilon, nothing beyond 2^(1/6) sigma
    def pair_energy(r):
        s = System(numpy.array([[0.0, 0.0, 0.0], [r, 0.0, 0.0]]), ['A', 'A'], Cell([10.0] * 3, [0.0] * 3),
                   model='wca', library=oracle_lib)
        return B200(s, 'wca', library=oracle_lib).system.potential_energy()
    assert abs(pair_energy(1.0) - 1.0) < 1e-12
    assert pair_energy(1.13) == 0.0


def test_tabulated_potentials(oracle_lib):
    """any pair potential of the reference's schema (func / params / species / cutoff /
    npoints, atooms/trajectory/hdf5.py:21-43) through lookup tables."""
    from atooms_b200 import models
    ka = models.get('kob_andersen')
    records = [dict(func='lennard_jones', params=p.params, species=p.species,
                    cutoff={'scheme': 'cs', 'radius': p.cutoff.radius}, npoints=20000) for p in ka.potentials()]
    assert isinstance(models.from_potentials(records), models.PairModel)          # representable natively
    tab = models.TabulatedModel(records)
    assert tab.tabulated and tab.nsp == 2 and numpy.allclose(tab.rcut, ka.rcut) and tab.npoints == 20000
    native = B200(make_system('ka_N256_rumd', oracle_lib, ka), ka, library=oracle_lib).system
    tabbed = B200(make_system('ka_N256_rumd', oracle_lib, tab), tab, library=oracle_lib).system
    assert abs(tabbed.potential_energy() - native.potential_energy()) < 1e-6 * abs(native.potential_energy())
    assert abs(tabbed.virial(False) - native.virial(False)) < 1e-5 * abs(native.virial(False))
    # a model the 12-6 table cannot hold: Gaussian core + inverse power cross term, force-shifted cut-off
    mixed = [dict(func='gaussian', params={'epsilon': 2.0, 'sigma': 1.0}, species=[1, 1], cutoff={'scheme': 'cs', 'radius': 3.0}),
             dict(func='inverse_power', params={'epsilon': 1.0, 'sigma': 0.9, 'exponent': 9}, species=[1, 2],
                  cutoff={'scheme': 'cf', 'radius': 2.2}),
             dict(func=lambda r, a=1.0: a * numpy.exp(-r) / r, params={'a': 0.5}, species=[2, 2],
                  cutoff={'scheme': 'cs', 'radius': 2.5})]
    model = models.from_potentials(mixed)
    assert isinstance(model, models.TabulatedModel) and model.nsp == 2
    u, w = model.tables()
    assert u.shape == (2, 2, 20000) and abs(u[0, 0, -1]) < 1e-12 and abs(u[0, 1, -1]) < 1e-12 and abs(w[0, 1, -1]) < 1e-12
    k = 12345
    r = numpy.sqrt(k * 9.0 / 19999)
    assert abs(u[0, 0, k] - (2.0 * numpy.exp(-r * r) - 2.0 * numpy.exp(-9.0))) < 1e-12
    bck = B200(make_system('ka_N256_rumd', oracle_lib, model), model, dt=0.002, library=oracle_lib)
    s = bck.system
    e0 = s.total_energy(per_particle=True)
    f = s.interaction.forces if s.interaction.forces is not None else None
    bck.run(300)
    assert abs(s.total_energy(per_particle=True) - e0) < 2e-4          # NVE on the tabulated forces
    # numerical check of the forces against the tabulated energy: dU = -F.dx
    s2 = B200(make_system('ka_N256_rumd', oracle_lib, model), model, library=oracle_lib).system
    s2.compute_interaction('forces')
    frc = s2.interaction.forces.T.copy()
    x = s2.view('position')
    h = 1e-3            # (several table cells wide: u and w are interpolated independently)
    x[7, 1] += h
    ep = s2.potential_energy()
    x[7, 1] -= 2 * h
    em = s2.potential_energy()
    assert abs(-(ep - em) / (2 * h) - frc[7, 1]) < 1e-3 * max(1.0, abs(frc[7, 1]))
    with pytest.raises(ValueError):
        models.from_potentials([dict(mixed[0], func='no_such_function')])


def test_observers_accessors(oracle_lib, pins):
    """What write_thermo reads (simulation/observers.py:247-256) + the energy pins."""
    model = atooms_b200.lennard_jones(2.5, shifted=False)
    with pytest.warns(UserWarning, match='nearest image'):        # L = 4.76 < 2 r_cut: said, not silent
        s = B200(make_system('lj_fcc_N108', oracle_lib, model, velocities=False), model, library=oracle_lib).system
    assert s._engine.stats()['small_box']
    assert round(abs(s.potential_energy(per_particle=True) - pins['lj_fcc_N108_unshifted_U_per_particle']['value']), 7) == 0
    assert round(abs(s.pressure - pins['lj_fcc_N108_unshifted_pressure']['value']), 7) == 0   # T = 0
    assert round(abs(s.force_norm(per_particle=True) - 4.2e-11), 2) == 0
    assert s.temperature == 0.0 and s.kinetic_energy() == 0.0
    assert s.total_energy(per_particle=True, cache=True) == s.potential_energy(per_particle=True)
    s = B200(make_system('lj_N3_inline', oracle_lib, 'lennard_jones'), 'lennard_jones', library=oracle_lib).system
    assert round(abs(s.potential_energy() - pins['lj_N3_shifted_U']['value']), 4) == 0
    assert round(abs(s.temperature - 1.0), 7) == 0        # tests/test_backends.py:91-99
    ka3 = atooms_b200.PairModel(numpy.ones((2, 2)), numpy.ones((2, 2)), 2.5 * numpy.ones((2, 2)))
    s = B200(make_system('lj_N4_inline', oracle_lib, ka3), ka3, library=oracle_lib).system
    assert round(abs(s.temperature - 10.0 / 9.0), 7) == 0  # mass-aware, tests/test_backends.py:101-105


def test_sync_contract(oracle_lib):
    """Host views handed out are pushed before the next run; the device is pulled lazily."""
    bck = B200(make_system('ka_N256_rumd', oracle_lib, 'kob_andersen'), 'kob_andersen', library=oracle_lib)
    s = bck.system
    e0 = s.potential_energy()
    pos = s.view('position')                      # docs/simulations.org:302-308: modify views in place
    assert pos.shape == (256, 3) and s.dump('position', order='F', view=True).shape == (3, 256)
    pos[0, 0] += 0.05
    assert abs(s.potential_energy() - e0) > 1e-6
    pos[0, 0] -= 0.05
    assert abs(s.potential_energy() - e0) < 1e-9
    s.particle[1].position[1] += 0.05             # through the materialised Particle list
    assert abs(s.potential_energy() - e0) > 1e-6
    s.particle[1].position[1] -= 0.05
    # tracked mode: only what was written through numpy is pushed
    s.sync = 'tracked'
    s.potential_energy()                     # (pushes the edit above)
    calls = []
    update = s._engine.update
    s._engine.update = lambda position=None, velocity=None: (calls.append((position is not None, velocity is not None)),
                                                           update(position, velocity))[1]
    assert abs(s.potential_energy() - e0) < 1e-9 and calls == []          # nothing written: nothing pushed
    vel = s.view('velocity')
    vel *= 1.0                                                             # in-place operator on the view
    s.kinetic_energy()
    assert calls == [(False, True)]
    pos[3] = pos[3] + 0.01
    numpy.add(pos[4], 0.01, out=pos[4])
    assert abs(s.potential_energy() - e0) > 1e-6 and calls[-1] == (True, False)
    pos[3] -= 0.01
    pos[4] -= 0.01
    assert abs(s.potential_energy() - e0) < 1e-9
    raw = pos.view(numpy.ndarray)                                          # a writer numpy cannot see
    raw[5, 0] += 0.05
    assert abs(s.potential_energy() - e0) < 1e-9
    s.mark_modified(velocity=False)
    assert abs(s.potential_energy() - e0) > 1e-6
    raw[5, 0] -= 0.05
    s.mark_modified()
    s._engine.update = update
    s.sync = 'always'
    bck.run(5)
    assert not numpy.allclose(s.dump('position'), golden('ka_N256_rumd')['position'])
    # velocities: set_temperature hits the target exactly, scale_velocities runs on the device
    s.set_temperature(1.5, seed=1)
    assert abs(s.temperature - 1.5) < 1e-10
    assert numpy.abs(s.cm_velocity).max() < 1e-12      # a property, as in the reference (system.py:361-364)
    assert numpy.allclose(s.cm_position, (s.mass[:, None] * s.dump('position')).sum(axis=0) / s.mass.sum())
    s.scale_velocities(2.0)
    assert abs(s.temperature - 6.0) < 1e-9
    assert abs(numpy.sum(s.dump('velocity') ** 2) / (3 * 255) - 6.0) < 1e-9


def test_thermostat_switches_the_backend_to_nvt(oracle_lib, pins):
    """tests/test_lammps.py:80-98: `sim.system.thermostat = Thermostat(2.0)` turns the run into NVT."""
    from atooms_b200 import Thermostat
    model = atooms_b200.lennard_jones(2.5, shifted=False)
    bck = B200(make_system('lj_N1000', oracle_lib, model), model, dt=0.002, library=oracle_lib)
    bck.system.set_temperature(1.4, seed=1)
    bck.system.thermostat = Thermostat(2.0, relaxation_time=0.2)
    T = []
    for _ in range(8):
        bck.run(500)
        T.append(bck.system.temperature)
    assert round(abs(sum(T[3:]) / len(T[3:]) - pins['lj_N1000_nvt_temperature']['value']), 1) == 0
    assert abs(bck.system.pressure - (1000 * 2.0 + bck.system.virial(per_particle=False) / 3) / 1000.0) < 1e-9
    with pytest.raises(ValueError):
        B200(bck.system, model, thermostat='langevin', library=oracle_lib)


def test_barostat_switches_the_backend_to_npt(oracle_lib):
    """`system.thermostat` and `system.barostat` together turn the run into NPT, as in the reference's
    backends (atooms/backends/lammps.py:229-233): the pressure relaxes to the target, the box follows."""
    from atooms_b200 import Thermostat, Barostat
    model = atooms_b200.lennard_jones(2.5, shifted=True)
    bck = B200(make_system('lj_N1000', oracle_lib, model), model, dt=0.004, library=oracle_lib)
    s = bck.system
    s.set_temperature(2.0, seed=1)
    s.thermostat = Thermostat(2.0, relaxation_time=0.2)
    bck.run(300)
    p_nvt, v0 = s.pressure, s.cell.volume
    assert p_nvt > 10.0 and abs(s.density - 1.0) < 1e-12
    s.barostat = Barostat(4.0, relaxation_time=0.5)
    P = []
    for _ in range(24):
        bck.run(100)
        P.append(s.pressure)
    assert abs(numpy.mean(P[12:]) - 4.0) < 0.4                    # <P> -> target
    assert s.cell.volume > 1.2 * v0 and abs(s.density - 1000 / s.cell.volume) < 1e-12
    pos = s.dump('position')
    assert numpy.abs(pos).max() <= 0.5 * s.cell.side[0] * (1 + 1e-9)   # the particles followed the box
    # the affine move itself: fractional coordinates kept, energy is that of the scaled configuration
    frac = s.dump('position') / s.cell.side
    s.set_density(0.9)
    assert abs(s.density - 0.9) < 1e-12 and numpy.abs(s.dump('position') / s.cell.side - frac).max() < 1e-9
    ref = System(s.dump('position'), s.species, Cell(s.cell.side), model=model, library=oracle_lib)
    ref.interaction = Interaction(model, library=oracle_lib)
    assert abs(s.potential_energy() - ref.potential_energy()) < 1e-6 * abs(ref.potential_energy())
    s.thermostat = None
    with pytest.raises(ValueError):
        bck.run(10)                                              # a barostat alone is not an ensemble we run


def test_small_systems_run_in_one_launch(oracle_lib):
    """configs[0] (LJ N = 256): B200.run hands NVE runs of small systems to the one-block replica kernel
    (one launch per run() instead of one per step); the trajectory, the box-crossing counters and the rmsd
    are those of the step-by-step path."""
    d = golden('lj_N256')
    vel = numpy.random.default_rng(4).normal(0.0, 1.4, size=(256, 3))
    vel -= vel.mean(axis=0)

    def backend(small):
        s = System(d['position'], d['species_names'][d['species']], Cell(d['side']), velocity=vel, model='lennard_jones',
                   library=oracle_lib)
        return B200(s, 'lennard_jones', dt=0.004, library=oracle_lib, small_system_kernel=small)
    fast, slow = backend(True), backend(False)
    assert fast._small_ok(100) and not slow._small_ok(100) and not fast._small_ok(5)
    for _ in range(3):
        fast.run(100)
        slow.run(100)
    assert fast._small is not None and slow._small is None
    assert numpy.abs(fast.system.dump('position') - slow.system.dump('position')).max() < 1e-8
    assert numpy.abs(fast.system.dump('velocity') - slow.system.dump('velocity')).max() < 1e-7
    assert numpy.array_equal(fast.system.image, slow.system.image) and numpy.abs(slow.system.image).max() >= 1
    assert abs(fast.rmsd - slow.rmsd) < 1e-8 and fast.steps == 300
    assert abs(fast.system.total_energy() - slow.system.total_energy()) < 1e-6


def test_deepcopy_and_reassign(oracle_lib):
    """tests/test_rumd.py:178-194."""
    bck = B200(make_system('ka_N256_rumd', oracle_lib, 'kob_andersen'), 'kob_andersen', library=oracle_lib)
    pos0 = bck.system.particle[0].position[0]
    s = copy.deepcopy(bck.system)
    assert round(abs(bck.system.potential_energy(per_particle=True) - (-6.4014371344819665)), 3) == 0
    bck.run(100)
    pos1 = bck.system.particle[0].position[0]
    bck.system = s
    assert bck.system is s
    bck.run(100)
    pos2 = bck.system.particle[0].position[0]
    assert abs(pos1 - pos0) > 1e-2
    assert abs(pos1 - pos2) < 1e-4


def test_checkpoint_roundtrip(oracle_lib, tmp_path):
    bck = B200(make_system('ka_N256_rumd', oracle_lib, 'kob_andersen'), 'kob_andersen', library=oracle_lib)
    bck.run(10)
    path = str(tmp_path / 'run')
    bck.write_checkpoint(path)
    ref = bck.system.dump('position').copy()
    bck.run(10)
    after = bck.system.dump('position').copy()
    bck.read_checkpoint(path)
    assert numpy.array_equal(bck.system.dump('position'), ref)
    assert bck.steps == 10
    bck.run(10)
    assert numpy.abs(bck.system.dump('position') - after).max() < 1e-9


def test_checkpoint_restores_images_rmsd_and_thermostat(oracle_lib, tmp_path):
    """run(a), checkpoint, restore into a FRESH backend, run(b) == run(a + b): unfolded
    coordinates (box-crossing counters), backend.rmsd (targeted-rmsd runs) and the Nose-Hoover friction."""
    from atooms_b200 import Thermostat

    def fresh():
        bck = B200(make_system('ka_N256_rumd', oracle_lib, 'kob_andersen'), 'kob_andersen', dt=0.004, library=oracle_lib)
        bck.system.set_temperature(2.0, seed=3)
        bck.system.thermostat = Thermostat(1.5, relaxation_time=0.1)
        return bck
    one = fresh()
    one.run(300)
    two = fresh()
    two.run(120)
    path = str(tmp_path / 'restart')
    two.write_checkpoint(path)
    assert numpy.abs(two.system.image).max() >= 1                 # particles did cross the box
    three = fresh()                                               # a new process would start like this
    three.read_checkpoint(path)
    assert three.steps == 120 and abs(three.rmsd - two.rmsd) < 1e-12
    three.run(180)
    assert numpy.abs(three.system.dump('position') - one.system.dump('position')).max() < 1e-8
    assert numpy.array_equal(three.system.image, one.system.image)
    assert abs(three.rmsd - one.rmsd) < 1e-9 and abs(three.system.temperature - one.system.temperature) < 1e-9
    other = B200(make_system('lj_N256', oracle_lib, 'lennard_jones'), 'lennard_jones', library=oracle_lib)
    with pytest.raises(ValueError):
        other.read_checkpoint(path)                               # not the same particles


def test_minimisation_backend(oracle_lib, pins):
    """tests/test_optimization.py:20-35 without the driver: reached_steps semantics."""
    pin = pins['lj_N256_minimised']
    bck = EnergyMinimization(make_system('lj_N256', oracle_lib, 'lennard_jones', velocities=False),
                             'lennard_jones', method='fire', ftol=1e-4, library=oracle_lib)
    assert bck.reached_steps is True              # lammps.py:307-309
    bck.run(50)
    assert bck.reached_steps is True and bck.iterations == 50     # stopped by the budget
    bck.run(100000)
    assert bck.reached_steps is False             # stopped by the tolerance
    assert bck.system.potential_energy(per_particle=True) < pin['U_per_particle_below']
    assert bck.system.force_norm_square(per_particle=True) < pin['fn2_per_particle_below']
    with pytest.raises(ValueError):
        EnergyMinimization(bck.system, 'lennard_jones', method='cg', library=oracle_lib)


def test_generic_interaction_protocol(oracle_lib):
    """Interaction.compute(observable, **variables) with host arrays (interaction.py:57-82)."""
    d = golden('ka_N256_rumd')

    class P(object):
        def __init__(self, species, mass):
            self.species, self.mass = species, mass
    particle = [P(str(s), m) for s, m in zip(d['species'], d['mass'])]
    position = numpy.asfortranarray(d['position'].T)          # (3, N) F order, as system.dump(order='F')
    inter = Interaction('kob_andersen', library=oracle_lib)
    assert inter.variables == {'position': 'particle.position', 'particle': 'particle', 'cell': 'cell'}
    assert inter.order == 'F' and inter.energy is None and inter.forces is None
    inter.compute('energy', position, particle, Cell(d['side']))
    assert round(abs(inter.energy / 256 - (-6.4014371344819665)), 3) == 0 and inter.forces is None
    inter.compute('forces', position, particle, Cell(d['side']))
    assert inter.forces.shape == (3, 256) and inter.forces.flags.f_contiguous
    fref = numpy.c_[d['fx'], d['fy'], d['fz']]
    assert numpy.abs(inter.forces.T - fref).max() / numpy.abs(fref).max() < 1e-5
    inter.compute('stress', position, particle, Cell(d['side']))
    assert inter.stress.shape == (3, 3) and abs(numpy.trace(inter.stress) - inter.virial) < 1e-8
    with pytest.raises(AssertionError):
        inter.compute('bogus', position, particle, Cell(d['side']))
    # hessian (interaction.py:33,76-82): (d, N, d, N), symmetric, rows sum to zero, and equal to minus the
    # derivative of the forces (central differences on one coordinate)
    inter.compute('hessian', position, particle, Cell(d['side']))
    H = inter.hessian
    assert H.shape == (3, 256, 3, 256)
    M = H.reshape(768, 768)
    assert numpy.abs(M - M.T).max() < 1e-9 * numpy.abs(M).max() and numpy.abs(H.sum(axis=3)).max() < 1e-8 * numpy.abs(M).max()
    h, (i, a) = 1e-5, (17, 2)
    shifted = position.copy(order='F')
    shifted[a, i] += h
    inter.compute('forces', shifted, particle, Cell(d['side']))
    fp = inter.forces.copy()
    shifted[a, i] -= 2 * h
    inter.compute('forces', shifted, particle, Cell(d['side']))
    numeric = -(fp - inter.forces) / (2 * h)                       # d2U / dx_jb dx_ia = -dF_jb / dx_ia
    assert numpy.abs(numeric - H[:, :, a, i]).max() < 1e-5 * numpy.abs(H[:, :, a, i]).max()
    total = inter + inter                                      # interaction.py:37-55
    assert abs(total.energy - 2 * inter.energy) < 1e-9


def test_trajectory_writer(oracle_lib, tmp_path):
    s = make_system('ka_N256_rumd', oracle_lib, 'kob_andersen')
    path = str(tmp_path / 'out.xyz')
    with Trajectory(path, 'w') as th:
        th.timestep = 0.002
        th.precision = 9
        th.variables = ['species', 'position', 'velocity']
        th.write(s, 0)
        th.write(s, 10)
    lines = open(path).read().splitlines()
    assert lines[0] == '256' and lines[1].startswith('step:0 columns:species,position,velocity dt:0.002 cell:6.0,6.0,6.0')
    assert len(lines) == 2 * 258
    if HAVE_REF:
        from atooms.trajectory import TrajectoryXYZ
        th = TrajectoryXYZ(path)
        assert th.steps == [0, 10]
        back = th[1]
        th.close()
        assert numpy.abs(numpy.array([p.position for p in back.particle]) - s.dump('position')).max() < 1e-8
        assert [p.species for p in back.particle] == list(s.species)


def test_trajectory_native_formatter(oracle_lib, tmp_path, monkeypatch):
    """b2_format_xyz (frame text on all host cores) writes byte for byte what the `%` formatter writes,
    in both dialects, with and without velocities, and answers its size query."""
    from atooms_b200 import TrajectoryRUMD, capi
    s = make_system('ka_N256_rumd', oracle_lib, 'kob_andersen')
    s.view('position')[7] = [-2.9999999996, 1e-12, 2.5]
    monkeypatch.setattr(capi, 'load', lambda path=None: oracle_lib)
    calls, native = [], [True]
    real_format, real_write = oracle_lib.format_xyz, Trajectory._write_native
    monkeypatch.setattr(oracle_lib, 'format_xyz', lambda *a, **k: calls.append(1) or real_format(*a, **k), raising=False)
    monkeypatch.setattr(Trajectory, '_write_native', lambda self, c, b: native[0] and real_write(self, c, b))
    for cls, variables, precision in ((Trajectory, ['species', 'position'], 6), (Trajectory, ['species', 'pos', 'vel'], 12),
                                      (TrajectoryRUMD, ['species', 'position', 'velocity'], 9)):
        texts = []
        for native[0] in (True, False):
            path = str(tmp_path / ('f%d.xyz' % native[0]))
            with cls(path, 'w') as th:
                th.timestep, th.precision, th.variables = 0.002, precision, variables
                th.write(s, 0)
                th.write(s, 3)
            texts.append(open(path, 'rb').read())
        assert texts[0] == texts[1] and len(texts[0]) > 256 * 20
    native[0] = True
    assert len(calls) == 6                                  # the native path did run, twice per file
    # columns it does not cover fall back to the block formatter
    with Trajectory(str(tmp_path / 'r.xyz'), 'w') as th:
        th.variables = ['species', 'position', 'radius']
        th.write(s, 0)
    assert len(calls) == 6
    # errors: capacity too small, bad precision
    import ctypes
    ids = numpy.zeros(4, dtype=numpy.int32)
    pos = numpy.zeros((4, 3))
    names = (ctypes.c_char_p * 1)(b'A')
    need = ctypes.c_int64()
    dp = lambda a: a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))
    ip = ids.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))
    assert oracle_lib.dll.b2_format_xyz(None, 0, ctypes.byref(need), 4, ip, 1, names, dp(pos), None, 6) == 0
    assert need.value >= 4 * len('A 0.000000 0.000000 0.000000\n')
    buf = ctypes.create_string_buffer(8)
    assert oracle_lib.dll.b2_format_xyz(buf, 8, ctypes.byref(need), 4, ip, 1, names, dp(pos), None, 6) != 0
    assert oracle_lib.dll.b2_format_xyz(None, 0, ctypes.byref(need), 4, ip, 1, names, dp(pos), None, 40) != 0


def test_trajectory_fast_path(oracle_lib, tmp_path):
    """block-formatted writer, read mode, RUMD dialect, in-memory snapshots — none of
    them builds Particle objects."""
    from atooms_b200 import TrajectoryRUMD, TrajectoryRam, Snapshot
    s = make_system('ka_N256_rumd', oracle_lib, 'kob_andersen')
    path = str(tmp_path / 'fast.xyz')
    with Trajectory(path, 'w') as th:
        th.timestep, th.precision, th.variables = 0.002, 9, ['species', 'position', 'velocity']
        th.write(s, 0)
        th.write(s, 25)
    with Trajectory(path) as th:                       # read mode
        assert th.steps == [0, 25] and len(th) == 2 and th.timestep == 0.002
        back = th[-1]
        assert isinstance(back, Snapshot) and len(back) == 256 and back.step == 25
        assert numpy.abs(back.position - s.dump('position')).max() < 1e-9
        assert numpy.abs(back.velocity - s.dump('velocity')).max() < 1e-9
        assert list(back.species) == list(s.species) and numpy.allclose(back.cell.side, s.cell.side)
        assert back.particle[3].species == s.species[3] and len(back.particle) == 256
        assert [f.step for f in th] == [0, 25]
    # a frame read back starts a backend like any other system
    again = B200(back, 'kob_andersen', library=oracle_lib)
    assert abs(again.system.potential_energy() - B200(s, 'kob_andersen', library=oracle_lib).system.potential_energy()) < 1e-5
    # RUMD dialect: integer types, ioformat header
    rpath = str(tmp_path / 'conf.xyz')
    with TrajectoryRUMD(rpath, 'w') as th:
        th.timestep = 0.005
        th.write(s, 7)
    head = open(rpath).read().splitlines()[:3]
    assert head[0] == '256' and head[1].startswith('ioformat=1 dt=0.005 timeStepIndex=7 boxLengths=6.0')
    assert 'numTypes=2 mass=1,1 columns=type,x,y,z,vx,vy,vz' in head[1] and head[2].split()[0] in ('0', '1')
    with TrajectoryRUMD(rpath) as th:
        b = th[0]
        assert th.steps == [7] and numpy.abs(b.position - s.dump('position')).max() < 1e-8
        assert list(b.species.astype(int)) == list(s._ids)
    if HAVE_REF:                                        # and the reference's own readers parse both files
        from atooms.trajectory import TrajectoryXYZ, TrajectoryRUMD as RefRUMD
        for cls, p, species in ((TrajectoryXYZ, path, s.species[5]), (RefRUMD, rpath, str(s._ids[5]))):
            th = cls(p)
            ref = th[0]
            th.close()
            assert numpy.abs(ref.particle[5].position - s.dump('position')[5]).max() < 1e-8
            assert str(ref.particle[5].species) == species
    # in-memory trajectory (ram.py's interface)
    ram = TrajectoryRam()
    ram.write(s, 0)
    ram.write(s, 10)
    ram[1] = s                                           # rewrites step 10
    ram[5] = s                                           # appends step 11
    assert ram.steps == [0, 10, 11] and len(ram) == 3
    f = ram[0]
    f.position[0, 0] += 1.0
    assert ram[0].position[0, 0] != f.position[0, 0]    # copies, as TrajectoryRam returns


# ---- the reference's own drivers on top of our backend -------------------------------
@needs_ref
def test_reference_simulation_driver(oracle_lib, pins, tmp_path):
    from atooms.simulation import Simulation, write_thermo, Scheduler
    pin = pins['lj_N1000_nve_x0']
    model = atooms_b200.lennard_jones(2.5, shifted=False)
    bck = B200(ref_data('lj_N1000_rho1.0.xyz', tmp_path), model, dt=pin['dt'], library=oracle_lib)
    out = str(tmp_path / 'sim')
    sim = Simulation(bck, output_path=out, checkpoint_interval=10, enable_speedometer=False)
    sim.add(write_thermo, Scheduler(5))
    assert sim.system is bck.system
    sim.run(10)
    assert round(abs(sim.system.particle[0].position[0] - pin['value'][1]), 5) == 0
    sim.run(10)
    assert round(abs(sim.system.particle[0].position[0] - pin['value'][2]), 5) == 0
    assert sim.current_step == 20
    thermo = numpy.loadtxt(out + '.thermo')
    assert thermo.shape[0] == 5 and list(thermo[:, 0]) == [0, 5, 10, 15, 20]
    assert numpy.abs(thermo[:, 4] - thermo[0, 4]).max() < 2e-2          # total energy (unshifted cutoff, dt 0.005)
    assert os.path.exists(out + '.chk.npz') and os.path.exists(out + '.chk.step')
    # restart from the checkpoint (simulation/core.py:240-273)
    bck2 = B200(ref_data('lj_N1000_rho1.0.xyz', tmp_path), model, dt=pin['dt'], library=oracle_lib)
    sim2 = Simulation(bck2, output_path=out, restart=True, enable_speedometer=False)
    sim2.run(30)
    assert sim2.current_step == 30
    bck.run(10)
    assert numpy.abs(bck2.system.dump('position') - bck.system.dump('position')).max() < 1e-9


@needs_ref
def test_integration_shim(oracle_lib, pins, tmp_path):
    """integration/b200.py under the name a maintainer gives it, atooms.backends.b200, driven by the
    reference's Simulation."""
    import importlib.util
    from atooms.simulation import Simulation
    spec = importlib.util.spec_from_file_location('atooms.backends.b200', os.path.join(ROOT, 'integration', 'b200.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    assert mod.B200 is B200 and mod.EnergyMinimization is EnergyMinimization
    pin = pins['lj_N1000_nve_x0']
    bck = mod.B200(ref_data('lj_N1000_rho1.0.xyz', tmp_path), atooms_b200.lennard_jones(2.5, shifted=False),
                   dt=pin['dt'], library=oracle_lib)
    Simulation(bck, enable_speedometer=False).run(10)
    assert round(abs(bck.system.particle[0].position[0] - pin['value'][1]), 5) == 0


@needs_ref
def test_reference_optimization_driver(oracle_lib, pins, tmp_path):
    """tests/test_optimization.py:20-35 with our backend in place of LAMMPS."""
    from atooms.optimization import Optimization
    pin = pins['lj_N256_minimised']
    bck = EnergyMinimization(ref_data('lj_N256_rho1.0.xyz', tmp_path), 'lennard_jones', library=oracle_lib)
    opt = Optimization(bck)
    opt.tolerance = 1e-4
    opt.run()
    assert bck.system.potential_energy(per_particle=True) < pin['U_per_particle_below']
    assert bck.system.force_norm_square(per_particle=True) < pin['fn2_per_particle_below']
    assert bck.reached_steps is False


@needs_ref
def test_interaction_with_reference_system(oracle_lib, pins, tmp_path):
    """system.interaction = Interaction(...) on a plain atooms System (system.py:245-335)."""
    from atooms.trajectory import TrajectoryXYZ
    th = TrajectoryXYZ(ref_data('lj_N1000_rho1.0.xyz', tmp_path))
    system = th[0]
    th.close()
    system.interaction = Interaction(atooms_b200.lennard_jones(2.5, shifted=False), library=oracle_lib)
    assert float('%.7f' % system.potential_energy(per_particle=True)) == pins['lj_N1000_unshifted_U_per_particle']['value']
    assert system.interaction.forces.shape == (3, 1000)
    assert system.force_norm_square(per_particle=True) > 0
    assert abs(system.virial(per_particle=False) - system.interaction.virial) < 1e-9
    # the reference's composite Interaction fans out to terms and sums them (interaction.py:85-115)
    from atooms.system.interaction import Interaction as Composite
    both = Composite(Interaction(atooms_b200.lennard_jones(2.5, False), library=oracle_lib),
                     Interaction(atooms_b200.lennard_jones(2.5, False), library=oracle_lib))
    e1 = system.interaction.energy
    system.interaction = both
    assert abs(system.potential_energy() - 2 * e1) < 1e-8


@needs_ref
def test_neighbors_export_round_trips_through_the_reference_reader(oracle_lib, tmp_path):
    """the engine's Verlet list, written in the reference's neighbours format
    (atooms/trajectory/xyz.py:484-506), comes back through the reference's own TrajectoryNeighbors."""
    from atooms.trajectory import TrajectoryXYZ
    from atooms.trajectory.xyz import TrajectoryNeighbors as RefNeighbors
    from oracle import oracle_np as onp
    th = TrajectoryXYZ(ref_data('lj_N256_rho1.0.xyz', tmp_path))
    system = th[0]
    th.close()
    inter = Interaction(atooms_b200.lennard_jones(2.5, True), skin=0.3, library=oracle_lib)
    system.interaction = inter
    system.potential_energy()
    rows = inter.neighbors()
    pos = numpy.array([p.position for p in system.particle])
    brute = onp.neighbor_list(pos, numpy.zeros(len(pos), dtype=numpy.int32), numpy.array(system.cell.side),
                              onp.lennard_jones(2.5, True), 0.3)
    assert len(rows) == 256 and all(numpy.array_equal(a, b) for a, b in zip(rows, brute))
    path = str(tmp_path / 'neighbors.xyz')
    with atooms_b200.TrajectoryNeighbors(path, 'w') as out:
        out.timestep = 0.002
        system.neighbors = rows
        out.write(system, 7)
    with RefNeighbors(path) as back:                  # offset 1 on both sides
        assert back.steps == [7] and back.timestep == 0.002
        got = back[0]
    for a, p in zip(rows, got.particle):
        assert list(a) == [int(j) for j in p.neighbors]


@needs_ref
def test_command_line_driver(oracle_lib, tmp_path, monkeypatch):
    """bin/md_b200.py, the bin/md.py look-alike (bin/md.py:12-81).  The engine is swapped for
    the oracle by patching the loader IN THIS TEST ONLY; the script itself has no such switch."""
    import argparse
    import importlib.util
    from atooms_b200 import capi
    monkeypatch.setattr(capi, 'load', lambda: oracle_lib)
    spec = importlib.util.spec_from_file_location('md_b200', os.path.join(hp_root(), 'bin', 'md_b200.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    out = str(tmp_path / 'run')
    params = argparse.Namespace(T=0.8, dt=0.002, input_file=ref_data('ka_N256_rho1.185_rumd.xyz.gz', tmp_path),
                                integrator='nve', model='kob_andersen', tau=0.2, skin=0.3, precision='fp64',
                                device=0, steps=60, thermo_interval=20, config_interval=30, checkpoint_interval=0,
                                restart=False, verbose=False, debug=False, output_dir=out)
    mod.main(params)
    thermo = numpy.loadtxt(os.path.join(out, 'trajectory.thermo'))
    assert list(thermo[:, 0]) == [0, 20, 40, 60]
    lines = open(os.path.join(out, 'trajectory')).read().splitlines()
    assert lines[0] == '256' and lines[1].startswith('step:0 ') and len(lines) == 3 * 258


def hp_root():
    return os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@needs_ref
def test_reference_trajectory_callbacks(oracle_lib, tmp_path):
    """the reference's trajectory callbacks (atooms/trajectory/decorators.py:24-136)
    work on particle objects — in-place edits of p.position / p.velocity, assignments to p.species, and a new
    list assigned to system.particle.  They must do to the array-backed System (device copy included) and to a
    Snapshot read back from a file what they do to the reference's System."""
    from atooms.trajectory import decorators as deco
    from atooms_b200 import Snapshot

    def backend():
        return B200(make_system('ka_N256_rumd', oracle_lib, 'kob_andersen'), 'kob_andersen', library=oracle_lib)
    u0 = backend().system.potential_energy(per_particle=True)
    # rigid moves and folds leave the energy alone; the device sees what the callback wrote
    s = backend().system
    deco.fold(deco.center(s))
    assert abs(s.potential_energy(per_particle=True) - u0) < 1e-10
    assert numpy.all(numpy.abs(s.dump('position')) <= 0.5 * s.cell.side + 1e-12)
    deco.fix_cm(s)
    assert numpy.abs(s.cm_position).max() < 1e-12
    deco.set_temperature(s, 0.7)
    assert abs(s.temperature - 0.7) < 0.1 and numpy.abs(s.cm_velocity).max() < 1e-12
    s = backend().system
    deco.set_density(s, 1.0)
    assert abs(s.density - 1.0) < 1e-12 and abs(s.potential_energy(per_particle=True) - u0) > 1e-3
    # relabelling: A, B -> 1, 2 -> 0, 1 -> A, B keeps every particle on its row of the pair table
    s = backend().system
    for layout, names in (('F', ['1', '2']), ('C', ['0', '1']), ('A', ['A', 'B'])):
        deco.change_species(s, layout)
        assert s.distinct_species == names
        assert abs(s.potential_energy(per_particle=True) - u0) < 1e-12
    # a new particle list: sorted by species, then the minority species alone (still species 1 of the model)
    ids = s._ids.copy()
    deco.sort(s)
    assert numpy.array_equal(s._ids, numpy.sort(ids)) and abs(s.potential_energy(per_particle=True) - u0) < 1e-10
    deco.filter_species(s, 'B')
    d = golden('ka_N256_rumd')
    only_b = System(d['position'][d['species'] == 1], ['B'] * int((d['species'] == 1).sum()), Cell(d['side'], d['center']),
                    model='kob_andersen', library=oracle_lib, species_order=['A', 'B'])
    ref_b = B200(only_b, 'kob_andersen', library=oracle_lib).system.potential_energy(per_particle=True)
    assert len(s) == len(only_b) and set(s._ids.tolist()) == {1}
    assert abs(s.potential_energy(per_particle=True) - ref_b) < 1e-10
    # a species the model has no row for cannot be taken over
    s.particle[0].species = 'Q'
    with pytest.raises(ValueError):
        s.potential_energy()
    # the same callbacks on frames read back from a file (array snapshots)
    path = str(tmp_path / 'cb.xyz')
    full = backend().system
    with Trajectory(path, 'w') as th:
        th.variables = ['species', 'position', 'velocity']
        th.write(full, 0)
    with Trajectory(path) as th:
        th.add_callback(deco.change_species, 'F')
        th.add_callback(deco.filter_species, '2')
        th.add_callback(deco.set_density, 0.5)
        frame = th[0]
    assert isinstance(frame, Snapshot) and len(frame) == len(only_b) and frame.distinct_species == ['2']
    assert abs(frame.density - 0.5) < 1e-12
    x = (len(only_b) / numpy.prod(d['side']) / 0.5) ** (1.0 / 3)
    assert numpy.abs(frame.position - x * d['position'][d['species'] == 1]).max() < 1e-5


@needs_ref
def test_reference_class_decorators_on_the_trajectory(oracle_lib, tmp_path):
    """`Sliced` and `Unfolded` (atooms/trajectory/decorators.py:147-241) subclass whatever trajectory they are given
    and override its read_init / read_system: the reader here follows that template.  Unfolding the frames must
    reproduce what the engine's own box-crossing counters say."""
    from atooms.trajectory.decorators import Sliced, Unfolded
    s = make_system('ka_N256_rumd', oracle_lib, 'kob_andersen')
    bck = B200(s, 'kob_andersen', dt=0.004, library=oracle_lib)
    bck.system.set_temperature(3.0, seed=5)
    path = str(tmp_path / 'u.xyz')
    unfolded = []
    with Trajectory(path, 'w') as th:
        th.precision = 12
        for _ in range(8):
            th.write(bck.system, bck.steps)
            unfolded.append(bck.system.dump('position') + bck.system.image * bck.system.cell.side)
            bck.run(100)
    assert numpy.abs(bck.system.image).sum() > 20              # particles did leave the cell
    with Trajectory(path) as th:
        sl = Sliced(th, slice(1, 7, 2))
        assert sl.steps == [100, 300, 500] and len(sl) == 3
        assert [sl[i].step for i in range(3)] == [100, 300, 500] and sl[-1].step == 500
        un = Unfolded(th)
        first = un[0].position.copy()        # (folded as written: compare displacements from it)
        for k in range(1, 8):
            assert numpy.abs((un[k].position - first) - (unfolded[k] - unfolded[0])).max() < 1e-9
        with pytest.raises(ValueError):
            un[2]                                            # no unfolding backwards (decorators.py:204-206)


@needs_ref
def test_system_members_of_the_reference(oracle_lib):
    """The members of atooms.system.System that change the particles themselves (system.py:160-162, 210-220,
    381-396, 569-628): replicate, set_composition, species_layout, packing_fraction — on the arrays, with the
    device copy following, against the reference's own System doing the same particle by particle."""
    from atooms.system import System as RefSystem, Particle as RefParticle, Cell as RefCell
    d = golden('ka_N256_rumd')
    names = d['species_names'][d['species']]
    s = B200(make_system('ka_N256_rumd', oracle_lib, 'kob_andersen'), 'kob_andersen', library=oracle_lib).system
    ref = RefSystem([RefParticle(position=x.copy(), velocity=v.copy(), species=n, mass=m)
                     for x, v, n, m in zip(d['position'], d['velocity'], names, d['mass'])], RefCell(d['side'].copy()))
    u0, k0 = s.potential_energy(per_particle=True), s.kinetic_energy(per_particle=True)
    assert abs(s.packing_fraction - ref.packing_fraction) < 1e-12
    # names only: nothing moves on the device
    for layout, expect in (('F', ['1', '2']), ('C', ['0', '1']), ('A', ['A', 'B'])):
        s.species_layout = layout
        ref.species_layout = layout
        assert s.species_layout == ref.species_layout == layout and s.distinct_species == expect
        assert s.particle[7].species == ref.particle[7].species
        assert abs(s.potential_energy(per_particle=True) - u0) < 1e-12
    # two periodic copies along x: same energy per particle, the reference's coordinates
    s.replicate(2, 0)
    ref.replicate(2, 0)
    assert len(s) == 512 and numpy.allclose(s.cell.side, ref.cell.side)
    theirs = numpy.array([p.position for p in ref.particle])
    assert numpy.abs(numpy.sort(theirs, axis=0) - numpy.sort(s.dump('position'), axis=0)).max() < 1e-12
    assert abs(s.potential_energy(per_particle=True) - u0) < 1e-12 and abs(s.kinetic_energy(per_particle=True) - k0) < 1e-12
    assert abs(s.packing_fraction - ref.packing_fraction) < 1e-12
    with pytest.raises(ValueError):
        s.replicate(1, 0)
    # a new composition: the device copy follows (the energy changes, the run goes on)
    s.set_composition({'A': 500, 'B': 12})
    assert s.composition == {'A': 500, 'B': 12} and abs(s.concentration['B'] - 12 / 512) < 1e-12
    assert abs(s.potential_energy(per_particle=True) - u0) > 1e-3
    with pytest.raises(ValueError):
        s.set_composition({'A': 3})
    with pytest.raises(ValueError):
        s.set_composition({'A': 500, 'B': 6, 'C': 6})       # the model has two species
    with pytest.raises(NotImplementedError):
        s.show()


def test_trajectory_base_members(oracle_lib, tmp_path):
    """What analysis code written against atooms/trajectory/base.py:212-527 uses besides read and write: a default
    step, append, the time axis, the block size of exponentially spaced frames, the read_* hooks, copy."""
    from atooms_b200 import TrajectoryRam
    from atooms_b200.trajectory import block_size_of
    s = make_system('ka_N256_rumd', oracle_lib, 'kob_andersen')
    path = str(tmp_path / 'b.xyz')
    steps = [0, 1, 2, 4, 8, 16, 17, 18, 20, 24, 32]
    with Trajectory(path, 'w') as th:
        th.timestep = 0.002
        for step in steps:
            th.write(s, step)
        th.append(s)                                         # one past the last step
        assert th.steps == steps + [33]
    with Trajectory(path) as th:
        assert th.read_len() == 12 and th.read_steps() == steps + [33] and th.read_timestep() == 0.002
        assert th.times[4] == 8 * 0.002 and abs(th.total_time - 33 * 0.002) < 1e-15
        assert th.block_size == 5 and not th.grandcanonical
        out = th.copy(cls=TrajectoryRam, steps=[0, 8, 32])
        assert out.steps == [0, 8, 32] and numpy.abs(out[1].position - th[4].position).max() == 0.0
        twin = th.copy(fout=str(tmp_path / 'c.xyz'))
    with Trajectory(twin.filename) as th:
        assert th.steps == steps + [33] and th.timestep == 0.002
    ram = TrajectoryRam()
    ram.write(s)
    ram.append(s)
    assert ram.steps == [0, 1] and ram.block_size == 1
    for seq, size in (([0, 1, 2, 3], 1), ([1, 2, 4, 11, 12, 14, 21, 22, 24], 3), ([0, 1, 3, 7], 4), ([5], 1),
                      ([0, 1, 10, 11, 20, 21, 30], 2)):
        assert block_size_of(seq) == size
        if HAVE_REF:
            from atooms.trajectory.utils import get_block_size
            assert get_block_size(seq) == size


def test_standalone_particle_and_reservoirs():
    """The stand-ins used when atooms itself is not importable carry the members of the originals
    (atooms/system/particle.py:32-107, reservoir.py:22-40)."""
    from atooms_b200.system import Particle as Own, Cell as OwnCell
    from atooms_b200 import Thermostat
    cell = OwnCell([4.0, 4.0, 4.0])
    a = Own(numpy.array([1.9, 0.0, -1.9]), numpy.zeros(3), 'A', mass=2.0, radius=0.6)
    b = Own(numpy.array([-1.9, 0.5, 1.9]), numpy.zeros(3), 'B')
    assert a.diameter == 1.2
    assert numpy.allclose(a.distance(b), [3.8, -0.5, -3.8]) and numpy.allclose(a.distance(b, cell), [-0.2, -0.5, 0.2])
    image = a.nearest_image(b, cell, copy=True)
    assert numpy.allclose(image.position, [-2.1, 0.0, 2.1]) and numpy.allclose(a.position, [1.9, 0.0, -1.9])
    a.nearest_image(b, cell)
    assert numpy.allclose(a.position, [-2.1, 0.0, 2.1])
    a.maxwellian(1.0)
    assert numpy.abs(a.velocity).max() > 0
    if HAVE_REF:
        from atooms.system.particle import Particle as Ref
        from atooms.system.cell import Cell as RefCell
        ra, rb = Ref(position=[1.9, 0.0, -1.9], radius=0.6), Ref(position=[-1.9, 0.5, 1.9])
        assert numpy.allclose(ra.distance(rb, RefCell([4.0, 4.0, 4.0])), [-0.2, -0.5, 0.2]) and ra.diameter == 1.2
        assert numpy.allclose(ra.nearest_image(rb, RefCell([4.0, 4.0, 4.0]), copy=True).position, image.position)
    th = Thermostat(2.0, collision_period=0.3)
    assert th.relaxation_time == 0.3 and th.collision_period == 0.3 and Thermostat(1.0, relaxation_time=0.5).collision_period == 0.5


def test_rumd_configurations_gzipped_and_second_header_format(oracle_lib, tmp_path):
    """RUMD keeps configurations as .xyz.gz and has two header dialects (atooms/trajectory/rumd.py:56-67, 79-98):
    `boxLengths= dt=` and `sim_box=<type>,Lx,Ly,Lz integrator=<name>,dt,...`."""
    import gzip
    from atooms_b200 import TrajectoryRUMD
    s = make_system('ka_N256_rumd', oracle_lib, 'kob_andersen')
    path = str(tmp_path / 'conf.xyz.gz')
    with TrajectoryRUMD(path, 'w') as th:                   # written compressed, by the native formatter too
        th.timestep = 0.005
        th.write(s, 0)
        th.write(s, 40)
    assert gzip.open(path, 'rt').readline().strip() == '256'
    with TrajectoryRUMD(path) as th:
        assert th.steps == [0, 40] and th.timestep == 0.005
        assert numpy.abs(th[1].position - s.dump('position')).max() < 1e-8
    second = str(tmp_path / 'second.xyz.gz')
    with gzip.open(second, 'wt') as fh:
        for step in (0, 8):
            fh.write('3\nioformat=2 timeStepIndex=%d numTypes=2 integrator=IntegratorNVT,0.004,0.5,0.4,0 '
                     'sim_box=RectangularSimulationBox,10.03,10.03,12.5 mass=1.0,1.0 columns=type,x,y,z,imx,imy,imz,pe,vir\n' % step)
            fh.write('0 4.2019 -0.8430 -0.7399 0 0 0 -5.3925 28.4504\n1 1.0 2.0 3.0 0 1 0 -4.0 20.0\n0 -1.5 0.25 4.0 -1 0 0 -6.0 30.0\n')
    with TrajectoryRUMD(second) as th:
        frame = th[1]
        assert th.steps == [0, 8] and th.timestep == 0.004 and numpy.allclose(frame.cell.side, [10.03, 10.03, 12.5])
        assert frame.distinct_species == ['0', '1'] and numpy.allclose(frame.position[2], [-1.5, 0.25, 4.0])
        assert numpy.allclose(frame.extra['pe'][:, 0], [-5.3925, -4.0, -6.0]) and frame.extra['imy'][1, 0] == 1
    data = '/root/reference/data/ka_rumd_N1200.xyz.gz'      # the reference's own file, where the tree is at hand
    if HAVE_REF and os.path.exists(data):
        from atooms.trajectory import TrajectoryRUMD as RefRUMD
        ref = RefRUMD(data)
        with TrajectoryRUMD(data) as th:
            assert th.steps == ref.steps and th.timestep == ref.timestep and th.block_size == ref.block_size
            mine, theirs = th[3], ref[3]
            assert numpy.allclose(mine.cell.side, theirs.cell.side)
            assert numpy.abs(mine.position - numpy.array([p.position for p in theirs.particle])).max() == 0.0
        ref.close()


def test_snapshot_host_side_members(oracle_lib):
    """A frame read back from a trajectory answers the host-side questions a System answers
    (atooms/system/system.py:160-230, 330-379), without a device context."""
    from atooms_b200 import Snapshot
    s = make_system('ka_N256_rumd', oracle_lib, 'kob_andersen')
    f = Snapshot.of(s)
    assert abs(f.temperature - s.temperature) < 1e-12 and abs(f.kinetic_energy() - s.kinetic_energy()) < 1e-9
    assert numpy.allclose(f.cm_position, s.cm_position) and numpy.allclose(f.cm_velocity, s.cm_velocity)
    assert numpy.allclose(f.cm('position'), s.cm_position) and abs(f.packing_fraction - s.packing_fraction) < 1e-12
    assert f.concentration == s.concentration and f.composition == s.composition
    f.position += 3.0 * f.cell.side
    f.fold()
    assert numpy.abs(f.position - s.dump('position')).max() < 1e-9
    f.set_density(1.0)
    assert abs(f.density - 1.0) < 1e-12
    f.set_temperature(0.8, seed=3)
    assert abs(f.temperature - 0.8) < 1e-12 and numpy.abs(f.cm_velocity).max() < 1e-12
    f.scale_velocities(2.0)
    assert abs(f.temperature - 3.2) < 1e-12
    f.temperature = 1.0
    assert abs(f.temperature - 1.0) < 1e-12
    # and starts a backend like any other system
    assert len(B200(f, 'kob_andersen', library=oracle_lib).system) == 256
