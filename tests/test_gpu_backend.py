"""
The Python drop-in classes on the real CUDA engine (B200).  The host-protocol
tests of test_host_protocol.py are re-run with the product library in place
of the oracle, plus a chunked driver loop shaped like Simulation.run
(atooms/simulation/core.py:386-404) at a size where the Python object model
of the reference would already hurt.
"""
import inspect

import numpy
import pytest

import test_host_protocol as hp
from conftest import golden
from test_gpu_parity import replicated
from atooms_b200 import B200, EnergyMinimization, System, Cell

pytestmark = pytest.mark.gpu

PROTOCOL_TESTS = [hp.test_backend_surface, hp.test_nve_pin_through_the_backend, hp.test_observers_accessors,
                  hp.test_sync_contract, hp.test_thermostat_switches_the_backend_to_nvt, hp.test_barostat_switches_the_backend_to_npt, hp.test_deepcopy_and_reassign, hp.test_small_systems_run_in_one_launch, hp.test_checkpoint_roundtrip, hp.test_checkpoint_restores_images_rmsd_and_thermostat,
                  hp.test_minimisation_backend, hp.test_generic_interaction_protocol, hp.test_trajectory_writer,
                  hp.test_model_registry_and_reference_schema, hp.test_trajectory_fast_path, hp.test_tabulated_potentials]
# the reference's OWN drivers (Simulation, Optimization, composite Interaction, its trajectory readers,
# the md.py look-alike) on the CUDA engine; the reference comes from baseline/_ref on the GPU box
REFERENCE_DRIVER_TESTS = [hp.test_reference_simulation_driver, hp.test_integration_shim, hp.test_reference_optimization_driver,
                          hp.test_interaction_with_reference_system,
                          hp.test_neighbors_export_round_trips_through_the_reference_reader,
                          hp.test_command_line_driver, hp.test_reference_trajectory_callbacks,
                          hp.test_reference_class_decorators_on_the_trajectory, hp.test_system_members_of_the_reference]


def _call(fn, cuda_lib, pins, tmp_path, monkeypatch):
    available = {'oracle_lib': cuda_lib, 'pins': pins, 'tmp_path': tmp_path, 'monkeypatch': monkeypatch}
    fn(**{name: available[name] for name in inspect.signature(fn).parameters})


@pytest.mark.parametrize('fn', PROTOCOL_TESTS, ids=lambda f: f.__name__)
def test_protocol_on_cuda(fn, cuda_lib, pins, tmp_path, monkeypatch):
    _call(fn, cuda_lib, pins, tmp_path, monkeypatch)


@pytest.mark.parametrize('fn', REFERENCE_DRIVER_TESTS, ids=lambda f: f.__name__)
def test_reference_drivers_on_cuda(fn, cuda_lib, pins, tmp_path, monkeypatch):
    assert hp.HAVE_REF, 'the reference must be installed under baseline/_ref (tools/install_reference.py)'
    _call(fn, cuda_lib, pins, tmp_path, monkeypatch)


def test_default_library_is_cuda():
    """No `library=` argument: the product path, which must be the CUDA engine."""
    d = golden('ka_N256_rumd')
    s = System(d['position'], d['species_names'][d['species']], Cell(d['side']), velocity=d['velocity'],
               model='kob_andersen')
    assert s._engine.lib.name == 'cuda-sm_100a'
    bck = B200(s, 'kob_andersen')
    assert round(abs(bck.system.potential_energy(per_particle=True) - (-6.4014371344819665)), 3) == 0


@pytest.mark.parametrize('precision', ['fp64', 'fp32'])
def test_chunked_driver_loop(precision):
    d = replicated('ka_N1200', 3, 0.0, 3)                      # N = 32400
    names = numpy.array(['A', 'B'])[d['species']]
    system = System(d['position'], names, Cell(d['side']), velocity=d['velocity'], model='kob_andersen',
                    precision=32 if precision == 'fp32' else 64)
    bck = B200(system, 'kob_andersen', dt=0.002, precision=precision)
    n = len(bck.system.particle)                               # O(1): no Particle objects are built
    assert n == 32400 and not bck.system._views_out
    energies, rmsd = [], []
    for chunk in range(10):                                    # observers every 50 steps
        bck.run(50)
        s = bck.system
        energies.append(s.total_energy(per_particle=True, cache=True))
        rmsd.append(bck.rmsd)
        assert abs(s.temperature - 2 * s.kinetic_energy() / (3 * (n - 1))) < 1e-12
    energies = numpy.array(energies)
    assert numpy.abs(energies - energies[0]).max() < (2e-4 if precision == 'fp64' else 1e-3)
    assert all(b > a for a, b in zip(rmsd, rmsd[1:]))
    assert bck.steps == 500
    # pressure from the device virial equals the definition (system.py:337-348)
    s = bck.system
    assert abs(s.pressure - (n * s.temperature + s.virial(per_particle=False) / 3) / s.cell.volume) < 1e-12


def test_fire_through_chunks_equals_one_call():
    """Optimization hands the backend 1000-step chunks (optimization/core.py:56): the result
    must not depend on the chunking."""
    d = golden('lj_N256')
    names = d['species_names'][d['species']]

    def fresh():
        return EnergyMinimization(System(d['position'], names, Cell(d['side']), model='lennard_jones'),
                                  'lennard_jones', method='fire', ftol=1e-4)
    one = fresh()
    one.run(100000)
    chunks = fresh()
    calls = 0
    while chunks.reached_steps:
        chunks.run(300)
        calls += 1
    assert calls > 2 and chunks.iterations == one.iterations
    assert abs(chunks.system.potential_energy() - one.system.potential_energy()) < 1e-9
