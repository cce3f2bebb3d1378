"""
`B200` and `EnergyMinimization` — simulation backends in the sense of
atooms/backends/__init__.py:7-11: duck-typed objects with `.system` and
`.run(steps)`, consumed unchanged by `atooms.simulation.Simulation`
(atooms/simulation/core.py:65-130, :308-315) and
`atooms.optimization.Optimization` (atooms/optimization/core.py:39-74).

The optional members `Simulation` probes with hasattr are all provided:
`trajectory_class` (core.py:89-92), `write_checkpoint` / `read_checkpoint`
(:225-227, :263-265), `rmsd` (:276-280), `version` (:443-444), `timestep`
(observers.py:198-199), `__str__` (core.py:126).
"""
import os

import numpy

from . import __version__, capi, models
from .interaction import Interaction
from .system import System
from .trajectory import Trajectory

__all__ = ['B200', 'EnergyMinimization']

_PRECISION = {'fp32': 32, 'f32': 32, 'float32': 32, 'single': 32, 32: 32,
              'fp64': 64, 'f64': 64, 'float64': 64, 'double': 64, 64: 64}


def _save_arrays(path, **arrays):
    """Named arrays in one file: a JSON header line, then the raw buffers.  (numpy.savez goes through
    zipfile, whose CRC-32 pass alone takes ~0.1 s on the 130 MB of a million-particle checkpoint — as much
    as 400 MD steps — and Simulation writes one at the end of every run(), simulation/core.py:409-415.)"""
    import json
    meta, blobs = [], []
    for name, a in arrays.items():
        a = numpy.ascontiguousarray(a)
        m = {'name': name, 'dtype': a.dtype.str, 'shape': list(a.shape), 'bytes': int(a.nbytes)}
        # an array of one repeated value (masses, radii, image counters of a short run: a quarter of the file)
        # is stored as that value
        if a.size > 1024 and a.dtype.kind in 'fiu' and (a.reshape(-1)[::1024] == a.flat[0]).all() and (a == a.flat[0]).all():
            m['fill'], m['bytes'] = a.flat[0].item(), 0
        else:
            blobs.append(a)
        meta.append(m)
    with open(path, 'wb') as fh:
        fh.write((json.dumps(meta) + '\n').encode())
        for a in blobs:
            fh.write(a.data if a.nbytes else b'')


def _load_arrays(path):
    import json
    with open(path, 'rb') as fh:
        first = fh.readline()
        if first[:2] == b'PK':                      # a round-1 checkpoint (numpy.savez)
            return numpy.load(path)
        out = {}
        for m in json.loads(first.decode()):
            if 'fill' in m:
                out[m['name']] = numpy.full(m['shape'], m['fill'], dtype=numpy.dtype(m['dtype']))
            else:
                out[m['name']] = numpy.frombuffer(fh.read(m['bytes']), dtype=numpy.dtype(m['dtype'])).reshape(m['shape']).copy()
    return out


def _last_frame(inp):
    """A System, a Trajectory (last frame) or a path, as LAMMPS.__init__ accepts
    (atooms/backends/lammps.py:151-188)."""
    if isinstance(inp, System) or (hasattr(inp, 'particle') and hasattr(inp, 'cell')):
        return inp
    if hasattr(inp, '__getitem__') and hasattr(inp, 'steps'):
        return inp[-1]
    if isinstance(inp, str) and os.path.exists(inp):
        # xyz and RUMD configurations: the readers of this package (arrays, no Particle objects)
        if inp.endswith('.xyz.gz') or inp.endswith('.xyz'):
            from .trajectory import TrajectoryRUMD
            with (TrajectoryRUMD if inp.endswith('.gz') else Trajectory)(inp) as th:
                if len(th):
                    return th[-1]
        try:
            from atooms import trajectory
        except ImportError:
            raise ValueError('reading %s needs the atooms trajectory readers' % inp)
        if not hasattr(numpy, 'set_string_function'):      # numpy >= 2: close() of the xyz reader
            numpy.set_string_function = lambda *a, **k: None   # (atooms/trajectory/xyz.py:472-474)
        try:
            th = trajectory.Trajectory(inp)
        except ValueError:
            # the factory goes by suffix (trajectory/factory.py:93-114): RUMD's .xyz.gz
            # configurations need their own reader
            th = trajectory.TrajectoryRUMD(inp) if inp.endswith('.xyz.gz') else trajectory.TrajectoryXYZ(inp)
        system = th[-1]
        th.close()
        return system
    raise ValueError('could not initialize system from {}'.format(inp))


class B200(object):

    """NVE molecular dynamics (velocity-Verlet, the LAMMPS backend's default
    `fix nve`, atooms/backends/lammps.py:236-238) on one B200."""

    version = __version__

    def __init__(self, inp, model='lennard_jones', dt=0.002, skin=0.3, precision='fp64', device=0,
                 library=None, thermostat='nose-hoover', compressibility=0.1, barostat_interval=10,
                 small_system_kernel=True):
        """`thermostat` names the scheme used WHEN `system.thermostat` is set (the reference's
        backends switch to NVT on that attribute, atooms/backends/lammps.py:229-241):
        'nose-hoover' (LAMMPS `fix nvt`, RUMD IntegratorNVT) or 'berendsen'."""
        if thermostat not in ('nose-hoover', 'berendsen'):
            raise ValueError("thermostat must be 'nose-hoover' or 'berendsen'")
        self.thermostat_kind = capi.NVT_NOSE_HOOVER if thermostat == 'nose-hoover' else capi.NVT_BERENDSEN
        # NPT (system.barostat set as well): isotropic weak-coupling volume moves every `barostat_interval`
        # steps, mu^3 = 1 - compressibility * (interval * dt / relaxation_time) * (P0 - P)
        self.compressibility = float(compressibility)
        self.barostat_interval = int(barostat_interval)
        # NVE of a system of at most SMALL particles: the whole run() in ONE kernel launch (one thread block,
        # configuration in shared memory, all pairs: b2_batch_run_nve) instead of one launch per step, which
        # at this size is pure launch latency (LJ N = 256: 13.6 against 23 us per step)
        self.small_system_kernel = bool(small_system_kernel)
        self._small = None
        self.model = models.get(model)
        self.timestep = float(dt)
        self.skin = float(skin)
        self.precision = _PRECISION[precision]
        self.device = device
        self._library = library
        self.trajectory_class = Trajectory
        self.verbose = False
        self.steps = 0
        self._system = None
        self.system = _last_frame(inp)

    def __str__(self):
        return 'B200 (atooms_b200 %s, fp%d, %s)' % (__version__, self.precision, self.model)

    # `Simulation` reads and assigns backend.system (simulation/core.py:111-123,
    # tests/test_simulation.py:179-187; tests/test_rumd.py:178-194 assigns a deepcopy)
    @property
    def system(self):
        return self._system

    @system.setter
    def system(self, value):
        if not isinstance(value, System):
            value = System.from_atooms(value, model=self.model, precision=self.precision, skin=self.skin,
                                       device=self.device, library=self._library)
        if value.interaction is None:
            value.interaction = Interaction(self.model, precision=self.precision, skin=self.skin,
                                            device=self.device, library=self._library)
        self._system = value
        value._push()
        value._engine.mark_reference()       # rmsd origin (rumd.py:117-121)

    @property
    def rmsd(self):
        """Root mean square displacement since the system was attached, from the
        unfolded coordinates (rumd.py:163-177)."""
        s = self._system
        return float((s._get_thermo()['msd_sum'] / max(len(s), 1)) ** 0.5)

    def run(self, steps):
        s = self._system
        s._push()
        th = s.thermostat
        ba = getattr(s, 'barostat', None)
        if ba is not None and getattr(ba, 'pressure', None) is not None:
            # NPT: `fix 1 all npt temp T T tau_T iso P P tau_P` in the reference (lammps.py:229-233)
            if th is None or getattr(th, 'temperature', None) is None:
                raise ValueError('a barostat needs a thermostat as well (NPT), as in atooms/backends/lammps.py:229-233')
            self._run_npt(int(steps), th, ba)
        elif th is not None and getattr(th, 'temperature', None) is not None:
            # NVT: `fix 1 all nvt temp T T relaxation_time` (lammps.py:233-235)
            s._engine.run_nvt(int(steps), self.timestep, th.temperature,
                              getattr(th, 'relaxation_time', None) or 100 * self.timestep, self.thermostat_kind)
        elif self._small_ok(steps):
            self._run_small(int(steps))
        else:
            s._engine.run_nve(int(steps), self.timestep)
        if steps > 0:
            s._advanced()
        self.steps += int(steps)

    SMALL = 1024

    def _small_ok(self, steps):
        s = self._system
        return (self.small_system_kernel and steps >= 20 and 2 <= len(s) <= self.SMALL
                and not getattr(self.model, 'tabulated', False)
                and numpy.all(s.cell.side >= 2.0 * self.model.rcut.max())           # one minimum image, all pairs
                and numpy.allclose(s.cell.center, 0.0))

    def _run_small(self, steps):
        """The run as one launch of the replica kernel on a copy of the state (a few KB each way), then the
        engine takes positions, velocities and box crossings back: observables, rmsd and later runs see one
        continuous trajectory."""
        s = self._system
        s._pull_image()
        if self._small is None:
            self._small = capi.Batch(s.cell.side, self.model.eps, self.model.sigma, self.model.rcut,
                                     self.model.shifted, precision=self.precision, device=self.device,
                                     library=s._engine.lib)
        b = self._small
        b.upload(s._pos[None], s._ids[None], s._mass[None])
        b.set_velocities(s._vel[None])
        b.run_nve(steps, self.timestep)
        pos = b.download()[0]
        vel, img = b.state()
        image = s._img + img[0]
        s._engine.update(pos, vel[0])
        s._engine.set_images(image)

    def _run_npt(self, steps, th, ba):
        """Thermostatted dynamics (the fused NVT step kernel) with an isotropic Berendsen barostat: every
        `barostat_interval` steps the instantaneous pressure (device reduction) rescales the box and, affinely,
        the particles (b2_set_box on resident particles; the list is rebuilt).  Weak coupling relaxes <P> to
        the target like the reference's `fix npt` does, with smaller volume fluctuations than its
        Nose-Hoover barostat: the same state point, not the same fluctuation spectrum."""
        s = self._system
        tau_t = getattr(th, 'relaxation_time', None) or 100 * self.timestep
        tau_p = getattr(ba, 'relaxation_time', None) or 1000 * self.timestep
        done = 0
        while done < steps:
            n = min(self.barostat_interval, steps - done)
            s._engine.run_nvt(n, self.timestep, th.temperature, tau_t, self.thermostat_kind)
            s._advanced()
            done += n
            mu3 = 1.0 - self.compressibility * (n * self.timestep / tau_p) * (ba.pressure - s.pressure)
            mu3 = min(max(mu3, 0.97), 1.03)                 # at most 1 % in length per move
            s.set_box(s.cell.side * mu3 ** (1.0 / 3.0))

    # binary checkpoints: the reference's fallback goes through TrajectoryXYZ and
    # breaks under numpy >= 2 (simulation/core.py:225-234)
    def write_checkpoint(self, output_path):
        """Everything a restart needs: positions, velocities, box-crossing counters, the rmsd origin, the
        thermostat's friction, and what identifies the system (species, masses, cell)."""
        s = self._system
        s._pull_image()
        _save_arrays(output_path + '.chk.npz', position=s._pos, velocity=s._vel, image=s._img, ids=s._ids,
                     names=numpy.array(s._names), mass=s._mass, radius=s._radius, side=s.cell.side,
                     center=s.cell.center, steps=numpy.array(self.steps), reference=s._engine.reference(),
                     thermostat_xi=numpy.array(s._engine.thermostat_state()))

    def read_checkpoint(self, output_path):
        """Restore the state written by write_checkpoint into THIS system: run(a), checkpoint, restore,
        run(b) continues as run(a + b) would, rmsd and NVT included.  The checkpoint must describe the same
        particles (species, count) in the same cell shape; the cell's size is taken from it (NPT)."""
        path = output_path + '.chk.npz'
        if not os.path.exists(path):
            return
        d = _load_arrays(path)
        s = self._system
        if len(d['ids']) != len(s):
            raise ValueError('checkpoint %s holds %d particles, the system %d' % (path, len(d['ids']), len(s)))
        if not numpy.array_equal(numpy.array(d['names'])[d['ids']], numpy.array(s._names)[s._ids]):
            raise ValueError('checkpoint %s holds other species than the system' % path)
        if not numpy.allclose(d['mass'], s._mass):
            raise ValueError('checkpoint %s holds other masses than the system' % path)
        if not numpy.allclose(d['side'], s.cell.side):
            s.set_box(d['side'])
        s._pull()
        s._pos[:], s._vel[:] = d['position'], d['velocity']
        s._host_pos_dirty = s._host_vel_dirty = True
        s._thermo = None
        s._push()
        if 'image' in d:
            s._engine.set_images(d['image'])           # after the push: b2_update counts folds on top
            s._img[:] = d['image']
            s._img_stale = False
        if 'reference' in d:
            s._engine.set_reference(d['reference'])
        if 'thermostat_xi' in d:
            s._engine.set_thermostat_state(float(numpy.asarray(d['thermostat_xi']).reshape(-1)[0]))
        self.steps = int(numpy.asarray(d['steps']).reshape(-1)[0])


class EnergyMinimization(B200):

    """Energy minimisation backend with the surface of the LAMMPS one
    (atooms/backends/lammps.py:291-363): `method`, `ftol`, `steps`,
    `reached_steps`; driven by `Optimization` (optimization/core.py:69-74).

    `ftol` follows LAMMPS `minimize 0.0 ftol ...`: stop when the 2-norm of the
    global force vector drops below it, i.e. sum F^2 / N < ftol^2 / N.
    `tolerance`, if given, is instead the bound on sum F^2 / N used by
    `Optimization._target_force_norm_square` (optimization/core.py:34-36).
    """

    def __init__(self, inp, model='lennard_jones', method='fire', ftol=1e-4, steps=100000, dt=None,
                 tolerance=None, skin=0.3, precision='fp64', device=0, library=None):
        method = method.lower()
        if method not in ('fire', 'sd'):
            raise ValueError("method must be 'fire' or 'sd', not %r" % method)
        B200.__init__(self, inp, model=model, dt=dt if dt is not None else (0.002 if method == 'fire' else 0.001),
                      skin=skin, precision=precision, device=device, library=library)
        self.method = method
        self.ftol = ftol
        self.tolerance = tolerance
        self.steps = steps
        self.max_evaluations = 100000
        # "Optimization backends must set a boolean reached_steps attribute. It is True at
        # the beginning." (lammps.py:307-309); read at step 0 by optimization/core.py:30-32
        self.reached_steps = True
        self.iterations = 0

    def __str__(self):
        return 'B200 energy minimization (%s)' % self.method

    def run(self, steps=None):
        if steps is not None:
            self.steps = steps
        s = self._system
        s._push()
        n = max(len(s), 1)
        fn2_tol = self.tolerance if self.tolerance is not None else self.ftol ** 2 / n
        done, converged = s._engine.minimize(capi.MIN_FIRE if self.method == 'fire' else capi.MIN_SD,
                                             int(min(100000000, self.steps)), self.timestep, fn2_tol)
        self.iterations += done
        # True when the iteration budget stopped us, False when the tolerance did (lammps.py:350-354)
        self.reached_steps = not converged
        s._advanced()
