"""
Array-backed `System` for the B200 backend.

atooms' `System` is a list of `Particle` objects, each owning 3-element numpy
arrays (atooms/system/system.py:38-41, particle.py:15-30); at N >= 1e5 every
accessor costs seconds and `System(N=10**6)` raises.  Like the
RUMD backend's wrapper (atooms/backends/rumd.py:225-357) this class keeps the
state in flat arrays — here mirrored on the GPU — and materialises
`system.particle` lazily, only when somebody asks for the list.

Synchronisation contract
  * the device is authoritative after `backend.run()`; host arrays are
    refreshed lazily, in place, on access (`b2_download`).
  * observers' accessors (temperature, kinetic / potential energy, virial,
    force norm, pressure, len(system.particle)) are device reductions
    (`b2_thermo`) or O(1): no N-sized copy, no Python loop.
  * once a MUTABLE handle on the host arrays has been handed out —
    `view()` / `dump(view=True)`, `system.position`, an element of
    `system.particle` — in-place edits by arbitrary code (the f90 kernels of
    docs/simulations.org:302-308) cannot be detected, so from then on the host
    arrays are pushed (`b2_update`) before every device operation and pulled
    right after every run; the handles stay valid because downloads write
    into the same arrays.  `release_views()` ends that mode.  With
    `system.sync = 'tracked'` the handles report numpy writes (item
    assignment, in-place operators, `out=`) and only arrays that were written
    are pushed; a writer numpy cannot see must then call `mark_modified()`.
    Host mirrors live in page-locked memory (b2_host_alloc).
"""
import copy

import numpy

from . import capi, models

__all__ = ['System', 'Cell', 'Particle', 'Thermostat', 'Barostat']


class Cell(object):

    """Orthorhombic periodic cell (atooms/system/cell.py:11-39)."""

    def __init__(self, side, center=None):
        self.side = numpy.asarray(side, dtype=numpy.float64)
        self.center = numpy.zeros_like(self.side) if center is None else numpy.asarray(center, dtype=numpy.float64)
        self.shape = 'cubic'
        self.periodic = numpy.ones(self.side.size, dtype=bool)

    @property
    def volume(self):
        return float(numpy.prod(self.side))

    @property
    def origin(self):                        # (atooms/system/cell.py:36-39)
        import warnings
        warnings.warn('cell origin is deprecated', DeprecationWarning)
        return numpy.zeros(self.side.size)


class Particle(object):

    """Minimal stand-in for atooms.system.particle.Particle (particle.py:15-30),
    used for the lazily materialised `system.particle` when atooms itself is not
    importable.  Attributes are views on the system's host arrays."""

    def __init__(self, position, velocity, species, mass=1.0, radius=0.5):
        self.position = position
        self.velocity = velocity
        self.species = species
        self.mass = mass
        self.radius = radius

    @property
    def kinetic_energy(self):
        return 0.5 * self.mass * numpy.dot(self.velocity, self.velocity)

    @property
    def diameter(self):
        return 2 * self.radius

    def distance(self, particle, cell=None, folded=True):
        """Separation vector from `particle`, of the nearest periodic image when a cell is given
        (atooms/system/particle.py:63-82; `folded` only selects between two equivalent reductions there)."""
        r = self.position - particle.position
        return r if cell is None else r - numpy.rint(r / cell.side) * cell.side

    def nearest_image(self, particle, cell, copy=False, folded=False):
        """Move this particle (or a copy of it) to its periodic image nearest to `particle` (particle.py:37-61)."""
        where = particle.position + self.distance(particle, cell)
        if copy:
            import copy as _copy
            image = _copy.deepcopy(self)
            image.position = where
            return image
        self.position[:] = where
        return self

    def maxwellian(self, T):
        """Velocity components drawn from a Gaussian of variance T / m (particle.py:101-107)."""
        import random
        for k in range(len(self.velocity)):
            self.velocity[k] = random.gauss(0, numpy.sqrt(T / self.mass))

    def fold(self, cell):
        # atooms/system/particle.py:90-99
        p = self.position - cell.center
        self.position[:] = p - numpy.rint(p / cell.side) * cell.side + cell.center
        return self

    def __repr__(self):
        return 'Particle(species={0.species}, mass={0.mass}, position={0.position}, ' \
               'velocity={0.velocity}, radius={0.radius})'.format(self)


class Thermostat(object):
    """Target temperature and coupling time, the two attributes of
    atooms.system.reservoir.Thermostat (reservoir.py:22-40) that backends read
    (lammps.py:233-235).  Assigning one to `system.thermostat` turns `B200.run`
    into NVT dynamics; `pressure` then uses this temperature (system.py:343-346)."""

    def __init__(self, temperature=None, name='', mass=1.0, coordinate=1.0, momentum=1.0, relaxation_time=1.0,
                 collision_period=None):
        self.name = name
        self.temperature = temperature
        self.coordinate = numpy.asarray(coordinate)
        self.momentum = numpy.asarray(momentum)
        self.mass = numpy.asarray(mass)
        self.relaxation_time = relaxation_time if collision_period is None else collision_period

    @property
    def collision_period(self):
        return self.relaxation_time


class Barostat(object):
    """Target pressure and coupling time, the attributes of atooms.system.reservoir.Barostat
    (reservoir.py:43-54) that backends read (lammps.py:229-233).  With `system.thermostat` AND
    `system.barostat` set, `B200.run` samples the NPT ensemble, as the reference's backends do."""

    def __init__(self, pressure, name='', mass=1.0, coordinate=1.0, momentum=1.0, relaxation_time=1.0):
        self.name = name
        self.pressure = pressure
        self.coordinate = numpy.asarray(coordinate)
        self.momentum = numpy.asarray(momentum)
        self.mass = numpy.asarray(mass)
        self.relaxation_time = relaxation_time


class _ParticleList(object):

    """`system.particle`: a sequence that builds Particle objects on demand.
    `len()` is free (Simulation.wall_time calls it, simulation/core.py:296-297);
    indexing or iterating creates Particle objects whose position / velocity are
    views on the system's host arrays and switches the system to the
    "views handed out" mode (see the module docstring)."""

    def __init__(self, system):
        self._system = system
        self._cache = {}

    def __len__(self):
        return len(self._system)

    def _make(self, i):
        s = self._system
        p = self._cache.get(i)
        if p is None:
            p = _bound_particle_class()(position=s._pos_user[i], velocity=s._vel_user[i],
                                        species=s._pending_species.get(i, s._names[s._ids[i]]),
                                        mass=s._mass[i], radius=s._radius[i])
            # the reference's Particle takes numpy.asarray of its arguments (particle.py:24-30), which drops
            # the tracking subclass: hand it the tracked row views again
            p.position, p.velocity = s._pos_user[i], s._vel_user[i]
            p.__dict__['_b2_owner'] = (s, i)
            # numpy.asarray keeps the row views (particle.py:24-30)
            s._pull_image()
            s._img_out = True                # a view of the counters is in user hands from now on
            p.periodic_image = s._img[i]
            self._cache[i] = p
        return p

    def __getitem__(self, i):
        self._system._hand_out()
        if isinstance(i, slice):
            return [self._make(k) for k in range(*i.indices(len(self)))]
        n = len(self)
        if i < 0:
            i += n
        if not 0 <= i < n:
            raise IndexError('particle index out of range')
        return self._make(i)

    def __iter__(self):
        self._system._hand_out()
        return (self._make(i) for i in range(len(self)))


def _particle_class():
    try:
        from atooms.system.particle import Particle as P   # the real one when atooms is installed
        return P
    except Exception:
        return Particle


_BOUND = {}


def _bound_particle_class():
    """The particle class with `species`, `mass` and `radius` that write through to the System they belong to (position and
    velocity are views already): the reference's trajectory callbacks assign them per particle
    (atooms/trajectory/decorators.py:33-89 `change_species`)."""
    base = _particle_class()
    cls = _BOUND.get(base)
    if cls is None:
        def _get(name):
            return lambda self: self.__dict__['_b2_' + name]

        def _set(name):
            def setter(self, value):
                self.__dict__['_b2_' + name] = value
                owner = self.__dict__.get('_b2_owner')
                if owner is not None:
                    owner[0]._assign(owner[1], name, value)
            return setter
        def _set_vector(name):
            def setter(self, value):
                owner = self.__dict__.get('_b2_owner')
                if owner is None:
                    self.__dict__['_b2_' + name] = value
                else:           # p.position = array: the values go into the row this particle is a view of
                    self.__dict__['_b2_' + name][...] = value
                    owner[0]._assign(owner[1], name, None)
            return setter
        cls = type('Particle', (base,), {'position': property(_get('position'), _set_vector('position')),
                                         'velocity': property(_get('velocity'), _set_vector('velocity')),
                                         'species': property(_get('species'), _set('species')),
                                         'mass': property(_get('mass'), _set('mass')),
                                         'radius': property(_get('radius'), _set('radius')),
                                         '__module__': base.__module__})
        _BOUND[base] = cls
    return cls


class _Tracked(numpy.ndarray):

    """Host array handed out to the user (`view()`, `system.position`, `particle[i].position`): it
    raises a flag on its System whenever it is written through numpy — item assignment, in-place
    operators, ufuncs with `out=`, `fill`, `sort`, `numpy.copyto` — so that only arrays that were
    really edited are pushed to the device.  Views of it stay tracked.  A write that bypasses numpy
    (a C extension filling the buffer) is not seen: follow it with `system.mark_modified()`."""

    _flag = None

    def __array_finalize__(self, obj):
        self._flag = getattr(obj, '_flag', None)

    def _touch(self):
        if self._flag is not None:
            self._flag[0] = True

    def __setitem__(self, key, value):
        self._touch()
        numpy.ndarray.__setitem__(self, key, value)

    def __array_ufunc__(self, ufunc, method, *inputs, out=None, **kwargs):
        plain = [x.view(numpy.ndarray) if isinstance(x, _Tracked) else x for x in inputs]
        if out is not None:
            for o in out:
                if isinstance(o, _Tracked):
                    o._touch()
            kwargs['out'] = tuple(o.view(numpy.ndarray) if isinstance(o, _Tracked) else o for o in out)
        if method == 'at' and isinstance(inputs[0], _Tracked):
            inputs[0]._touch()
        result = getattr(ufunc, method)(*plain, **kwargs)
        if out is not None:                       # in-place operators must hand the same object back
            return out[0] if len(out) == 1 else out
        return result

    def __array_function__(self, func, types, args, kwargs):
        if func in (numpy.copyto, numpy.put, numpy.place, numpy.putmask, numpy.fill_diagonal) and args and \
                isinstance(args[0], _Tracked):
            args[0]._touch()
        strip = lambda x: x.view(numpy.ndarray) if isinstance(x, _Tracked) else x
        args = tuple([strip(y) for y in x] if isinstance(x, (list, tuple)) else strip(x) for x in args)
        return func(*args, **{k: strip(v) for k, v in kwargs.items()})

    def fill(self, value):
        self._touch()
        self.view(numpy.ndarray).fill(value)

    def sort(self, *args, **kwargs):
        self._touch()
        self.view(numpy.ndarray).sort(*args, **kwargs)

    def put(self, *args, **kwargs):
        self._touch()
        self.view(numpy.ndarray).put(*args, **kwargs)


def _tracked(array, flag):
    out = array.view(_Tracked)
    out._flag = flag
    return out


class System(object):

    def __init__(self, position, species, cell, velocity=None, mass=None, radius=None,
                 model='lennard_jones', precision=64, skin=0.3, device=0, library=None, species_order=None):
        position = numpy.array(position, dtype=numpy.float64)
        if position.ndim != 2 or position.shape[1] != 3:
            raise ValueError('position must be (N, 3); only three dimensions are supported')
        n = position.shape[0]
        # host mirrors in page-locked memory: the push / pull around run() with views handed out is DMA
        lib = library if library is not None else capi.load()
        self._pos = lib.host_empty((n, 3))
        self._pos[:] = position
        self._vel = lib.host_empty((n, 3))
        self._vel[:] = 0.0 if velocity is None else numpy.asarray(velocity, dtype=numpy.float64)
        self._mass = numpy.ones(n) if mass is None else numpy.array(mass, dtype=numpy.float64)
        self._radius = numpy.full(n, 0.5) if radius is None else numpy.array(radius, dtype=numpy.float64)
        self._img = lib.host_empty((n, 3), numpy.int32)
        self._img[:] = 0
        self.cell = cell if hasattr(cell, 'side') else Cell(cell)
        self.model = models.get(model)
        # species names -> 0-based ids in table order
        names = [str(s) for s in species]
        if species_order is not None:         # (a system rebuilt from a subset keeps the ids of its species)
            order = list(species_order) + sorted(set(names) - set(species_order))
        else:
            order = self.model.species if self.model.species is not None else sorted(set(names))
        if len(order) > self.model.nsp:
            raise ValueError('system has %d species but the model %s has %d'
                             % (len(order), self.model, self.model.nsp))
        lookup = {s: i for i, s in enumerate(order)}
        try:
            self._ids = numpy.array([lookup[s] for s in names], dtype=numpy.int32)
        except KeyError as err:
            raise ValueError('species %s not in the model species %s' % (err, order))
        self._names = list(order)
        self.thermostat = None
        self.barostat = None
        self.reservoir = None
        self.interaction = None      # set by the backend (atooms_b200.interaction.Interaction)
        self.precision = int(precision)
        self._skin_value = float(skin)
        self._engine = capi.Engine(self.precision, device=device, library=library)
        self._engine.set_box(self.cell.side, self.cell.center)
        self._engine.set_model(self.model)
        self._engine.set_skin(skin)
        self._engine.upload(self._pos, self._ids, self._vel, self._mass)
        if numpy.any(self.cell.side < 2.0 * self.model.rcut.max()):
            import warnings
            warnings.warn('cell side %g is below 2 r_cut = %g: only the nearest image of each pair interacts '
                          '(minimum-image convention, include/atooms_b200.h)'
                          % (self.cell.side.min(), 2.0 * self.model.rcut.max()), stacklevel=2)
        self._host_pos_dirty = False     # host positions changed since the last push
        self._host_vel_dirty = False
        self._device_newer = False       # device moved on since the last pull
        self._views_out = False          # a mutable handle on the host arrays is in user hands
        self._img_stale = False
        self._img_out = False
        self._identity_dirty = False     # species or masses were assigned through particle objects
        self._pending_species = {}       # ... particle -> name, until the next push settles them
        self._touched = {'pos': [False], 'vel': [False]}      # ... and was written through (see _Tracked)
        # 'always': while handles are out, both arrays are pushed before every device operation — right for any
        # writer.  'tracked': only arrays written through numpy since the last push (see _Tracked) are pushed.
        self.sync = 'always'
        self._pos_user = _tracked(self._pos, self._touched['pos'])
        self._vel_user = _tracked(self._vel, self._touched['vel'])
        self._particle = _ParticleList(self)
        self._thermo = None              # cached b2_thermo of the current state

    # -- construction from / conversion to reference objects ------------------
    @classmethod
    def from_atooms(cls, system, **kwargs):
        """Build from an `atooms.system.System` (or anything with .particle / .cell)."""
        if isinstance(system, cls):
            return system
        from .trajectory import Snapshot
        if isinstance(system, Snapshot):                  # array-backed frame: no Particle objects
            return cls(position=system.position, species=system.species.tolist(),
                       cell=Cell(system.cell.side, getattr(system.cell, 'center', None)),
                       velocity=system.velocity, mass=system.mass, radius=system.radius, **kwargs)
        particle = system.particle
        if system.cell is None:
            raise ValueError('the B200 backend needs a periodic cell')
        return cls(position=[p.position for p in particle],
                   species=[p.species for p in particle],
                   cell=Cell(system.cell.side, getattr(system.cell, 'center', None)),
                   velocity=[p.velocity for p in particle],
                   mass=[p.mass for p in particle],
                   radius=[getattr(p, 'radius', 0.5) for p in particle], **kwargs)

    # -- synchronisation ---------------------------------------------------------
    def _push(self):
        """Host -> device, if the host side was (or may have been) touched."""
        if self._views_out:
            # handles on the host arrays are in user hands
            if self.sync == 'tracked':        # push what was written through them (numpy writes only)
                self._host_pos_dirty |= self._touched['pos'][0]
                self._host_vel_dirty |= self._touched['vel'][0]
            else:                             # any writer, f2py / C kernels included: push everything
                self._host_pos_dirty = self._host_vel_dirty = True
            self._touched['pos'][0] = self._touched['vel'][0] = False
        if self._identity_dirty:
            # species / masses changed: the device copy is rebuilt from the host arrays (which the holder of the
            # particle objects owns anyway), the box-crossing counters are kept
            self._resolve_species()
            self._engine.upload(self._pos, self._ids, self._vel, self._mass)
            self._engine.set_images(self._img)
            self._identity_dirty = self._host_pos_dirty = self._host_vel_dirty = False
            self._thermo = None
            if self.interaction is not None:
                self.interaction._invalidate()
        if self._host_pos_dirty or self._host_vel_dirty:
            self._engine.update(self._pos if self._host_pos_dirty else None,
                                self._vel if self._host_vel_dirty else None)
            self._host_pos_dirty = self._host_vel_dirty = False
            self._thermo = None
            if self.interaction is not None:
                self.interaction._invalidate()

    def _pull(self):
        """Device -> host (in place), if the device moved on."""
        if self._device_newer:
            # (the box-crossing counters follow only when somebody looks at them: _pull_image)
            self._engine.download_into(self._pos, self._vel, None)
            self._device_newer = False
            self._img_stale = True

    def _pull_image(self):
        self._pull()
        if self._img_stale:
            self._engine.download_into(None, None, self._img)
            self._img_stale = False

    def _hand_out(self):
        """A mutable handle on the host arrays is about to leave our hands."""
        self._pull()
        self._views_out = True

    def mark_modified(self, position=True, velocity=True):
        """The host arrays were written by something numpy cannot see (a C extension, a raw buffer)."""
        self._touched['pos'][0] |= bool(position)
        self._touched['vel'][0] |= bool(velocity)

    def release_views(self):
        """Promise that host arrays handed out earlier are no longer edited:
        back to lazy synchronisation."""
        self._push()
        self._views_out = False

    def _advanced(self):
        """Called by the backend after the device state changed."""
        self._device_newer = True
        self._thermo = None
        if self.interaction is not None:
            self.interaction._invalidate()
        if self._views_out:
            self._pull()
            if self._img_out:
                self._pull_image()

    def _get_thermo(self):
        self._push()
        if self._thermo is None:
            self._thermo = self._engine.thermo()
        return self._thermo

    # -- array accessors -----------------------------------------------------------
    @property
    def position(self):
        """(N, 3) host array, mutable: editing it in place edits the system."""
        self._hand_out()
        return self._pos_user

    @property
    def velocity(self):
        self._hand_out()
        return self._vel_user

    @property
    def image(self):
        self._pull_image()
        return self._img

    @property
    def species(self):
        if self._pending_species:
            return numpy.array(self._labels().tolist())
        return numpy.array(self._names)[self._ids]

    @property
    def mass(self):
        return self._mass.copy()

    def __len__(self):
        return len(self._ids)

    # -- the reference's System interface (atooms/system/system.py) ----------------
    @property
    def particle(self):
        """Lazily materialised particle sequence (cf. rumd.py:326-352, which rebuilds
        the whole list on each access)."""
        return self._particle

    @particle.setter
    def particle(self, particles):
        """Replace the particles by a list of particle objects, as the reference's System allows and its trajectory
        callbacks do (`sort`, `filter_species`: atooms/trajectory/decorators.py:91-99).  The device context is
        rebuilt; cell, model, reservoirs, interaction and synchronisation mode stay."""
        particles = list(particles)
        if not particles:
            raise ValueError('a system needs at least one particle')
        n = len(particles)
        images = [getattr(q, 'periodic_image', None) for q in particles]
        self._rebuild(numpy.array([numpy.asarray(q.position, dtype=numpy.float64) for q in particles]).reshape(n, 3),
                      numpy.array([numpy.asarray(q.velocity, dtype=numpy.float64) for q in particles]).reshape(n, 3),
                      [str(q.species) for q in particles], [float(q.mass) for q in particles],
                      [float(getattr(q, 'radius', 0.5)) for q in particles],
                      numpy.array(images, dtype=numpy.int32).reshape(n, 3) if all(im is not None for im in images) else None)

    def _rebuild(self, position, velocity, names, mass, radius, image=None):
        """This system with another set of particles (any number): a new device context in place of the old one;
        cell, model, reservoirs, interaction and synchronisation mode stay."""
        self._resolve_species()
        new = System(position=position, species=names, cell=self.cell, velocity=velocity, mass=mass, radius=radius,
                     model=self.model, precision=self.precision, skin=self._skin(), device=self._engine.device,
                     library=self._engine.lib, species_order=self._names)
        if image is not None:
            new._img[:] = image
            new._engine.set_images(new._img)
        for name in ('thermostat', 'barostat', 'reservoir', 'interaction', 'sync'):
            setattr(new, name, getattr(self, name))
        old_engine = self._engine
        self.__dict__.update(new.__dict__)
        self._particle._system = self
        old_engine.close()
        new.__dict__.clear()                 # (its engine and arrays live on in self)
        if self.interaction is not None:
            self.interaction._invalidate()

    def replicate(self, times, axis):
        """`times` copies of the system side by side along `axis`, the cell grown with them, the result centred
        as before (atooms/system/system.py:381-396) — on the arrays, not particle by particle."""
        n = int(times)
        if n <= 1:
            raise ValueError('replicate: times must be larger than 1')
        self._pull_image()
        self._resolve_species()
        length = float(self.cell.side[axis])
        shift = (numpy.arange(n) - 0.5 * (n - 1)) * length
        pos = numpy.tile(self._pos, (n, 1))
        pos[:, axis] += numpy.repeat(shift, len(self))
        names = [self._names[i] for i in self._ids] * n
        side = numpy.array(self.cell.side, dtype=numpy.float64)
        side[axis] *= n
        self.cell.side = side
        self._rebuild(pos, numpy.tile(self._vel, (n, 1)), names, numpy.tile(self._mass, n), numpy.tile(self._radius, n),
                      numpy.tile(self._img, (n, 1)))

    @property
    def packing_fraction(self):
        return float(numpy.pi / 6 * ((2.0 * self._radius) ** 3).sum() / self.cell.volume)      # system.py:160-162

    @property
    def species_layout(self):
        """'A' (names), 'C' (integers from 0) or 'F' (integers from 1), as system.py:569-589."""
        try:
            lowest = min(int(name) for name in self.distinct_species)
        except ValueError:
            return 'A'
        if lowest not in (0, 1):
            raise ValueError('Numeric species should start from 0 or 1')
        return 'C' if lowest == 0 else 'F'

    @species_layout.setter
    def species_layout(self, layout):
        """Rename the species (system.py:591-628): a change of names only, every particle keeps its row of the
        pair table and nothing moves on the device."""
        if layout[0] not in 'ACF':
            raise ValueError('layout must be A, C or F (not %s)' % layout)
        self._resolve_species()
        current = self.species_layout
        if layout[0] == current:
            return
        import string
        used = self.distinct_species
        if layout[0] == 'A':
            lowest = min(int(name) for name in used)
            renamed = {name: string.ascii_uppercase[int(name) - lowest] for name in used}
        elif current == 'A':
            renamed = {name: str(k + (1 if layout[0] == 'F' else 0)) for k, name in enumerate(used)}
        else:
            renamed = {name: str(int(name) + (1 if layout[0] == 'F' else -1)) for name in used}
        self._names = [renamed.get(name, name) for name in self._names]
        self._particle._cache.clear()

    def set_composition(self, N):
        """Assign the species at random so that N[name] particles carry `name` (system.py:210-220)."""
        import random
        if sum(N.values()) != len(self):
            raise ValueError('composition %s does not match N = %d' % (N, len(self)))
        self._resolve_species()
        for name in N:
            if str(name) not in self._names:
                if len(self._names) >= self.model.nsp:
                    raise ValueError('species %r: the model %s has %d species (%s)' % (name, self.model, self.model.nsp, self._names))
                self._names.append(str(name))
        species = []
        for name in N:
            species += [self._names.index(str(name))] * N[name]
        random.shuffle(species)
        self._ids[:] = species
        self._identity_dirty = True
        self._thermo = None
        self._particle._cache.clear()

    def show(self, *args, **kwargs):
        raise NotImplementedError('visualisation backends (atooms/system/system.py:556-567) are unavailable on this box')

    def _assign(self, i, name, value):
        """`particle[i].species = ...` / `.mass = ...` (see _bound_particle_class)."""
        if name == 'species':
            self._pending_species[int(i)] = str(value)       # settled by _resolve_species
        elif name == 'radius':
            self._radius[i] = float(value)                   # (host side only: the engine has no use for it)
            return
        elif name in ('position', 'velocity'):
            return                                           # (written through the tracked row view already)
        else:
            self._mass[i] = float(value)
        self._identity_dirty = True
        self._thermo = None

    def _labels(self):
        """Species name of every particle, names assigned through particle objects included."""
        labels = numpy.array(self._names, dtype=object)[self._ids]
        for i, name in self._pending_species.items():
            labels[i] = name
        return labels

    def _resolve_species(self):
        """Names assigned particle by particle (`_assign`) -> rows of the model's table.  If every name is one the
        system has, each particle simply takes that species.  Otherwise the assignment must be a RENAMING: all
        particles of a species carry one and the same new name (`change_species`: 'A' -> '0', 'B' -> '1', whatever
        the names meant before), and the species keeps its row.  Anything else does not fit a table of fixed species."""
        if not self._pending_species:
            return
        labels = self._labels()
        self._pending_species = {}
        if set(labels.tolist()) <= set(self._names):
            lookup = {name: k for k, name in enumerate(self._names)}
            self._ids[:] = [lookup[name] for name in labels]
            return
        names = list(self._names)
        for k in numpy.unique(self._ids):
            mine = set(labels[self._ids == k].tolist())
            if len(mine) != 1:
                raise ValueError('species %s of the model %s (species %s) would be split into %s: only a renaming of '
                                 'all particles of a species can be taken over' % (names[k], self.model, self._names, sorted(mine)))
            names[k] = mine.pop()
        used = [names[k] for k in numpy.unique(self._ids)]
        if len(set(used)) != len(used):
            raise ValueError('two species of the model %s were given the same new name %s' % (self.model, used))
        self._names = names

    @property
    def number_of_dimensions(self):
        return 3

    @property
    def distinct_species(self):
        if self._pending_species:
            return sorted(set(self._labels().tolist()))
        return sorted(set(self._names[i] for i in numpy.unique(self._ids)))

    @property
    def composition(self):
        if self._pending_species:
            names, counts = numpy.unique(self._labels().astype(str), return_counts=True)
            return {str(k): int(c) for k, c in zip(names, counts)}
        ids, counts = numpy.unique(self._ids, return_counts=True)
        return {self._names[i]: int(c) for i, c in zip(ids, counts)}

    @property
    def concentration(self):
        n = float(len(self))
        return {k: v / n for k, v in self.composition.items()}

    @property
    def density(self):
        return len(self) / self.cell.volume if len(self) else 0.0

    def kinetic_energy(self, per_particle=False, normed=False):
        # system.py:232-243
        ekin = self._get_thermo()['kinetic'] if len(self) else 0.0
        return ekin / len(self) if (per_particle or normed) else ekin

    @property
    def temperature(self):
        # system.py:164-179: (N-1) d degrees of freedom
        n = len(self)
        if n > 1:
            return 2.0 / ((n - 1) * 3) * self.kinetic_energy()
        if n == 1:
            return 2.0 / 3 * self.kinetic_energy()
        return 0.0

    @temperature.setter
    def temperature(self, T):
        self.set_temperature(T)

    def set_temperature(self, temperature, seed=None):
        """Maxwell-Boltzmann velocities, centre of mass removed, rescaled to exactly
        `temperature` (system.py:185-198); numpy's generator replaces random.gauss."""
        rng = numpy.random.default_rng(seed)
        n = len(self)
        self._push()
        self._pull()
        self._vel[:] = rng.normal(0.0, 1.0, size=(n, 3)) * numpy.sqrt(temperature / self._mass)[:, None]
        self._vel -= (self._mass[:, None] * self._vel).sum(axis=0) / self._mass.sum()
        self._host_vel_dirty = True
        self._thermo = None
        t_old = self.temperature
        if t_old > 0.0:
            self.scale_velocities((temperature / t_old) ** 0.5)

    def set_box(self, side):
        """New cell sides with the particles following affinely (fractional coordinates kept): the volume
        move of a barostat, `System.set_density` of the reference (system.py:150-157).  On the device."""
        self._push()
        side = numpy.asarray(side, dtype=numpy.float64) * numpy.ones(3)
        self._engine.set_box(side, self.cell.center)
        self.cell.side = side.copy()
        if self.interaction is not None:
            self.interaction._invalidate()
        self._advanced()

    def set_density(self, rho):
        self.set_box(self.cell.side * (self.density / rho) ** (1.0 / 3.0))

    def scale_velocities(self, factor):
        # system.py:227-230, on the device
        self._push()
        self._engine.scale_velocities(factor)
        self._advanced()

    # properties, as in the reference (system.py:361-369)
    @property
    def cm_velocity(self):
        t = self._get_thermo()
        return t['momentum'] / t['mass']

    @property
    def cm_position(self):
        """Mass-weighted mean of the (folded) positions, system.py:350-369."""
        self._pull()
        return (self._mass[:, None] * self._pos).sum(axis=0) / self._mass.sum()

    def cm(self, what):
        if what not in ('position', 'velocity'):
            raise ValueError('cm of %r is not defined for the array-backed system' % what)
        return self.cm_position if what == 'position' else self.cm_velocity

    def fix_momentum(self):
        vcm = self.cm_velocity
        self._pull()
        self._vel -= vcm
        self._host_vel_dirty = True
        self._thermo = None

    def fold(self):
        """Positions on the device are always folded; expose them folded (system.py:376-379)."""
        self._pull()
        return self

    def compute_interaction(self, what=None):
        # system.py:245-259 — but the state is already resident: no marshalling
        if self.interaction is not None:
            self.interaction.compute_resident(what, self)

    def potential_energy(self, per_particle=False, normed=False, cache=False):
        if self.interaction is None:
            return 0.0
        if not cache or self.interaction.energy is None:
            self.compute_interaction('energy')
        e = self.interaction.energy
        return e / len(self) if (per_particle or normed) else e

    def total_energy(self, per_particle=False, normed=None, cache=False):
        if normed is not None:
            per_particle = normed
        return self.potential_energy(per_particle, cache=cache) + self.kinetic_energy(per_particle)

    def force_norm_square(self, per_particle=True, cache=False):
        # system.py:305-319; the sum is reduced on the device
        if self.interaction is None:
            return 0.0
        f2 = self._get_thermo()['fnorm2']
        return f2 / len(self) if per_particle else f2

    def force_norm(self, per_particle=True, cache=False):
        if self.interaction is None:
            return 0.0
        f = self._get_thermo()['fnorm2'] ** 0.5
        return f / len(self) if per_particle else f

    def virial(self, per_particle=True, cache=False):
        if self.interaction is None:
            return 0.0
        w = self._get_thermo()['virial']
        return w / len(self) if per_particle else w

    @property
    def pressure(self):
        # system.py:337-348
        T = self.thermostat.temperature if (self.thermostat and self.thermostat.temperature is not None) \
            else self.temperature
        return (len(self) * T + self._get_thermo()['virial'] / 3) / self.cell.volume

    def dump(self, what=None, order='C', dtype=None, view=False, clear=False, flat=False):
        """Array marshalling compatible with system.py:398-530 for the attributes
        the path uses.  `view=True` hands out the host array itself."""
        if what is None:
            return None
        aliases = {'box': 'cell.side', 'pos': 'particle.position', 'vel': 'particle.velocity',
                   'spe': 'particle.species', 'rad': 'particle.radius',
                   'position': 'particle.position', 'velocity': 'particle.velocity',
                   'species': 'particle.species', 'mass': 'particle.mass', 'radius': 'particle.radius'}
        what = aliases.get(what, what)
        if what == 'cell.side':
            data = self.cell.side
        elif what == 'cell.center':
            data = self.cell.center
        elif what == 'particle.position':
            data = self.position if view else self._synced('_pos')
        elif what == 'particle.velocity':
            data = self.velocity if view else self._synced('_vel')
        elif what == 'particle.species':
            data = self.species
        elif what == 'particle.mass':
            data = self._mass
        elif what == 'particle.radius':
            data = self._radius
        else:
            raise ValueError('Unknown attribute %s' % what)
        if not view:
            data = numpy.array(data, dtype=dtype)
        elif dtype is not None and numpy.dtype(dtype) != data.dtype:
            data = data.astype(dtype)
        if data.ndim == 2 and order == 'F':
            data = data.T          # (3, N) Fortran-ordered view of the same memory
        if flat:
            data = data.flatten(order=order)
        return data

    def view(self, what=None, order='C', dtype=None, clear=False, flat=False):
        return self.dump(what, order=order, dtype=dtype, view=True, clear=clear, flat=flat)

    def _synced(self, name):
        self._pull()
        return getattr(self, name)

    def update(self, other, full=False, exclude=None, only=None):
        """Take the state of `other` (same particles), as `System.update` does (system.py:72-101)."""
        other = System.from_atooms(other, model=self.model, precision=self.precision) \
            if not isinstance(other, System) else other
        other._pull()
        self._pull()
        self._pos[:], self._vel[:] = other._pos, other._vel
        self._host_pos_dirty = self._host_vel_dirty = True
        self._thermo = None

    def __deepcopy__(self, memo):
        """An independent system with its own device context (tests/test_rumd.py:178-194)."""
        self._pull()
        self._resolve_species()
        new = System(self._pos, [self._names[i] for i in self._ids], Cell(self.cell.side.copy(), self.cell.center.copy()),
                     velocity=self._vel, mass=self._mass, radius=self._radius, model=self.model,
                     precision=self.precision, skin=self._skin(), device=self._engine.device,
                     library=self._engine.lib)
        new.thermostat = copy.deepcopy(self.thermostat, memo)
        new.barostat = copy.deepcopy(self.barostat, memo)
        new.reservoir = copy.deepcopy(self.reservoir, memo)
        self._pull_image()
        new._img[:] = self._img
        if self.interaction is not None:
            it = self.interaction
            # same engine settings as the original, not the defaults
            new.interaction = it.__class__(self.model, precision=getattr(it, 'precision', self.precision),
                                           skin=getattr(it, 'skin', self._skin()),
                                           device=getattr(it, 'device', self._engine.device),
                                           species=getattr(it, 'species', None),
                                           library=getattr(it, '_library', None) or self._engine.lib)
        return new

    def _skin(self):
        return getattr(self, '_skin_value', 0.3)

    def __str__(self):
        # array-based: Simulation._report prints str(system) at start (simulation/core.py:348-351)
        txt = 'system composed by N={} particles\n'.format(len(self))
        txt += 'with chemical composition C={}\n'.format(self.composition)
        txt += 'with chemical concentration x={}\n'.format(self.concentration)
        txt += 'enclosed in a cubic box at number density rho={:.6f}\n'.format(self.density)
        txt += 'interacting via {} on the B200 engine (fp{})'.format(self.model, self.precision)
        return txt

    def close(self):
        self._engine.close()
