"""
Trajectory fast path for the array-backed System.

The reference's writers format one particle at a time with `str.format`
(atooms/trajectory/xyz.py:437-474, rumd.py:88-98) and `TrajectoryXYZ` relies on
`numpy.set_string_function`, which numpy >= 2 removed; its
in-memory trajectory deep-copies N Particle objects per frame
(atooms/trajectory/ram.py:22-33).  With an engine at ~1e10 particle-steps/s those
loops are the wall time at any realistic cadence.  Here

  Trajectory        xyz files in the reference's layout (`N`, then
                    `step:.. columns:.. dt:.. cell:..`, then one line per particle), written
                    from the flat arrays with ONE formatting call per block of 65536 particles,
                    and read back the same way (mode 'r': `steps`, `len`, indexing, iteration);
                    the reference's reader parses what this writes, and vice versa.
  TrajectoryRUMD    the same for RUMD's xyz dialect (integer types, `ioformat=1 ...` header,
                    atooms/trajectory/rumd.py:69-98).
  TrajectoryRam     frames kept in memory as array snapshots (ram.py's interface: write, update
                    of an existing step, `th[i]` returns a copy, `th[i] = system`).
  Snapshot          what reading returns: an array-backed frame with the members the reference's
                    System offers to analysis code (`particle` materialised lazily, `cell`,
                    `dump`, `density`, ...); `System.from_atooms` / `B200(...)` accept it.

All provide what `Simulation`'s observers use (atooms/simulation/observers.py:167-207):
`timestep`, `precision`, `variables`, `write(system, step)`, `close()`, context manager.
"""
import copy
import os

import numpy

__all__ = ['Trajectory', 'TrajectoryRUMD', 'TrajectoryRam', 'TrajectoryNeighbors', 'Snapshot']

_BLOCK = 65536        # particles formatted per call

_ALIASES = {'pos': 'position', 'vel': 'velocity', 'spe': 'species', 'rad': 'radius', 'particle.position': 'position',
            'particle.velocity': 'velocity', 'particle.species': 'species', 'particle.mass': 'mass',
            'particle.radius': 'radius'}
_WIDTH = {'position': 3, 'velocity': 3, 'species': 1, 'mass': 1, 'radius': 1}


class _Cell(object):

    def __init__(self, side, center=None):
        self.side = numpy.array(side, dtype=numpy.float64)
        self.center = numpy.zeros_like(self.side) if center is None else numpy.array(center, dtype=numpy.float64)
        self.shape = 'cubic'
        self.periodic = numpy.ones(self.side.size, dtype=bool)

    @property
    def volume(self):
        return float(numpy.prod(self.side))


class Snapshot(object):

    """One frame as flat arrays.  `particle` builds Particle objects only when indexed."""

    def __init__(self, position, species, cell, velocity=None, mass=None, radius=None, step=0):
        self.position = numpy.array(position, dtype=numpy.float64).reshape(-1, 3)
        n = len(self.position)
        self.species = numpy.array(species)
        self.velocity = numpy.zeros((n, 3)) if velocity is None else numpy.array(velocity, dtype=numpy.float64).reshape(n, 3)
        self.mass = numpy.ones(n) if mass is None else numpy.array(mass, dtype=numpy.float64)
        self.radius = numpy.full(n, 0.5) if radius is None else numpy.array(radius, dtype=numpy.float64)
        self.cell = cell if hasattr(cell, 'side') else _Cell(cell)
        self.step = step
        self.thermostat = self.barostat = self.reservoir = self.interaction = None

    @classmethod
    def of(cls, system, step=0):
        """Copy of the state of a System of this package, a Snapshot, or a reference System."""
        if isinstance(system, Snapshot):
            out = copy.deepcopy(system)
            out.step = step
            return out
        if hasattr(system, 'dump') and hasattr(system, '_ids'):          # atooms_b200.System: arrays, no objects
            return cls(system.dump('position'), system.species, _Cell(system.cell.side, system.cell.center),
                       system.dump('velocity'), system.dump('mass'), system.dump('radius'), step)
        p = system.particle
        return cls([q.position for q in p], [q.species for q in p],
                   _Cell(system.cell.side, getattr(system.cell, 'center', None)),
                   [q.velocity for q in p], [q.mass for q in p], [getattr(q, 'radius', 0.5) for q in p], step)

    def __len__(self):
        return len(self.position)

    number_of_dimensions = 3

    @property
    def particle(self):
        return _Particles(self)

    @particle.setter
    def particle(self, particles):
        """A new list of particle objects (the reference's callbacks `sort` / `filter_species` assign one,
        atooms/trajectory/decorators.py:91-99)."""
        particles = list(particles)
        n = len(particles)
        self.position = numpy.array([q.position for q in particles], dtype=numpy.float64).reshape(n, 3)
        self.velocity = numpy.array([q.velocity for q in particles], dtype=numpy.float64).reshape(n, 3)
        self.species = numpy.array([q.species for q in particles])
        self.mass = numpy.array([q.mass for q in particles], dtype=numpy.float64)
        self.radius = numpy.array([getattr(q, 'radius', 0.5) for q in particles], dtype=numpy.float64)

    def _assign(self, i, name, value):
        """`particle[i].species = ...` / `.mass = ...` write through (atooms_b200.system._bound_particle_class)."""
        if name == 'species':
            if self.species.dtype != object:
                self.species = self.species.astype(object)      # (a fixed-width string array would truncate)
            self.species[i] = value
        elif name == 'radius':
            self.radius[i] = float(value)
        elif name in ('position', 'velocity'):
            pass                                             # (written through the row view already)
        else:
            self.mass[i] = float(value)

    @property
    def distinct_species(self):
        return sorted(set(self.species.tolist()))

    @property
    def cm_position(self):
        return (self.mass[:, None] * self.position).sum(axis=0) / self.mass.sum()

    @property
    def cm_velocity(self):
        return (self.mass[:, None] * self.velocity).sum(axis=0) / self.mass.sum()

    def kinetic_energy(self, per_particle=False, normed=False):
        k = 0.5 * float((self.mass[:, None] * self.velocity ** 2).sum())
        return k / len(self) if (per_particle or normed) else k

    @property
    def temperature(self):
        n = len(self)
        return 2.0 * self.kinetic_energy() / (3 * (n - 1)) if n > 1 else 0.0      # (system.py: ndim (N - 1) degrees of freedom)

    @temperature.setter
    def temperature(self, T):
        self.set_temperature(T)

    # the host-side members of atooms/system/system.py:186-208, 227-230, 330-379 — on the arrays
    @property
    def concentration(self):
        return {k: v / float(len(self)) for k, v in self.composition.items()}

    @property
    def packing_fraction(self):
        return float(numpy.pi / 6 * ((2.0 * self.radius) ** 3).sum() / self.cell.volume)

    def cm(self, what):
        return (self.mass[:, None] * getattr(self, what)).sum(axis=0) / self.mass.sum()

    def scale_velocities(self, factor):
        self.velocity *= factor

    def fix_momentum(self):
        self.velocity -= self.cm_velocity

    def fold(self):
        center = getattr(self.cell, 'center', 0.0)
        r = self.position - center
        self.position[:] = r - numpy.rint(r / self.cell.side) * self.cell.side + center

    def set_density(self, rho):
        factor = (self.density / rho) ** (1.0 / 3)
        self.position *= factor
        self.cell.side = self.cell.side * factor

    def set_temperature(self, temperature, seed=None):
        """Maxwell-Boltzmann velocities at `temperature`, total momentum zero, rescaled to it exactly."""
        rng = numpy.random.default_rng(seed)
        self.velocity[:] = rng.normal(size=self.velocity.shape) * numpy.sqrt(temperature / self.mass)[:, None]
        self.fix_momentum()
        current = self.temperature
        if current > 0:
            self.velocity *= numpy.sqrt(temperature / current)

    @property
    def composition(self):
        names, counts = numpy.unique(self.species, return_counts=True)
        return {str(k): int(v) for k, v in zip(names, counts)}

    @property
    def density(self):
        return len(self) / self.cell.volume if len(self) else 0.0

    def dump(self, what=None, order='C', dtype=None, view=False, clear=False, flat=False):
        what = _ALIASES.get(what, what)
        if what in ('cell.side', 'box'):
            data = self.cell.side
        elif what in _WIDTH:
            data = getattr(self, what)
        else:
            raise ValueError('Unknown attribute %s' % what)
        data = data if view else numpy.array(data, dtype=dtype)
        if data.ndim == 2 and order == 'F':
            data = data.T
        return data.flatten(order=order) if flat else data

    view = dump

    def update(self, other, full=False, exclude=None, only=None):
        other = Snapshot.of(other, self.step)
        self.position[:], self.velocity[:] = other.position, other.velocity
        self.species, self.mass, self.radius = other.species, other.mass, other.radius
        self.cell = other.cell


class _Particles(object):

    def __init__(self, frame):
        self._f = frame

    def __len__(self):
        return len(self._f)

    def _make(self, i):
        from .system import _bound_particle_class
        f = self._f
        sp = f.species[i]
        p = _bound_particle_class()(position=f.position[i], velocity=f.velocity[i],
                                    species=sp.item() if hasattr(sp, 'item') else sp, mass=f.mass[i], radius=f.radius[i])
        p.__dict__['_b2_owner'] = (f, i)         # species / mass assigned to the object reach the frame's arrays
        return p

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self._make(k) for k in range(*i.indices(len(self)))]
        if i < 0:
            i += len(self)
        if not 0 <= i < len(self):
            raise IndexError('particle index out of range')
        return self._make(i)

    def __iter__(self):
        return (self._make(i) for i in range(len(self)))


def _columns(variables):
    return [_ALIASES.get(v, v).replace('particle.', '') for v in variables]


def _format_rows(handle, blocks, formats):
    """Write len(blocks[0]) lines; blocks are 1-d (strings / integers) or 2-d (floats) arrays and `formats`
    their per-value format.  One `%` call per _BLOCK lines: no per-particle Python."""
    n = len(blocks[0])
    row = ' '.join(' '.join([f] * (1 if b.ndim == 1 else b.shape[1])) for b, f in zip(blocks, formats)) + '\n'
    width = sum(1 if b.ndim == 1 else b.shape[1] for b in blocks)
    for lo in range(0, n, _BLOCK):
        hi = min(n, lo + _BLOCK)
        cells = numpy.empty((hi - lo, width), dtype=object)
        k = 0
        for b in blocks:
            w = 1 if b.ndim == 1 else b.shape[1]
            cells[:, k:k + w] = b[lo:hi].reshape(hi - lo, w)       # numpy scalars -> Python objects, in C
            k += w
        handle.write((row * (hi - lo)) % tuple(cells.ravel().tolist()))


def _arrays_of(system, columns):
    """The columns of `system` as arrays: a System of this package, a Snapshot, or a reference System."""
    if not (hasattr(system, '_ids') or isinstance(system, Snapshot)):
        system = Snapshot.of(system)                       # a reference System: one pass over its Particle list
    return [numpy.asarray(system.dump(c)) for c in columns]


def block_size_of(steps):
    """Frames per block: the period after which the increments between steps recur (1 for equally spaced frames,
    the whole trajectory when nothing recurs) — what atooms/trajectory/utils.py:98-133 finds for exponentially
    spaced samples like 1, 2, 4, 11, 12, 14, ..."""
    steps = list(steps)
    if len(steps) < 2:
        return 1
    inc = [b - a for a, b in zip(steps[:-1], steps[1:])]
    if all(d == inc[0] for d in inc):
        return 1
    for period in range(1, len(inc)):
        if all(d == inc[k % period] for k, d in enumerate(inc)):
            return period
    return len(steps)


class _BaseMembers(object):

    """The members of atooms/trajectory/base.py:212-527 that analysis code built on the reference uses besides
    read / write: time axis, block size, the read_* / write_* hooks, append and copy."""

    def append(self, system):
        self.write(system)

    def _next_step(self, step):
        steps = self.steps
        return (steps[-1] + 1 if len(steps) else 0) if step is None else step

    @property
    def times(self):
        dt = self.timestep if self.timestep is not None else 1.0
        return [step * dt for step in self.steps]

    @property
    def total_time(self):
        dt = self.timestep if self.timestep is not None else 1.0
        return (self.steps[-1] - self.steps[0]) * dt

    @property
    def block_size(self):
        fixed = self.__dict__.get('_block_size')
        return block_size_of(self.steps) if fixed is None else fixed

    @block_size.setter
    def block_size(self, value):
        self.__dict__['_block_size'] = value

    @property
    def grandcanonical(self):
        return False

    @property
    def fields(self):
        return self.variables

    @fields.setter
    def fields(self, value):
        self.variables = value

    def read_len(self):
        return len(self)

    def read_steps(self):
        return list(self.steps)

    def read_timestep(self):
        return self.timestep if self.timestep is not None else 1.0

    def read_block_size(self):
        return None

    def write_init(self, system):
        pass

    def write_system(self, system, step):
        self.write(system, step)

    def write_timestep(self, value):
        self.timestep = value

    def write_block_size(self, value):
        pass

    def copy(self, cls=None, fout=None, steps=None, **ignored):
        """The frames (those at `steps`, if given) written to another trajectory of class `cls`; returns it, closed
        (base.py:384-443)."""
        cls = self.__class__ if cls is None else cls
        if fout is None:
            root, ext = os.path.splitext(getattr(self, 'filename', None) or 'trajectory')
            fout = root + '.copy.' + getattr(cls, 'suffix', ext.lstrip('.') or 'xyz')
        if fout == getattr(self, 'filename', None):
            raise IOError('copy: trying to overwrite the input trajectory')
        out = cls(fout, 'w')
        out.variables, out.timestep = list(self.variables), self.timestep
        if getattr(self, 'precision', None) is not None:
            out.precision = self.precision
        for k, step in enumerate(self.steps):
            if steps is None or step in steps:
                out.write(self[k], step)
        out.close()
        return out


class Trajectory(_BaseMembers):

    """xyz trajectory in the reference's layout (atooms/trajectory/xyz.py:135-474)."""

    suffix = 'xyz'
    _names = {'id': 'species', 'name': 'species', 'type': 'species', 'spe': 'species', 'pos': 'position',
              'vel': 'velocity', 'rad': 'radius'}

    def __init__(self, filename, mode='r'):
        self.filename = filename
        self.mode = mode
        self.timestep = None
        self.precision = 6
        self.variables = ['species', 'position']
        self._fh = None
        self.callbacks = []                    # read mode: applied to every frame (atooms/trajectory/base.py:237-249,331-341)
        self._frames = []                      # read mode: (step, offset of the first particle line, n, header dict)
        self._written = []                     # write mode: steps written so far
        if mode in ('w', 'a'):
            base = os.path.dirname(filename)
            if base and not os.path.exists(base):
                os.makedirs(base)
            self._fh = self._open(filename, mode)
        elif mode == 'r':
            self._fh = self._open(filename, 'r')
            self._index()
        else:
            raise ValueError('unknown mode %r' % mode)

    @staticmethod
    def _open(filename, mode):
        """Plain text, or gzip when the name says so (RUMD keeps its configurations as .xyz.gz,
        atooms/trajectory/rumd.py:10-98): read in one go, the frame index then works on the text in memory."""
        if not filename.endswith('.gz'):
            return open(filename, mode)
        import gzip
        import io
        if mode == 'r':
            with gzip.open(filename, 'rt') as fh:
                return io.StringIO(fh.read())
        return gzip.open(filename, mode + 't')

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def close(self):
        if self._fh is not None:
            self._fh.close()
            self._fh = None

    # ---- writing ---------------------------------------------------------------------
    def _header(self, step, columns, system):
        line = 'step:{} columns:{} '.format(step, ','.join(columns))
        if self.timestep is not None:
            line += 'dt:{:g} '.format(self.timestep)
        return line + 'cell:{} '.format(','.join(repr(float(x)) for x in system.cell.side))

    def _species_column(self, names):
        return names.astype(object) if names.dtype.kind in 'US' else names

    def write(self, system, step=None):
        if self._fh is None or self.mode == 'r':
            raise IOError('trajectory %s is not open for writing' % self.filename)
        step = self._next_step(step)           # (base.py:267-275: one past the last step when none is given)
        self._written.append(step)
        columns = _columns(self.variables)
        blocks = _arrays_of(system, columns)
        n = len(blocks[0]) if blocks else 0
        formats = []
        for k, (c, b) in enumerate(zip(columns, blocks)):
            if c == 'species':
                blocks[k], fmt = self._species_column(b), '%s'
            else:
                fmt = '%.{}f'.format(self.precision) if b.dtype.kind == 'f' else '%s'
            formats.append(fmt)
        self._fh.write('%d\n%s\n' % (n, self._header(step, columns, system)))
        if n and not self._write_native(columns, blocks):
            _format_rows(self._fh, blocks, formats)

    def _write_native(self, columns, blocks):
        """species + position [+ velocity] in three dimensions go through b2_format_xyz (all host cores)."""
        if columns not in (['species', 'position'], ['species', 'position', 'velocity']):
            return False
        if any(b.ndim != 2 or b.shape[1] != 3 or b.dtype.kind != 'f' for b in blocks[1:]):
            return False
        try:
            from . import capi
            lib = capi.load()
        except Exception:                                 # library not built: the Python formatter below
            return False
        names, ids = numpy.unique(numpy.asarray(blocks[0]).astype(str), return_inverse=True)
        raw = lib.format_xyz(ids, list(names), blocks[1], blocks[2] if len(blocks) > 2 else None, self.precision)
        self._fh.flush()
        self._fh.buffer.write(raw)
        return True

    # ---- reading ---------------------------------------------------------------------
    def _parse_header(self, line):
        meta = {}
        for token in line.replace('=', ':').split():
            if ':' in token:
                key, value = token.split(':', 1)
                meta[key] = value
        # RUMD's second header format (atooms/trajectory/rumd.py:56-67): the time step is the second field of
        # `integrator=`, the cell follows the box type in `sim_box=`
        if 'integrator' in meta and 'dt' not in meta:
            meta['dt'] = meta['integrator'].split(',')[1]
        if 'sim_box' in meta and 'cell' not in meta:
            meta['cell'] = ','.join(meta['sim_box'].split(',')[1:])
        return meta

    def _index(self):
        fh = self._fh
        fh.seek(0)
        while True:
            first = fh.readline()
            if not first.strip():
                break
            n = int(first)
            meta = self._parse_header(fh.readline())
            start = fh.tell()
            for _ in range(n):
                fh.readline()
            step = int(meta.get('step', meta.get('timeStepIndex', len(self._frames))))
            self._frames.append((step, start, n, meta))
        if self._frames and 'dt' in self._frames[0][3]:
            self.timestep = float(self._frames[0][3]['dt'])

    @property
    def steps(self):
        override = self.__dict__.get('_steps')
        if override is not None:
            return override
        return [f[0] for f in self._frames] if self.mode == 'r' else self._written

    @steps.setter
    def steps(self, value):                  # (the reference's `Sliced` decorator narrows them, decorators.py:162-165)
        self.__dict__['_steps'] = list(value)

    def __len__(self):
        return len(self.steps)

    def __iter__(self):
        return (self[i] for i in range(len(self)))

    def _layout(self, meta, ncol):
        cols = [self._names.get(c, c) for c in meta['columns'].split(',')] if 'columns' in meta else \
            (['species', 'position'] if ncol < 7 else ['species', 'position', 'velocity'])
        return cols

    def _build(self, meta, names, table, step):
        """A Snapshot from the columns named in the header; `table` holds the numeric columns."""
        cols, k, got = self._layout(meta, table.shape[1] + 1), 0, {}
        for c in cols:
            if c == 'species':
                continue
            w = _WIDTH.get(c, 1)
            got[c] = table[:, k:k + w]
            k += w
        side = numpy.array([float(x) for x in meta.get('cell', meta.get('boxLengths', '0,0,0')).split(',')])
        snap = Snapshot(got['position'], names, _Cell(side), got.get('velocity'),
                        got['mass'][:, 0] if 'mass' in got else None,
                        got['radius'][:, 0] if 'radius' in got else None, step)
        snap.extra = {c: v for c, v in got.items() if c not in _WIDTH}
        return snap

    # The template of atooms/trajectory/base.py:209-251, so that the reference's class decorators (`Sliced`,
    # `Unfolded`: they override read_init / read_system of whatever class they wrap) work on this one:
    # read(i) = read_init() once, read_system(i), callbacks.
    def read_init(self):
        return None

    def __getitem__(self, i):
        if i < 0:
            i += len(self)
        if not 0 <= i < len(self):
            raise IndexError('frame %d of %d' % (i, len(self)))
        if not getattr(self, '_initialized_read', False):
            self.read_init()
            self._initialized_read = True
        frame = self.read_system(i)
        for cbk, args, kwargs in list(self.callbacks):
            frame = cbk(frame, *args, **kwargs)
        return frame

    def read_system(self, i):
        if self._fh is None or self.mode != 'r':
            raise IOError('trajectory %s is not open for reading' % self.filename)
        step, start, n, meta = self._frames[i]
        self._fh.seek(start)
        tokens = numpy.array(''.join(self._fh.readline() for _ in range(n)).split())
        ncol = len(tokens) // max(n, 1)
        tokens = tokens.reshape(n, ncol)
        return self._build(meta, self._read_species(tokens[:, 0]), tokens[:, 1:].astype(numpy.float64), step)

    def _read_species(self, column):
        return column

    read = __getitem__

    def register_callback(self, cbk, *args, **kwargs):
        """A function `cbk(system, *args, **kwargs) -> system` applied to every frame read, the reference's
        own callbacks included (atooms/trajectory/decorators.py: center, fold, sort, filter_species,
        change_species, set_density, set_temperature, fix_cm)."""
        if (cbk, args, kwargs) not in self.callbacks:
            self.callbacks.append((cbk, args, kwargs))

    add_callback = register_callback


class TrajectoryRUMD(Trajectory):

    """RUMD's dialect (atooms/trajectory/rumd.py:10-98): integer types in the first column (the position of
    the species in the sorted list of distinct species), header `ioformat=1 dt=.. timeStepIndex=..
    boxLengths=.. numTypes=.. mass=.. columns=type,x,y,z,vx,vy,vz`."""

    def __init__(self, filename, mode='r'):
        Trajectory.__init__(self, filename, mode)
        self.variables = ['species', 'position', 'velocity']
        self.precision = 9

    def _header(self, step, columns, system):
        names = numpy.asarray(system.dump('species'))
        mass = numpy.asarray(system.dump('mass'))
        distinct = sorted(set(names.tolist()))
        first = [float(mass[numpy.argmax(names == s)]) for s in distinct]
        names = {'position': 'x,y,z', 'velocity': 'vx,vy,vz', 'species': 'type'}
        return 'ioformat=1 dt=%g timeStepIndex=%d boxLengths=%s numTypes=%d mass=%s columns=%s' % (
            self.timestep if self.timestep is not None else 1.0, step,
            ','.join('%.12f' % x for x in system.cell.side), len(distinct), ','.join('%g' % m for m in first),
            ','.join(names.get(c, c) for c in columns))

    def _species_column(self, names):
        distinct = sorted(set(names.tolist()))
        return numpy.searchsorted(numpy.array(distinct), names).astype(object)

    def _layout(self, meta, ncol):
        out, skip = [], 0
        for c in meta.get('columns', 'type,x,y,z').split(','):
            if skip:
                skip -= 1
                continue
            if c == 'type':
                out.append('species')
            elif c == 'x':
                out.append('position')
                skip = 2
            elif c == 'vx':
                out.append('velocity')
                skip = 2
            else:
                out.append(c)
        return out


class TrajectoryRam(_BaseMembers):

    """Frames kept in memory as array snapshots: the interface of atooms/trajectory/ram.py:7-47 (write
    replaces the frame of a step already present, `th[i]` returns a copy, `th[i] = system` writes)
    without a deep copy of N Particle objects per frame."""

    def __init__(self, filename=None, mode='w'):
        self.filename, self.mode = filename, mode
        self.timestep = None
        self.precision = None
        self.variables = ['species', 'position', 'velocity']
        self.steps = []
        self._frames = []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def close(self):
        pass

    def write(self, system, step=None):
        step = self._next_step(step)
        frame = Snapshot.of(system, step)
        if step in self.steps:
            self._frames[self.steps.index(step)] = frame
        else:
            self.steps.append(step)
            self._frames.append(frame)

    def __len__(self):
        return len(self._frames)

    def read_init(self):
        return None

    def read_system(self, i):
        return copy.deepcopy(self._frames[i])

    def __getitem__(self, i):
        frame = self.read_system(i)
        for cbk, args, kwargs in list(self.__dict__.get('callbacks', ())):
            frame = cbk(frame, *args, **kwargs)
        return frame

    read = __getitem__

    def register_callback(self, cbk, *args, **kwargs):
        callbacks = self.__dict__.setdefault('callbacks', [])
        if (cbk, args, kwargs) not in callbacks:
            callbacks.append((cbk, args, kwargs))

    add_callback = register_callback

    def __setitem__(self, i, value):
        try:
            step = self.steps[i]
        except IndexError:
            step = self.steps[-1] + 1 if self.steps else 0
        self.write(value, step)

    def __iter__(self):
        return (self[i] for i in range(len(self)))


class TrajectoryNeighbors(Trajectory):

    """Writer of the reference's neighbours format (atooms/trajectory/xyz.py:484-506): an xyz
    layout whose single column holds each particle's neighbours, comma-separated, numbered from
    `offset` (1 by default, as there).  `write` takes anything with a `neighbors` attribute (a
    list of integer arrays, 0-based — what `Interaction.neighbors()` returns) or the list itself."""

    def __init__(self, filename, mode='w', offset=1):
        Trajectory.__init__(self, filename, mode)
        self.variables = ['neighbors']
        self.offset = int(offset)

    def write(self, system, step):
        if self._fh is None:
            raise IOError('trajectory %s is not open for writing' % self.filename)
        rows = getattr(system, 'neighbors', system)
        line = 'step:{} columns:neighbors '.format(step)
        if self.timestep is not None:
            line += 'dt:{:g} '.format(self.timestep)
        cell = getattr(system, 'cell', None)
        if cell is not None:
            line += 'cell:{} '.format(','.join(repr(float(x)) for x in cell.side))
        self._fh.write('%d\n%s\n' % (len(rows), line))
        off = self.offset
        self._fh.write('\n'.join(','.join(map(str, (numpy.asarray(r, dtype=numpy.int64) + off).tolist())) for r in rows))
        self._fh.write('\n')
