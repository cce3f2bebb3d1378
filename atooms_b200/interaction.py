"""
`Interaction` — the pair interaction evaluated on the B200.

Drop-in for the reference's interaction protocol
(atooms/system/interaction.py:16-82): inputs are declared in `variables` and
arrive as keyword arguments from `System.compute_interaction`
(atooms/system/system.py:245-259); results are left as attributes `energy`,
`virial`, `forces` ((ndim, N), Fortran order, like `zeros_like(position)`),
`stress` ((ndim, ndim)), `hessian` ((ndim, N, ndim, N), dense: small systems only).

Two ways in:
  * `compute(observable, position, particle, cell)` — the generic protocol,
    for a plain `atooms.system.System`: host arrays go through the C ABI
    (`b2_upload` + `b2_compute`) on every call, as the LAMMPS interaction
    re-sends the whole system (atooms/backends/lammps.py:96-140).
  * `compute_resident(observable, system)` — used by our array-backed System,
    whose state already lives on the device: no marshalling at all.
"""
import numpy

from . import capi, models

__all__ = ['Interaction']

try:   # subclass the real protocol class when atooms is importable, so that isinstance checks hold
    from atooms.system.interaction import InteractionBase as _Base
except Exception:
    class _Base(object):
        """Local mirror of atooms.system.interaction.InteractionBase (interaction.py:16-55)."""

        def __init__(self):
            self.variables = {'position': 'particle.position'}
            self.order = 'F'
            self.observable = ['energy', 'forces', 'virial', 'stress', 'hessian']
            for name in self.observable:
                setattr(self, name, None)

        def __add__(self, other):
            total = _Base()
            for attr in self.observable:
                a, b = getattr(self, attr), getattr(other, attr)
                if a is None and b is None:
                    continue
                if a is not None and b is not None:
                    setattr(total, attr, a + b)
                else:
                    raise ValueError('attribute {} not set in {} or {}'.format(attr, self, other))
            return total

        def __radd__(self, other):
            return self if other == 0 else self.__add__(other)


class Interaction(_Base):

    def __init__(self, model='lennard_jones', precision=64, skin=0.3, device=0, species=None, library=None):
        _Base.__init__(self)
        self.model = models.get(model)
        self.precision = int(precision)
        self.skin = float(skin)
        self.device = device
        self.species = species if species is not None else self.model.species
        self._library = library
        # (3, N) Fortran-ordered positions are N x 3 interleaved in memory: passed as is
        self.variables = {'position': 'particle.position', 'particle': 'particle', 'cell': 'cell'}
        self.order = 'F'
        self._engine = None
        self._key = None
        self._valid_for = None

    def __str__(self):
        return 'b200 %s' % self.model

    # what the reference's HDF5 trajectory stores about an interaction (trajectory/hdf5.py:208-237)
    name = 'atomic'

    @property
    def potential(self):
        return self.model.potentials()

    def _invalidate(self):
        self.energy = self.virial = self.forces = self.stress = None

    # -- generic protocol ---------------------------------------------------------
    def compute(self, observable, position, particle, cell):
        """Host buffers in, attributes out; semantics of interaction.py:57-82."""
        if observable is None:
            observable = 'forces'
        assert observable in self.observable, 'unsupported observable {}'.format(observable)
        ndim, n = position.shape
        if ndim != 3:
            raise ValueError('the B200 engine is three-dimensional')
        names = [str(p.species) for p in particle]
        order = self.species if self.species is not None else sorted(set(names))
        lookup = {s: i for i, s in enumerate(order)}
        ids = numpy.array([lookup[s] for s in names], dtype=numpy.int32)
        mass = numpy.array([p.mass for p in particle], dtype=numpy.float64)
        key = (n, tuple(numpy.asarray(cell.side, dtype=float)), tuple(numpy.asarray(cell.center, dtype=float)))
        if self._engine is None or key != self._key:
            if self._engine is not None:
                self._engine.close()
            self._engine = capi.Engine(self.precision, device=self.device, library=self._library)
            self._engine.set_box(cell.side, cell.center)
            self._engine.set_model(self.model)
            self._engine.set_skin(self.skin)
            self._key = key
        # host -> device: the whole configuration, every call (the generic protocol is stateless)
        self._engine.upload(position, ids, None, mass)
        if observable == 'hessian':                  # interaction.py:76-82: shape (ndim, N, ndim, N)
            self.hessian = self._engine.hessian()
            return
        self._fill(observable, self._engine, position)

    def _fill(self, observable, engine, like=None):
        want_forces = observable in ('forces', 'stress')
        out = engine.compute(forces=want_forces, stress=(observable == 'stress'))
        self.energy = out['energy']
        if observable == 'energy':
            return
        self.virial = out['virial']
        forces = out['forces']                       # (N, 3) C order == (3, N) F order in memory
        self.forces = forces.T if like is None or like.flags.f_contiguous else numpy.ascontiguousarray(forces.T)
        if observable == 'stress':
            self.stress = out['stress']

    def neighbors(self, system=None):
        """The engine's Verlet list (pairs within r_cut + skin) of the configuration last
        computed — or of `system`, an array-backed System bound to a backend — as one sorted
        int32 array per particle, 0-based, in the caller's numbering.  This is what the
        reference stores in `particle.neighbors` / `system.neighbors` and writes with
        TrajectoryNeighbors (atooms/trajectory/xyz.py:476-506)."""
        engine = system._engine if system is not None and getattr(system, '_engine', None) is not None else self._engine
        if engine is None:
            raise ValueError('no configuration resident: call compute() first')
        if system is not None and hasattr(system, '_push'):
            system._push()
        return engine.neighbors()

    # -- resident state -------------------------------------------------------------
    def compute_resident(self, observable, system):
        if observable is None:
            observable = 'forces'
        assert observable in self.observable, 'unsupported observable {}'.format(observable)
        if observable == 'hessian':
            system._push()
            self.hessian = system._engine.hessian()
            return
        th = system._get_thermo()            # one device reduction, cached until the state changes
        self.energy = th['potential']
        if observable == 'energy':
            # accessors called with cache=True expect the virial of the same state
            self.virial = th['virial']
            return
        self._fill(observable, system._engine)
