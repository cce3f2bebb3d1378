"""
Ensembles of small independent replicas (BASELINE.json configs[3]).

The reference minimises one system per `Optimization` object, each chunk a
LAMMPS subprocess (atooms/optimization/core.py:39-74,
atooms/backends/lammps.py:314-363).  `ReplicaEnsemble` holds thousands of
small configurations that share cell and model and minimises them in one
kernel launch (one thread block per replica).  Replicas are independent: over
several GPUs they are dealt out round-robin, one process per GPU, with no
communication ("replicas only").
"""
import numpy

from . import capi, models

__all__ = ['ReplicaEnsemble']


class ReplicaEnsemble(object):

    def __init__(self, position, species, side, model='kob_andersen', mass=None, precision=64, device=0,
                 library=None, rank=0, world=1):
        """position (nrep, n, 3); species: names or ids, (n,) shared or (nrep, n).
        With world > 1 this rank keeps replicas rank, rank + world, ..."""
        self.model = models.get(model)
        pos = numpy.asarray(position, dtype=numpy.float64)
        if pos.ndim != 3 or pos.shape[2] != 3:
            raise ValueError('position must be (nrep, n, 3)')
        spe = numpy.asarray(species)
        if spe.dtype.kind not in 'iu':
            order = self.model.species if self.model.species is not None else sorted(set(spe.ravel().tolist()))
            lookup = {str(s): i for i, s in enumerate(order)}
            spe = numpy.vectorize(lambda s: lookup[str(s)])(spe)
        spe = numpy.broadcast_to(spe.astype(numpy.int32), pos.shape[:2])
        self.index = numpy.arange(pos.shape[0])[rank::world]      # global replica numbers held here
        self.side = numpy.asarray(side, dtype=numpy.float64)
        self.batch = capi.Batch(self.side, self.model.eps, self.model.sigma, self.model.rcut, self.model.shifted,
                                precision=precision, device=device, library=library)
        m = None if mass is None else numpy.broadcast_to(numpy.asarray(mass, dtype=numpy.float64), pos.shape[:2])[self.index]
        self.batch.upload(pos[self.index], spe[self.index], m)
        self.nrep, self.n = self.batch.nrep, self.batch.n
        self.result = None

    def minimize(self, method='fire', tolerance=1e-10, steps=100000, dt=None):
        """Minimise every replica until sum F^2 / N < tolerance
        (atooms/optimization/core.py:34-36; dryrun.py:140 default 1e-10) or `steps` iterations."""
        method = method.lower()
        if method not in ('fire', 'sd'):
            raise ValueError("method must be 'fire' or 'sd'")
        if dt is None:
            dt = 0.002 if method == 'fire' else 0.001
        self.result = self.batch.minimize(capi.MIN_FIRE if method == 'fire' else capi.MIN_SD, steps, dt, tolerance)
        return self.result

    def set_velocities(self, velocity):
        """velocity (nrep, n, 3) of ALL replicas (this rank keeps its share)."""
        self.batch.set_velocities(numpy.asarray(velocity, dtype=numpy.float64)[self.index])

    def run_nve(self, steps, dt=0.002):
        """`steps` velocity-Verlet steps of every replica in one kernel launch; returns
        per-replica kinetic / potential energy and virial of the final configuration."""
        self.result = self.batch.run_nve(steps, dt)
        return self.result

    def positions(self):
        return self.batch.download()

    def velocities(self):
        return self.batch.state()[0]

    def close(self):
        self.batch.close()
