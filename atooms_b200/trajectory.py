"""
Vectorised xyz trajectory writer for the array-backed System.

The reference's `TrajectoryXYZ.write_system` formats one particle at a time
with `str.format` and relies on `numpy.set_string_function`, which numpy >= 2
removed (atooms/trajectory/xyz.py:437-474).  This class writes
the same layout (`N`, then `step:.. columns:.. dt:.. cell:..`, then one line
per particle) from the flat arrays, so the reference's reader parses it back.
It provides the members `Simulation`'s observers use
(atooms/simulation/observers.py:167-207): `timestep`, `precision`,
`variables`, `write(system, step)`, `close()`, context manager.
"""
import os

import numpy

__all__ = ['Trajectory', 'TrajectoryNeighbors']


class Trajectory(object):

    suffix = 'xyz'

    def __init__(self, filename, mode='r'):
        self.filename = filename
        self.mode = mode
        self.timestep = None
        self.precision = 6
        self.variables = ['species', 'position']
        self._fh = None
        if mode in ('w', 'a'):
            base = os.path.dirname(filename)
            if base and not os.path.exists(base):
                os.makedirs(base)
            self._fh = open(filename, mode)
        elif mode != 'r':
            raise ValueError('unknown mode %r' % mode)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def write(self, system, step):
        if self._fh is None:
            raise IOError('trajectory %s is not open for writing' % self.filename)
        n = len(system.particle) if not hasattr(system, '_ids') else len(system)
        columns = [v.replace('particle.', '') for v in self.variables]
        blocks = []
        for var in columns:
            data = numpy.asarray(system.dump(var))
            if data.dtype.kind in 'fc':
                fmt = '%.{}f'.format(self.precision)
                data = data.reshape(n, -1)
                blocks.append([' '.join(fmt % x for x in row) for row in data])
            else:
                blocks.append([str(x) for x in data])
        line = 'step:{} columns:{} '.format(step, ','.join(columns))
        if self.timestep is not None:
            line += 'dt:{:g} '.format(self.timestep)
        line += 'cell:{} '.format(','.join(repr(float(x)) for x in system.cell.side))
        self._fh.write('%d\n%s\n' % (n, line))
        self._fh.write('\n'.join(' '.join(fields) for fields in zip(*blocks)))
        self._fh.write('\n')

    def close(self):
        if self._fh is not None:
            self._fh.close()
            self._fh = None


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
