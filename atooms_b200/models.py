"""
Pair-potential tables for the engine.

The reference defines its potentials inside the external engines
(`rumd.Pot_LJ_12_6(...).SetParams(i, j, Epsilon, Sigma, Rcut)`,
tests/test_rumd.py:26-31; LAMMPS `pair_style lj/cut`, `pair_coeff`,
`pair_modify shift yes`, tests/test_lammps.py:24-27).  Here a model is a plain
table of 12-6 Lennard-Jones parameters per species pair.
"""
import numpy

__all__ = ['PairModel', 'lennard_jones', 'kob_andersen', 'get']


class PairModel(object):

    """u(r) = 4 eps [(sigma/r)^12 - (sigma/r)^6] (- u(rcut) if shifted) for r < rcut.

    `rcut` is an absolute distance per pair (RUMD's `Rcut` is in units of
    sigma: pass rcut = Rcut * sigma).  `species` lists the species names in
    table order; None means "sorted distinct species of the system"
    (atooms/system/system.py:117-122).
    """

    def __init__(self, eps, sigma, rcut, shifted=True, species=None, name='lj'):
        self.eps = numpy.atleast_2d(numpy.asarray(eps, dtype=numpy.float64))
        self.sigma = numpy.atleast_2d(numpy.asarray(sigma, dtype=numpy.float64))
        self.rcut = numpy.atleast_2d(numpy.asarray(rcut, dtype=numpy.float64))
        self.nsp = self.eps.shape[0]
        if not (self.eps.shape == self.sigma.shape == self.rcut.shape == (self.nsp, self.nsp)):
            raise ValueError('eps, sigma and rcut must be square matrices of equal size')
        if not (numpy.allclose(self.eps, self.eps.T) and numpy.allclose(self.sigma, self.sigma.T)
                and numpy.allclose(self.rcut, self.rcut.T)):
            raise ValueError('pair tables must be symmetric')
        self.shifted = bool(shifted)
        self.species = None if species is None else [str(s) for s in species]
        self.name = name

    def __str__(self):
        return '%s(nsp=%d, rcut_max=%g, %s)' % (self.name, self.nsp, self.rcut.max(),
                                                'cut&shifted' if self.shifted else 'cut')


def lennard_jones(rcut=2.5, shifted=True, epsilon=1.0, sigma=1.0):
    """One-component LJ (`pair_style lj/cut 2.5`, tests/test_lammps.py:24-27)."""
    return PairModel([[epsilon]], [[sigma]], [[rcut]], shifted, name='lennard_jones')


def kob_andersen(shifted=True, rcut=2.5, nsp=2):
    """Kob-Andersen 80:20 binary mixture, cut at `rcut` sigma_ab (tests/test_rumd.py:26-31,
    data/ka_rumd.ff).  nsp=3 appends a species with A parameters (for the
    three-type fixture data/ka_rumd_N1200.xyz.gz)."""
    eps = numpy.array([[1.0, 1.5], [1.5, 0.5]])
    sig = numpy.array([[1.0, 0.8], [0.8, 0.88]])
    if nsp == 3:
        e3, s3 = numpy.ones((3, 3)), numpy.ones((3, 3))
        e3[:2, :2], s3[:2, :2] = eps, sig
        e3[2, 1] = e3[1, 2] = 1.5
        s3[2, 1] = s3[1, 2] = 0.8
        eps, sig = e3, s3
    elif nsp != 2:
        raise ValueError('kob_andersen: nsp must be 2 or 3')
    return PairModel(eps, sig, rcut * sig, shifted, name='kob_andersen')


_REGISTRY = {'lennard_jones': lennard_jones, 'lj': lennard_jones,
             'kob_andersen': kob_andersen, 'ka': kob_andersen}


def get(model):
    """Accept a PairModel, a registered name, or a dict(eps=, sigma=, rcut=, shifted=)."""
    if isinstance(model, PairModel):
        return model
    if isinstance(model, str):
        try:
            return _REGISTRY[model]()
        except KeyError:
            raise ValueError('unknown model %r (known: %s)' % (model, ', '.join(sorted(_REGISTRY))))
    if isinstance(model, dict):
        return PairModel(**model)
    raise ValueError('cannot build a pair model from %r' % (model,))
