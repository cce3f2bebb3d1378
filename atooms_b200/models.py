"""
Pair-potential tables for the engine.

The reference defines its potentials inside the external engines
(`rumd.Pot_LJ_12_6(...).SetParams(i, j, Epsilon, Sigma, Rcut)`,
tests/test_rumd.py:26-31; LAMMPS `pair_style lj/cut`, `pair_coeff`,
`pair_modify shift yes`, tests/test_lammps.py:24-27).  Here a model is a plain
table of 12-6 Lennard-Jones parameters per species pair.
"""
import numpy

__all__ = ['PairModel', 'TabulatedModel', 'lennard_jones', 'kob_andersen', 'wca', 'from_potentials', 'get',
           'FUNCTIONS']


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

    def potentials(self):
        """The table in the reference's own description of a pair interaction: one record per
        species pair with `func`, `params`, `species` (1-based), `cutoff.scheme` ('cs' cut and
        shifted, 'c' cut) and `cutoff.radius` — the fields of `_PairPotential` / `_CutOff` that the
        reference's HDF5 trajectories store (atooms/trajectory/hdf5.py:21-43, 208-237)."""
        out = []
        for a in range(self.nsp):
            for b in range(a, self.nsp):
                out.append(_Potential('lennard_jones', {'epsilon': float(self.eps[a, b]), 'sigma': float(self.sigma[a, b])},
                                      [a + 1, b + 1], _CutOff('cs' if self.shifted else 'c', float(self.rcut[a, b]))))
        return out

    def __str__(self):
        return '%s(nsp=%d, rcut_max=%g, %s)' % (self.name, self.nsp, self.rcut.max(),
                                                'cut&shifted' if self.shifted else 'cut')


def _lj(r, epsilon=1.0, sigma=1.0):
    x = (sigma / r) ** 6
    return 4 * epsilon * (x * x - x), 24 * epsilon * (2 * x * x - x)


def _inverse_power(r, epsilon=1.0, sigma=1.0, exponent=12):
    u = epsilon * (sigma / r) ** exponent
    return u, exponent * u


def _gaussian(r, epsilon=1.0, sigma=1.0):
    u = epsilon * numpy.exp(-(r / sigma) ** 2)
    return u, 2 * (r / sigma) ** 2 * u


def _harmonic(r, epsilon=1.0, sigma=1.0):
    x = numpy.clip(1.0 - r / sigma, 0.0, None)           # soft spheres: 0.5 eps (1 - r/sigma)^2 inside sigma
    return 0.5 * epsilon * x * x, epsilon * x * r / sigma


def _yukawa(r, epsilon=1.0, sigma=1.0, kappa=1.0):
    u = epsilon * sigma / r * numpy.exp(-kappa * (r - sigma))
    return u, u * (1.0 + kappa * r)


# name -> f(r, **params) returning (u, w = -r du/dr): what `func` of the reference's schema may name
FUNCTIONS = {'lennard_jones': _lj, 'lj': _lj, 'inverse_power': _inverse_power, 'gaussian': _gaussian,
             'harmonic': _harmonic, 'yukawa': _yukawa}


class TabulatedModel(object):

    """Any pair potential, through lookup tables (`b2_set_pair_tabulated`).

    One record per species pair in the reference's schema (atooms/trajectory/hdf5.py:21-43): `func` — a name
    of FUNCTIONS or a callable f(r, **params) -> (u, w) with w = -r du/dr, or -> u alone (w is then taken by
    central differences) — `params`, `species` (1-based pair), `cutoff.scheme` ('c' cut, 'cs' cut and
    shifted, 'cf' cut and force-shifted: u - u(rc) - (r - rc) u'(rc)), `cutoff.radius`, `npoints`."""

    tabulated = True
    shifted = False

    def __init__(self, potentials, species=None, name='tabulated'):
        get_ = lambda o, k: o[k] if isinstance(o, dict) else getattr(o, k)
        self._pairs = {}
        npoints = []
        for phi in potentials:
            a, b = sorted(int(x) - 1 for x in get_(phi, 'species'))
            cut = get_(phi, 'cutoff')
            func = get_(phi, 'func')
            if not callable(func):
                if str(func) not in FUNCTIONS:
                    raise ValueError('unknown pair potential %r (known: %s; or pass a callable)'
                                     % (func, ', '.join(sorted(FUNCTIONS))))
                func = FUNCTIONS[str(func)]
            scheme = str(get_(cut, 'scheme')).lower()
            if scheme not in ('c', 'cs', 'cf'):
                raise ValueError("cut-off scheme %r is not supported ('c', 'cs', 'cf')" % scheme)
            self._pairs[(a, b)] = (func, dict(get_(phi, 'params')), scheme, float(get_(cut, 'radius')))
            try:
                npoints.append(int(get_(phi, 'npoints')))
            except (KeyError, AttributeError):
                npoints.append(20000)
        self.nsp = 1 + max(b for _, b in self._pairs)
        self.npoints = max(npoints)
        self.rcut = numpy.zeros((self.nsp, self.nsp))
        for a in range(self.nsp):
            for b in range(a, self.nsp):
                if (a, b) not in self._pairs:
                    raise ValueError('no potential for the species pair (%d, %d)' % (a + 1, b + 1))
                self.rcut[a, b] = self.rcut[b, a] = self._pairs[(a, b)][3]
        self.species = None if species is None else [str(s) for s in species]
        self.name = name
        self._tables = None

    @staticmethod
    def _evaluate(func, params, r):
        out = func(r, **params)
        if isinstance(out, tuple):
            return numpy.asarray(out[0], dtype=numpy.float64), numpy.asarray(out[1], dtype=numpy.float64)
        h = 1e-6 * r
        return numpy.asarray(out, dtype=numpy.float64), -r * (numpy.asarray(func(r + h, **params)) -
                                                              numpy.asarray(func(r - h, **params))) / (2 * h)

    def tables(self):
        """(u, w), each (nsp, nsp, npoints), at r^2 = k rcut^2 / (npoints - 1)."""
        if self._tables is None:
            u = numpy.zeros((self.nsp, self.nsp, self.npoints))
            w = numpy.zeros_like(u)
            for (a, b), (func, params, scheme, rc) in self._pairs.items():
                r = numpy.sqrt(numpy.arange(self.npoints) * (rc * rc / (self.npoints - 1)))
                r[0] = r[1]                                   # r = 0: any finite value will do
                with numpy.errstate(over='ignore', invalid='ignore'):
                    uu, ww = self._evaluate(func, params, r)
                    uc, wc = self._evaluate(func, params, numpy.array([rc]))
                if scheme in ('cs', 'cf'):
                    uu = uu - uc[0]
                if scheme == 'cf':                            # force-shifted: F(rc) = 0 as well
                    uu = uu + (r - rc) * wc[0] / rc           # u'(rc) = -w(rc) / rc
                    ww = ww - wc[0] * r / rc
                big = numpy.finfo(numpy.float32).max / 16     # keep the float copy of w finite
                u[a, b] = u[b, a] = numpy.clip(numpy.nan_to_num(uu, nan=big, posinf=big), -big, big)
                w[a, b] = w[b, a] = numpy.clip(numpy.nan_to_num(ww, nan=big, posinf=big), -big, big)
            self._tables = (u, w)
        return self._tables

    def potentials(self):
        return [_Potential(getattr(f, '__name__', str(f)).lstrip('_'), p, [a + 1, b + 1], _CutOff(s, rc), self.npoints)
                for (a, b), (f, p, s, rc) in sorted(self._pairs.items())]

    def __str__(self):
        return '%s(nsp=%d, rcut_max=%g, tabulated with %d points)' % (self.name, self.nsp, self.rcut.max(), self.npoints)


class _CutOff(object):
    def __init__(self, scheme, radius):
        self.scheme, self.radius = scheme, radius


class _Potential(object):
    interacting_bodies = 2

    def __init__(self, func, params, species, cutoff, npoints=20000):
        self.func, self.params, self.species, self.cutoff, self.npoints = func, params, species, cutoff, npoints

    def __str__(self):
        return self.func


def from_potentials(potentials, species=None):
    """Build the model from records in the reference's schema (see PairModel.potentials): objects
    or dicts with func / params / species / cutoff.  12-6 Lennard-Jones with one cut-off scheme
    ('c' or 'cs') for all pairs becomes the native table; anything else — other functions of
    FUNCTIONS, callables, mixed or force-shifted cut-offs — a TabulatedModel."""
    get_ = lambda o, k: o[k] if isinstance(o, dict) else getattr(o, k)
    potentials = list(potentials)
    native = all(str(get_(phi, 'func')) in ('lennard_jones', 'lj') for phi in potentials) and \
        len(set(str(get_(get_(phi, 'cutoff'), 'scheme')).lower() for phi in potentials)) == 1 and \
        str(get_(get_(potentials[0], 'cutoff'), 'scheme')).lower() in ('c', 'cs')
    if not native:
        return TabulatedModel(potentials, species=species)
    pairs, schemes = {}, set()
    for phi in potentials:
        func = str(get_(phi, 'func'))
        if func not in ('lennard_jones', 'lj'):
            raise ValueError('only lennard_jones pair potentials are supported, not %r' % func)
        a, b = (int(x) - 1 for x in get_(phi, 'species'))
        cut = get_(phi, 'cutoff')
        scheme = str(get_(cut, 'scheme')).lower()
        if scheme not in ('c', 'cs'):
            raise ValueError('cut-off scheme %r is not supported (c: cut, cs: cut and shifted)' % scheme)
        schemes.add(scheme)
        par = get_(phi, 'params')
        pairs[(min(a, b), max(a, b))] = (float(par['epsilon']), float(par['sigma']), float(get_(cut, 'radius')))
    if len(schemes) != 1:
        raise ValueError('one cut-off scheme for all pairs is required')
    nsp = 1 + max(b for _, b in pairs)
    eps, sig, rc = numpy.zeros((nsp, nsp)), numpy.ones((nsp, nsp)), numpy.zeros((nsp, nsp))
    for a in range(nsp):
        for b in range(a, nsp):
            if (a, b) not in pairs:
                raise ValueError('no potential for the species pair (%d, %d)' % (a + 1, b + 1))
            e, s_, r = pairs[(a, b)]
            eps[a, b] = eps[b, a] = e
            sig[a, b] = sig[b, a] = s_
            rc[a, b] = rc[b, a] = r
    return PairModel(eps, sig, rc, schemes.pop() == 'cs', species=species, name='lennard_jones')


def wca(epsilon=1.0, sigma=1.0):
    """Weeks-Chandler-Andersen: the repulsive part of LJ, cut at its minimum 2^(1/6) sigma and
    shifted up by epsilon."""
    return PairModel([[epsilon]], [[sigma]], [[2.0 ** (1.0 / 6.0) * sigma]], True, name='wca')


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
             'kob_andersen': kob_andersen, 'ka': kob_andersen, 'wca': wca}


def get(model):
    """Accept a PairModel, a registered name, or a dict(eps=, sigma=, rcut=, shifted=)."""
    if isinstance(model, (PairModel, TabulatedModel)):
        return model
    if isinstance(model, (list, tuple)):                 # records in the reference's schema
        return from_potentials(model)
    if isinstance(model, str):
        try:
            return _REGISTRY[model]()
        except KeyError:
            raise ValueError('unknown model %r (known: %s)' % (model, ', '.join(sorted(_REGISTRY))))
    if isinstance(model, dict):
        return PairModel(**model)
    raise ValueError('cannot build a pair model from %r' % (model,))
