"""
Oracle-A: NumPy fp64 O(N^2) restatement of the pair-interaction hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is imported by the product
package `atooms_b200`; only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may use it, and only as the checker.

Parity status: the reference tree (atooms 3.14.1) ships NO force code (removed
in 3.0.0, CHANGELOG.md:151-152); its backends shell out to LAMMPS / RUMD, which
are absent here.  This file therefore restates the *published* algorithm the
reference's call sites assume, and is pinned by every known-answer value the
reference's tests and fixtures hold for the path (tests/test_oracle_pins.py;
values in tests/golden/pins.json with their file:line).  Beyond those pins
(8-9 significant digits on energies/pressure, 5e-6 on per-particle forces)
fp64 parity is "parity unpinned": kernel-vs-oracle only.

Conventions followed (citations are into /root/reference):
  * outputs & zeroing ......... atooms/system/interaction.py:57-82
      energy (float), virial (float), forces (ndim, N), stress (ndim, ndim)
  * inputs .................... atooms/system/system.py:245-259 (position (ndim,N) F-order)
  * pressure / virial sign .... atooms/system/system.py:337-348  P = (N T + W/d) / V,
                                atooms/backends/lammps.py:128-133 W = (P - rho T) d V
      => W = sum_{i<j} r_ij . F_ij
  * minimum image ............. atooms/system/particle.py:129-130  d - rint(d/L) L
  * fold ...................... atooms/system/particle.py:90-99
  * kinetic T ................. atooms/system/system.py:164-179  dof = (N-1) d
  * LJ units, NVE = vel.Verlet  atooms/backends/lammps.py:110-133,236-238
  * KA table, cut&shift 2.5 s   tests/test_rumd.py:26-31, data/ka_rumd.ff:4-7
  * minimiser stop rule ....... atooms/optimization/core.py:30-36 (sum F^2 / N < tol)
      FIRE / SD parameters are NOT in the reference (LAMMPS defaults implied by
      backends/lammps.py:336).  We use Bitzek et al., PRL 97, 170201 (2006):
      N_min 5, f_inc 1.1, f_dec 0.5, alpha0 0.1, f_alpha 0.99, dt_max 10 dt.
      Minimiser parity is unpinned beyond the inequalities of
      tests/test_optimization.py:20-35.
"""
import numpy

__all__ = ['PairTable', 'lennard_jones', 'kob_andersen', 'compute', 'neighbor_list', 'quantize32',
           'fold', 'kinetic_energy', 'temperature', 'nve', 'fire', 'steepest_descent',
           'maxwell_boltzmann']


class PairTable(object):
    """12-6 Lennard-Jones parameters for `nsp` species.

    u(r) = 4 eps [(s/r)^12 - (s/r)^6] - shift,  r < rcut (strict), else 0.
    `rcut` is an absolute distance.  `shifted` => shift = u_LJ(rcut).
    """

    def __init__(self, eps, sigma, rcut, shifted=True):
        self.eps = numpy.atleast_2d(numpy.asarray(eps, dtype=numpy.float64))
        self.sigma = numpy.atleast_2d(numpy.asarray(sigma, dtype=numpy.float64))
        self.rcut = numpy.atleast_2d(numpy.asarray(rcut, dtype=numpy.float64))
        self.nsp = self.eps.shape[0]
        assert self.eps.shape == self.sigma.shape == self.rcut.shape == (self.nsp, self.nsp)
        self.shifted = bool(shifted)
        # Derived coefficients, evaluated in this exact order everywhere
        # (the C oracle and the CUDA library use the same expressions).
        s2 = self.sigma * self.sigma
        s6 = s2 * s2 * s2
        self.c6 = 4.0 * self.eps * s6
        self.c12 = 4.0 * self.eps * s6 * s6
        self.rc2 = self.rcut * self.rcut
        ir2 = 1.0 / self.rc2
        ir6 = ir2 * ir2 * ir2
        self.ushift = ir6 * (self.c12 * ir6 - self.c6) if self.shifted else numpy.zeros_like(self.eps)


def lennard_jones(rcut=2.5, shifted=True, eps=1.0, sigma=1.0):
    return PairTable([[eps]], [[sigma]], [[rcut]], shifted)


def kob_andersen(shifted=True, nsp=2):
    """KA 80:20 table (tests/test_rumd.py:26-31), rcut = 2.5 sigma_ab.

    nsp=3 adds a third species carrying A parameters (used for the N=1212
    three-type fixture, stated explicitly in SURVEY.md 8d C2(iii))."""
    eps = numpy.array([[1.0, 1.5], [1.5, 0.5]])
    sig = numpy.array([[1.0, 0.8], [0.8, 0.88]])
    if nsp == 3:
        e3 = numpy.ones((3, 3))
        s3 = numpy.ones((3, 3))
        e3[:2, :2], s3[:2, :2] = eps, sig
        e3[2, 1] = e3[1, 2] = 1.5
        s3[2, 1] = s3[1, 2] = 0.8
        eps, sig = e3, s3
    return PairTable(eps, sig, 2.5 * sig, shifted)


def _min_image(d, side):
    # atooms/system/particle.py:129-130
    return d - numpy.rint(d / side) * side


def fold(position, side, center=None):
    """atooms/system/particle.py:90-99 on an (N, 3) array; returns (folded, image shift)."""
    c = numpy.zeros(3) if center is None else numpy.asarray(center)
    p = position - c
    n = numpy.rint(p / side)
    return p - n * side + c, n.astype(numpy.int32)


def compute(position, species, side, table, chunk=512, per_particle=False):
    """Energy, forces (N,3), virial, stress (3,3) [+ pe_i, vir_i] by direct O(N^2) sum.

    Pair expressions (identical in oracle_c.c and the CUDA kernels):
        r2inv = 1/r2 ; r6inv = r2inv^3
        u  = r6inv (c12 r6inv - c6) - ushift
        w  = r6inv (12 c12 r6inv - 6 c6)          ( = r . F, the pair virial )
        F_i += (w r2inv) d ,  d = r_i - r_j (minimum image)
    Totals: U = sum_{i<j} u, W = sum_{i<j} w, stress_ab = sum_{i<j} d_a F_b;
    per particle: pe_i = 1/2 sum_j u_ij, vir_i = 1/2 sum_j w_ij (RUMD columns).
    """
    pos = numpy.ascontiguousarray(position, dtype=numpy.float64)
    spe = numpy.asarray(species, dtype=numpy.int64)
    side = numpy.asarray(side, dtype=numpy.float64)
    n = pos.shape[0]
    forces = numpy.zeros((n, 3))
    pe = numpy.zeros(n)
    vir = numpy.zeros(n)
    stress = numpy.zeros((3, 3))
    for start in range(0, n, chunk):
        stop = min(n, start + chunk)
        d = pos[start:stop, None, :] - pos[None, :, :]
        d = _min_image(d, side)
        r2 = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]
        ti = spe[start:stop, None]
        tj = spe[None, :]
        mask = r2 < table.rc2[ti, tj]
        mask[numpy.arange(stop - start), numpy.arange(start, stop)] = False
        r2s = numpy.where(mask, r2, 1.0)
        r2inv = 1.0 / r2s
        r6inv = r2inv * r2inv * r2inv
        c12 = table.c12[ti, tj]
        c6 = table.c6[ti, tj]
        u = numpy.where(mask, r6inv * (c12 * r6inv - c6) - table.ushift[ti, tj], 0.0)
        w = numpy.where(mask, r6inv * (12.0 * c12 * r6inv - 6.0 * c6), 0.0)
        fpair = w * r2inv
        fvec = fpair[..., None] * d
        forces[start:stop] = fvec.sum(axis=1)
        pe[start:stop] = 0.5 * u.sum(axis=1)
        vir[start:stop] = 0.5 * w.sum(axis=1)
        stress += 0.5 * numpy.einsum('ija,ijb->ab', d, fvec)
    out = {'energy': float(pe.sum()), 'forces': forces, 'virial': float(vir.sum()),
           'stress': stress}
    if per_particle:
        out['pe'] = pe
        out['vir'] = vir
    return out


def quantize32(position, side, center=None):
    """Positions as the fp32 engine holds them: 32-bit fixed-point fractions of
    the cell, x_int = round(x / L * 2^32) (wrapping), returned as float64
    coordinates together with the integers."""
    c = numpy.zeros(3) if center is None else numpy.asarray(center, dtype=numpy.float64)
    side = numpy.asarray(side, dtype=numpy.float64)
    f = (numpy.asarray(position, dtype=numpy.float64) - c) / side
    f -= numpy.rint(f)
    k = numpy.rint(f * 4294967296.0).astype(numpy.int64)
    k = ((k + 2 ** 31) % 2 ** 32 - 2 ** 31)
    return k * (side / 4294967296.0) + c, k.astype(numpy.int32)


def neighbor_list(position, species, side, table, skin, dtype=numpy.float64, chunk=512, center=None):
    """Brute-force full Verlet list: j in list(i) iff r2 < (rcut_ab + skin)^2, strict.

    The predicate is evaluated exactly as the engine of that precision does,
    so that membership is bit-comparable:
      float64: d = x_i - x_j ; d = d - rint(d / L) * L   (particle.py:129-130)
      float32: d = float32(wrapped int32 difference of the fixed-point
               coordinates) * float32(L / 2^32)
      r2 = (dx*dx + dy*dy) + dz*dz in `dtype`, no FMA contraction.
    Returns a list of sorted int32 arrays (original numbering).
    """
    spe = numpy.asarray(species, dtype=numpy.int64)
    rl = (table.rcut + skin).astype(dtype)
    rl2 = rl * rl
    fixed = numpy.dtype(dtype) == numpy.float32
    if fixed:
        _, ipos = quantize32(position, side, center)
        scale = (numpy.asarray(side, dtype=numpy.float64) / 4294967296.0).astype(numpy.float32)
    else:
        pos = numpy.ascontiguousarray(position, dtype=numpy.float64)
        side = numpy.asarray(side, dtype=numpy.float64)
    n = len(spe)
    out = []
    for start in range(0, n, chunk):
        stop = min(n, start + chunk)
        if fixed:
            w = (ipos[start:stop, None, :].astype(numpy.uint32) - ipos[None, :, :].astype(numpy.uint32))
            d = w.astype(numpy.int32).astype(numpy.float32) * scale
        else:
            d = pos[start:stop, None, :] - pos[None, :, :]
            d = d - numpy.rint(d / side) * side
        r2 = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]
        assert r2.dtype == dtype
        mask = r2 < rl2[spe[start:stop, None], spe[None, :]]
        mask[numpy.arange(stop - start), numpy.arange(start, stop)] = False
        for k in range(stop - start):
            out.append(numpy.nonzero(mask[k])[0].astype(numpy.int32))
    return out


def kinetic_energy(velocity, mass):
    # atooms/system/particle.py:109-112 summed, system.py:232-243
    return float(0.5 * numpy.sum(numpy.asarray(mass)[:, None] * velocity * velocity))


def temperature(velocity, mass):
    # atooms/system/system.py:164-179
    n, d = velocity.shape
    if n > 1:
        return 2.0 / ((n - 1) * d) * kinetic_energy(velocity, mass)
    return 2.0 / d * kinetic_energy(velocity, mass) if n == 1 else 0.0


def maxwell_boltzmann(n, T, mass, seed):
    """MB velocities, COM removed, rescaled to exactly T (system.py:185-198),
    with numpy's default_rng instead of random.gauss."""
    rng = numpy.random.default_rng(seed)
    mass = numpy.broadcast_to(numpy.asarray(mass, dtype=numpy.float64), (n,))
    v = rng.normal(0.0, 1.0, size=(n, 3)) * numpy.sqrt(T / mass)[:, None]
    v -= (mass[:, None] * v).sum(axis=0) / mass.sum()
    v *= numpy.sqrt(T / temperature(v, mass))
    return v


def nve(position, velocity, mass, species, side, table, dt, steps, center=None,
        callback=None):
    """Velocity-Verlet NVE (LAMMPS `fix nve`, backends/lammps.py:236-238).

    Written in the kick-drift form the CUDA path uses: after a half kick the
    velocities live at half steps; each step is F(x) -> v += dt F/m -> x += dt v,
    and a closing half kick returns v at the integer step.  Algebraically this
    is velocity-Verlet.  Positions are folded after every drift (the fused
    kernel does the same); `image` counts the box crossings.
    """
    pos = numpy.array(position, dtype=numpy.float64)
    vel = numpy.array(velocity, dtype=numpy.float64)
    m = numpy.broadcast_to(numpy.asarray(mass, dtype=numpy.float64), (pos.shape[0],))[:, None]
    image = numpy.zeros(pos.shape, dtype=numpy.int32)
    if steps <= 0:
        return pos, vel, image
    f = compute(pos, species, side, table)['forces']
    vel = vel - (0.5 * dt) * f / m          # v(t - dt/2)
    for step in range(steps):
        if step > 0:
            f = compute(pos, species, side, table)['forces']
        vel = vel + dt * f / m
        pos = pos + dt * vel
        pos, shift = fold(pos, side, center)
        image += shift
        if callback is not None:
            callback(step + 1, pos, vel)
    f = compute(pos, species, side, table)['forces']
    vel = vel + (0.5 * dt) * f / m
    return pos, vel, image


FIRE_DEFAULTS = dict(n_min=5, f_inc=1.1, f_dec=0.5, alpha0=0.1, f_alpha=0.99, dtmax_factor=10.0)


def fire(position, mass, species, side, table, dt, max_steps, fn2_tol, center=None,
         **params):
    """FIRE (Bitzek et al. 2006) with a semi-implicit Euler step.

    Per iteration, with x, v, F(x) at the same time level:
        fn2 = sum F^2 / N ; stop if fn2 < tol       (optimization/core.py:34-36)
        P = F.v
        if P > 0: npos += 1; if npos > n_min: dt = min(dt f_inc, dt_max), alpha *= f_alpha
                  v = (1 - alpha) v + alpha |v|/|F| F      [with alpha as it was BEFORE this update]
        else:     npos = 0; dt *= f_dec; alpha = alpha0; v = 0
        v += dt F / m ; x += dt v ; fold ; F = F(x)
    Returns (pos, steps_done, converged, energy, fn2).
    """
    p = dict(FIRE_DEFAULTS)
    p.update(params)
    pos = numpy.array(position, dtype=numpy.float64)
    n = pos.shape[0]
    m = numpy.broadcast_to(numpy.asarray(mass, dtype=numpy.float64), (n,))[:, None]
    vel = numpy.zeros_like(pos)
    dt_now, dt_max, alpha, npos = dt, dt * p['dtmax_factor'], p['alpha0'], 0
    out = compute(pos, species, side, table)
    it = 0
    converged = False
    while True:
        f = out['forces']
        ff = float(numpy.sum(f * f))
        if ff / n < fn2_tol:
            converged = True
            break
        if it >= max_steps:
            break
        power = float(numpy.sum(f * vel))
        vv = float(numpy.sum(vel * vel))
        if power > 0.0:
            vel = (1.0 - alpha) * vel + (alpha * numpy.sqrt(vv / ff)) * f
            npos += 1
            if npos > p['n_min']:
                dt_now = min(dt_now * p['f_inc'], dt_max)
                alpha *= p['f_alpha']
        else:
            npos = 0
            dt_now *= p['f_dec']
            alpha = p['alpha0']
            vel = numpy.zeros_like(pos)
        vel = vel + dt_now * f / m
        pos = pos + dt_now * vel
        pos, _ = fold(pos, side, center)
        out = compute(pos, species, side, table)
        it += 1
    return pos, it, converged, out['energy'], ff / n


def steepest_descent(position, species, side, table, step, max_steps, fn2_tol,
                     max_disp=0.1, center=None):
    """Steepest descent with a capped step that adapts to gradient agreement.

    Per iteration, with F = F(x) and F_prev the force of the previous iterate:
        fn2 = sum F^2 / N ; stop if fn2 < tol            (optimization/core.py:34-36)
        if it > 0:  h_now *= 1.1 if F . F_prev > 0 else 0.5   (overshoot => halve)
        h = min(h_now, max_disp / max_i |F_i|) ;  x += h F ; fold
    No energy evaluations and no rejected moves, so one force evaluation per
    iteration.  Returns (pos, steps_done, converged, energy, fn2).
    """
    pos = numpy.array(position, dtype=numpy.float64)
    n = pos.shape[0]
    prev = numpy.zeros_like(pos)
    h_now = step
    it = 0
    converged = False
    while True:
        out = compute(pos, species, side, table)
        f = out['forces']
        ff = float(numpy.sum(f * f))
        if ff / n < fn2_tol:
            converged = True
            break
        if it >= max_steps:
            break
        if it > 0:
            h_now *= 1.1 if float(numpy.sum(f * prev)) > 0.0 else 0.5
        fmax = float(numpy.sqrt(numpy.max(numpy.sum(f * f, axis=1))))
        h = min(h_now, max_disp / fmax)
        pos, _ = fold(pos + h * f, side, center)
        prev = f
        it += 1
    return pos, it, converged, out['energy'], ff / n
