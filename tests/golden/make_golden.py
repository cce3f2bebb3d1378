#!/usr/bin/env python
"""
Generate the golden fixtures under tests/golden/ from the reference tree.

Run HERE (the build container), never on the GPU box:

    python tests/golden/make_golden.py [/root/reference]

It imports the *unmodified* reference package from /root/reference, reads its
data fixtures with the reference's own trajectory readers and stores the raw
arrays as .npz files, plus `pins.json`, which holds the known-answer values
quoted from the reference's tests (each with its file:line).  The .npz files
and pins.json are committed; /root/reference does not exist on the GPU box.

numpy >= 2 removed numpy.set_string_function, which the reference's
TrajectoryXYZ.close() still calls (atooms/trajectory/xyz.py:472-474).  We
install a no-op shim in THIS process only; the reference is not touched.
"""
import json
import os
import sys
import tempfile

import numpy

REF = sys.argv[1] if len(sys.argv) > 1 else '/root/reference'
HERE = os.path.dirname(os.path.abspath(__file__))

if not hasattr(numpy, 'set_string_function'):
    numpy.set_string_function = lambda *a, **k: None
sys.path.insert(0, REF)

from atooms.trajectory import TrajectoryXYZ, TrajectoryRUMD  # noqa: E402


def species_ids(system):
    """0-based ids by position in the sorted list of distinct species
    (atooms/system/system.py:117-122)."""
    distinct = system.distinct_species
    return numpy.array([distinct.index(p.species) for p in system.particle],
                       dtype=numpy.int32), [str(s) for s in distinct]


def arrays(system, extra=()):
    out = {}
    out['position'] = numpy.array([p.position for p in system.particle], dtype=numpy.float64)
    out['velocity'] = numpy.array([p.velocity for p in system.particle], dtype=numpy.float64)
    out['mass'] = numpy.array([p.mass for p in system.particle], dtype=numpy.float64)
    out['species'], names = species_ids(system)
    out['species_names'] = numpy.array(names)
    out['side'] = numpy.array(system.cell.side, dtype=numpy.float64)
    out['center'] = numpy.array(system.cell.center, dtype=numpy.float64)
    for name in extra:
        out[name] = numpy.array([float(getattr(p, name)) for p in system.particle], dtype=numpy.float64)
    return out


def save(name, **data):
    path = os.path.join(HERE, name + '.npz')
    numpy.savez_compressed(path, **data)
    n = data['position'].shape
    print('wrote %-28s position%s side=%s' % (name + '.npz', n, data['side']))


def first_frame(cls, fname, frame=0, extra=()):
    th = cls(os.path.join(REF, 'data', fname))
    system = th[frame]
    meta = {'steps': list(th.steps), 'timestep': th.timestep}
    th.close()
    return arrays(system, extra), meta


def main():
    # --- single-frame fixtures -------------------------------------------
    for fname, key in [('lj_N256_rho1.0.xyz', 'lj_N256'),
                       ('lj_N1000_rho1.0.xyz', 'lj_N1000'),
                       ('lj_fcc_N108.xyz', 'lj_fcc_N108'),
                       ('trajectory.xyz', 'ka_N1200')]:
        data, _ = first_frame(TrajectoryXYZ, fname)
        save(key, **data)

    # RUMD file with per-particle forces / pe / vir columns (fp32 golden vector)
    data, meta = first_frame(TrajectoryRUMD, 'ka_N256_rho1.185_rumd.xyz.gz',
                             extra=('fx', 'fy', 'fz', 'pe', 'vir'))
    data['timestep'] = numpy.float64(meta['timestep'])
    save('ka_N256_rumd', **data)

    # N=1212, three types: positions + species only (pe/vir unusable)
    data, _ = first_frame(TrajectoryRUMD, 'ka_rumd_N1200.xyz.gz')
    save('ka_N1212_rumd', **data)

    # --- 80 frames of KA N=150 (replica ensemble seeds) --------------------
    th = TrajectoryXYZ(os.path.join(REF, 'data', 'ka_N150_rho1.200.xyz'))
    frames = []
    for i in range(len(th.steps)):
        system = th[i]
        frames.append(numpy.array([p.position for p in system.particle]))
    spe, names = species_ids(system)
    side = numpy.array(system.cell.side)
    th.close()
    numpy.savez_compressed(os.path.join(HERE, 'ka_N150_frames.npz'),
                           position=numpy.array(frames), species=spe,
                           species_names=numpy.array(names), side=side,
                           center=numpy.zeros(3))
    print('wrote ka_N150_frames.npz', numpy.array(frames).shape, side)

    # --- inline 3- and 4-particle RUMD configurations -----------------------
    # text quoted from the reference test file (tests/test_backends.py:25-40)
    src = open(os.path.join(REF, 'tests', 'test_backends.py')).read()
    ns = {}
    start = src.index('xyz = """')
    stop = src.index('xyz_io2 = """')
    exec(src[start:stop], ns)
    for key, name in [('xyz', 'lj_N3_inline'), ('xyz_2', 'lj_N4_inline')]:
        tmp = tempfile.mkdtemp()
        path = os.path.join(tmp, 'c.xyz')
        with open(path, 'w') as fh:
            fh.write(ns[key])
        th = TrajectoryRUMD(path)
        system = th[0]
        th.close()
        save(name, **arrays(system))

    # --- pins ----------------------------------------------------------------
    pins = {
        'lj_N1000_unshifted_U_per_particle': {
            'value': -4.2446836, 'source': 'tests/test_lammps.py:109', 'rel_tol': 2e-8,
            'setup': 'lj_N1000, LJ eps=sigma=1 rc=2.5 unshifted'},
        'lj_fcc_N108_unshifted_U_per_particle': {
            'value': -7.7615881, 'source': 'tests/test_lammps.py:119', 'places': 7},
        'lj_fcc_N108_unshifted_pressure': {
            'value': -3.3935748, 'source': 'tests/test_lammps.py:121', 'places': 7,
            'setup': 'T=0 (no velocities) so P = W/(3V)'},
        'lj_fcc_N108_force_norm_per_particle': {
            'value': 4.2e-11, 'source': 'tests/test_lammps.py:120', 'places': 2},
        'lj_N1000_nve_x0': {
            'value': [3.62635, 3.64526, 3.675987], 'steps': [0, 10, 20], 'dt': 0.005,
            'source': 'tests/test_lammps.py:33-40', 'places': 5,
            'setup': 'lj_N1000 with file velocities, unshifted LJ 2.5, velocity-Verlet NVE'},
        'pair_r1.122_unshifted_U': {
            'value': -0.99999388, 'source': 'docs/basics.org:166-183', 'places': 7},
        'lj_N3_shifted_U': {
            'value': 36.9236726612, 'source': 'tests/test_backends.py:95-98', 'places': 4},
        'lj_N3_temperature': {
            'value': 1.0, 'source': 'tests/test_backends.py:96', 'places': 7},
        'lj_N4_temperature_mass': {
            'value': 10.0 / 9.0, 'source': 'tests/test_backends.py:101-105', 'places': 7},
        'ka_N256_shifted_U_per_particle': {
            'value': -6.4014371344819665, 'source': 'tests/test_rumd.py:187', 'places': 3},
        'lj_N256_minimised': {
            'U_per_particle_below': -6.8, 'fn2_per_particle_below': 1e-4,
            'source': 'tests/test_optimization.py:20-35'},
        'lj_N1000_nvt_temperature': {
            'value': 2.0, 'places': 1, 'source': 'tests/test_lammps.py:42-98',
            'setup': 'lj_N1000, unshifted LJ 2.5, start at T=1.5 (or 1.4), fix nvt temp 2.0 2.0 0.2, dt=0.002, '
                     'mean of T sampled every 500 steps after 1500 steps'},
        'ka_table': {
            'eps': [[1.0, 1.5], [1.5, 0.5]], 'sigma': [[1.0, 0.8], [0.8, 0.88]],
            'rcut_over_sigma': 2.5, 'source': 'tests/test_rumd.py:26-31; data/ka_rumd.ff:4-7'},
    }
    with open(os.path.join(HERE, 'pins.json'), 'w') as fh:
        json.dump(pins, fh, indent=1, sort_keys=True)
    print('wrote pins.json')


if __name__ == '__main__':
    main()
