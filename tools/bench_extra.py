"""Timings of the non-headline configurations of BASELINE.json (for DESIGN.md / profiles/):
   C1  LJ N=256 rho=1.0 NVE 1e4 steps        (configs[0])
   C4  4096 x KA N=150 replicas, FIRE / SD   (configs[3])
"""
import os, sys, time, json
import numpy
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
from atooms_b200 import capi
from atooms_b200.batch import ReplicaEnsemble
from oracle import oracle_np as onp

out = {}
g = lambda n: numpy.load(os.path.join(ROOT, 'tests', 'golden', n + '.npz'))
# ---- C1 ----
d = g('lj_N256')
vel = onp.maxwell_boltzmann(256, 2.0, 1.0, seed=1)
for prec in (64, 32):
    e = capi.Engine(prec); e.set_box(d['side']); e.set_pair_table([[1.0]], [[1.0]], [[2.5]], True)
    e.upload(d['position'], d['species'], vel)
    t0 = e.thermo(); e.run_nve(1000, 0.002)
    t = time.perf_counter(); e.run_nve(10000, 0.002); wall = time.perf_counter() - t
    ms, launches = e.last_timing(); t1 = e.thermo()
    out['C1_fp%d' % prec] = {'steps': 10000, 'device_ms': ms, 'wall_s': wall, 'us_per_step': ms * 100 / 1000,
                             'particle_steps_per_s': 256 * 10000 / (ms * 1e-3), 'launches': launches,
                             'dE_per_particle': (t1['kinetic'] + t1['potential'] - t0['kinetic'] - t0['potential']) / 256}
    e.close()
# ---- C1 through the one-block-per-system kernel (whole run in one launch) ----
for prec in (64, 32):
    ens = ReplicaEnsemble(d['position'][None], d['species'], d['side'], 'lennard_jones', precision=prec)
    ens.set_velocities(vel[None])
    r0 = ens.run_nve(1000, 0.002)
    t = time.perf_counter(); r1 = ens.run_nve(10000, 0.002); wall = time.perf_counter() - t
    out['C1_single_block_fp%d' % prec] = {'steps': 10000, 'wall_s': wall, 'us_per_step': wall * 100,
                                          'particle_steps_per_s': 256 * 10000 / wall,
                                          'dE_per_particle': float(r1['energy'][0] + r1['kinetic'][0] - r0['energy'][0] - r0['kinetic'][0]) / 256}
    ens.close()
    many = ReplicaEnsemble(numpy.repeat(d['position'][None], 1184, axis=0), d['species'], d['side'], 'lennard_jones', precision=prec)
    many.set_velocities(numpy.repeat(vel[None], 1184, axis=0))
    many.run_nve(100, 0.002)
    t = time.perf_counter(); many.run_nve(2000, 0.002); wall = time.perf_counter() - t
    out['C1_x1184_replicas_fp%d' % prec] = {'replicas': 1184, 'steps': 2000, 'wall_s': wall,
                                            'particle_steps_per_s': 1184 * 256 * 2000 / wall}
    many.close()
# ---- C4 ----
fr = g('ka_N150_frames')
nrep = 4096
pos = numpy.empty((nrep, 150, 3))
for k in range(nrep):
    pos[k] = fr['position'][k % 80] + numpy.random.default_rng(100 + k).uniform(-0.05, 0.05, size=(150, 3))
for prec in (64, 32):
    for method, tol in (('fire', 1e-10 if prec == 64 else 1e-7), ('sd', 1e-10 if prec == 64 else 1e-7)):
        ens = ReplicaEnsemble(pos, fr['species'], fr['side'], 'kob_andersen', precision=prec)
        t = time.perf_counter(); r = ens.minimize(method, tolerance=tol, steps=100000); wall = time.perf_counter() - t
        it = int(r['steps'].sum())
        out['C4_%s_fp%d' % (method, prec)] = {
            'replicas': nrep, 'tolerance': tol, 'wall_s': wall, 'converged': int(r['converged'].sum()),
            'iterations_total': it, 'iterations_mean': it / nrep, 'iterations_max': int(r['steps'].max()),
            'particle_iterations_per_s': it * 150 / wall, 'pair_evaluations_per_s': it * 150 * 149 / wall,
            'mean_U_per_particle': float(r['energy'].mean() / 150)}
        ens.close()
print(json.dumps(out, indent=1))
