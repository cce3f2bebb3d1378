"""Small end-to-end exercise of every kernel family (for compute-sanitizer runs)."""
import os, sys
import numpy
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
from atooms_b200 import capi
from atooms_b200.batch import ReplicaEnsemble
g = lambda n: numpy.load(os.path.join(ROOT, 'tests', 'golden', n + '.npz'))
eps = numpy.array([[1.0, 1.5], [1.5, 0.5]]); sig = numpy.array([[1.0, 0.8], [0.8, 0.88]])
d = g('ka_N1200')
rng = numpy.random.default_rng(0)
# 2x2x2 tiling: cell-list mode with M = 2
sh = numpy.array([[i, j, k] for i in range(2) for j in range(2) for k in range(2)])
pos = (d['position'][None] + sh[:, None, :] * d['side']).reshape(-1, 3)
spe = numpy.tile(d['species'], 8)
vel = rng.normal(0, 1.0, pos.shape); vel -= vel.mean(0)
for prec in (32, 64):
    e = capi.Engine(prec); e.set_box(d['side'] * 2, d['side'] * 0.5); e.set_pair_table(eps, sig, 2.5 * sig, True)
    e.upload(pos, spe, vel)
    r = e.compute(per_particle=True, stress=True)
    e.run_nve(120, 0.002)
    nl = e.neighbors()
    e.update(position=e.download()[0])
    print(prec, r['energy'] / len(spe), e.minimize(capi.MIN_FIRE, 60, 0.002, 1e-30), e.minimize(capi.MIN_SD, 40, 0.001, 1e-30),
          e.thermo()['potential'] / len(spe), len(nl), e.stats()['rebuilds'])
    e.close()
# small box: all-pairs list
d2 = g('lj_N256')
e = capi.Engine(64); e.set_box(d2['side']); e.set_pair_table([[1.0]], [[1.0]], [[2.5]], True)
e.upload(d2['position'], d2['species']); e.run_nve(50, 0.002); print('N256', e.thermo()['potential'] / 256); e.close()
fr = g('ka_N150_frames')
ens = ReplicaEnsemble(fr['position'][:6], fr['species'], fr['side'], 'kob_andersen', precision=32)
print(ens.minimize('fire', 1e-30, 50)['energy'][:2]); ens.set_velocities(rng.normal(0, 0.5, (6, 150, 3))); print(ens.run_nve(50)['kinetic'][:2])
print('sanity ok')
