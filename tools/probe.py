"""Quick device-time probe of the NVE loop (not the bench): KA replicated, fp32 + fp64."""
import os, sys, time
import numpy
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from atooms_b200 import capi

def ka_system(reps, seed=2):
    d = numpy.load(os.path.join(ROOT, 'tests', 'golden', 'ka_N1200.npz'))
    side = d['side']; n0 = len(d['species'])
    scale = (n0 / 1.2 / numpy.prod(side)) ** (1 / 3.)
    pos0 = d['position'] * scale; side = side * scale
    sh = numpy.stack(numpy.meshgrid(*[numpy.arange(reps)] * 3, indexing='ij'), -1).reshape(-1, 3)
    pos = (pos0[None] + sh[:, None, :] * side).reshape(-1, 3) - 0.5 * (reps - 1) * side
    rng = numpy.random.default_rng(seed)
    pos += rng.uniform(-0.01, 0.01, pos.shape)
    n = len(pos)
    vel = rng.normal(0, numpy.sqrt(0.8), (n, 3)); vel -= vel.mean(0)
    return pos, numpy.tile(d['species'], reps ** 3), side * reps, vel

if __name__ == '__main__':
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 200
    pos, spe, side, vel = ka_system(reps)
    n = len(pos)
    eps = numpy.array([[1.0, 1.5], [1.5, 0.5]]); sig = numpy.array([[1.0, 0.8], [0.8, 0.88]])
    for prec in (32, 64):
        e = capi.Engine(prec)
        e.set_box(side); e.set_pair_table(eps, sig, 2.5 * sig, True)
        t = time.time(); e.upload(pos, spe, vel); t_up = time.time() - t
        t = time.time(); r = e.compute(forces=False); t_c = time.time() - t
        e.run_nve(50, 0.002)
        th0 = e.thermo()
        e.run_nve(steps, 0.002); ms, launches = e.last_timing()
        th1 = e.thermo(); st = e.stats()
        print('N=%d fp%d upload %.3fs first-compute %.3fs U/N=%.6f | %d steps: %.3f ms/step, %.3e p-ts/s, launches %d, rebuilds %d, maxnb %d cap %d, dE/N=%.2e'
              % (n, prec, t_up, t_c, r['energy'] / n, steps, ms / steps, n * steps / (ms * 1e-3), launches,
                 st['rebuilds'], st['max_neighbors'], st['neighbor_capacity'],
                 (th1['kinetic'] + th1['potential'] - th0['kinetic'] - th0['potential']) / n), flush=True)
        e.close()
