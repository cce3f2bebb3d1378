"""One FIRE minimisation of 4096 KA N=150 replicas (BASELINE.json configs[3]) for profiling k_batch_minimize."""
import os
import sys

import numpy

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from atooms_b200.batch import ReplicaEnsemble   # noqa: E402

precision = int(sys.argv[1]) if len(sys.argv) > 1 else 64
fr = numpy.load(os.path.join(ROOT, 'tests', 'golden', 'ka_N150_frames.npz'))
nrep = 4096
pos = numpy.empty((nrep, 150, 3))
for k in range(nrep):
    pos[k] = fr['position'][k % 80] + numpy.random.default_rng(100 + k).uniform(-0.05, 0.05, size=(150, 3))
ens = ReplicaEnsemble(pos, fr['species'], fr['side'], 'kob_andersen', precision=precision)
r = ens.minimize('fire', tolerance=1e-10 if precision == 64 else 1e-7, steps=100000)
print(int(r['converged'].sum()), int(r['steps'].sum()))
