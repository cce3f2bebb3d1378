"""FIRE minima of LJ N=256 (data/lj_N256_rho1.0.xyz) from ten starts jittered by 1e-4, both
precisions, on the CUDA engine: the spread behind tests/test_gpu_parity.py::
test_fire_reaches_the_reference_inequalities.  Writes gpurun_out/fire_energies.json."""
import json
import os
import sys

import numpy

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'tests'))
sys.path.insert(0, ROOT)
from conftest import make_engine, golden   # noqa: E402
from oracle import oracle_np as onp        # noqa: E402
from atooms_b200 import capi               # noqa: E402

lib = capi.load()
d = dict(golden('lj_N256').items())
lj = onp.lennard_jones(2.5, True)
out = {64: [], 32: []}
for seed in range(10):
    dd = dict(d)
    if seed:
        dd['position'] = d['position'] + numpy.random.default_rng(seed).uniform(-1e-4, 1e-4, size=d['position'].shape)
    for prec, tol in ((64, 1e-8 / 256), (32, 1e-7)):
        eng = make_engine(lib, dd, lj, prec, velocities=False)
        steps, conv = eng.minimize(capi.MIN_FIRE, 100000, 0.002, tol)
        t = eng.thermo()
        out[prec].append([steps, bool(conv), t['potential'] / 256, t['fnorm2'] / 256])
        eng.close()
os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, 'gpurun_out', 'fire_energies.json'), 'w'), indent=1)
