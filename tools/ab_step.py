"""A/B of the step kernels at the bench shape (KA N = 1.2M): windowed list (default) against the
global-index list (B2_LIST_GLOBAL=1, read once per process).  Prints one JSON line per call.
    python tools/ab_step.py [precision] [md_steps]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                      # noqa: E402
from atooms_b200 import capi     # noqa: E402

precision = int(sys.argv[1]) if len(sys.argv) > 1 else 32
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
pos, spe, side, vel = bench.ka_workload()
lib = capi.load()
eng = bench.make_engine(lib, precision, pos, spe, side, vel)
eng.compute(forces=False)
eng.set_profile(True)
eng.run_nve(200, bench.DT)
t0 = eng.thermo()
prof = {'ms_steps': 0.0, 'ms_refresh': 0.0, 'steps': 0, 'refreshes': 0}
import time
tw = time.perf_counter()
for _ in range(steps // 100):
    eng.run_nve(100, bench.DT)
    p = eng.profile()
    for k in prof:
        prof[k] += p[k]
wall = time.perf_counter() - tw
t1 = eng.thermo()
st = eng.stats()
n = len(spe)
print(json.dumps({'list': 'global' if os.environ.get('B2_LIST_GLOBAL') else 'windowed', 'precision': precision,
                  'list_mode': st['list_mode'], 'ncells': st['ncells'], 'cap': st['neighbor_capacity'],
                  'step_kernel_ms': prof['ms_steps'] / max(prof['steps'], 1),
                  'refresh_ms_each': prof['ms_refresh'] / max(prof['refreshes'], 1), 'refreshes': prof['refreshes'],
                  'p_ts_per_s_wall': n * steps / wall,
                  'e0': (t0['kinetic'] + t0['potential']) / n, 'e1': (t1['kinetic'] + t1['potential']) / n}))
