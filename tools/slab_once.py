"""Step / refresh times of ONE context whose particles fill a slab of the box (what a rank of the slab
decomposition holds): the bench system placed in a box twice as long along x.  List format by environment:
B2_LIST_GLOBAL=1 | B2_LIST_WINDOWED=1 (all super-blocks) | B2_REGION_X=lo,hi (the slab's super-blocks only).
    python tools/slab_once.py [md_steps] [box length along x in units of the system's]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy                      # noqa: E402
import bench                      # noqa: E402
from atooms_b200 import capi     # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
pos, spe, side, vel = bench.ka_workload()
stretch = float(sys.argv[2]) if len(sys.argv) > 2 else 2.0
side = numpy.array(side, dtype=float) * numpy.array([stretch, 1.0, 1.0])
lib = capi.load()
eng = bench.make_engine(lib, 32, pos, spe, side, vel)
eng.compute(forces=False)
eng.set_profile(True)
eng.run_nve(100, bench.DT)
prof = {'ms_steps': 0.0, 'ms_refresh': 0.0, 'steps': 0, 'refreshes': 0}
for _ in range(steps // 100):
    eng.run_nve(100, bench.DT)
    p = eng.profile()
    for k in prof:
        prof[k] += p[k]
st = eng.stats()
print(json.dumps({'env': {k: os.environ[k] for k in ('B2_LIST_GLOBAL', 'B2_LIST_WINDOWED', 'B2_REGION_X') if k in os.environ},
                  'list_mode': st['list_mode'], 'cap': st['neighbor_capacity'],
                  'step_kernel_ms': prof['ms_steps'] / max(prof['steps'], 1),
                  'refresh_ms_each': prof['ms_refresh'] / max(prof['refreshes'], 1), 'refreshes': prof['refreshes']}))
