"""Compare step-kernel tuning variants (separately compiled .so files) on the bench workload."""
import glob, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from atooms_b200 import capi
pos, spe, side, vel = bench.ka_workload()
n = len(spe)
for path in sorted(glob.glob(os.path.join(ROOT, 'atooms_b200', 'lib', 'variant_*.so'))):
    lib = capi.Library(path)
    for prec in (32, 64):
        eng = bench.make_engine(lib, prec, pos, spe, side, vel)
        eng.set_profile(True)
        eng.run_nve(100, bench.DT)
        ms = ks = kr = 0.0; st = 0
        for _ in range(3):
            eng.run_nve(100, bench.DT)
            t, _l = eng.last_timing(); p = eng.profile()
            ms += t; ks += p['ms_steps']; kr += p['ms_refresh']; st += p['steps']
        print('%-22s fp%d  %.3f ms/step  kernel %.3f  refresh/step %.3f  -> %.3e p-ts/s' %
              (os.path.basename(path), prec, ms / st, ks / st, kr / st, n * st / ms * 1e3), flush=True)
        eng.close()
