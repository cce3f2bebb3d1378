"""Where the plugin leg of bench.py spends host time: cProfile of Simulation(B200(System)).run() with resident
state (no host views).    python tools/plugin_profile.py [chunks]"""
import cProfile
import os
import pstats
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                      # noqa: E402
from atooms_b200 import capi     # noqa: E402

chunks = int(sys.argv[1]) if len(sys.argv) > 1 else 20
views = len(sys.argv) > 2 and sys.argv[2] == 'views'
pos, spe, side, vel = bench.ka_workload()
lib = capi.load()
bench.plugin_e2e(lib, 32, pos, spe, side, vel, 20, views=views)          # warm
pr = cProfile.Profile()
pr.enable()
out = bench.plugin_e2e(lib, 32, pos, spe, side, vel, chunks, views=views)
pr.disable()
print(out)
pstats.Stats(pr).sort_stats('cumulative').print_stats(45)
