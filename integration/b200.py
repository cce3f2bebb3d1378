"""B200-native backend (CUDA, sm_100a) for atooms.

Drop this file next to atooms/backends/lammps.py as atooms/backends/b200.py:

    from atooms.simulation import Simulation
    from atooms.backends.b200 import B200
    Simulation(B200('input.xyz', model='kob_andersen', dt=0.002)).run(10000)

Everything lives in the top-level package atooms_b200; this module only gives it the place the
reference's other backends have (atooms/backends/__init__.py:7-11).
"""
from atooms_b200 import (B200, EnergyMinimization, Interaction, System, Trajectory,     # noqa: F401
                         TrajectoryRUMD, TrajectoryRam, Thermostat, Barostat)
