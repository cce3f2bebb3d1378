"""
atooms_b200 — B200-native pair-interaction / MD backend for atooms.

    from atooms_b200 import B200, EnergyMinimization, Interaction, System

Host code is Python; all numerics run in hand-written CUDA (sm_100a) behind
the C ABI of include/atooms_b200.h.  There is no CPU fallback.
"""
__version__ = '0.1.0'

from . import capi, models                                   # noqa: E402,F401
from .models import PairModel, lennard_jones, kob_andersen   # noqa: E402,F401
from .system import System, Cell, Particle, Thermostat, Barostat   # noqa: E402,F401
from .interaction import Interaction                          # noqa: E402,F401
from .trajectory import (Trajectory, TrajectoryRUMD, TrajectoryRam, TrajectoryNeighbors,  # noqa: E402,F401
                         Snapshot)
from .backend import B200, EnergyMinimization                 # noqa: E402,F401
