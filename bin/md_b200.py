#!/usr/bin/env python
"""
Molecular dynamics driver for the B200 backend.

Same command line as the reference's bin/md.py (which drives the RUMD
backend, bin/md.py:12-81): input file, -T (switches to NVT), -P (with -T: NPT), --dt, -n steps,
-t thermo interval, -c config interval, -C checkpoint interval, -r restart,
output directory.  In place of RUMD's `.ff` potential script, --model names a
registered pair model (lennard_jones, kob_andersen; atooms_b200/models.py).

    python bin/md_b200.py -i data/ka_N256_rho1.185_rumd.xyz.gz --model kob_andersen \\
                          -T 0.8 -n 10000 -t 100 -c 1000 /tmp/run
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main(params):
    import numpy
    if not hasattr(numpy, 'set_string_function'):      # numpy >= 2 and the reference's xyz reader
        numpy.set_string_function = lambda *a, **k: None
    from atooms.simulation import Simulation, Scheduler, write_thermo, write_config
    from atooms.core.utils import setup_logging, mkdir
    from atooms_b200 import B200, Thermostat, Barostat

    if params.verbose:
        setup_logging(level=20)
    if params.debug:
        setup_logging(level=10)
    output_base = os.path.join(params.output_dir, 'trajectory')
    mkdir(params.output_dir)
    backend = B200(params.input_file, model=params.model, dt=params.dt, skin=params.skin,
                   precision=params.precision, device=params.device)
    if params.T is not None or params.integrator == 'nvt':
        if params.T is None:
            raise SystemExit('-I nvt needs a temperature (-T)')
        backend.system.thermostat = Thermostat(params.T, relaxation_time=params.tau)
    if getattr(params, 'P', None) is not None:
        if params.T is None:
            raise SystemExit('-P (NPT) needs a temperature (-T) as well')
        backend.system.barostat = Barostat(params.P, relaxation_time=getattr(params, 'tau_p', 1.0))
    sim = Simulation(backend, output_path=output_base, steps=params.steps,
                     checkpoint_interval=params.checkpoint_interval, restart=params.restart)
    if params.thermo_interval > 0:
        sim.add(write_thermo, Scheduler(params.thermo_interval))
    if params.config_interval > 0:
        sim.add(write_config, Scheduler(params.config_interval))
    sim.run()
    tsp = sim.wall_time(per_step=True, per_particle=True)
    print('# %d particles, %d steps, %.3e particle-timesteps/s' %
          (len(backend.system.particle), sim.current_step, 1.0 / tsp if tsp > 0 else float('nan')))


if __name__ == '__main__':
    parser = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    parser.add_argument('-T', dest='T', type=float, help='temperature (switches to NVT)')
    parser.add_argument('-P', dest='P', type=float, help='pressure (with -T: NPT)')
    parser.add_argument('--tau-p', dest='tau_p', type=float, default=1.0, help='barostat relaxation time')
    parser.add_argument('--dt', dest='dt', type=float, default=0.002, help='time step')
    parser.add_argument('-i', dest='input_file', required=True, help='input file (any atooms trajectory format)')
    parser.add_argument('-I', dest='integrator', default='nve', choices=['nve', 'nvt'], help='integrator')
    parser.add_argument('--model', dest='model', default='lennard_jones', help='pair model name')
    parser.add_argument('--tau', dest='tau', type=float, default=0.2, help='thermostat relaxation time')
    parser.add_argument('--skin', dest='skin', type=float, default=0.3, help='Verlet skin')
    parser.add_argument('--precision', dest='precision', default='fp32', choices=['fp32', 'fp64'])
    parser.add_argument('--device', dest='device', type=int, default=0, help='CUDA device')
    parser.add_argument('-n', dest='steps', type=int, default=0, help='number of steps')
    parser.add_argument('-t', '--thermo-interval', dest='thermo_interval', type=int, default=0, help='energy interval')
    parser.add_argument('-c', '--config-interval', dest='config_interval', type=int, default=0, help='config interval')
    parser.add_argument('-C', '--checkpoint-interval', dest='checkpoint_interval', type=int, default=0,
                        help='checkpoint interval')
    parser.add_argument('-r', dest='restart', action='store_true', help='restart')
    parser.add_argument('-v', dest='verbose', action='store_true', help='verbose output')
    parser.add_argument('-d', dest='debug', action='store_true', help='debug output')
    parser.add_argument(dest='output_dir', type=str, help='output directory')
    main(parser.parse_args())
