#!/usr/bin/env python
"""
bench.py — particle-timesteps/s of the NVE hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One bench "step" = one `run(INNER)` chunk of the backend: INNER velocity-Verlet
timesteps of the whole system, the unit atooms' Simulation hands to a backend
between observer calls (atooms/simulation/core.py:308-315).

Workload at N = 1 (BASELINE.json configs[2]): Kob-Andersen
80:20, N = 1 200 000 — frame 0 of the reference's data/trajectory.xyz
(tests/golden/ka_N1200.npz) rescaled to rho = 1.2, tiled 10^3, jitter +-0.01
(seed 2), Maxwell-Boltzmann velocities T = 0.8 (seed 3); cut & shifted at
2.5 sigma_ab, dt = 0.002, skin 0.3.  N > 1 ranks: ONE periodic box of N x 1.2M
particles, slab-decomposed along x, one slab per GPU, with a halo exchange every
step and particle migration at every list refresh ("weak": fixed work per GPU).
After the timed region the N-rank run is checked against a single-domain run of
the same engine (`parity` block of the JSON line); out of tolerance => rc != 0.

value : device-resident throughput (CUDA events on the launching stream).
e2e   : the same metric through the plugin, as a user of the reference runs it:
        atooms.simulation.Simulation(B200(System(...))).run() with the reference's
        own driver (baseline/_ref), speedometer and write_thermo on, and an observer
        holding host views of positions and velocities, which the System pushes
        from page-locked host memory before and pulls back after every
        backend.run().  detail.e2e_capi: the same through the C ABI alone.
roofline : the fused step kernel k_step_nve against the measured HBM peak,
        algorithmic bytes = 7 P + 4 nbar per particle-step.
cpu_baseline : the C/OpenMP oracle (oracle/oracle_c.c) on this box's cores, on
        a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

INNER = 100          # MD timesteps per bench step (one backend.run(INNER) chunk)
DT = 0.002
SKIN = 0.3
RHO = 1.2
TEMPERATURE = 0.8
KA_EPS = numpy.array([[1.0, 1.5], [1.5, 0.5]])
KA_SIG = numpy.array([[1.0, 0.8], [0.8, 0.88]])
# algorithmic bytes per particle-step, SURVEY.md 8(d): 7 P + 4 nbar, nbar = 93.3 (KA, rho 1.2, skin 0.3)
NBAR = 93.3
B_ALG = {32: 7 * 16 + 4 * NBAR, 64: 7 * 32 + 4 * NBAR}


def ka_workload(reps=10, seed=2):
    """SURVEY.md 8d C3.  Returns position (N,3), species, side, velocity."""
    d = numpy.load(os.path.join(ROOT, 'tests', 'golden', 'ka_N1200.npz'))
    n0 = len(d['species'])
    scale = (n0 / RHO / numpy.prod(d['side'])) ** (1.0 / 3.0)
    pos0, side0 = d['position'] * scale, d['side'] * scale
    shifts = numpy.stack(numpy.meshgrid(*[numpy.arange(reps)] * 3, indexing='ij'), -1).reshape(-1, 3)
    pos = (pos0[None, :, :] + shifts[:, None, :] * side0).reshape(-1, 3) - 0.5 * (reps - 1) * side0
    rng = numpy.random.default_rng(seed)
    pos += rng.uniform(-0.01, 0.01, size=pos.shape)
    n = pos.shape[0]
    # Maxwell-Boltzmann, COM removed, rescaled to exactly T with (N-1)d dof (system.py:164-198)
    rng = numpy.random.default_rng(seed + 1)
    vel = rng.normal(0.0, numpy.sqrt(TEMPERATURE), size=(n, 3))
    vel -= vel.mean(axis=0)
    vel *= numpy.sqrt(TEMPERATURE / (numpy.sum(vel * vel) / (3 * (n - 1))))
    return pos, numpy.tile(d['species'], reps ** 3).astype(numpy.int32), side0 * reps, vel


def make_engine(lib, precision, pos, spe, side, vel, device=0):
    from atooms_b200 import capi
    eng = capi.Engine(precision, device=device, library=lib)
    eng.set_box(side)
    eng.set_pair_table(KA_EPS, KA_SIG, 2.5 * KA_SIG, True)
    eng.set_skin(SKIN)
    eng.upload(pos, spe, vel)
    return eng


class ClockSampler(object):
    """nvidia-smi clocks and throttle reasons DURING the timed region."""
    QUERY = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
             'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, device):
        self.device = device
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix='.csv')
            os.close(fd)
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.device), '--query-gpu=' + self.QUERY,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=open(self.path, 'w'), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': []}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, reasons = [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(',')]
                if len(f) < 9:
                    continue
                sm.append(float(f[1]))
                out['sm_max_mhz'] = float(f[2])
                for name, val in zip(names, f[5:9]):
                    if val.lower().startswith('active'):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out['sm_mhz'] = float(numpy.median(sm))
        out['reasons'] = sorted(reasons)
        return out


def measured_peak():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        try:
            return float(json.load(open(path))['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
        except Exception:
            pass
    return 6650.0, 'fallback (B200_PROFILING.md: 6.65 TB/s)'


def ncu_traffic(kernel, precision, n):
    """DRAM bytes per launch of the step kernel from the committed ncu capture (same workload), else None."""
    try:
        t = json.load(open(os.path.join(ROOT, 'profiles', 'ncu_traffic.json')))['%s_fp%d' % (kernel, precision)]
        return float(t['bytes_per_launch']) if int(t['particles']) == int(n) else None
    except Exception:
        return None


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count()


def oracle_library(threads=None):
    """The CPU oracle: used by the cpu_baseline leg and by --impl reference only.
    `threads` sets the OpenMP team through the oracle's own entry point: OMP_NUM_THREADS is read
    once per process and torch.distributed.run pre-sets it to 1.  Returns (library, threads in use)."""
    import ctypes
    from atooms_b200 import capi
    path = os.path.join(ROOT, 'oracle', '_build', 'liboracle_b2.so')
    if not os.path.exists(path):
        subprocess.check_call(['make', '-s', '-C', os.path.join(ROOT, 'oracle')])
    lib = capi.Library(path)
    lib.dll.b2_oracle_set_threads.argtypes = [ctypes.c_int]
    lib.dll.b2_oracle_set_threads.restype = ctypes.c_int
    lib.dll.b2_oracle_threads.restype = ctypes.c_int
    if threads:
        lib.dll.b2_oracle_set_threads(int(threads))
    return lib, int(lib.dll.b2_oracle_threads())


def time_oracle(pos, spe, side, vel, md_steps, target_s=None):
    """particle-timesteps/s of the C/OpenMP oracle over NVE steps (list builds included, as in
    the GPU number).  With `target_s` the sample grows from `md_steps` until about that many
    seconds of CPU work are timed (at most 300 steps): hosts differ by more than 10x.
    Returns (rate, seconds, steps, threads)."""
    lib, threads = oracle_library(host_cores())
    eng = make_engine(lib, 64, pos, spe, side, vel)
    eng.compute(forces=False)                  # first list build + forces: not timed (as on the GPU)
    t0 = time.perf_counter()
    eng.run_nve(md_steps, DT)
    spent, done = time.perf_counter() - t0, md_steps
    if target_s is not None and spent < target_s:
        more = min(int(md_steps * (target_s - spent) / max(spent, 1e-3)), 300 - md_steps)
        if more > 0:
            t0 = time.perf_counter()
            eng.run_nve(more, DT)
            spent, done = spent + time.perf_counter() - t0, done + more
    eng.close()
    return len(spe) * done / spent, spent, done, threads


def run_reference(args, rank, world):
    """--impl reference: the reference has no CPU engine for this path;
    the timed stand-in is the oracle port on all host cores, same workload, bounded sample."""
    if rank != 0:
        return
    lib, cores = oracle_library(host_cores())      # cores = the OpenMP team actually used
    pos, spe, side, vel = ka_workload()
    n = len(spe)
    inner = 5                                   # MD timesteps per bench step (bounded sample)
    eng = make_engine(lib, 64, pos, spe, side, vel)
    eng.compute(forces=False)
    t0 = time.perf_counter()
    eng.run_nve(inner, DT)
    per_md_step = (time.perf_counter() - t0) / inner
    # bounded: the whole --steps K --warmup W run stays within ~3 minutes on any host
    inner = int(min(inner, max(1, 180.0 / (per_md_step * max(args.steps + args.warmup, 1)))))
    for _ in range(args.warmup):
        eng.run_nve(inner, DT)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        eng.run_nve(inner, DT)
    elapsed = time.perf_counter() - t0
    value = n * inner * args.steps / elapsed
    line = {
        'impl': 'reference', 'metric': 'particle-timesteps/s', 'value': value, 'unit': 'particle-timesteps/s',
        'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': 1e3 * elapsed / args.steps, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': workload_config(n, inner),
        'cpu_baseline': {'value': value, 'unit': 'particle-timesteps/s', 'cores': cores, 'threads': cores,
                         'kind': 'port', 'sample': '%d bench steps x %d NVE timesteps of the full N=%d system, '
                                   'oracle/oracle_c.c (C + OpenMP, fp64, Verlet list)' % (args.steps, inner, n)},
        'e2e': {'value': value, 'unit': 'particle-timesteps/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    print(json.dumps(line), flush=True)


def workload_config(n, inner, world=1, halo_bytes=0, comm=None):
    cfg = {'workload': 'Kob-Andersen 80:20 N=%d rho=1.2 NVE velocity-Verlet dt=0.002, LJ cut&shifted 2.5 sigma_ab, '
                       'Verlet skin 0.3 (BASELINE.json configs[2], SURVEY.md 8d C3)' % (n * world),
           'particles_per_gpu': n, 'md_timesteps_per_step': inner,
           'cache': 'inputs larger than L2 (neighbour list ~0.8 GB per GPU vs 126 MB L2)'}
    if world > 1:
        cfg['decomposition'] = ('one periodic box of %d x 1.2M particles, %d slabs along x, ring halo exchange '
                                '+ 4-byte max-reduction per step, migration at list refresh'
                                % (world, world))
        cfg['halo_bytes_per_step_per_gpu'] = int(halo_bytes)
        cfg['comm'] = comm
    return cfg


def parity_system(world, seed=5):
    """The system of the multi-GPU parity check: KA 80:20 at rho = 1.2 in a box of
    (world x 10) x 20 x 20 (4800 particles per slab; the slab is 10 wide, r_list = 2.8), jittered,
    T = 1.5 so that the list is refreshed several times within 120 steps and particles migrate."""
    d = numpy.load(os.path.join(ROOT, 'tests', 'golden', 'ka_N1200.npz'))
    n0 = len(d['species'])
    scale = (n0 / RHO / numpy.prod(d['side'])) ** (1.0 / 3.0)
    pos0, side0 = d['position'] * scale, d['side'] * scale
    reps = numpy.array([world, 2, 2])
    shifts = numpy.stack(numpy.meshgrid(*[numpy.arange(r) for r in reps], indexing='ij'), -1).reshape(-1, 3)
    pos = (pos0[None, :, :] + shifts[:, None, :] * side0).reshape(-1, 3) - 0.5 * (reps - 1) * side0
    rng = numpy.random.default_rng(seed)
    pos += rng.uniform(-0.02, 0.02, size=pos.shape)
    n = len(pos)
    vel = rng.normal(0, numpy.sqrt(1.5), size=(n, 3))
    vel -= vel.mean(axis=0)
    return pos, vel, numpy.tile(d['species'], len(shifts)).astype(numpy.int32), side0 * reps


def dd_parity(lib, rank, world, steps=120):
    """Correctness of the multi-GPU product path, where the driver can see it (its test box has
    one GPU): the slab-decomposed run over all ranks (peer-memory halo, refresh on the device when
    available) against ONE engine holding the whole system on rank 0's GPU, both precisions, the
    scenario of tests/test_gpu_dd.py::_compare.  Returns the `parity` block (rank 0) or None."""
    from atooms_b200.dd import SlabDomain
    pos, vel, spe, side = parity_system(world)
    slab = numpy.clip(numpy.floor((pos[:, 0] / side[0] + 0.5) * world).astype(int), 0, world - 1)
    mine = numpy.nonzero(slab == rank)[0]
    block = {'steps': steps, 'particles': int(len(spe)), 'ok': True}
    for precision in (64, 32):
        dom = SlabDomain(pos[mine], vel[mine], spe[mine], side, model=atooms_model(), dt=DT, skin=SKIN,
                         precision=precision, gid=mine, library=lib)
        t0 = dom.thermo()
        dom.run(steps)
        t1 = dom.thermo()
        gpos, gvel, gimg = dom.gather()
        comm = ('peer-memory mailboxes, refresh on the device' if dom._ids_resident else 'peer-memory mailboxes') \
            if dom._peer else ('NCCL enqueued from C' if dom._fast else 'torch.distributed')
        refreshes = dom.refreshes
        dom.close()
        if rank != 0:
            continue
        ref = make_engine(lib, precision, pos, spe, side, vel, device=int(os.environ.get('LOCAL_RANK', '0')))
        e0 = ref.thermo()
        ref.run_nve(steps, DT)
        rpos, rvel, rimg = ref.download(image=True)
        e1 = ref.thermo()
        ref.close()
        tol_x, tol_e = (1e-8, 1e-8) if precision == 64 else (2e-4, 1e-5)
        dx = float(numpy.abs((gpos + gimg * side) - (rpos + rimg * side)).max())
        dv = float(numpy.abs(gvel - rvel).max())
        de0 = abs((t0['kinetic'] + t0['potential']) - (e0['kinetic'] + e0['potential'])) / abs(e0['potential'])
        de1 = abs((t1['kinetic'] + t1['potential']) - (e1['kinetic'] + e1['potential'])) / abs(e1['potential'])
        ok = bool(t1['n'] == len(spe) and refreshes >= 3 and dx < tol_x and dv < 10 * tol_x and de0 < 1e-9
                  and de1 < tol_e)
        block['fp%d' % precision] = {'max_abs_dx': dx, 'max_abs_dv': dv, 'rel_dE_start': float(de0),
                                     'rel_dE_end': float(de1), 'refreshes': int(refreshes), 'comm': comm,
                                     'tol_dx': tol_x, 'tol_dE': tol_e, 'ok': ok}
        block['ok'] = block['ok'] and ok
    return block if rank == 0 else None


def reference_drivers():
    """atooms.simulation from the unmodified reference installed under baseline/_ref (it travels to the
    GPU box; tools/install_reference.py), or None.  The DRIVER only: the engine underneath is ours."""
    ref = os.path.join(ROOT, 'baseline', '_ref')
    if os.path.isdir(os.path.join(ref, 'atooms')) and ref not in sys.path:
        sys.path.insert(0, ref)
    try:
        if not hasattr(numpy, 'set_string_function'):          # numpy >= 2 (atooms/trajectory/xyz.py:472-474)
            numpy.set_string_function = lambda *a, **k: None
        from atooms import simulation
        return simulation
    except Exception:
        return None


def plugin_e2e(lib, precision, pos, spe, side, vel, chunks, device=0, views=True):
    """The metric measured END TO END THROUGH THE PLUGIN, the call a user of the reference makes:
        Simulation(B200(System(...))).run(steps)         (atooms/simulation/core.py:308-425)
    with the reference's own driver, its speedometer on (which cuts the run into steps / 20 chunks,
    core.py:107-109), write_thermo every INNER steps, and — `views` — an observer that works on the
    host arrays the way docs/simulations.org:274-308 teaches (system.view(...)): it reads the positions and
    rescales the velocities in place.  While such views are in user hands the System pushes what was written
    through them host -> device before the next device operation and pulls both arrays back after every
    backend.run().  Host mirrors are page-locked (b2_host_alloc).  Copies are counted where they
    happen (Engine.update / download_into).  Returns rate, seconds, bytes per INNER steps, driver name."""
    import shutil
    from atooms_b200 import B200, System, Cell
    model = atooms_model()
    names = numpy.array(['A', 'B'])[spe]
    system = System(pos, names, Cell(side), velocity=vel, model=model, precision=precision, skin=SKIN,
                    device=device, library=lib)
    bck = B200(system, model, dt=DT, skin=SKIN, precision=precision, device=device, library=lib)
    system.sync = 'tracked'          # push only what numpy wrote through the views (atooms_b200/system.py)
    eng = system._engine
    count = {'h2d': 0, 'd2h': 0, 'runs': 0}
    upd, dwn, run = eng.update, eng.download_into, bck.run

    def update(position=None, velocity=None):
        count['h2d'] += sum(a.nbytes for a in (position, velocity) if a is not None)
        return upd(position, velocity)

    def download_into(position, velocity=None, image=None):
        count['d2h'] += sum(a.nbytes for a in (position, velocity, image) if a is not None)
        return dwn(position, velocity, image)

    def counted_run(steps):
        count['runs'] += 1
        return run(steps)
    eng.update, eng.download_into, bck.run = update, download_into, counted_run
    seen = []

    def host_observer(sim):
        # host-side work on the live arrays, the way docs/simulations.org:274-308 teaches: read the positions,
        # edit the velocities of some particles IN PLACE (factor 1: the run stays NVE); the edit must reach the
        # device, which takes the whole velocity array (what was written where is not tracked)
        x, v = sim.system.view('position'), sim.system.view('velocity')
        seen.append(float(x[::4096, 0].sum()))
        v[:4096] *= 1.0
    simulation = reference_drivers()
    n = len(spe)
    tmp = tempfile.mkdtemp()
    try:
        if simulation is not None:
            sim = simulation.Simulation(bck, output_path=os.path.join(tmp, 'run'), enable_speedometer=True)
            sim.add(simulation.write_thermo, simulation.Scheduler(INNER))
            if views:
                sim.add(host_observer, simulation.Scheduler(INNER))
            sim.run(INNER)                                   # warm-up chunk(s)
            for k in count:
                count[k] = 0
            t0 = time.perf_counter()
            sim.run(INNER * chunks)
            elapsed = time.perf_counter() - t0
            driver = 'atooms.simulation.Simulation (reference driver, speedometer on, write_thermo)'
        else:
            class _Sim(object):
                pass
            sim = _Sim()
            sim.system = bck.system
            bck.run(INNER)
            for k in count:
                count[k] = 0
            t0 = time.perf_counter()
            for _ in range(chunks):
                bck.run(INNER)
                s = bck.system
                seen.append(s.temperature + s.potential_energy(per_particle=True) + s.pressure)
                count['d2h'] += 9 * 8
                if views:
                    host_observer(sim)
            elapsed = time.perf_counter() - t0
            driver = 'chunk loop (reference not installed under baseline/_ref)'
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    out = {'value': n * INNER * chunks / elapsed, 'seconds': elapsed, 'h2d': count['h2d'] // chunks,
           'd2h': count['d2h'] // chunks + 9 * 8, 'backend_run_calls': count['runs'], 'driver': driver,
           'energy': float(bck.system.total_energy(per_particle=True, cache=True))}
    bck.system.close()
    return out


def c5_strong(lib, rank, world, local_rank, steps=200, warmup=100, precision=32, cells=126):
    """BASELINE.json configs[4]: LJ FCC N = 4 * 126^3 = 8 001 504, rho = 0.85, cut &
    shifted 2.5, NVE dt = 0.005, skin 0.3, jitter +-0.05 a, T = 1.0 — ONE system over all ranks (strong
    scaling: the total is fixed), slabs along x; with one rank the plain engine holds all 8M particles.
    The configuration north_star quotes the 8-GPU target on.  Returns the detail block (rank 0)."""
    import torch
    import torch.distributed as dist
    from atooms_b200 import capi, models
    a = (4.0 / 0.85) ** (1.0 / 3.0)
    L = cells * a
    basis = numpy.array([[0, 0, 0], [0.5, 0.5, 0], [0.5, 0, 0.5], [0, 0.5, 0.5]])
    lo, hi = (rank * cells) // world, ((rank + 1) * cells) // world
    grid = numpy.stack(numpy.meshgrid(numpy.arange(lo, hi), numpy.arange(cells), numpy.arange(cells),
                                      indexing='ij'), -1).reshape(-1, 3)
    pos = ((grid[:, None, :] + basis[None, :, :]) * a).reshape(-1, 3) - 0.5 * L
    gid = (((grid[:, 0] * cells + grid[:, 1]) * cells + grid[:, 2])[:, None] * 4 + numpy.arange(4)[None, :]).reshape(-1)
    pos += numpy.random.default_rng(5 + 1000 * rank).uniform(-0.05 * a, 0.05 * a, size=pos.shape)
    vel = numpy.random.default_rng(6 + 1000 * rank).normal(0.0, 1.0, size=pos.shape)
    vel -= vel.mean(axis=0)
    vel *= numpy.sqrt(1.0 / (numpy.sum(vel * vel) / (3 * (len(vel) - 1))))
    spe = numpy.zeros(len(pos), dtype=numpy.int32)
    side = numpy.array([L, L, L])
    model = models.lennard_jones(2.5, True)
    dt = 0.005
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        from atooms_b200.dd import SlabDomain
        dom = SlabDomain(pos, vel, spe, side, model=model, dt=dt, skin=SKIN, precision=precision, gid=gid, library=lib)
        t0 = dom.thermo()
        dom.run(warmup)
        r0 = dom.refreshes
        torch.cuda.synchronize()
        dist.barrier()
        ev0.record()
        dom.run(steps)
        ev1.record()
        torch.cuda.synchronize()
        dist.barrier()
        ms = torch.tensor([ev0.elapsed_time(ev1)], device='cuda')
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms = float(ms)
        t1 = dom.thermo()
        ntot, refreshes = t1['n'], dom.refreshes - r0
        comm = 'peer-memory mailboxes' if dom._peer else ('NCCL from C' if dom._fast else 'torch.distributed')
        dom.close()
    else:
        eng = capi.Engine(precision, device=local_rank, library=lib)
        eng.set_box(side)
        eng.set_pair_table(model.eps, model.sigma, model.rcut, True)
        eng.set_skin(SKIN)
        eng.upload(pos, spe, vel)
        t0 = eng.thermo()
        eng.run_nve(warmup, dt)
        r0 = eng.stats()['rebuilds']
        ms = 0.0
        for _ in range(max(1, steps // 100)):
            eng.run_nve(100, dt)
            ms += eng.last_timing()[0]
        steps = max(1, steps // 100) * 100
        t1 = eng.thermo()
        ntot, refreshes, comm = len(spe), eng.stats()['rebuilds'] - r0, 'none'
        eng.close()
    if rank != 0:
        return None
    return {'config': 'LJ FCC N=%d rho=0.85 NVE dt=0.005 skin 0.3, cut&shifted 2.5 (BASELINE.json configs[4])' % ntot,
            'scaling': 'strong', 'value': ntot * steps / (ms * 1e-3), 'unit': 'particle-timesteps/s',
            'ms_per_md_step': ms / steps, 'md_steps': steps, 'refreshes': int(refreshes), 'comm': comm,
            'dtype': 'f%d' % precision,
            'E_per_particle_start': (t0['kinetic'] + t0['potential']) / ntot,
            'E_per_particle_end': (t1['kinetic'] + t1['potential']) / ntot}


def c4_replicas(lib, rank, world, local_rank, nrep=4096, check=16):
    """BASELINE.json configs[3]: FIRE minimisation of 4096 Kob-Andersen N=150 replicas
    (frame k mod 80 of the reference's data/ka_N150_rho1.200.xyz, jitter +-0.05, seed 100 + k) to
    sum F^2 / N < 1e-10, dealt round-robin over the ranks with no communication ("replicas only").  The
    first `check` replicas of rank 0 are minimised by the CPU oracle as well and compared."""
    import torch
    import torch.distributed as dist
    from atooms_b200 import capi
    from atooms_b200.batch import ReplicaEnsemble
    fr = numpy.load(os.path.join(ROOT, 'tests', 'golden', 'ka_N150_frames.npz'))
    mine = numpy.arange(nrep)[rank::world]
    pos = numpy.empty((len(mine), 150, 3))
    for q, k in enumerate(mine):
        pos[q] = fr['position'][k % 80] + numpy.random.default_rng(100 + int(k)).uniform(-0.05, 0.05, size=(150, 3))
    out = {}
    for precision, tol in ((64, 1e-10), (32, 1e-7)):
        ens = ReplicaEnsemble(pos, fr['species'], fr['side'], 'kob_andersen', precision=precision, device=local_rank,
                              library=lib)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        r = ens.minimize('fire', tolerance=tol, steps=100000)
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        tot = torch.tensor([float(r['steps'].sum()), float(r['converged'].sum()), float(r['energy'].sum())],
                           dtype=torch.float64, device='cuda')
        w = torch.tensor([wall], dtype=torch.float64, device='cuda')
        if world > 1:
            dist.all_reduce(tot)
            dist.all_reduce(w, op=dist.ReduceOp.MAX)
        wall = float(w)
        it = float(tot[0])
        block = {'tolerance': tol, 'seconds': wall, 'replicas_per_s': nrep / wall, 'converged': int(tot[1]),
                 'iterations_total': int(it), 'particle_iterations_per_s': it * 150 / wall,
                 'pair_evaluations_per_s': it * 150 * 149 / wall, 'mean_U_per_particle': float(tot[2]) / nrep / 150}
        if rank == 0 and precision == 64 and check:
            olib, _ = oracle_library(host_cores())
            ref = ReplicaEnsemble(pos[:check], fr['species'], fr['side'], 'kob_andersen', precision=64, library=olib)
            ro = ref.minimize('fire', tolerance=tol, steps=100000)
            block['parity'] = {'replicas_checked': check,
                               'max_abs_dU_per_particle': float(numpy.abs(r['energy'][:check] - ro['energy']).max() / 150),
                               'iterations_equal': bool(numpy.array_equal(r['steps'][:check], ro['steps']))}
            ref.close()
        out['fp%d' % precision] = block
        ens.close()
    if rank != 0:
        return None
    out['config'] = '%d x Kob-Andersen N=150 (rho 1.2, L=5) FIRE to sum F^2/N < tol (BASELINE.json configs[3]), ' \
                    'replica k on rank k mod %d, no communication' % (nrep, world)
    return out


def atooms_model():
    from atooms_b200 import models
    return models.PairModel(KA_EPS, KA_SIG, 2.5 * KA_SIG, True, name='kob_andersen')


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--precision', type=int, default=32, choices=[32, 64])
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--headline-only', action='store_true', help='skip the C4 / C5 legs (detail.c4_replicas, detail.c5_strong)')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'b200' else args.warmup

    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))

    if args.impl == 'reference':
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from atooms_b200 import capi

    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device; the product has no CPU path '
                         '(use --impl reference for the CPU stand-in)')
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    lib = capi.load()
    pos, spe, side, vel = ka_workload(seed=2 + 10 * rank)
    n = len(spe)
    stream = torch.cuda.Stream()
    results = {}
    clocks = None
    if world > 1:
        # ONE periodic system of world x 1.2M particles, slab-decomposed along x: rank r owns
        # tile r of the (world L) x L x L box; per step one ring halo exchange (NCCL send/recv)
        # and one 4-byte all_reduce; particles migrate at every list refresh (atooms_b200/dd.py)
        from atooms_b200.dd import SlabDomain
        gside = side * numpy.array([world, 1, 1])
        gpos = pos + numpy.array([(rank - 0.5 * (world - 1)) * side[0], 0.0, 0.0])
    for precision in ([args.precision] + [p for p in (32, 64) if p != args.precision]):
        primary = precision == args.precision
        if world > 1:
            if not primary:
                continue
            dom = SlabDomain(gpos, vel, spe, gside, model=atooms_model(), dt=DT, skin=SKIN, precision=precision,
                             gid=numpy.arange(n, dtype=numpy.int64) + rank * n, library=lib)
            dom.thermo()
            for _ in range(args.warmup):
                dom.run(INNER)
            l0 = dom.engine.stats()['launches']
            r0 = dom.refreshes
            c0 = dict(dom.counters)
            sampler = ClockSampler(local_rank) if rank == 0 else None
            barrier()
            if sampler:
                sampler.start()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            for _ in range(args.steps):
                dom.run(INNER)
            ev1.record()
            barrier()
            ms = max_over_ranks(ev0.elapsed_time(ev1))
            if sampler:
                clocks = sampler.stop()
            launches = dom.engine.stats()['launches'] - l0
            refreshes_timed = dom.refreshes - r0
            cnt = {k: dom.counters[k] - c0[k] for k in c0}
            th = dom.thermo()
            ntot = th['n']
            value = ntot * INNER * args.steps / (ms * 1e-3)
            # end to end: host buffers in and out around every chunk
            cap = int(n * 1.2)
            hpos = torch.empty((cap, 3), dtype=torch.float64, pin_memory=True)
            hvel = torch.empty((cap, 3), dtype=torch.float64, pin_memory=True)
            e2e_steps = max(2, args.steps // 2)
            m = dom.store_host(hpos, hvel)

            def e2e_step(m):
                dom.load_host(hpos[:m], hvel[:m])
                dom.run(INNER)
                th = dom.thermo()
                return dom.store_host(hpos, hvel), th
            m, th = e2e_step(m)
            barrier()
            t0 = time.perf_counter()
            h2d = d2h = 0
            for _ in range(e2e_steps):
                h2d += 2 * 3 * m * 8
                m, th = e2e_step(m)
                d2h += 2 * 3 * m * 8 + 9 * 8
            barrier()
            e2e_s = max_over_ranks(time.perf_counter() - t0)
            results[precision] = dict(value=value, ms=ms, launches=launches,
                                      prof={'ms_steps': ms - cnt['refresh_ms'], 'ms_refresh': cnt['refresh_ms'],
                                            'step_launches': cnt['enqueued'], 'rounds': cnt['rounds'],
                                            'steps': INNER * args.steps, 'refreshes': refreshes_timed},
                                      e2e=ntot * INNER * e2e_steps / e2e_s, e2e_steps=e2e_steps,
                                      stats=dom.engine.stats(), energy=(th['kinetic'] + th['potential']) / ntot,
                                      h2d=h2d // e2e_steps, d2h=d2h // e2e_steps, ntot=ntot,
                                      halo_bytes=dom.halo_bytes_per_step,
                                      comm='peer-memory mailboxes over NVLink (b2_peer_run)' if dom._peer else
                                      ('NCCL enqueued from C (b2_nccl_run)' if dom._fast else 'torch.distributed'))
            dom.close()
            continue
        eng = make_engine(lib, precision, pos, spe, side, vel, device=local_rank)
        eng.set_stream(stream.cuda_stream)
        eng.compute(forces=False)
        eng.set_profile(True)
        # ---- device-resident ------------------------------------------------
        for _ in range(args.warmup):
            eng.run_nve(INNER, DT)
        l0 = eng.stats()['launches']
        prof = {'ms_steps': 0.0, 'ms_refresh': 0.0, 'step_launches': 0, 'steps': 0, 'refreshes': 0}
        sampler = ClockSampler(local_rank) if (primary and rank == 0) else None
        barrier()
        if sampler:
            sampler.start()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            ev0.record(stream)
            for _ in range(args.steps):
                eng.run_nve(INNER, DT)
                p = eng.profile()
                for k in prof:
                    prof[k] += p[k]
            ev1.record(stream)
        barrier()
        ms = max_over_ranks(ev0.elapsed_time(ev1))
        if sampler:
            clocks = sampler.stop()
        launches = eng.stats()['launches'] - l0
        value = world * n * INNER * args.steps / (ms * 1e-3)
        # ---- end to end through the C ABI with host buffers --------------------
        hpos = torch.empty((n, 3), dtype=torch.float64, pin_memory=True).numpy()
        hvel = torch.empty((n, 3), dtype=torch.float64, pin_memory=True).numpy()
        p0, v0, _ = eng.download()
        hpos[:], hvel[:] = p0, v0
        e2e_steps = max(2, args.steps // 2)
        out_pos = torch.empty((n, 3), dtype=torch.float64, pin_memory=True).numpy()
        out_vel = torch.empty((n, 3), dtype=torch.float64, pin_memory=True).numpy()

        bufs = [(hpos, hvel), (out_pos, out_vel)]           # pinned; what one step wrote is the next one's input

        def e2e_step(k):
            src, dst = bufs[k & 1], bufs[(k + 1) & 1]
            eng.update(src[0], src[1])                        # host -> device: 2 x 3N x 8 bytes
            eng.run_nve(INNER, DT)
            eng.download_into(dst[0], dst[1])                 # device -> host
            return eng.thermo()                               # + the scalars an observer reads
        e2e_step(0)
        barrier()
        t0 = time.perf_counter()
        for k in range(1, e2e_steps + 1):
            th = e2e_step(k)
        barrier()
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        e2e_value = world * n * INNER * e2e_steps / e2e_s
        st = eng.stats()
        results[precision] = dict(value=value, ms=ms, launches=launches, prof=prof, e2e=e2e_value,
                                  e2e_steps=e2e_steps, stats=st, h2d=2 * 3 * n * 8, d2h=2 * 3 * n * 8 + 16 * 8,
                                  energy=(th['kinetic'] + th['potential']) / n, ntot=n, halo_bytes=0)
        eng.close()
        if primary:
            # the headline end-to-end number: through the plugin, driven by the reference's Simulation
            # (at least 20 chunks: the reference's driver writes a checkpoint of the whole system at the end
            # of every Simulation.run(), simulation/core.py:409-415, which a 5-chunk sample would be charged for)
            pchunks = max(20, e2e_steps)
            results[precision]['plugin'] = plugin_e2e(lib, precision, pos, spe, side, vel, pchunks,
                                                      device=local_rank, views=True)
            results[precision]['plugin_resident'] = plugin_e2e(lib, precision, pos, spe, side, vel, pchunks,
                                                               device=local_rank, views=False)
            results[precision]['plugin']['chunks'] = pchunks

    parity = dd_parity(lib, rank, world) if world > 1 else None
    extra = {}
    if not args.headline_only:
        extra['c5_strong'] = c5_strong(lib, rank, world, local_rank)
        extra['c4_replicas'] = c4_replicas(lib, rank, world, local_rank)

    if rank == 0:
        r = results[args.precision]
        peak, peak_src = measured_peak()
        # average duration of one k_step_nve launch that did work (no-op launches of a
        # speculative batch are inside the bracket, so this is an upper bound).  With
        # several GPUs the step kernel is not bracketed separately: the whole step
        # (kernel + halo exchange + refreshes) is charged to it, per GPU.
        if world == 1:
            kernel_ms = r['prof']['ms_steps'] / max(r['prof']['steps'], 1)
        else:
            kernel_ms = r['ms'] / (args.steps * INNER)
        achieved = B_ALG[args.precision] * (r['ntot'] / world) / (kernel_ms * 1e-3) / 1e9
        # list_mode 2: the windowed list and its kernels (csrc/b2_win.cuh); else the global-index list
        step_kernel = 'k_step_win' if r['stats'].get('list_mode') == 2 else 'k_step_nve'
        line = {
            'metric': 'particle-timesteps/s', 'value': r['value'], 'unit': 'particle-timesteps/s',
            'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
            'ms_per_step': r['ms'] / args.steps, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32' if args.precision == 32 else 'f64', 'data': 'synthetic',
            'config': workload_config(n, INNER, world, r['halo_bytes'], r.get('comm')),
            'e2e': ({'value': r['plugin']['value'], 'unit': 'particle-timesteps/s',
                     'h2d_bytes_per_step': int(r['plugin']['h2d']), 'd2h_bytes_per_step': int(r['plugin']['d2h']),
                     'steps': r['plugin']['chunks'], 'backend_run_calls': r['plugin']['backend_run_calls'],
                     'through': 'Simulation(B200(System)).run(): ' + r['plugin']['driver'] +
                                ', an observer holds host views, reads positions and edits velocities '
                                'in place every chunk (sync = tracked: what was written is pushed before, both '
                                'arrays are pulled after every backend.run)'}
                    if 'plugin' in r else
                    {'value': r['e2e'], 'unit': 'particle-timesteps/s',
                     'h2d_bytes_per_step': int(r['h2d']), 'd2h_bytes_per_step': int(r['d2h']),
                     'steps': r['e2e_steps'], 'through': 'SlabDomain.load_host / run / store_host'}),
            'gpu_launches': int(r['launches']),
            'clocks': clocks,
            'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                         'frac': achieved / peak,
                         'traffic': ncu_traffic(step_kernel, args.precision, n) if world == 1 else None,
                         'kernel': step_kernel,
                         'kernel_ms': kernel_ms, 'bytes_per_particle_step': B_ALG[args.precision],
                         'peak_source': peak_src,
                         'step_kernel_share': (r['prof']['ms_steps'] / r['ms']) if world == 1 else None,
                         'refresh_share': (r['prof']['ms_refresh'] / r['ms']) if world == 1 else None},
            'detail': {('fp%d' % p): {'value': results[p]['value'], 'e2e': results[p]['e2e'],
                                      'ms_per_md_step': results[p]['ms'] / (args.steps * INNER),
                                      'step_kernel_ms': results[p]['prof']['ms_steps'] / max(results[p]['prof']['steps'], 1),
                                      'refresh_ms_each': results[p]['prof']['ms_refresh'] / max(results[p]['prof']['refreshes'], 1),
                                      'refreshes': results[p]['prof']['refreshes'],
                                      'noop_step_launches': results[p]['prof']['step_launches'] - results[p]['prof']['steps'],
                                      'max_neighbors': results[p]['stats']['max_neighbors'],
                                      'total_energy_per_particle': results[p]['energy']}
                       for p in results},
        }
        if 'plugin' in r:
            # the same chunk through the C ABI alone (push, b2_run_nve, pull, thermo), and through the plugin
            # with the state left resident (observers read scalars only)
            line['detail']['e2e_capi'] = {'value': r['e2e'], 'h2d_bytes_per_step': int(r['h2d']),
                                          'd2h_bytes_per_step': int(r['d2h'])}
            pr = r['plugin_resident']
            line['detail']['e2e_plugin_resident'] = {'value': pr['value'], 'h2d_bytes_per_step': int(pr['h2d']),
                                                     'd2h_bytes_per_step': int(pr['d2h']),
                                                     'backend_run_calls': pr['backend_run_calls']}
        for key, block in extra.items():
            if block is not None:
                line['detail'][key] = block
        if parity is not None:
            line['parity'] = parity
        if not args.no_cpu_baseline:
            cpu_value, cpu_s, md, cores = time_oracle(pos, spe, side, vel, 10, target_s=15.0)
            line['cpu_baseline'] = {'value': cpu_value, 'unit': 'particle-timesteps/s', 'cores': cores,
                                    'threads': cores, 'kind': 'port',
                                    'sample': '%d NVE timesteps of the full N=%d system (%.1f s), oracle/oracle_c.c '
                                              '(C + OpenMP, fp64, Verlet list)' % (md, n, cpu_s)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if parity is not None and not parity['ok']:
        raise SystemExit('bench.py: the %d-GPU run does not reproduce the single-domain trajectory: %s'
                         % (world, json.dumps(parity)))


if __name__ == '__main__':
    main()
