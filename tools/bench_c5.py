"""BASELINE.json configs[4]: LJ FCC N = 4 * 126^3 = 8 001 504, rho = 0.85, cut & shifted
r_c = 2.5, NVE dt = 0.005, skin 0.3, jitter +-0.05 a (seed 5), T = 1.0 (seed 6); ONE system, slab
decomposition along x over the ranks of torchrun (strong scaling: the total is fixed).
    torchrun --nproc-per-node N tools/bench_c5.py [--steps 300] [--precision 32]
With one rank it runs the plain single-GPU engine on all 8M particles."""
import argparse, json, os, sys
import numpy, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from atooms_b200 import capi, models

ap = argparse.ArgumentParser()
ap.add_argument('--steps', type=int, default=300)
ap.add_argument('--warmup', type=int, default=100)
ap.add_argument('--precision', type=int, default=32)
ap.add_argument('--cells', type=int, default=126)
args = ap.parse_args()
rank, world = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
local = int(os.environ.get('LOCAL_RANK', 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))

nc = args.cells
a = (4.0 / 0.85) ** (1.0 / 3.0)
L = nc * a
basis = numpy.array([[0, 0, 0], [0.5, 0.5, 0], [0.5, 0, 0.5], [0, 0.5, 0.5]])
# this rank's share of lattice planes along x (the first refresh would migrate strays anyway)
lo, hi = (rank * nc) // world, ((rank + 1) * nc) // world
ix = numpy.arange(lo, hi)
grid = numpy.stack(numpy.meshgrid(ix, numpy.arange(nc), numpy.arange(nc), indexing='ij'), -1).reshape(-1, 3)
pos = ((grid[:, None, :] + basis[None, :, :]) * a).reshape(-1, 3) - 0.5 * L
gid = ((grid[:, 0] * nc + grid[:, 1]) * nc + grid[:, 2])[:, None] * 4 + numpy.arange(4)[None, :]
gid = gid.reshape(-1)
rng = numpy.random.default_rng(5 + 1000 * rank)
pos += rng.uniform(-0.05 * a, 0.05 * a, size=pos.shape)
rng = numpy.random.default_rng(6 + 1000 * rank)
vel = rng.normal(0.0, 1.0, size=pos.shape)
vel -= vel.mean(axis=0)
vel *= numpy.sqrt(1.0 / (numpy.sum(vel * vel) / (3 * (len(vel) - 1))))
spe = numpy.zeros(len(pos), dtype=numpy.int32)
side = numpy.array([L, L, L])
dt = 0.005
model = models.lennard_jones(2.5, True)

if world > 1:
    from atooms_b200.dd import SlabDomain
    dom = SlabDomain(pos, vel, spe, side, model=model, dt=dt, skin=0.3, precision=args.precision, gid=gid)
    t0 = dom.thermo()
    dom.run(args.warmup)
    r0 = dom.refreshes
    torch.cuda.synchronize(); dist.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(); dom.run(args.steps); ev1.record(); torch.cuda.synchronize(); dist.barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], device='cuda'); dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms)
    t1 = dom.thermo()
    ntot, refreshes, comm = t1['n'], dom.refreshes - r0, ('peer' if dom._peer else 'nccl')
    halo = dom.halo_bytes_per_step
else:
    eng = capi.Engine(args.precision, device=local)
    eng.set_box(side); eng.set_pair_table(model.eps, model.sigma, model.rcut, True); eng.set_skin(0.3)
    eng.upload(pos, spe, vel)
    t0 = eng.thermo(); t0['n'] = len(spe)
    eng.run_nve(args.warmup, dt)
    r0 = eng.stats()['rebuilds']
    ms = 0.0
    for _ in range(max(1, args.steps // 100)):
        eng.run_nve(100, dt); ms += eng.last_timing()[0]
    args.steps = max(1, args.steps // 100) * 100
    t1 = eng.thermo(); ntot, refreshes, comm, halo = len(spe), eng.stats()['rebuilds'] - r0, 'none', 0
if rank == 0:
    e0 = (t0['kinetic'] + t0['potential']) / ntot
    e1 = (t1['kinetic'] + t1['potential']) / ntot
    print(json.dumps({'config': 'LJ FCC N=%d rho=0.85 NVE dt=0.005 skin 0.3 (BASELINE.json configs[4])' % ntot,
                      'n_gpus': world, 'precision': args.precision, 'steps': args.steps,
                      'ms_per_md_step': ms / args.steps, 'value': ntot * args.steps / (ms * 1e-3),
                      'unit': 'particle-timesteps/s', 'scaling': 'strong', 'refreshes': refreshes, 'comm': comm,
                      'halo_bytes_per_step_per_gpu': halo, 'E_per_particle_start': e0, 'E_per_particle_end': e1}))
if world > 1:
    dist.barrier(); dist.destroy_process_group()
