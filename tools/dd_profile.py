"""Time the phases of SlabDomain.refresh / run on 2+ GPUs (torchrun)."""
import os, sys, time
import numpy, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from atooms_b200.dd import SlabDomain

rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
dist.init_process_group('nccl', device_id=torch.device('cuda', int(os.environ['LOCAL_RANK'])))
pos, spe, side, vel = bench.ka_workload(seed=2 + 10 * rank)
n = len(spe)
gside = side * numpy.array([world, 1, 1])
gpos = pos + numpy.array([(rank - 0.5 * (world - 1)) * side[0], 0.0, 0.0])
dom = SlabDomain(gpos, vel, spe, gside, model=bench.atooms_model(), dt=bench.DT, skin=bench.SKIN, precision=32,
                 gid=numpy.arange(n, dtype=numpy.int64) + rank * n)
dom.thermo(); dom.run(100)
def timed(f, reps=5):
    torch.cuda.synchronize(); dist.barrier(); t = time.perf_counter()
    for _ in range(reps): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t) / reps * 1e3
t_ref = timed(dom.refresh)
eng = dom.engine
eng.dd_begin(10**9, dom.dt, False)
t_step = timed(lambda: (eng.peer_run(10) if dom._peer else eng.nccl_run(10)), 10) / 10
def only_kernels():
    for _ in range(10):
        eng.dd_step(dom._vmax.data_ptr()); eng.dd_commit(dom._vmax.data_ptr())
t_kern = timed(only_kernels, 10) / 10
t_poll = timed(lambda: eng.dd_poll(), 20)
if rank == 0:
    print('refresh %.2f ms | dd step with comm %.3f ms | step+commit only %.3f ms | poll %.3f ms | batch %d' % (t_ref, t_step, t_kern, t_poll, dom._batch))
# refresh breakdown
dom.profile = {}
for _ in range(5): dom.refresh()
if rank == 0:
    print('peer %s fast %s' % (dom._peer, dom._fast))
    for k, v in dom.profile.items(): print('   %-12s %.3f ms' % (k, v / 5))
dom.profile = None
def phase(name, f):
    torch.cuda.synchronize(); t = time.perf_counter(); r = f(); torch.cuda.synchronize()
    if rank == 0: print('   %-28s %.2f ms' % (name, (time.perf_counter() - t) * 1e3))
    return r
phase('dd_begin (list build)', lambda: eng.dd_begin(100, dom.dt, False))
dist.barrier(); dist.destroy_process_group()
