"""
The N > 1 path on CPU: slab decomposition over 2 ranks with the gloo backend and
the CPU oracle as the engine behind the C ABI (the product uses NCCL + the CUDA
library through the very same atooms_b200/dd.py).  The decomposed run must
reproduce the single-domain run of the same global system: migration, ghost
selection, the per-step halo exchange and the global refresh decision are all
on this path.
"""
import os
import sys

import numpy
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))


def global_system(reps_x=2, seed=5):
    d = numpy.load(os.path.join(ROOT, 'tests', 'golden', 'ka_N1200.npz'))
    side0 = d['side']
    pos = numpy.concatenate([d['position'] + numpy.array([k * side0[0], 0, 0]) for k in range(reps_x)])
    pos[:, 0] -= 0.5 * (reps_x - 1) * side0[0]
    rng = numpy.random.default_rng(seed)
    pos += rng.uniform(-0.02, 0.02, size=pos.shape)
    n = len(pos)
    vel = rng.normal(0, numpy.sqrt(1.5), size=(n, 3))
    vel -= vel.mean(axis=0)
    side = side0 * numpy.array([reps_x, 1, 1])
    return pos, vel, numpy.tile(d['species'], reps_x).astype(numpy.int32), side


def _worker(rank, world, port, steps, out, reps_x=2):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from atooms_b200 import capi
    from atooms_b200.dd import SlabDomain
    lib = capi.Library(os.path.join(ROOT, 'oracle', '_build', 'liboracle_b2.so'))
    pos, vel, spe, side = global_system(reps_x)
    n = len(spe)
    # deliberately NOT pre-sorted into slabs: rank r starts with every world-th particle that
    # lies in its own or an adjacent slab; the first refresh migrates the strays
    slab = numpy.clip(numpy.floor((pos[:, 0] / side[0] + 0.5) * world).astype(int), 0, world - 1)
    mine = numpy.nonzero(slab == rank)[0]
    dom = SlabDomain(pos[mine], vel[mine], spe[mine], side, model='kob_andersen', dt=0.002, skin=0.3,
                     precision=64, gid=mine, library=lib)
    t0 = dom.thermo()
    dom.run(steps)
    t1 = dom.thermo()
    gpos, gvel, gimg = dom.gather()
    # host round trip (bench.py's end-to-end leg): same particles back in, same state, and on it goes
    hp, hv = torch.empty((dom.n_own + 8, 3), dtype=torch.float64), torch.empty((dom.n_own + 8, 3), dtype=torch.float64)
    m = dom.store_host(hp, hv)
    dom.load_host(hp[:m], hv[:m])
    t2 = dom.thermo()
    assert abs((t2['kinetic'] + t2['potential']) - (t1['kinetic'] + t1['potential'])) < 1e-9 * abs(t1['potential'])
    dom.run(20)
    t3 = dom.thermo()
    assert t3['n'] == t1['n']
    assert abs((t3['kinetic'] + t3['potential']) - (t1['kinetic'] + t1['potential'])) < 1e-4 * abs(t1['potential'])
    if rank == 0:
        numpy.savez(out, pos=gpos, vel=gvel, img=gimg, e0=t0['kinetic'] + t0['potential'],
                    e1=t1['kinetic'] + t1['potential'], n=t1['n'], refreshes=dom.refreshes,
                    momentum=t1['momentum'], virial=t1['virial'])
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize('world', [2, 3])
def test_slab_run_matches_single_domain(world, tmp_path, oracle_lib):
    """world = 2: both ring neighbours are the same rank; world = 3: a genuine ring (the hop
    arithmetic, distinct left / right peers), on a box three fixtures long."""
    steps = 120
    reps_x = 2 if world == 2 else 3
    out = str(tmp_path / 'dd.npz')
    port = 29500 + (os.getpid() + 17 * world) % 2000
    mp.spawn(_worker, args=(world, port, steps, out, reps_x), nprocs=world, join=True)
    got = numpy.load(out)
    # the same global system on ONE engine
    from conftest import make_engine
    from oracle import oracle_np as onp
    pos, vel, spe, side = global_system(reps_x)
    ref = make_engine(oracle_lib, {'position': pos, 'velocity': vel, 'species': spe, 'side': side,
                                   'center': numpy.zeros(3)}, onp.kob_andersen(), 64)
    e0 = ref.thermo()
    ref.run_nve(steps, 0.002)
    rpos, rvel, rimg = ref.download(image=True)
    e1 = ref.thermo()
    assert int(got['n']) == len(spe)
    assert int(got['refreshes']) >= 3                      # the hot system forces several list refreshes
    assert abs(got['e0'] - (e0['kinetic'] + e0['potential'])) < 1e-9 * abs(got['e0'])
    assert abs(got['e1'] - (e1['kinetic'] + e1['potential'])) < 1e-8 * abs(got['e1'])
    assert abs(got['virial'] - e1['virial']) < 1e-7 * abs(e1['virial'])
    unf_got = got['pos'] + got['img'] * side
    unf_ref = rpos + rimg * side
    assert numpy.abs(unf_got - unf_ref).max() < 1e-8
    assert numpy.abs(got['vel'] - rvel).max() < 1e-7
    assert numpy.abs(got['momentum']).max() < 1e-9
