"""
Slab decomposition on the real engine: two ranks, CUDA tensors, one process per
rank.  With two GPUs the ranks talk NCCL (the product path); on a single-GPU box
both ranks share GPU 0 and the same code runs over gloo — the engine-side
primitives (ghost flag, pack / unpack, global-max commit, device upload) are the
CUDA ones either way.  The decomposed run must match one engine holding the
whole system.
"""
import os

import numpy
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from test_dd_gloo import global_system

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, steps, precision, backend, mode, out, reps_x=2):
    # variants of the peer-memory step (switches of the CUDA library, read once per process): the commit as a kernel of
    # its own instead of the step kernel's last block; the halo pushed by the step kernel instead of k_peer_push
    if mode == 'peer/commit-kernel':
        os.environ['B2_PEER_COMMIT_KERNEL'] = '1'
    if mode == 'peer/push-fused':
        os.environ['B2_PEER_PUSH_FUSED'] = '1'
    mode = mode.split('/')[0]
    fast, peer = mode != 'torch', mode.startswith('peer')
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    ngpu = torch.cuda.device_count()
    torch.cuda.set_device(rank % ngpu)
    dist.init_process_group(backend, rank=rank, world_size=world)
    from atooms_b200.dd import SlabDomain
    pos, vel, spe, side = global_system(reps_x)
    slab = numpy.clip(numpy.floor((pos[:, 0] / side[0] + 0.5) * world).astype(int), 0, world - 1)
    mine = numpy.nonzero(slab == rank)[0]
    dom = SlabDomain(pos[mine], vel[mine], spe[mine], side, model='kob_andersen', dt=0.002, skin=0.3,
                     precision=precision, gid=mine, fast_comm=fast, peer=peer, device_refresh=(mode == 'peer'))
    assert dom.engine.lib.name == 'cuda-sm_100a' and dom._fast == fast
    t0 = dom.thermo()
    dom.run(steps)
    t1 = dom.thermo()
    assert dom._peer == peer, 'the peer-memory halo fell back to NCCL'
    assert dom._ids_resident == (mode == 'peer'), 'the refresh did not run on the device'
    gpos, gvel, gimg = dom.gather()
    # host round trip (what bench.py's end-to-end leg does), then on with the device-side refresh
    cap = int(1.2 * dom.n_own) + 64
    hp = torch.empty((cap, 3), dtype=torch.float64, pin_memory=True)
    hv = torch.empty((cap, 3), dtype=torch.float64, pin_memory=True)
    m = dom.store_host(hp, hv)
    dom.load_host(hp[:m], hv[:m])
    t2 = dom.thermo()
    assert abs((t2['kinetic'] + t2['potential']) - (t1['kinetic'] + t1['potential'])) < 1e-9 * abs(t1['potential'])
    dom.run(40)
    t3 = dom.thermo()
    assert t3['n'] == t1['n']
    assert abs((t3['kinetic'] + t3['potential']) - (t1['kinetic'] + t1['potential'])) < 2e-4 * abs(t1['potential'])
    if rank == 0:
        numpy.savez(out, pos=gpos, vel=gvel, img=gimg, e0=t0['kinetic'] + t0['potential'],
                    e1=t1['kinetic'] + t1['potential'], n=t1['n'], refreshes=dom.refreshes, virial=t1['virial'])
    dist.barrier()
    dist.destroy_process_group()


def test_four_rank_ring_on_cuda(tmp_path, cuda_lib):
    """Four slabs: left and right neighbours are different ranks and particles leave through
    both faces (peer-memory halo, refresh on the device).  Needs four GPUs."""
    if torch.cuda.device_count() < 4:
        pytest.skip('needs four GPUs')
    _compare(64, 'peer', tmp_path, cuda_lib, world=4, reps_x=3)


@pytest.mark.parametrize('mode', ['torch', 'nccl', 'peer-host', 'peer', 'peer/commit-kernel', 'peer/push-fused'])
@pytest.mark.parametrize('precision', [64, 32])
def test_two_rank_slab_run_on_cuda(precision, mode, tmp_path, cuda_lib):
    _compare(precision, mode, tmp_path, cuda_lib)


def _compare(precision, mode, tmp_path, cuda_lib, world=2, reps_x=2):
    fast = mode != 'torch'                    # ('peer/...' variants: see _worker)
    from conftest import make_engine
    from oracle import oracle_np as onp
    steps = 120
    backend = 'nccl' if torch.cuda.device_count() >= 2 else 'gloo'
    if fast and backend != 'nccl':
        pytest.skip('the C-side communication paths need two GPUs')
    out = str(tmp_path / 'dd.npz')
    port = 29500 + (os.getpid() + precision + 7 * len(mode)) % 2000
    mp.spawn(_worker, args=(world, port, steps, precision, backend, mode, out, reps_x), nprocs=world, join=True)
    got = numpy.load(out)
    pos, vel, spe, side = global_system(reps_x)
    ref = make_engine(cuda_lib, {'position': pos, 'velocity': vel, 'species': spe, 'side': side,
                                 'center': numpy.zeros(3)}, onp.kob_andersen(), precision)
    e0 = ref.thermo()
    ref.run_nve(steps, 0.002)
    rpos, rvel, rimg = ref.download(image=True)
    e1 = ref.thermo()
    tol = 1e-8 if precision == 64 else 2e-4
    assert int(got['n']) == len(spe) and int(got['refreshes']) >= 3
    assert abs(got['e0'] - (e0['kinetic'] + e0['potential'])) < 1e-9 * abs(got['e0'])
    assert abs(got['e1'] - (e1['kinetic'] + e1['potential'])) < (1e-8 if precision == 64 else 1e-5) * abs(got['e1'])
    unf_got, unf_ref = got['pos'] + got['img'] * side, rpos + rimg * side
    assert numpy.abs(unf_got - unf_ref).max() < tol
    assert numpy.abs(got['vel'] - rvel).max() < 10 * tol
