"""
b2_kinetic_static: the entry point that allocates nothing.

It exists for hosts whose own allocator owns the GPU (here: torch tensors from
the caching allocator) and which therefore hand this library raw device
pointers and cannot let it cudaMalloc.  The test drives it through the C ABI on
torch-owned memory and checks it against the existing, allocating entry point
(b2_thermo on an engine holding the same particles), twice in a row — the
second call on a smaller system, so rows left in the shared scratch by the
first call would show up if they were read.

Kept in a file of its own that pytest collects last.
"""
import numpy
import pytest

from conftest import golden, make_engine
from oracle import oracle_np as onp
from atooms_b200 import capi

pytestmark = pytest.mark.gpu

# same tolerances as tests/test_gpu_parity.py: totals 1e-12 (fp64) / 1e-6 (fp32) relative
TOL = {64: 1e-12, 32: 1e-6}


def _system(reps, seed):
    d = golden('ka_N1200')
    n = 1200 * reps ** 3
    rng = numpy.random.default_rng(seed)
    mass = rng.uniform(0.5, 2.0, size=n)
    vel = onp.maxwell_boltzmann(n, 0.8, mass, seed + 1) + 0.01     # non-zero total momentum
    shifts = numpy.array([[i, j, k] for i in range(reps) for j in range(reps) for k in range(reps)])
    pos = (d['position'][None] + shifts[:, None, :] * d['side']).reshape(-1, 3)
    return {'position': pos, 'species': numpy.tile(d['species'], reps ** 3), 'side': d['side'] * reps,
            'center': 0.5 * (reps - 1) * d['side'], 'velocity': vel, 'mass': mass}


def _existing(cuda_lib, data, precision):
    th = make_engine(cuda_lib, data, onp.kob_andersen(), precision).thermo()
    return numpy.array([th['kinetic'], *th['momentum'], th['mass']])


def _static(data, precision, out):
    import torch
    dtype = torch.float64 if precision == 64 else torch.float32
    vm = torch.tensor(numpy.c_[data['velocity'], data['mass']], dtype=dtype, device='cuda').contiguous()
    stream = torch.cuda.current_stream().cuda_stream
    capi.kinetic_static(vm.data_ptr(), vm.shape[0], precision, out.data_ptr(), stream)
    torch.cuda.synchronize()
    return out.cpu().numpy()


@pytest.mark.parametrize('precision', [64, 32])
def test_static_matches_thermo_two_calls_in_a_row(cuda_lib, precision):
    import torch
    out = torch.full((8,), float('nan'), dtype=torch.float64, device='cuda')
    for reps, seed in ((2, 11), (1, 23)):          # 9600 particles (3 scratch rows), then 1200 (1 row)
        data = _system(reps, seed)
        ref = _existing(cuda_lib, data, precision)
        got = _static(data, precision, out)
        scale = numpy.array([ref[0], *([numpy.abs(ref[1:4]).max()] * 3), ref[4]])
        assert numpy.all(numpy.abs(got[:5] - ref) <= TOL[precision] * scale), (got, ref)
        assert numpy.all(got[5:] == 0.0)
        # and against plain NumPy sums over the very array the kernel read (an independent check:
        # b2_thermo is another CUDA path)
        dtype = numpy.float64 if precision == 64 else numpy.float32
        v, m = data['velocity'].astype(dtype).astype(numpy.float64), data['mass'].astype(dtype).astype(numpy.float64)
        host = numpy.array([0.5 * numpy.sum(m * numpy.sum(v * v, axis=1)), *(m[:, None] * v).sum(axis=0), m.sum()])
        assert numpy.all(numpy.abs(got[:5] - host) <= 1e-12 * scale), (got, host)


def test_static_rejects_oversize_and_bad_arguments(cuda_lib):
    with pytest.raises(ValueError, match='B2_STATIC_MAX_PARTICLES'):
        capi.kinetic_static(0, capi.STATIC_MAX_PARTICLES + 1, 64, 0)
    with pytest.raises(ValueError):
        capi.kinetic_static(0, 10, 64, 0)           # NULL pointers
    with pytest.raises(ValueError):
        capi.kinetic_static(0, 0, 16, 0)            # bad precision


def test_static_rejects_misaligned_pointers(cuda_lib):
    """A tensor view that does not start on a 16-byte boundary must be refused on the host: the
    vector loads would otherwise fault inside the caller's own CUDA context."""
    import torch
    buf = torch.zeros(4 * 64 + 1, dtype=torch.float32, device='cuda')
    out = torch.zeros(9, dtype=torch.float64, device='cuda')
    stream = torch.cuda.current_stream().cuda_stream
    with pytest.raises(ValueError, match='aligned'):
        capi.kinetic_static(buf[1:].data_ptr(), 64, 32, out.data_ptr(), stream)
    with pytest.raises(ValueError, match='aligned'):
        capi.kinetic_static(buf.data_ptr(), 64, 32, out.data_ptr() + 4, stream)
    capi.kinetic_static(buf.data_ptr(), 64, 32, out.data_ptr(), stream)      # the aligned call still works
    torch.cuda.synchronize()
    assert numpy.all(out.cpu().numpy() == 0.0)


def test_static_empty_input(cuda_lib):
    import torch
    out = torch.full((8,), 7.0, dtype=torch.float64, device='cuda')
    capi.kinetic_static(0, 0, 64, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert numpy.all(out.cpu().numpy() == 0.0)
