"""
ctypes binding of the C ABI declared in include/atooms_b200.h.

`load()` returns the product library (hand-written CUDA, sm_100a).  There is
no CPU fallback: if the shared object is missing or no GPU is usable the
call raises.  `Library(path)` binds any shared object exporting the ABI; the
test-suite uses it to drive the CPU oracle through the very same wrapper.
"""
import ctypes
import os

import numpy

__all__ = ['Library', 'Engine', 'Batch', 'load', 'library_path', 'B2Error',
           'FP32', 'FP64', 'ENERGY', 'FORCES', 'VIRIAL', 'STRESS', 'MIN_SD', 'MIN_FIRE',
           'SYMBOLS', 'kinetic_static', 'STATIC_MAX_PARTICLES']

ABI_VERSION = 1
FP32, FP64 = 32, 64
ENERGY, FORCES, VIRIAL, STRESS = 1, 2, 4, 8
GHOST = 256
MIN_SD, MIN_FIRE = 0, 1
NVT_BERENDSEN, NVT_NOSE_HOOVER = 0, 1
T_KINETIC, T_POTENTIAL, T_VIRIAL, T_FNORM2, T_PX, T_PY, T_PZ, T_MASS, T_MSD, T_LEN = \
    0, 1, 2, 3, 4, 5, 6, 7, 8, 16
S_STEPS, S_REBUILDS, S_LAUNCHES, S_MAX_NEIGH, S_NEIGH_CAP, S_LIST_MODE, S_NCELLS, S_SMALL_BOX, S_WIN_BLOCKS, S_LEN = \
    0, 1, 2, 3, 4, 5, 6, 7, 8, 16

_ERRORS = {-1: ValueError, -5: NotImplementedError}

_c_dp = ctypes.POINTER(ctypes.c_double)
_c_ip = ctypes.POINTER(ctypes.c_int32)
_c_lp = ctypes.POINTER(ctypes.c_int64)
_vp = ctypes.c_void_p
_i64 = ctypes.c_int64
_int = ctypes.c_int
_dbl = ctypes.c_double

# name -> (restype, argtypes); the single source of truth for the loader and
# for the "every declared symbol is exported" test.
SYMBOLS = {
    'b2_abi_version': (_int, []),
    'b2_backend_name': (ctypes.c_char_p, []),
    'b2_last_error': (ctypes.c_char_p, [_vp]),
    'b2_create': (_int, [_int, _int, ctypes.POINTER(_vp)]),
    'b2_destroy': (_int, [_vp]),
    'b2_set_stream': (_int, [_vp, _vp]),
    'b2_set_box': (_int, [_vp, _c_dp, _c_dp]),
    'b2_set_pair_table': (_int, [_vp, _int, _c_dp, _c_dp, _c_dp, _int]),
    'b2_set_pair_tabulated': (_int, [_vp, _int, _int, _c_dp, _c_dp, _c_dp]),
    'b2_set_skin': (_int, [_vp, _dbl]),
    'b2_upload': (_int, [_vp, _i64, _c_dp, _c_dp, _c_ip, _c_dp]),
    'b2_update': (_int, [_vp, _c_dp, _c_dp]),
    'b2_download': (_int, [_vp, _c_dp, _c_dp, _c_ip]),
    'b2_scale_velocities': (_int, [_vp, _dbl]),
    'b2_mark_reference': (_int, [_vp]),
    'b2_compute': (_int, [_vp, _int, _c_dp, _c_dp, _c_dp, _c_dp, _c_dp, _c_dp]),
    'b2_thermo': (_int, [_vp, _c_dp]),
    'b2_neighbors': (_int, [_vp, _c_ip, _c_ip, ctypes.c_int32, _c_ip]),
    'b2_run_nve': (_int, [_vp, _i64, _dbl]),
    'b2_minimize': (_int, [_vp, _int, _i64, _dbl, _dbl, _c_lp, ctypes.POINTER(_int)]),
    'b2_run_nvt': (_int, [_vp, _i64, _dbl, _dbl, _dbl, _int]),
    'b2_stats': (_int, [_vp, _c_lp]),
    'b2_last_timing': (_int, [_vp, _c_dp, _c_lp]),
    'b2_set_profile': (_int, [_vp, _int]),
    'b2_profile': (_int, [_vp, _c_dp]),
    'b2_set_images': (_int, [_vp, _c_ip]),
    'b2_get_reference': (_int, [_vp, _c_dp]),
    'b2_set_reference': (_int, [_vp, _c_dp]),
    'b2_get_thermostat': (_int, [_vp, _c_dp]),
    'b2_set_thermostat': (_int, [_vp, _dbl]),
    'b2_hessian': (_int, [_vp, _c_dp]),
    'b2_format_xyz': (_int, [ctypes.c_char_p, _i64, ctypes.POINTER(_i64), _i64, _c_ip, _int, ctypes.POINTER(ctypes.c_char_p),
                      _c_dp, _c_dp, _int]),
    'b2_host_alloc': (_int, [ctypes.POINTER(_vp), _i64]),
    'b2_host_free': (_int, [_vp]),
    'b2_kinetic_static': (_int, [_vp, _i64, _int, _vp, _vp]),
    'b2_upload_device': (_int, [_vp, _i64, _vp, _vp, _vp, _vp]),
    'b2_download_device': (_int, [_vp, _vp, _vp, _vp]),
    'b2_dd_begin': (_int, [_vp, _i64, _dbl, _int]),
    'b2_dd_step': (_int, [_vp, _vp]),
    'b2_dd_commit': (_int, [_vp, _vp]),
    'b2_dd_poll': (_int, [_vp, _c_lp, ctypes.POINTER(_int)]),
    'b2_dd_end': (_int, [_vp]),
    'b2_dd_pack': (_int, [_vp, _vp, _i64, _vp]),
    'b2_dd_unpack': (_int, [_vp, _i64, _i64, _vp]),
    'b2_batch_create': (_int, [_int, _int, ctypes.POINTER(_vp)]),
    'b2_batch_destroy': (_int, [_vp]),
    'b2_batch_set_model': (_int, [_vp, _c_dp, _int, _c_dp, _c_dp, _c_dp, _int]),
    'b2_batch_upload': (_int, [_vp, _i64, _int, _c_dp, _c_ip, _c_dp]),
    'b2_batch_minimize': (_int, [_vp, _int, _i64, _dbl, _dbl, _c_lp, ctypes.POINTER(_int), _c_dp, _c_dp, _c_dp]),
    'b2_batch_download': (_int, [_vp, _c_dp, _c_dp]),
    'b2_batch_set_velocities': (_int, [_vp, _c_dp]),
    'b2_batch_run_nve': (_int, [_vp, _i64, _dbl, _c_dp, _c_dp, _c_dp]),
    'b2_batch_get_state': (_int, [_vp, _c_dp, _c_ip]),
}


class B2Error(RuntimeError):
    """Failure reported by the engine (negative B2_E* code)."""

    def __init__(self, code, message):
        RuntimeError.__init__(self, '[b2 error %d] %s' % (code, message))
        self.code = code


def library_path():
    """Where build() puts the product library (in-tree, so it travels to the GPU box)."""
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), 'lib', 'libatooms_b200.so')


class Library(object):

    """A shared object exporting the b2_* ABI."""

    def __init__(self, path):
        if not os.path.exists(path):
            raise OSError('atooms_b200: shared library not found: %s '
                          '(build it with `python -c "import __graft_entry__ as g; g.build()"`; '
                          'there is no CPU fallback)' % path)
        self.path = path
        # RTLD_LOCAL: two libraries exporting the same b2_* names (product + oracle in the
        # tests) must never see each other's symbols; both are also linked -Bsymbolic
        self.dll = ctypes.CDLL(path, mode=ctypes.RTLD_LOCAL)
        for name, (restype, argtypes) in SYMBOLS.items():
            fn = getattr(self.dll, name)   # AttributeError if the symbol is missing
            fn.restype = restype
            fn.argtypes = argtypes
        # optional NCCL fast path (include/atooms_b200_nccl.h): CUDA library only
        self.has_nccl = hasattr(self.dll, 'b2_nccl_run')
        if self.has_nccl:
            self.dll.b2_nccl_unique_id.argtypes = [ctypes.c_char_p]
            self.dll.b2_nccl_init.argtypes = [_vp, ctypes.c_char_p, _int, _int]
            self.dll.b2_nccl_set_halo.argtypes = [_vp, _vp, _i64, _vp, _i64, _i64, _i64]
            self.dll.b2_nccl_run.argtypes = [_vp, _i64]
            self.dll.b2_peer_export.argtypes = [_vp, _i64, _int, _int, ctypes.c_char_p]
            self.dll.b2_peer_connect.argtypes = [_vp, ctypes.c_char_p]
            self.dll.b2_peer_run.argtypes = [_vp, _i64]
            self.dll.b2_peer_set_ids.argtypes = [_vp, _vp, _vp]
            self.dll.b2_peer_refresh.argtypes = [_vp, _dbl, _dbl, _dbl, ctypes.POINTER(ctypes.c_int64)]
            self.dll.b2_peer_state.argtypes = [_vp, _vp, _vp, _vp, _vp]
            self.dll.b2_peer_update.argtypes = [_vp, _i64, _vp, _vp]
            for name in ('b2_nccl_unique_id', 'b2_nccl_init', 'b2_nccl_set_halo', 'b2_nccl_run',
                         'b2_peer_export', 'b2_peer_connect', 'b2_peer_run', 'b2_peer_set_ids',
                         'b2_peer_refresh', 'b2_peer_state', 'b2_peer_update'):
                getattr(self.dll, name).restype = _int
        version = self.dll.b2_abi_version()
        if version != ABI_VERSION:
            raise OSError('atooms_b200: %s has ABI version %d, expected %d' % (path, version, ABI_VERSION))
        self.name = self.dll.b2_backend_name().decode()

    def format_xyz(self, ids, names, position, velocity=None, precision=6):
        """The particle lines of an xyz frame as bytes (b2_format_xyz: formatted on all host cores)."""
        ids = numpy.ascontiguousarray(ids, dtype=numpy.int32)
        pos = numpy.ascontiguousarray(position, dtype=numpy.float64)
        vel = None if velocity is None else numpy.ascontiguousarray(velocity, dtype=numpy.float64)
        n = len(ids)
        raw = [str(s).encode() for s in names]
        arr = (ctypes.c_char_p * len(raw))(*raw)
        written = _i64()
        args = (n, _ip(ids), len(raw), arr, _dp(pos), _dp(vel), int(precision))
        self.check(self.dll.b2_format_xyz(None, 0, ctypes.byref(written), *args))          # size query
        buf = numpy.empty(max(written.value, 1), dtype=numpy.uint8)
        self.check(self.dll.b2_format_xyz(buf.ctypes.data_as(ctypes.c_char_p), buf.size, ctypes.byref(written), *args))
        return memoryview(buf[:written.value])

    def host_empty(self, shape, dtype=numpy.float64):
        """A numpy array in page-locked host memory (b2_host_alloc), freed with the array: host
        mirrors in such memory are copied to and from the device by DMA, asynchronously."""
        import weakref
        shape = tuple(int(x) for x in (shape if hasattr(shape, '__len__') else (shape,)))
        nbytes = int(numpy.prod(shape, dtype=numpy.int64)) * numpy.dtype(dtype).itemsize
        ptr = _vp()
        self.check(self.dll.b2_host_alloc(ctypes.byref(ptr), nbytes))
        buf = (ctypes.c_char * max(nbytes, 1)).from_address(ptr.value)
        arr = numpy.frombuffer(buf, dtype=dtype, count=int(numpy.prod(shape, dtype=numpy.int64))).reshape(shape)
        weakref.finalize(buf, self.dll.b2_host_free, ptr.value)    # buf lives as long as any view of it
        return arr

    def check(self, code, handle=None):
        if code == 0:
            return
        msg = self.dll.b2_last_error(handle)
        msg = msg.decode() if msg else ''
        if code in _ERRORS:
            raise _ERRORS[code]('[b2 error %d] %s' % (code, msg))
        raise B2Error(code, msg)


STATIC_MAX_PARTICLES = 16777216


def kinetic_static(vel_mass_ptr, n, precision, out_ptr, stream=0, library=None):
    """Kinetic energy, momentum and mass of `n` particles from a device array the
    caller owns (n x (vx, vy, vz, mass), float4 or double4), written to 8 doubles
    at the device address `out_ptr`; asynchronous on `stream`.

    This is the allocation-free entry point (b2_kinetic_static): meant for hosts
    whose own pool owns the GPU memory (e.g. torch tensors' data_ptr()), where a
    cudaMalloc from this library would fail or stall the device.  At most
    STATIC_MAX_PARTICLES per call; overlapping calls on one GPU race on the
    shared scratch; make the first call on a GPU before any stream capture."""
    lib = library if library is not None else load()
    lib.check(lib.dll.b2_kinetic_static(_vp(int(vel_mass_ptr) if vel_mass_ptr else 0), int(n), int(precision),
                                        _vp(int(stream) if stream else 0), _vp(int(out_ptr) if out_ptr else 0)))


_product = None


def load():
    """The product library (CUDA).  Raises if it has not been built."""
    global _product
    if _product is None:
        _product = Library(library_path())
    return _product


def _dp(a):
    return None if a is None else a.ctypes.data_as(_c_dp)


def _ip(a):
    return None if a is None else a.ctypes.data_as(_c_ip)


def _as_xyz(a, n=None, name='array'):
    """Accept (N,3) C-order or (3,N) F-order float64; return an array whose
    memory is N x 3 interleaved (no copy if it already is)."""
    a = numpy.asarray(a)
    if a.ndim != 2:
        raise ValueError('%s must be 2-dimensional' % name)
    if a.shape[0] == 3 and a.flags.f_contiguous and not (a.shape[1] == 3 and a.flags.c_contiguous):
        a = a.T
    if a.shape[1] != 3:
        raise ValueError('%s must be (N, 3) or Fortran-ordered (3, N)' % name)
    a = numpy.ascontiguousarray(a, dtype=numpy.float64)
    if n is not None and a.shape[0] != n:
        raise ValueError('%s has %d rows, expected %d' % (name, a.shape[0], n))
    return a


class Engine(object):

    """One b2_ctx: a particle system resident on one GPU."""

    def __init__(self, precision=FP64, device=0, library=None, stream=None):
        self.lib = library if library is not None else load()
        self.precision = int(precision)
        self.device = int(device)
        self.n = 0
        handle = _vp()
        self.lib.check(self.lib.dll.b2_create(self.device, self.precision, ctypes.byref(handle)))
        self._h = handle
        if stream is not None:
            self.set_stream(stream)

    # -- lifetime ---------------------------------------------------------
    def close(self):
        h, self._h = getattr(self, '_h', None), None
        if h:
            self.lib.dll.b2_destroy(h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _call(self, name, *args):
        if not self._h:
            raise B2Error(-4, 'engine is closed')
        self.lib.check(getattr(self.lib.dll, name)(self._h, *args), self._h)

    # -- model ------------------------------------------------------------
    def set_stream(self, stream):
        self._call('b2_set_stream', _vp(int(stream) if stream else 0))

    def set_box(self, side, center=None):
        side = numpy.ascontiguousarray(side, dtype=numpy.float64)
        center = numpy.zeros(3) if center is None else numpy.ascontiguousarray(center, dtype=numpy.float64)
        if side.shape != (3,) or center.shape != (3,):
            raise ValueError('only 3-dimensional orthorhombic cells are supported')
        self._call('b2_set_box', _dp(side), _dp(center))

    def set_pair_table(self, eps, sigma, rcut, shifted=True):
        eps = numpy.atleast_2d(numpy.ascontiguousarray(eps, dtype=numpy.float64))
        sigma = numpy.atleast_2d(numpy.ascontiguousarray(sigma, dtype=numpy.float64))
        rcut = numpy.atleast_2d(numpy.ascontiguousarray(rcut, dtype=numpy.float64))
        nsp = eps.shape[0]
        if not (eps.shape == sigma.shape == rcut.shape == (nsp, nsp)):
            raise ValueError('eps, sigma, rcut must be square matrices of the same size')
        self._call('b2_set_pair_table', nsp, _dp(eps), _dp(sigma), _dp(rcut), int(bool(shifted)))

    def set_pair_tabulated(self, rcut, u, w):
        """rcut (nsp, nsp); u, w (nsp, nsp, npoints) sampled uniformly in r^2 over [0, rcut^2] (b2_set_pair_tabulated)."""
        rcut = numpy.atleast_2d(numpy.ascontiguousarray(rcut, dtype=numpy.float64))
        u = numpy.ascontiguousarray(u, dtype=numpy.float64)
        w = numpy.ascontiguousarray(w, dtype=numpy.float64)
        nsp = rcut.shape[0]
        if rcut.shape != (nsp, nsp) or u.shape != w.shape or u.shape[:2] != (nsp, nsp) or u.ndim != 3:
            raise ValueError('rcut must be (nsp, nsp), u and w (nsp, nsp, npoints)')
        self._call('b2_set_pair_tabulated', nsp, int(u.shape[2]), _dp(rcut), _dp(u), _dp(w))

    def set_model(self, model):
        """Bind a model of atooms_b200.models: the 12-6 table or lookup tables."""
        if getattr(model, 'tabulated', False):
            u, w = model.tables()
            self.set_pair_tabulated(model.rcut, u, w)
        else:
            self.set_pair_table(model.eps, model.sigma, model.rcut, model.shifted)

    def set_skin(self, skin):
        self._call('b2_set_skin', float(skin))

    # -- state ------------------------------------------------------------
    def upload(self, position, species, velocity=None, mass=None):
        pos = _as_xyz(position, name='position')
        n = pos.shape[0]
        vel = None if velocity is None else _as_xyz(velocity, n, 'velocity')
        spe = numpy.ascontiguousarray(species, dtype=numpy.int32)
        if spe.shape != (n,):
            raise ValueError('species must have one entry per particle')
        m = None if mass is None else numpy.ascontiguousarray(mass, dtype=numpy.float64)
        if m is not None and m.shape != (n,):
            raise ValueError('mass must have one entry per particle')
        self._call('b2_upload', n, _dp(pos), _dp(vel), _ip(spe), _dp(m))
        self.n = n

    def update(self, position=None, velocity=None):
        pos = None if position is None else _as_xyz(position, self.n, 'position')
        vel = None if velocity is None else _as_xyz(velocity, self.n, 'velocity')
        self._call('b2_update', _dp(pos), _dp(vel))

    def download(self, position=True, velocity=True, image=False):
        pos = numpy.empty((self.n, 3)) if position else None
        vel = numpy.empty((self.n, 3)) if velocity else None
        img = numpy.empty((self.n, 3), dtype=numpy.int32) if image else None
        self._call('b2_download', _dp(pos), _dp(vel), _ip(img))
        return pos, vel, img

    def download_into(self, position=None, velocity=None, image=None):
        """Copy state into caller-owned (N,3) C-contiguous arrays (e.g. pinned host memory)."""
        for a, dt in ((position, numpy.float64), (velocity, numpy.float64), (image, numpy.int32)):
            if a is not None and (a.dtype != dt or a.shape != (self.n, 3) or not a.flags.c_contiguous):
                raise ValueError('download_into: arrays must be C-contiguous (N, 3) float64 / int32')
        self._call('b2_download', _dp(position), _dp(velocity), _ip(image))

    def scale_velocities(self, factor):
        self._call('b2_scale_velocities', float(factor))

    def mark_reference(self):
        self._call('b2_mark_reference')

    def hessian(self):
        """(3, N, 3, N) second derivatives of the potential energy (b2_hessian)."""
        out = numpy.empty((3, self.n, 3, self.n))
        self._call('b2_hessian', _dp(out))
        return out

    # -- restart state (see include/atooms_b200.h) ------------------------------------------------
    def set_images(self, image):
        image = numpy.ascontiguousarray(image, dtype=numpy.int32).reshape(self.n, 3)
        self._call('b2_set_images', _ip(image))

    def reference(self):
        ref = numpy.empty((self.n, 3))
        self._call('b2_get_reference', _dp(ref))
        return ref

    def set_reference(self, ref):
        ref = numpy.ascontiguousarray(ref, dtype=numpy.float64).reshape(self.n, 3)
        self._call('b2_set_reference', _dp(ref))

    def thermostat_state(self):
        xi = ctypes.c_double()
        self._call('b2_get_thermostat', ctypes.byref(xi))
        return xi.value

    def set_thermostat_state(self, xi):
        self._call('b2_set_thermostat', float(xi))

    # -- single point -----------------------------------------------------
    def compute(self, what=ENERGY | FORCES | VIRIAL, forces=True, per_particle=False, stress=False):
        """Returns a dict with energy, virial and optionally forces (N,3),
        stress (3,3), pe (N), vir (N)."""
        e, w = _dbl(), _dbl()
        f = numpy.empty((self.n, 3)) if forces else None
        s = numpy.empty((3, 3)) if stress else None
        pe = numpy.empty(self.n) if per_particle else None
        vir = numpy.empty(self.n) if per_particle else None
        if stress:
            what |= STRESS
        self._call('b2_compute', int(what), ctypes.byref(e), ctypes.byref(w), _dp(s), _dp(f), _dp(pe), _dp(vir))
        out = {'energy': e.value, 'virial': w.value}
        if forces:
            out['forces'] = f
        if stress:
            out['stress'] = s
        if per_particle:
            out['pe'], out['vir'] = pe, vir
        return out

    def thermo(self):
        out = numpy.zeros(T_LEN)
        self._call('b2_thermo', _dp(out))
        return {'kinetic': out[T_KINETIC], 'potential': out[T_POTENTIAL], 'virial': out[T_VIRIAL],
                'fnorm2': out[T_FNORM2], 'momentum': out[T_PX:T_PZ + 1].copy(), 'mass': out[T_MASS],
                'msd_sum': out[T_MSD]}

    def neighbors(self):
        """List of sorted int32 arrays (original numbering) for the current configuration."""
        count = numpy.zeros(self.n, dtype=numpy.int32)
        mx = ctypes.c_int32()
        self._call('b2_neighbors', _ip(count), None, 0, ctypes.byref(mx))
        stride = max(int(mx.value), 1)
        nbr = numpy.empty((self.n, stride), dtype=numpy.int32)
        self._call('b2_neighbors', _ip(count), _ip(nbr), stride, ctypes.byref(mx))
        return [nbr[i, :count[i]].copy() for i in range(self.n)]

    def neighbors_matrix(self):
        """The same list as (count[n], nbr[n, max_count]) arrays, rows sorted and padded with -1:
        no per-particle Python objects, for systems of millions of particles."""
        count = numpy.zeros(self.n, dtype=numpy.int32)
        mx = ctypes.c_int32()
        self._call('b2_neighbors', _ip(count), None, 0, ctypes.byref(mx))
        stride = max(int(mx.value), 1)
        nbr = numpy.empty((self.n, stride), dtype=numpy.int32)
        self._call('b2_neighbors', _ip(count), _ip(nbr), stride, ctypes.byref(mx))
        nbr[numpy.arange(stride)[None, :] >= count[:, None]] = -1
        return count, nbr

    # -- dynamics ---------------------------------------------------------
    def run_nve(self, steps, dt):
        self._call('b2_run_nve', int(steps), float(dt))

    def run_nvt(self, steps, dt, temperature, relaxation_time, kind=NVT_NOSE_HOOVER):
        self._call('b2_run_nvt', int(steps), float(dt), float(temperature), float(relaxation_time), int(kind))

    def minimize(self, method, max_steps, dt, fn2_tol):
        done, conv = _i64(), _int()
        self._call('b2_minimize', int(method), int(max_steps), float(dt), float(fn2_tol),
                   ctypes.byref(done), ctypes.byref(conv))
        return done.value, bool(conv.value)

    # -- introspection ----------------------------------------------------
    def stats(self):
        out = numpy.zeros(S_LEN, dtype=numpy.int64)
        self._call('b2_stats', out.ctypes.data_as(_c_lp))
        return {'steps': int(out[S_STEPS]), 'rebuilds': int(out[S_REBUILDS]),
                'launches': int(out[S_LAUNCHES]), 'max_neighbors': int(out[S_MAX_NEIGH]),
                'neighbor_capacity': int(out[S_NEIGH_CAP]), 'list_mode': int(out[S_LIST_MODE]),
                'ncells': int(out[S_NCELLS]), 'small_box': bool(out[S_SMALL_BOX]),
                'window_blocks': int(out[S_WIN_BLOCKS])}

    def last_timing(self):
        ms, launches = _dbl(), _i64()
        self._call('b2_last_timing', ctypes.byref(ms), ctypes.byref(launches))
        return ms.value, launches.value

    def set_profile(self, on=True):
        self._call('b2_set_profile', int(bool(on)))

    def profile(self):
        out = numpy.zeros(8)
        self._call('b2_profile', _dp(out))
        return {'ms_steps': out[0], 'ms_refresh': out[1], 'step_launches': int(out[2]),
                'steps': int(out[3]), 'refreshes': int(out[4])}

    # -- domain decomposition primitives: raw addresses in the engine's memory space ----
    # (device pointers for the CUDA library, host pointers for the CPU oracle)
    def upload_device(self, n, pos_ptr, species_ptr, vel_ptr=0, mass_ptr=0):
        self._call('b2_upload_device', int(n), _vp(pos_ptr or 0), _vp(vel_ptr or 0), _vp(species_ptr or 0),
                   _vp(mass_ptr or 0))
        self.n = int(n)

    def download_device(self, pos_ptr=0, vel_ptr=0, image_ptr=0):
        self._call('b2_download_device', _vp(pos_ptr or 0), _vp(vel_ptr or 0), _vp(image_ptr or 0))

    def dd_begin(self, steps, dt, opening_kick=True):
        self._call('b2_dd_begin', int(steps), float(dt), int(bool(opening_kick)))

    def dd_step(self, vmax_ptr):
        self._call('b2_dd_step', _vp(vmax_ptr))

    def dd_commit(self, gvmax_ptr):
        self._call('b2_dd_commit', _vp(gvmax_ptr))

    def dd_poll(self):
        done, need = _i64(), _int()
        self._call('b2_dd_poll', ctypes.byref(done), ctypes.byref(need))
        return done.value, bool(need.value)

    def dd_end(self):
        self._call('b2_dd_end')

    def dd_pack(self, ids_ptr, n, out_ptr):
        self._call('b2_dd_pack', _vp(ids_ptr or 0), int(n), _vp(out_ptr or 0))

    def dd_unpack(self, first_id, n, in_ptr):
        self._call('b2_dd_unpack', int(first_id), int(n), _vp(in_ptr or 0))

    # -- NCCL fast path (CUDA library only) ----------------------------------------------------
    def nccl_unique_id(self):
        buf = ctypes.create_string_buffer(128)
        self.lib.check(self.lib.dll.b2_nccl_unique_id(buf))
        return buf.raw

    def nccl_init(self, unique_id, rank, nranks):
        self._call('b2_nccl_init', ctypes.c_char_p(bytes(unique_id)), int(rank), int(nranks))

    def nccl_set_halo(self, send_l_ptr, n_send_l, send_r_ptr, n_send_r, n_ghost_l, n_ghost_r):
        self._call('b2_nccl_set_halo', _vp(send_l_ptr or 0), int(n_send_l), _vp(send_r_ptr or 0), int(n_send_r),
                   int(n_ghost_l), int(n_ghost_r))

    def nccl_run(self, count):
        self._call('b2_nccl_run', int(count))


    # -- peer-memory halo (CUDA library only, one rank per GPU) ---------------------------------
    def peer_export(self, halo_capacity, rank, nranks):
        buf = ctypes.create_string_buffer(64)
        self._call('b2_peer_export', int(halo_capacity), int(rank), int(nranks), buf)
        return buf.raw

    def peer_connect(self, handles):
        self._call('b2_peer_connect', ctypes.c_char_p(b''.join(handles)))

    def peer_run(self, count):
        self._call('b2_peer_run', int(count))

    def peer_set_ids(self, gid_ptr, image_ptr=None):
        self._call('b2_peer_set_ids', gid_ptr, image_ptr)

    def peer_refresh(self, x_lo, x_hi, margin):
        """Device-side slab refresh; returns (owned, ghosts_l, ghosts_r, send_l, send_r, out, in)."""
        out = (ctypes.c_int64 * 7)()
        self._call('b2_peer_refresh', float(x_lo), float(x_hi), float(margin), out)
        self.n = int(out[0] + out[1] + out[2])
        return tuple(int(v) for v in out)

    def peer_update(self, n_first, pos_ptr, vel_ptr):
        self._call('b2_peer_update', int(n_first), pos_ptr, vel_ptr)

    def peer_state(self, gid_ptr=None, image_ptr=None, species_ptr=None, mass_ptr=None):
        self._call('b2_peer_state', gid_ptr, image_ptr, species_ptr, mass_ptr)


class Batch(object):

    """An ensemble of small independent replicas sharing cell and pair table
    (b2_batch_*): one thread block per replica, all pairs, whole minimisation
    in one launch."""

    def __init__(self, side, eps, sigma, rcut, shifted=True, precision=FP64, device=0, library=None):
        self.lib = library if library is not None else load()
        self.precision = int(precision)
        handle = _vp()
        self.lib.check(self.lib.dll.b2_batch_create(int(device), self.precision, ctypes.byref(handle)))
        self._h = handle
        side = numpy.ascontiguousarray(side, dtype=numpy.float64)
        eps = numpy.atleast_2d(numpy.ascontiguousarray(eps, dtype=numpy.float64))
        sigma = numpy.atleast_2d(numpy.ascontiguousarray(sigma, dtype=numpy.float64))
        rcut = numpy.atleast_2d(numpy.ascontiguousarray(rcut, dtype=numpy.float64))
        self._call('b2_batch_set_model', _dp(side), eps.shape[0], _dp(eps), _dp(sigma), _dp(rcut), int(bool(shifted)))
        self.nrep = self.n = 0

    def _call(self, name, *args):
        if not self._h:
            raise B2Error(-4, 'batch is closed')
        self.lib.check(getattr(self.lib.dll, name)(self._h, *args))

    def close(self):
        h, self._h = getattr(self, '_h', None), None
        if h:
            self.lib.dll.b2_batch_destroy(h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, position, species, mass=None):
        """position (nrep, n, 3); species (n,) shared by all replicas or (nrep, n)."""
        pos = numpy.ascontiguousarray(position, dtype=numpy.float64)
        if pos.ndim != 3 or pos.shape[2] != 3:
            raise ValueError('position must be (nrep, n, 3)')
        nrep, n = pos.shape[:2]
        spe = numpy.asarray(species, dtype=numpy.int32)
        spe = numpy.ascontiguousarray(numpy.broadcast_to(spe, (nrep, n)))
        m = None if mass is None else numpy.ascontiguousarray(numpy.broadcast_to(numpy.asarray(mass, dtype=numpy.float64), (nrep, n)))
        self._call('b2_batch_upload', nrep, n, _dp(pos), _ip(spe), _dp(m))
        self.nrep, self.n = nrep, n

    def minimize(self, method, max_steps, dt, fn2_tol):
        steps = numpy.zeros(self.nrep, dtype=numpy.int64)
        conv = numpy.zeros(self.nrep, dtype=numpy.int32)
        energy, fn2, virial = numpy.zeros(self.nrep), numpy.zeros(self.nrep), numpy.zeros(self.nrep)
        self._call('b2_batch_minimize', int(method), int(max_steps), float(dt), float(fn2_tol),
                   steps.ctypes.data_as(_c_lp), conv.ctypes.data_as(ctypes.POINTER(_int)), _dp(energy), _dp(fn2), _dp(virial))
        return {'steps': steps, 'converged': conv.astype(bool), 'energy': energy, 'fn2': fn2, 'virial': virial}

    def set_velocities(self, velocity):
        vel = numpy.ascontiguousarray(velocity, dtype=numpy.float64)
        if vel.shape != (self.nrep, self.n, 3):
            raise ValueError('velocity must be (nrep, n, 3)')
        self._call('b2_batch_set_velocities', _dp(vel))

    def run_nve(self, steps, dt):
        kin, energy, virial = numpy.zeros(self.nrep), numpy.zeros(self.nrep), numpy.zeros(self.nrep)
        self._call('b2_batch_run_nve', int(steps), float(dt), _dp(kin), _dp(energy), _dp(virial))
        return {'kinetic': kin, 'energy': energy, 'virial': virial}

    def state(self):
        vel = numpy.empty((self.nrep, self.n, 3))
        img = numpy.empty((self.nrep, self.n, 3), dtype=numpy.int32)
        self._call('b2_batch_get_state', _dp(vel), _ip(img))
        return vel, img

    def download(self, forces=False):
        pos = numpy.empty((self.nrep, self.n, 3))
        frc = numpy.empty((self.nrep, self.n, 3)) if forces else None
        self._call('b2_batch_download', _dp(pos), _dp(frc))
        return (pos, frc) if forces else pos
