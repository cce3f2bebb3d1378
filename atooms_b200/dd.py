"""
Spatial slab decomposition of one large periodic system over several GPUs.

The reference has no domain decomposition: its only
multi-process route is `mpirun -n lammps_mpi_tasks` inside the LAMMPS
subprocess (atooms/backends/lammps.py:61-62).  Here each rank (one process per
GPU) owns the particles of one slab along x and keeps read-only copies
("ghosts") of the neighbours' particles within r_list of its two faces.

  every step     fused step kernel on the owned particles (b2_dd_step)
                 -> all_reduce(MAX) of ONE float, the max |v|^2, so that every
                    rank takes the same list-refresh decision (b2_dd_commit)
                 -> halo: pack boundary positions, send/recv with the two ring
                    neighbours (NCCL over NVLink through torch.distributed),
                    unpack into the ghosts
  every refresh  (displacement bound > skin/2, ~ every 15-20 steps; host-
                 synchronous) migrate particles that left the slab, re-select
                 the ghosts, re-upload the domain, rebuild the Verlet list

Coordinates stay global on every rank (same box), so packed positions need no
translation.  The engine primitives do no communication; everything that
moves between ranks is a torch tensor, so the same code runs on CUDA tensors
with the NCCL backend (product) and on CPU tensors with gloo (tests, with the
CPU oracle standing in as the engine).
"""
import time

import numpy
import torch
import torch.distributed as dist

from . import capi, models

__all__ = ['SlabDomain']


class SlabDomain(object):

    def __init__(self, position, velocity, species, side, model='lennard_jones', dt=0.002, skin=0.3,
                 precision=64, mass=None, gid=None, center=None, device=None, library=None, group=None,
                 fast_comm=None, peer=None, device_refresh=None):
        """`position` ... are the particles this rank holds initially (global
        coordinates, any x: the first refresh sends strays to the ring neighbour
        that owns them; they must not be more than one slab away)."""
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.model = models.get(model)
        self.dt, self.skin, self.precision = float(dt), float(skin), int(precision)
        self.side = numpy.asarray(side, dtype=numpy.float64)
        self.center = numpy.zeros(3) if center is None else numpy.asarray(center, dtype=numpy.float64)
        self.rlist = float(self.model.rcut.max()) + self.skin
        self.width = self.side[0] / self.world
        if self.world > 1 and self.width < 2.0 * self.rlist * (1 + 1e-9):
            raise ValueError('slab width %g is below 2 r_list = %g: use fewer ranks' % (self.width, 2 * self.rlist))
        lib = library if library is not None else capi.load()
        on_gpu = lib.name.startswith('cuda')
        if device is None:
            device = torch.device('cuda', torch.cuda.current_device()) if on_gpu else torch.device('cpu')
        self.dev = torch.device(device)
        self.engine = capi.Engine(self.precision, device=self.dev.index or 0, library=lib)
        if on_gpu:
            self.engine.set_stream(torch.cuda.current_stream(self.dev).cuda_stream)
        self.engine.set_box(self.side, self.center)
        self.engine.set_model(self.model)
        self.engine.set_skin(self.skin)
        f64 = dict(dtype=torch.float64, device=self.dev)
        n = len(species)
        self.pos = torch.as_tensor(numpy.ascontiguousarray(position, dtype=numpy.float64).reshape(n, 3)).to(**f64)
        self.vel = torch.zeros((n, 3), **f64) if velocity is None else \
            torch.as_tensor(numpy.ascontiguousarray(velocity, dtype=numpy.float64).reshape(n, 3)).to(**f64)
        self.mass = torch.ones(n, **f64) if mass is None else torch.as_tensor(numpy.asarray(mass, dtype=numpy.float64)).to(**f64)
        self.spe = torch.as_tensor(numpy.asarray(species, dtype=numpy.int32)).to(self.dev)
        self.gid = torch.arange(n, dtype=torch.int64, device=self.dev) if gid is None else \
            torch.as_tensor(numpy.asarray(gid, dtype=numpy.int64)).to(self.dev)
        self.img = torch.zeros((n, 3), dtype=torch.int32, device=self.dev)
        self._vmax = torch.zeros(2, dtype=torch.float32, device=self.dev)
        # NCCL fast path: the per-step sequence (step, all-reduce, commit, pack, send/recv,
        # unpack) enqueued from C (b2_nccl_run) instead of one Python round trip per collective
        if fast_comm is None:
            fast_comm = on_gpu and lib.has_nccl and self.world > 1 and dist.get_backend(group) == 'nccl'
        self._fast = bool(fast_comm)
        # a backend that cannot move device memory (gloo) gets host-staged messages
        self._stage = on_gpu and self.world > 1 and dist.get_backend(group) != 'nccl'
        self._cdev = torch.device('cpu') if self._stage else self.dev
        if self._fast:
            box = [self.engine.nccl_unique_id() if self.rank == 0 else None]
            dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
            self.engine.nccl_init(box[0], self.rank, self.world)
        # peer-memory halo (b2_peer_*): boundary positions are stored straight into the neighbours'
        # HBM over NVLink and the all-reduce is a mailbox wait: no library call per step.
        # Needs the C path's set-up (NCCL backend, one rank per GPU); connected at the first refresh.
        if peer is None:
            peer = self._fast
        self._peer = bool(peer) and self._fast
        self._peer_cap = 0
        # refresh on the device (b2_peer_refresh): migration, ghost selection and re-ingestion in
        # five kernels, exchanged through the same mailboxes; the engine then owns the global ids
        # and image counters too, and the tensors below are fetched back only on demand
        self._dev_refresh = self._peer if device_refresh is None else (bool(device_refresh) and self._peer)
        self._ids_resident = False
        self._n_own_res = 0
        self._uploads = 0
        self._stale = False              # owned particles were overwritten: ghosts and list are out of date
        self.profile = None
        self._resident = False
        self._batch = 8
        self.refreshes = 0
        self.steps = 0
        self.halo_bytes_per_step = 0
        self.counters = {'rounds': 0, 'enqueued': 0, 'refreshes': 0, 'refresh_ms': 0.0}

    # -- helpers ---------------------------------------------------------------------
    @property
    def n_own(self):
        return self._n_own_res if self._ids_resident else self.pos.shape[0]

    def _left(self):
        return (self.rank - 1) % self.world

    def _right(self):
        return (self.rank + 1) % self.world

    def _exchange(self, to_left, to_right):
        """The refresh exchange: each argument is (migrants (m, 12), ghosts (g, 4)); both go to
        that neighbour as one flat message.  Returns ((migrants, ghosts) from the left,
        (migrants, ghosts) from the right).  Posting order [to-left, to-right] / [from-right,
        from-left] keeps the pairing unambiguous when both neighbours are the same rank
        (world size 2)."""
        counts = torch.tensor([to_left[0].shape[0], to_left[1].shape[0],
                               to_right[0].shape[0], to_right[1].shape[0]], dtype=torch.int64, device=self._cdev)
        all_counts = [torch.empty_like(counts) for _ in range(self.world)]
        dist.all_gather(all_counts, counts, group=self.group)
        mr, gr = (int(v) for v in all_counts[self._right()][0:2])     # the right neighbour's to-left message
        ml, gl = (int(v) for v in all_counts[self._left()][2:4])
        flat = lambda pair: torch.cat([pair[0].reshape(-1), pair[1].reshape(-1),
                                       pair[0].new_zeros(1)]).to(self._cdev).contiguous()
        from_right = torch.empty(12 * mr + 4 * gr + 1, dtype=torch.float64, device=self._cdev)
        from_left = torch.empty(12 * ml + 4 * gl + 1, dtype=torch.float64, device=self._cdev)
        ops = [dist.P2POp(dist.isend, flat(to_left), self._left(), self.group),
               dist.P2POp(dist.isend, flat(to_right), self._right(), self.group),
               dist.P2POp(dist.irecv, from_right, self._right(), self.group),
               dist.P2POp(dist.irecv, from_left, self._left(), self.group)]
        for req in dist.batch_isend_irecv(ops):
            req.wait()
        from_left, from_right = from_left.to(self.dev), from_right.to(self.dev)
        split = lambda t, m, g: (t[:12 * m].reshape(m, 12), t[12 * m:12 * m + 4 * g].reshape(g, 4))
        return split(from_left, ml, gl), split(from_right, mr, gr)

    def _tick(self, name):
        """Phase timer of refresh() (only when self.profile is a dict; synchronises)."""
        if self.profile is None:
            return
        import time
        if self.dev.type == 'cuda':
            torch.cuda.synchronize(self.dev)
        now = time.perf_counter()
        if name is not None:
            self.profile[name] = self.profile.get(name, 0.0) + (now - self._t_last) * 1e3
        self._t_last = now

    # -- refresh: migration + ghost selection + re-upload ---------------------------------
    def _refresh_on_device(self):
        lo, hi = self.rank / self.world, (self.rank + 1) / self.world
        margin = self.rlist / self.side[0] * (1 + 1e-6) + 1e-12
        c = self.engine.peer_refresh(lo, hi, margin)
        self._stale = False
        self._n_own_res, self._n_gl, self._n_gr = c[0], c[1], c[2]
        self.halo_bytes_per_step = 24 * (c[3] + c[4])
        self.pos = self.vel = self.mass = self.spe = self.gid = self.img = None     # stale: see _pull_state
        self.refreshes += 1

    def _pull_state(self):
        """Fetch the owned particles' state back from the engine (device-refresh mode)."""
        n, m = self._n_own_res, self.engine.n
        i32 = dict(dtype=torch.int32, device=self.dev)
        pos = torch.empty((m, 3), dtype=torch.float64, device=self.dev)
        vel = torch.empty((m, 3), dtype=torch.float64, device=self.dev)
        gid = torch.empty(m, dtype=torch.int64, device=self.dev)
        img, spe = torch.empty((m, 3), **i32), torch.empty(m, **i32)
        mass = torch.empty(m, dtype=torch.float64, device=self.dev)
        self.engine.download_device(pos.data_ptr(), vel.data_ptr(), 0)
        self.engine.peer_state(gid.data_ptr(), img.data_ptr(), spe.data_ptr(), mass.data_ptr())
        self.pos, self.vel, self.gid, self.img = pos[:n], vel[:n], gid[:n], img[:n]       # img: total crossings
        self.spe, self.mass = spe[:n], mass[:n]

    def refresh(self):
        if self._ids_resident and self._resident:
            return self._refresh_on_device()
        self._tick(None)
        side = torch.as_tensor(self.side, dtype=torch.float64, device=self.dev)
        ctr = torch.as_tensor(self.center, dtype=torch.float64, device=self.dev)
        if self._resident:
            n = self.n_own
            img = torch.empty((self.engine.n, 3), dtype=torch.int32, device=self.dev)
            pos = torch.empty((self.engine.n, 3), dtype=torch.float64, device=self.dev)
            vel = torch.empty((self.engine.n, 3), dtype=torch.float64, device=self.dev)
            self.engine.download_device(pos.data_ptr(), vel.data_ptr(), img.data_ptr())
            self.pos, self.vel = pos[:n], vel[:n]          # already folded into the global cell
            self.img += img[:n]
        else:
            # fold the initial coordinates (atooms/system/particle.py:90-99); like b2_upload,
            # this fold is not counted as a box crossing
            rel = self.pos - ctr
            self.pos = rel - torch.round(rel / side) * side + ctr
        self._tick('download')
        # which slab owns each particle now; only the few that changed slab are looked at further
        u = (self.pos[:, 0] - ctr[0]) / side[0] + 0.5
        slab = torch.clamp(torch.floor(u * self.world).to(torch.int64), 0, self.world - 1)
        moved = torch.nonzero(slab != self.rank).flatten()
        hop = slab[moved] - self.rank
        if self.world > 2:
            hop = (hop + self.world // 2) % self.world - self.world // 2
            if moved.numel() and int(hop.abs().max()) > 1:
                raise RuntimeError('a particle moved more than one slab between two refreshes')
        else:
            hop = torch.ones_like(hop)                    # two ranks: either way round the ring
        idx_l, idx_r = moved[hop == -1], moved[hop == 1]
        self._tick('classify')

        def payload(idx):
            return torch.cat([self.pos[idx], self.vel[idx], self.mass[idx, None],
                              self.gid[idx, None].to(torch.float64), self.spe[idx, None].to(torch.float64),
                              self.img[idx].to(torch.float64)], dim=1)

        def ghost(idx):
            return torch.cat([self.pos[idx], self.spe[idx, None].to(torch.float64)], dim=1)
        if self.world > 1:
            # One exchange carries both the migrants and the ghost layer.  The ghost message
            # holds the particles that STAY here and lie within r_list of the face; the ones
            # that just crossed that face are ghosts of the rank they left as well, but that
            # rank already has their coordinates, so it appends them itself (same order on
            # both sides: [staying near-face particles, then the migrants in sending order]).
            gone = idx_l.numel() + idx_r.numel()
            keep = torch.nonzero(slab == self.rank).flatten() if gone else None
            x_in = u if keep is None else u[keep]
            lo, hi = self.rank / self.world, (self.rank + 1) / self.world
            margin = self.rlist / self.side[0] * (1 + 1e-6) + 1e-12
            near_l = torch.nonzero(x_in < lo + margin).flatten()          # numbering after the removal
            near_r = torch.nonzero(x_in >= hi - margin).flatten()
            src_l = near_l if keep is None else keep[near_l]
            src_r = near_r if keep is None else keep[near_r]
            mig_l, mig_r = payload(idx_l), payload(idx_r)
            (from_l, g_from_l), (from_r, g_from_r) = self._exchange(
                (mig_l, ghost(src_l)), (mig_r, ghost(src_r)))
            n_keep = self.n_own - gone
            n_fl, n_fr = from_l.shape[0], from_r.shape[0]
            if gone or n_fl or n_fr:
                new = torch.cat([from_l, from_r], dim=0)
                sel = (lambda t: t[keep]) if gone else (lambda t: t)
                self.pos = torch.cat([sel(self.pos), new[:, 0:3]], dim=0)
                self.vel = torch.cat([sel(self.vel), new[:, 3:6]], dim=0)
                self.mass = torch.cat([sel(self.mass), new[:, 6]], dim=0)
                self.gid = torch.cat([sel(self.gid), new[:, 7].round().to(torch.int64)], dim=0)
                self.spe = torch.cat([sel(self.spe), new[:, 8].round().to(torch.int32)], dim=0)
                self.img = torch.cat([sel(self.img), new[:, 9:12].round().to(torch.int32)], dim=0)
            if n_fl:
                near_l = torch.cat([near_l, torch.arange(n_keep, n_keep + n_fl, device=self.dev)])
            if n_fr:
                near_r = torch.cat([near_r, torch.arange(n_keep + n_fl, n_keep + n_fl + n_fr, device=self.dev)])
            if idx_l.numel():
                g_from_l = torch.cat([g_from_l, mig_l[:, [0, 1, 2, 8]]], dim=0)
            if idx_r.numel():
                g_from_r = torch.cat([g_from_r, mig_r[:, [0, 1, 2, 8]]], dim=0)
            self._send_l, self._send_r = near_l.to(torch.int32).contiguous(), near_r.to(torch.int32).contiguous()
        else:
            g_from_l = g_from_r = self.pos.new_zeros((0, 4))
            self._send_l = self._send_r = torch.zeros(0, dtype=torch.int32, device=self.dev)
        n = self.n_own
        self._tick('exchange')
        self._n_gl, self._n_gr = g_from_l.shape[0], g_from_r.shape[0]
        ghosts = torch.cat([g_from_l, g_from_r], dim=0)
        ng = ghosts.shape[0]
        pos = torch.cat([self.pos, ghosts[:, 0:3]], dim=0).contiguous()
        vel = torch.cat([self.vel, self.vel.new_zeros((ng, 3))], dim=0).contiguous()
        mass = torch.cat([self.mass, self.mass.new_ones(ng)], dim=0).contiguous()
        spe = torch.cat([self.spe, ghosts[:, 3].round().to(torch.int32) | capi.GHOST], dim=0).contiguous()
        self.engine.upload_device(n + ng, pos.data_ptr(), spe.data_ptr(), vel.data_ptr(), mass.data_ptr())
        self._uploads += 1
        self._tick('upload')
        if not self._fast:          # halo buffers of the generic (torch.distributed) path
            f64 = dict(dtype=torch.float64, device=self.dev)
            self._out_l = torch.zeros((max(len(self._send_l), 1), 3), **f64)
            self._out_r = torch.zeros((max(len(self._send_r), 1), 3), **f64)
            self._in_l = torch.zeros((max(self._n_gl, 1), 3), **f64)
            self._in_r = torch.zeros((max(self._n_gr, 1), 3), **f64)
        self.halo_bytes_per_step = 24 * (len(self._send_l) + len(self._send_r))
        if self._fast:
            self.engine.nccl_set_halo(self._send_l.data_ptr(), len(self._send_l), self._send_r.data_ptr(),
                                      len(self._send_r), self._n_gl, self._n_gr)
        if self._peer and self._peer_cap == 0:
            # mailboxes are sized once, from the largest halo of any rank plus 50 % headroom for
            # density fluctuations (b2_peer_run refuses a halo that outgrows them); the IPC
            # handles travel by all_gather
            need = torch.tensor([max(len(self._send_l), len(self._send_r), self._n_gl, self._n_gr)],
                                dtype=torch.int64, device=self.dev)
            dist.all_reduce(need, op=dist.ReduceOp.MAX, group=self.group)
            self._peer_cap = int(need) + int(need) // 2 + 4096
            try:
                mine = self.engine.peer_export(self._peer_cap, self.rank, self.world)
                handles = [None] * self.world
                dist.all_gather_object(handles, mine, group=self.group)
                self.engine.peer_connect(handles)
                ok = 1
            except capi.B2Error:
                ok = 0
            flag = torch.tensor([ok], dtype=torch.int64, device=self.dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
            if int(flag) == 0:
                self._peer = False              # no peer access on this box: stay on NCCL
        if self._peer and self._dev_refresh:
            # hand the ids over; the engine's first upload has no headroom for a domain that grows,
            # a repeated one has (12 %)
            if self._uploads == 1:
                self._uploads += 1
                self.engine.upload_device(n + ng, pos.data_ptr(), spe.data_ptr(), vel.data_ptr(), mass.data_ptr())
            none = torch.full((ng,), -1, dtype=torch.int64, device=self.dev)
            gid_all = torch.cat([self.gid, none]).contiguous()
            img_all = torch.cat([self.img, torch.zeros((ng, 3), dtype=torch.int32, device=self.dev)]).contiguous()
            self.engine.peer_set_ids(gid_all.data_ptr(), img_all.data_ptr())
            self._ids_resident, self._n_own_res = True, n
        self._tick('halo setup')
        self._resident = True
        self._stale = False
        self.refreshes += 1

    def _halo(self):
        """Ghost positions of the new time level: pack -> ring exchange -> unpack."""
        if self.world == 1:
            return
        eng, n = self.engine, self.n_own
        eng.dd_pack(self._send_l.data_ptr(), len(self._send_l), self._out_l.data_ptr())
        eng.dd_pack(self._send_r.data_ptr(), len(self._send_r), self._out_r.data_ptr())
        out_l, out_r = self._out_l.to(self._cdev), self._out_r.to(self._cdev)
        in_l = self._in_l if not self._stage else torch.empty_like(self._in_l, device=self._cdev)
        in_r = self._in_r if not self._stage else torch.empty_like(self._in_r, device=self._cdev)
        ops = [dist.P2POp(dist.isend, out_l, self._left(), self.group),
               dist.P2POp(dist.isend, out_r, self._right(), self.group),
               dist.P2POp(dist.irecv, in_r, self._right(), self.group),
               dist.P2POp(dist.irecv, in_l, self._left(), self.group)]
        for req in dist.batch_isend_irecv(ops):
            req.wait()
        if self._stage:
            self._in_l.copy_(in_l)
            self._in_r.copy_(in_r)
        eng.dd_unpack(n, self._n_gl, self._in_l.data_ptr())
        eng.dd_unpack(n + self._n_gl, self._n_gr, self._in_r.data_ptr())

    # -- dynamics -------------------------------------------------------------------------
    def run(self, steps):
        """`steps` velocity-Verlet NVE steps of the whole decomposed system."""
        steps = int(steps)
        if steps <= 0:
            return
        if not self._resident or self._stale:
            self.refresh()
        eng = self.engine
        eng.dd_begin(steps, self.dt, True)
        base, since_refresh = 0, 0
        while base < steps:
            m = min(self._batch, steps - base)
            for _ in range(0 if self._fast else m):
                eng.dd_step(self._vmax.data_ptr())
                if self.world > 1 and not self._stage:
                    dist.all_reduce(self._vmax[:1], op=dist.ReduceOp.MAX, group=self.group)
                elif self.world > 1:
                    v = self._vmax[:1].to(self._cdev)
                    dist.all_reduce(v, op=dist.ReduceOp.MAX, group=self.group)
                    self._vmax[:1].copy_(v)
                eng.dd_commit(self._vmax.data_ptr())
                self._halo()
            if self._peer:
                eng.peer_run(m)
            elif self._fast:
                eng.nccl_run(m)
            done, need = eng.dd_poll()
            since_refresh += done
            base += done
            c = self.counters
            c['rounds'] += 1
            c['enqueued'] += m
            if need:
                # every rank sees the same flag (the bound comes from the global max)
                t0 = time.perf_counter()
                self._batch = max(1, min(since_refresh, 64))
                self.refresh()
                eng.dd_begin(steps - base, self.dt, False)
                since_refresh = 0
                c['refreshes'] += 1
                c['refresh_ms'] += (time.perf_counter() - t0) * 1e3     # both ends are synchronous
            elif done == m and base < steps:
                self._batch = min(self._batch + 2, 64)
            if base < steps and not need:
                eng.dd_begin(steps - base, self.dt, False)     # new budget for the next round
        eng.dd_end()
        self.steps += steps

    # -- observables -----------------------------------------------------------------------
    def thermo(self):
        """Totals over the whole system (owned particles of every rank, all_reduce SUM)."""
        if not self._resident or self._stale:
            self.refresh()
        t = self.engine.thermo()
        vec = torch.tensor([t['kinetic'], t['potential'], t['virial'], t['fnorm2'], *t['momentum'], t['mass'],
                            float(self.n_own)], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            vec = vec.to(self._cdev)
            dist.all_reduce(vec, op=dist.ReduceOp.SUM, group=self.group)
        v = vec.cpu().numpy()
        return {'kinetic': v[0], 'potential': v[1], 'virial': v[2], 'fnorm2': v[3], 'momentum': v[4:7],
                'mass': v[7], 'n': int(round(v[8]))}

    def gather(self):
        """(gid-ordered) global position, velocity, image on every rank — for tests."""
        if self._ids_resident and self._resident:
            self._pull_state()
            pos, vel, img = self.pos, self.vel, self.img
        elif self._resident:
            n = self.n_own
            pos = torch.empty((self.engine.n, 3), dtype=torch.float64, device=self.dev)
            vel = torch.empty((self.engine.n, 3), dtype=torch.float64, device=self.dev)
            img = torch.empty((self.engine.n, 3), dtype=torch.int32, device=self.dev)
            self.engine.download_device(pos.data_ptr(), vel.data_ptr(), img.data_ptr())
            pos, vel, img = pos[:n], vel[:n], self.img + img[:n]
        else:
            pos, vel, img = self.pos, self.vel, self.img
        mine = torch.cat([self.gid[:, None].to(torch.float64), pos, vel, img.to(torch.float64)], dim=1).cpu()
        if self.world > 1:
            parts = [None] * self.world
            dist.all_gather_object(parts, mine.numpy(), group=self.group)
            allp = numpy.concatenate(parts, axis=0)
        else:
            allp = mine.numpy()
        order = numpy.argsort(allp[:, 0])
        allp = allp[order]
        return allp[:, 1:4], allp[:, 4:7], numpy.round(allp[:, 7:10]).astype(numpy.int64)

    def load_host(self, position, velocity):
        """Replace the owned particles' coordinates and velocities from host arrays
        (pinned memory makes the copies asynchronous); same particles, same order."""
        if self._ids_resident and self._resident:
            # the engine owns the particles: overwrite the owned ones in place (pinned host or device
            # memory, read by the engine directly); the next refresh re-sends the ghost layer
            p, v = torch.as_tensor(position), torch.as_tensor(velocity)
            ok = all(t.dtype == torch.float64 and t.is_contiguous() and tuple(t.shape) == (self._n_own_res, 3)
                     and (t.is_cuda or t.is_pinned()) for t in (p, v))
            if ok:
                self.engine.peer_update(self._n_own_res, p.data_ptr(), v.data_ptr())
                self._stale = True
                return
        if self._ids_resident:
            self._pull_state()              # ids, species, masses in the engine's current order
            self._ids_resident = False
        self.pos = torch.as_tensor(position).to(self.dev, non_blocking=True)
        self.vel = torch.as_tensor(velocity).to(self.dev, non_blocking=True)
        self._resident = False

    def store_host(self, position, velocity):
        """Copy the owned particles' state into host arrays of shape (n_own, 3)."""
        n = self.n_own
        pos = torch.empty((self.engine.n, 3), dtype=torch.float64, device=self.dev)
        vel = torch.empty((self.engine.n, 3), dtype=torch.float64, device=self.dev)
        self.engine.download_device(pos.data_ptr(), vel.data_ptr(), 0)
        torch.as_tensor(position)[:n].copy_(pos[:n])
        torch.as_tensor(velocity)[:n].copy_(vel[:n])
        return n

    def close(self):
        self.engine.close()
