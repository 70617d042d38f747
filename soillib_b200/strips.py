"""Row-strip decomposition of the erosion step across GPUs (new work - the reference is
single-GPU).

The map is cut into contiguous row strips along dim 0 (the slow axis, so a strip and each of its halo rows is
one contiguous block).  Per step:

  1. every rank launches the particle kernel: all ranks draw all start positions from identical RNG streams
     and run the particles that start in their rows;
  2. particles that left a strip are handed to the neighbour (64-byte records) and continued there, in rounds,
     until none is in flight anywhere - the union over ranks is exactly the single-GPU particle set;
  3. every rank runs the fused dense pass on its rows (deposits are always local, so no flux reduction);
  4. halo rows are exchanged with the neighbours: one row of `layers` (the dense stencils), and `halo` rows of the
     maps the particles gather from (height, discharge, momentum, debris momentum).

Peer-mapped mode replaces 2: a particle that moves into a neighbour's rows keeps going inside the same kernel, its
deposits going to the neighbour's records over NVLink.  With `halo` >= the distance a particle can travel (one row
per crossing, so `maxage`) all of its gathers stay in local memory - only the posted reductions cross the link, and
the particle kernel is as fast as on one GPU; with a shallower halo it falls back to gathering from the owner.

`Transport` moves rows and records: `DistTransport` with torch.distributed point-to-point ops (NCCL over NVLink
on GPUs, gloo in the CPU tests), `LocalTransport` between several strips living in one process (how the
multi-rank path is tested on a single GPU).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _lib

RECORD_FLOATS = 16  # one particle in flight = 64 bytes (transport.cu pstate)


def bind_host_to_gpu(device_index: int) -> bool:
    """Pins the calling thread (and the threads it starts later: NCCL's proxy, pinned-memory first touch) to the CPUs
    next to GPU `device_index`, as NVML reports them.  One process per GPU on a two-socket box otherwise lands on
    either socket, and a strip's upload / download then crosses the socket interconnect.  Returns False (and changes
    nothing) where NVML or the affinity call is unavailable."""
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            pr = torch.cuda.get_device_properties(device_index)
            bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
            handle = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        except Exception:  # noqa: BLE001 - older torch: no PCI ids; NVML order == CUDA order without CUDA_VISIBLE_DEVICES
            handle = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        pynvml.nvmlDeviceSetCpuAffinity(handle)
        return True
    except Exception:  # noqa: BLE001
        return False


def partition_rows(nx: int, world: int):
    """Contiguous, balanced row ranges [(lo, hi), ...] for `world` ranks."""
    if world < 1 or nx < world:
        raise ValueError("need at least one row per rank")
    base, extra = divmod(nx, world)
    out, lo = [], 0
    for r in range(world):
        hi = lo + base + (1 if r < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


class Strip:
    """Device state of one rank: its rows of every map plus halo rows, flux records, RNG streams, outboxes."""

    def __init__(self, global_nx, ny, lo, hi, scale, n_particles, albedo=False, seed=0, with_creep=True,
                 device=None, outbox_capacity=None, halo=1, view=False):
        lib = _lib.load()
        self.lib = lib
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.global_nx, self.ny, self.lo, self.hi = int(global_nx), int(ny), int(lo), int(hi)
        if halo < 1:
            raise ValueError("halo must be at least one row")
        self.halo = int(halo)
        self.halo_lo = min(self.halo, lo)
        self.halo_hi = min(self.halo, global_nx - hi)
        self.row0 = lo - self.halo_lo
        self.rows = (hi - lo) + self.halo_lo + self.halo_hi
        self.scale = tuple(float(v) for v in scale)
        self.n_particles = int(n_particles)
        self.albedo, self.with_creep = bool(albedo), bool(with_creep)
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=self.device)  # noqa: E731
        r, n = self.rows, self.ny
        self.layers = [z(r, n, 2), z(r, n, 2)]
        self.cur = 0
        self.height, self.uplift = z(r, n), z(r, n)
        self.rainfall = torch.ones((r, n), dtype=torch.float32, device=self.device)
        self.discharge, self.mass, self.debris = z(r, n), z(r, n), z(r, n)
        self.momentum, self.debris_momentum = z(r, n, 2), z(r, n, 2)
        self.albedo_bedrock = torch.full((r, n, 3), 0.5, device=self.device) if albedo else None
        self.albedo_surface = torch.full((r, n, 3), 0.5, device=self.device) if albedo else None
        nflux = lib.soilb200_erode_flux_bytes(r, n, int(albedo))
        self.flux = [torch.zeros(nflux, dtype=torch.uint8, device=self.device) for _ in range(2)]
        self.tags = [torch.zeros(lib.soilb200_erode_ws(r, n), dtype=torch.uint8, device=self.device) for _ in range(2)]
        self.fcur = 0
        self.pre = torch.empty(lib.soilb200_erode_pre_bytes(r, n), dtype=torch.uint8, device=self.device)
        # packed particle view (soilb200_model_t.view): the particles gather from it and the dense pass writes it instead of
        # height / discharge / momentum / debris / debris_momentum.  Its rows come in pairs (2 x 2-cell quads), so the halo
        # blocks must be whole pairs: an even halo depth and an even number of owned rows.
        self.view = None
        if view:
            if self.halo_lo % 2 or self.halo_hi % 2 or (hi - lo) % 2:
                raise ValueError("a strip with a packed view needs an even halo depth and an even number of rows")
            self.rowsq, self.nyq = (r + 1) // 2, (n + 1) // 2
            raw = torch.zeros(lib.soilb200_erode_view_bytes(r, n), dtype=torch.uint8, device=self.device)
            self.view = raw.view(torch.float32).view(2, self.rowsq, self.nyq * 16)   # [system][row pair][quads of 4 x 4 floats]
        self.view_valid = False
        self.spawn = torch.empty(lib.soilb200_erode_spawn_bytes(self.n_particles), dtype=torch.uint8, device=self.device)
        self.rng_f = torch.empty((self.n_particles, 48), dtype=torch.uint8, device=self.device)
        self.rng_d = torch.empty((self.n_particles, 48), dtype=torch.uint8, device=self.device)
        s = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(lib.soilb200_seed_xorwow(self.rng_f.data_ptr(), self.n_particles, seed, 0, 0, s), "seed")
        _lib.check(lib.soilb200_seed_xorwow(self.rng_d.data_ptr(), self.n_particles, seed + 1, 0, 0, s), "seed")
        self.capacity = int(outbox_capacity or max(1024, 2 * self.n_particles))
        self.outbox = [torch.zeros((self.capacity, RECORD_FLOATS), dtype=torch.float32, device=self.device) for _ in range(2)]
        self.counts = torch.zeros(2, dtype=torch.int32, device=self.device)
        self.mass_sum = torch.zeros(2, dtype=torch.float64, device=self.device)
        self.c = _lib.Model()
        self.peer_tables = None   # [table for fcur == 0, table for fcur == 1] once connect()ed
        self.peer_self = 0
        self._peer_keepalive = []

    # ---- initial conditions: slices of global maps, halo rows included
    def own(self, t):
        """The owned rows of a local map."""
        return t[self.halo_lo:self.halo_lo + (self.hi - self.lo)]

    def set_from_global(self, bedrock=None, sediment=None, uplift=None, rainfall=None):
        sl = slice(self.row0, self.row0 + self.rows)
        L = self.layers[self.cur]
        if bedrock is not None:
            L[..., 0] = bedrock[sl].to(self.device)
        if sediment is not None:
            L[..., 1] = sediment[sl].to(self.device)
        if uplift is not None:
            self.uplift.copy_(uplift[sl].to(self.device))
        if rainfall is not None:
            self.rainfall.copy_(rainfall[sl].to(self.device))
        torch.add(L[..., 0], L[..., 1], out=self.height)
        self.view_valid = False

    def import_view(self):
        """maps -> packed view (all rows, halo rows included); needed before the first step and after the maps were edited."""
        if self.view is not None and not self.view_valid:
            self._fill(0, False)
            _lib.check(self.lib.soilb200_erode_import(C.byref(self.c), self._stream()), "erode import")
            self.view_valid = True

    def edit(self):
        """Call before writing into the strip's maps by hand: brings the maps up to date and schedules a re-import."""
        self.export_view()
        self.view_valid = False

    def export_view(self):
        """packed view -> height, discharge, momentum, debris, debris_momentum (for whoever wants to read them)."""
        if self.view is not None and self.view_valid:
            self._fill(0, False)
            _lib.check(self.lib.soilb200_erode_export(C.byref(self.c), self._stream()), "erode export")

    # ---- C-ABI plumbing
    def _fill(self, phases, want_sum):
        c, p = self.c, (lambda t: t.data_ptr())
        c.nx, c.ny = self.rows, self.ny
        c.scale[0], c.scale[1], c.scale[2] = self.scale
        c.with_creep = 1 if self.with_creep else 0
        c.layers, c.layers_next, c.height = p(self.layers[self.cur]), p(self.layers[1 - self.cur]), p(self.height)
        c.rainfall, c.uplift = p(self.rainfall), p(self.uplift)
        c.discharge, c.mass, c.momentum = p(self.discharge), p(self.mass), p(self.momentum)
        c.debris, c.debris_momentum = p(self.debris), p(self.debris_momentum)
        c.albedo_bedrock = p(self.albedo_bedrock) if self.albedo else None
        c.albedo_surface = p(self.albedo_surface) if self.albedo else None
        c.rng_fluvial, c.rng_debris, c.n_particles = p(self.rng_f), p(self.rng_d), self.n_particles
        c.flux, c.touched = p(self.flux[self.fcur]), p(self.tags[self.fcur])
        c.flux_prev, c.touched_prev = p(self.flux[1 - self.fcur]), p(self.tags[1 - self.fcur])
        c.mass_sum = p(self.mass_sum) if want_sum else None
        c.global_nx, c.global_row0, c.own_row_lo, c.own_row_hi = self.global_nx, self.row0, self.lo, self.hi
        c.outbox_lo, c.outbox_hi, c.outbox_counts = p(self.outbox[0]), p(self.outbox[1]), p(self.counts)
        c.outbox_capacity, c.phases = self.capacity, phases
        c.pre = p(self.pre)
        c.view = p(self.view) if self.view is not None else None
        c.dense_row_lo = c.dense_row_hi = 0
        c.spawn = None if os.environ.get("SOILB200_NO_SPAWN") else p(self.spawn)
        if self.peer_tables is not None:
            t = self.peer_tables[self.fcur]
            c.peers, c.n_peers, c.peer_self = p(t), t.numel() // C.sizeof(_lib.Peer), self.peer_self
        else:
            c.peers, c.n_peers, c.peer_self = None, 0, 0

    def shared_tensors(self):
        """The arrays a particle of another rank may gather from or deposit into."""
        return [self.height, self.discharge, self.momentum, self.debris_momentum, self.flux[0], self.flux[1],
                self.tags[0], self.tags[1]] + ([self.view] if self.view is not None else [])

    def set_peers(self, entries, self_index, keepalive=()):
        """entries[r] = (lo, hi, row0, rows, [8 tensors or raw device addresses as in shared_tensors(), valid on this GPU])."""
        tables = []
        for f in (0, 1):
            arr = (_lib.Peer * len(entries))()
            for r, (lo, hi, row0, rows, ts) in enumerate(entries):
                ptr = [t if isinstance(t, int) else t.data_ptr() for t in ts]
                arr[r].own_row_lo, arr[r].own_row_hi, arr[r].global_row0, arr[r].rows = lo, hi, row0, rows
                arr[r].height, arr[r].discharge = ptr[0], ptr[1]
                arr[r].momentum, arr[r].debris_momentum = ptr[2], ptr[3]
                arr[r].flux, arr[r].touched = ptr[4 + f], ptr[6 + f]
                arr[r].view = ptr[8] if len(ptr) > 8 else None
            raw = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
            tables.append(raw.to(self.device))
        self.peer_tables, self.peer_self = tables, int(self_index)
        self._peer_keepalive = list(keepalive)

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def launch_particles(self, param):
        self.counts.zero_()
        self._fill(1, False)
        _lib.check(self.lib.soilb200_erode_step(C.byref(self.c), param.c(), self._stream()), "erode particles")

    def resume(self, param, inbox, n_in):
        self.counts.zero_()
        if n_in == 0:
            return
        self._fill(1, False)
        _lib.check(self.lib.soilb200_erode_resume(C.byref(self.c), param.c(), inbox.data_ptr(), int(n_in),
                                                  self._stream()), "erode resume")

    def launch_dense(self, param, want_sum=False):
        self._fill(2, want_sum)
        self.c.dense_row_lo = self.c.dense_row_hi = 0
        _lib.check(self.lib.soilb200_erode_step(C.byref(self.c), param.c(), self._stream()), "erode dense")
        self.cur, self.fcur = 1 - self.cur, 1 - self.fcur

    # ---- the dense pass in two parts: the rows the neighbours keep as halo first, so that they can travel while
    # the interior is computed.  launch_dense_bands + launch_dense_interior == launch_dense.
    def dense_plan(self):
        """([(lo, hi), ...] boundary bands in global rows, (lo, hi) interior or None)."""
        own = self.hi - self.lo
        d_lo = min(self.halo, own) if self.lo > 0 else 0
        d_hi = min(self.halo, own - d_lo) if self.hi < self.global_nx else 0
        bands = ([(self.lo, self.lo + d_lo)] if d_lo else []) + ([(self.hi - d_hi, self.hi)] if d_hi else [])
        inner = (self.lo + d_lo, self.hi - d_hi)
        return bands, (inner if inner[0] < inner[1] else None)

    def _dense_part(self, param, rows, first, last, want_sum):
        self._fill(2 | (0 if last else 4) | (0 if first else 8), want_sum)
        self.c.dense_row_lo, self.c.dense_row_hi = rows
        _lib.check(self.lib.soilb200_erode_step(C.byref(self.c), param.c(), self._stream()), "erode dense")

    def launch_dense_bands(self, param, want_sum=False):
        """The owned rows next to a cut.  Afterwards layers[1 - cur] / height / ... hold their new values there."""
        bands, inner = self.dense_plan()
        for i, rows in enumerate(bands):
            self._dense_part(param, rows, i == 0, inner is None and i == len(bands) - 1, want_sum)
        return len(bands)

    def launch_dense_interior(self, param, want_sum=False):
        bands, inner = self.dense_plan()
        if inner is not None:
            self._dense_part(param, inner, not bands, True, want_sum)
        self.cur, self.fcur = 1 - self.cur, 1 - self.fcur

    # ---- halo rows.  Every block is a contiguous slice (rows are the slow axis): nothing is packed.
    def halo_maps(self, nxt=False):
        """(map, depth) pairs to refresh after the dense pass: the current layers one row deep (dense stencils), the
        maps the particles gather from `halo` rows deep.  nxt: between launch_dense_bands and launch_dense_interior
        the new layers are still the `next` buffer."""
        d = self.halo
        L = self.layers[1 - self.cur if nxt else self.cur]
        if self.view is not None:   # the view replaces the four particle maps; its rows go in pairs (handled by _rows)
            return [(L, 1), (self.view[0], d), (self.view[1], d)]
        return [(L, 1), (self.height, d), (self.discharge, d), (self.momentum, d), (self.debris_momentum, d)]

    def _rows(self, t, a, b):
        """Local rows [a, b) of a halo map; the two views are indexed by row pair."""
        if self.view is not None and t.data_ptr() in (self.view[0].data_ptr(), self.view[1].data_ptr()):
            return t[a // 2:b // 2]
        return t[a:b]

    def send_blocks(self, side, depth_of_receiver, nxt=False):
        """The owned rows next to `side` (0: first rows, 1: last rows) that the neighbour on that side keeps as halo."""
        own0, own1 = self.halo_lo, self.halo_lo + (self.hi - self.lo)
        out = []
        for t, d in self.halo_maps(nxt):
            d = min(d, depth_of_receiver)
            if d > self.hi - self.lo:
                raise ValueError("halo deeper than a neighbouring strip")
            out.append(self._rows(t, own0, own0 + d) if side == 0 else self._rows(t, own1 - d, own1))
        return out

    def recv_blocks(self, side, nxt=False):
        """The halo rows on `side`, innermost `depth` rows of each map (views to receive into)."""
        own0, own1 = self.halo_lo, self.halo_lo + (self.hi - self.lo)
        have = self.halo_lo if side == 0 else self.halo_hi
        out = []
        for t, d in self.halo_maps(nxt):
            d = min(d, have)
            out.append(self._rows(t, own0 - d, own0) if side == 0 else self._rows(t, own1, own1 + d))
        return out

    # the one-row packed form [layers row (ny*2) | height row (ny)] (kept for callers that move single rows)
    def boundary_row(self, side):
        """Packed first (side 0) or last (side 1) OWNED row of the current layers + height."""
        i = self.halo_lo if side == 0 else self.halo_lo + (self.hi - self.lo) - 1
        return torch.cat([self.layers[self.cur][i].reshape(-1), self.height[i]])

    def set_halo_row(self, side, packed):
        i = self.halo_lo - 1 if side == 0 else self.halo_lo + (self.hi - self.lo)
        self.layers[self.cur][i].copy_(packed[: 2 * self.ny].view(self.ny, 2))
        self.height[i].copy_(packed[2 * self.ny:])


class LocalTransport:
    """All strips live in this process (several logical ranks on one GPU): rows and records are copied."""

    def connect(self, strips):
        """Peer mode: every strip's particles may read / deposit into every other strip's arrays directly."""
        entries = [(s.lo, s.hi, s.row0, s.rows, s.shared_tensors()) for s in strips]
        for r, s in enumerate(strips):
            s.set_peers(entries, r)

    def barrier(self, strips):
        pass  # one stream: launches are already ordered

    def counts(self, strips):
        return [s.counts.cpu().tolist() for s in strips]

    def exchange_particles(self, strips, counts):
        inboxes = []
        for r, s in enumerate(strips):
            parts = []
            if r > 0 and counts[r - 1][1]:
                parts.append(strips[r - 1].outbox[1][: counts[r - 1][1]])
            if r + 1 < len(strips) and counts[r + 1][0]:
                parts.append(strips[r + 1].outbox[0][: counts[r + 1][0]])
            inboxes.append(torch.cat(parts).contiguous() if parts else None)
        return inboxes

    def exchange_halos(self, strips, nxt=False):
        for r, s in enumerate(strips):
            if r > 0:        # my low halo <- the last owned rows of the strip below
                for dst, src in zip(s.recv_blocks(0, nxt), strips[r - 1].send_blocks(1, s.halo_lo, nxt)):
                    dst.copy_(src)
            if r + 1 < len(strips):
                for dst, src in zip(s.recv_blocks(1, nxt), strips[r + 1].send_blocks(0, s.halo_hi, nxt)):
                    dst.copy_(src)

    def exchange_halos_begin(self, strips):
        """Between the boundary bands and the interior of the dense pass (one stream here: nothing overlaps)."""
        self.exchange_halos(strips, nxt=True)
        return None

    def exchange_halos_end(self, strips, handle):
        pass

    def reduce_sum(self, values):
        return sum(values)


class DistTransport:
    """One strip per process; neighbours talk through torch.distributed point-to-point ops."""

    def __init__(self, group=None, high_priority=True):
        import torch.distributed as dist
        self.dist = dist
        if group is None and high_priority and dist.get_backend() == "nccl":
            # NCCL's kernels on a high-priority stream: the halo send/recv that overlaps the interior dense pass must not
            # queue behind that pass's quarter of a million blocks
            try:
                opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)
                group = dist.new_group(backend="nccl", pg_options=opts)
            except Exception:  # noqa: BLE001 - older torch: fall back to the default group
                group = None
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)

    def connect(self, strips):
        """Peer mode over CUDA IPC: every rank publishes its strip's arrays, every rank maps all of them.

        One process per GPU on one NVSwitch box; the mapped pointers are used by the particle kernel for loads
        and reductions over NVLink."""
        s = strips[0]
        lib = s.lib
        exported = []
        with torch.cuda.device(s.device):
            for t in s.shared_tensors():
                handle, off = C.create_string_buffer(64), C.c_uint64(0)
                _lib.check(lib.soilb200_ipc_export(t.data_ptr(), handle, C.byref(off)), "ipc_export")
                exported.append((handle.raw, int(off.value)))
        mine = (s.lo, s.hi, s.row0, s.rows, exported)
        everyone = [None] * self.world
        self.dist.all_gather_object(everyone, mine, group=self.group)
        entries, opened = [], {}
        self._opened = opened
        with torch.cuda.device(s.device):
            for r, (lo, hi, row0, rows, exp) in enumerate(everyone):
                if r == self.rank:
                    ptrs = s.shared_tensors()
                else:
                    ptrs = []
                    for handle, off in exp:
                        if handle not in opened:     # an allocation can be opened once per process
                            base = C.c_void_p()
                            _lib.check(lib.soilb200_ipc_open(handle, C.byref(base)), "ipc_open")
                            opened[handle] = int(base.value)
                        ptrs.append(opened[handle] + off)
                entries.append((lo, hi, row0, rows, ptrs))
        s.set_peers(entries, self.rank)
        torch.cuda.synchronize(s.device)
        self.dist.barrier(group=self.group)

    def disconnect(self, strips):
        """Unmaps the peers' arrays (before any rank frees them) and drops the peer tables."""
        s = strips[0]
        torch.cuda.synchronize(s.device)
        self.dist.barrier(group=self.group)          # nobody's kernels still use a mapping
        with torch.cuda.device(s.device):
            for base in getattr(self, "_opened", {}).values():
                _lib.check(s.lib.soilb200_ipc_close(base), "ipc_close")
        self._opened = {}
        s.peer_tables = None
        self.dist.barrier(group=self.group)

    def barrier(self, strips):
        """Orders the phases across ranks on the device: a one-element all-reduce is stream-ordered after this
        rank's kernels and completes only when every rank has reached it - no host synchronisation."""
        if not hasattr(self, "_token"):
            self._token = torch.zeros(1, dtype=torch.float32, device=strips[0].device)
        self.dist.all_reduce(self._token, group=self.group)

    def all_counts(self, counts_tensor):
        """[world][2] outbox counts of every rank, on the host (one small all-gather per hand-off round)."""
        mine = counts_tensor.to(torch.int32)
        out = [torch.zeros_like(mine) for _ in range(self.world)]
        self.dist.all_gather(out, mine, group=self.group)
        return [t.cpu().tolist() for t in out]

    def counts(self, strips):
        return self.all_counts(strips[0].counts)

    def exchange_records(self, out_lo, out_hi, counts, device):
        """Ships out_lo[:n] to rank-1 and out_hi[:n] to rank+1; returns the records received, concatenated."""
        d, r = self.dist, self.rank
        ops, recv = [], []
        if r > 0:
            if counts[r][0]:
                ops.append(d.P2POp(d.isend, out_lo[: counts[r][0]].contiguous(), r - 1, self.group))
            if counts[r - 1][1]:
                buf = torch.empty((counts[r - 1][1], RECORD_FLOATS), dtype=torch.float32, device=device)
                ops.append(d.P2POp(d.irecv, buf, r - 1, self.group))
                recv.append(buf)
        if r + 1 < self.world:
            if counts[r][1]:
                ops.append(d.P2POp(d.isend, out_hi[: counts[r][1]].contiguous(), r + 1, self.group))
            if counts[r + 1][0]:
                buf = torch.empty((counts[r + 1][0], RECORD_FLOATS), dtype=torch.float32, device=device)
                ops.append(d.P2POp(d.irecv, buf, r + 1, self.group))
                recv.append(buf)
        if ops:
            for w in d.batch_isend_irecv(ops):
                w.wait()
        return torch.cat(recv).contiguous() if recv else None

    def exchange_particles(self, strips, counts):
        s = strips[0]
        return [self.exchange_records(s.outbox[0], s.outbox[1], counts, s.device)]

    def exchange_rows(self, first_row, last_row):
        """Sends my first owned row down-rank and my last owned row up-rank; returns (halo_lo, halo_hi) or None."""
        d, r = self.dist, self.rank
        ops, lo, hi = [], None, None
        if r > 0:
            lo = torch.empty_like(first_row)
            ops += [d.P2POp(d.isend, first_row.contiguous(), r - 1, self.group), d.P2POp(d.irecv, lo, r - 1, self.group)]
        if r + 1 < self.world:
            hi = torch.empty_like(last_row)
            ops += [d.P2POp(d.isend, last_row.contiguous(), r + 1, self.group), d.P2POp(d.irecv, hi, r + 1, self.group)]
        if ops:
            for w in d.batch_isend_irecv(ops):
                w.wait()
        return lo, hi

    def exchange_halos(self, strips, nxt=False):
        """One batch of point-to-point ops: every map's boundary rows go straight from the owner's array into the
        neighbour's halo rows (contiguous slices on both sides, no staging).  All strips use the same `halo`."""
        s, d, r = strips[0], self.dist, self.rank
        ops = []
        if r > 0:
            ops += [d.P2POp(d.isend, t, r - 1, self.group) for t in s.send_blocks(0, s.halo, nxt)]
            ops += [d.P2POp(d.irecv, t, r - 1, self.group) for t in s.recv_blocks(0, nxt)]
        if r + 1 < self.world:
            ops += [d.P2POp(d.isend, t, r + 1, self.group) for t in s.send_blocks(1, s.halo, nxt)]
            ops += [d.P2POp(d.irecv, t, r + 1, self.group) for t in s.recv_blocks(1, nxt)]
        if ops:
            for w in d.batch_isend_irecv(ops):
                w.wait()

    def exchange_halos_begin(self, strips):
        """Issues the halo exchange on a communication stream, ordered after what the caller's stream holds now (the
        boundary bands of the dense pass): the send/recv run while the caller's stream computes the interior."""
        s = strips[0]
        if not s.device.type == "cuda":
            self.exchange_halos(strips, nxt=True)
            return None
        if not hasattr(self, "_comm"):
            self._comm = torch.cuda.Stream(device=s.device, priority=-1)
        main = torch.cuda.current_stream(s.device)
        ready = torch.cuda.Event()
        ready.record(main)
        self._comm.wait_event(ready)
        with torch.cuda.stream(self._comm):
            self.exchange_halos(strips, nxt=True)
            done = torch.cuda.Event()
            done.record(self._comm)
        return done

    def exchange_halos_end(self, strips, handle):
        if handle is not None:
            torch.cuda.current_stream(strips[0].device).wait_event(handle)

    def reduce_sum(self, values):
        t = torch.tensor([float(sum(values))], dtype=torch.float64, device=values_device(values))
        self.dist.all_reduce(t, group=self.group)
        return float(t.item())


def values_device(values):
    return "cuda" if torch.cuda.is_available() else "cpu"


def _mark(strip):
    e = torch.cuda.Event(enable_timing=True)
    e.record(torch.cuda.current_stream(strip.device))
    return e


def erode_strips(strips, transport, param, steps=1, max_rounds=None, events=None, mass_sum=False, overlap=True, export=True):
    """`steps` erosion steps over the strips this process holds (one under DistTransport, all under LocalTransport).

    mass_sum: the last step's dense pass also leaves sum(bedrock), sum(sediment) of the owned rows in strip.mass_sum.
    overlap (peer-mapped strips): the dense pass runs the rows next to a cut first and ships them to the neighbours on
    a communication stream while the interior is computed.
    export: strips with a packed view write their height / discharge / momentum / debris maps from it before returning
    (pass False in a loop that does not look at them, and call strip.export_view() at the end).

    Returns the number of particle hand-off rounds of the last step.  A particle changes strips at most once per
    crossing, so maxage rounds always suffice (a particle zig-zagging along a strip boundary is the worst case);
    with strips of thousands of rows a handful of rounds is typical, and late rounds carry few particles.
    """
    if max_rounds is None:
        max_rounds = int(param.maxage) + 64
    rounds = 0
    for s in strips:
        s.import_view()          # (no-op without a view, or when it is current)
    try:
        return _erode_strips(strips, transport, param, steps, max_rounds, events, mass_sum, overlap)
    finally:
        if export:
            for s in strips:
                s.export_view()


def _erode_strips(strips, transport, param, steps, max_rounds, events, mass_sum, overlap):
    rounds = 0
    if strips[0].peer_tables is not None:
        # peer-mapped strips: particles roam across ranks inside ONE kernel per rank; only the phases are ordered
        mark = (lambda: events.append(_mark(strips[0]))) if events is not None else (lambda: None)
        for _ in range(int(steps)):
            mark()                         # (events: 6 CUDA events per step, for the phase breakdown of bench.py)
            for s in strips:
                s.launch_particles(param)
            mark()
            transport.barrier(strips)      # every rank's deposits are in before any dense pass reads them
            mark()
            want_sum = bool(mass_sum) and _ == int(steps) - 1
            if overlap:
                # the rows the neighbours keep as halo first; they travel while the interior is computed
                for s in strips:
                    s.launch_dense_bands(param, want_sum)
                handle = transport.exchange_halos_begin(strips)
                for s in strips:
                    s.launch_dense_interior(param, want_sum)
                mark()
                transport.exchange_halos_end(strips, handle)
            else:
                for s in strips:
                    s.launch_dense(param, want_sum)
                mark()
                transport.exchange_halos(strips)
            mark()
            transport.barrier(strips)      # every rank's maps are final before any particle reads them
            mark()
        return 0
    for _ in range(int(steps)):
        for s in strips:
            s.launch_particles(param)
        rounds = 0
        while True:
            counts = transport.counts(strips)          # [rank][lo, hi] for every rank
            for c, cap in ((c, strips[0].capacity) for c in counts):
                if max(c) > cap:
                    raise RuntimeError(f"particle outbox overflow: {max(c)} > {cap}; raise outbox_capacity")
            if sum(sum(c) for c in counts) == 0:
                break
            rounds += 1
            if rounds > max_rounds:
                raise RuntimeError("particle hand-off did not terminate")
            inboxes = transport.exchange_particles(strips, counts)
            for s, box in zip(strips, inboxes):
                s.resume(param, box, 0 if box is None else box.shape[0])
        for s in strips:
            s.launch_dense(param, bool(mass_sum) and _ == int(steps) - 1)
        transport.exchange_halos(strips)
    return rounds
