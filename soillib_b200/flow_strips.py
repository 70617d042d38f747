"""The radius-1/2 stencil ops of the flow path over row strips (first row): direction, steepest
random_weighted and normal shard naturally - a rank holds its rows of the height map plus `halo` rows of each neighbour (2 for the
normal map's 4th-order stencil, 1 for the others), runs the single-GPU kernel on that block and keeps its own rows.
Results are bit-identical to the undivided map; `steepest` returns GLOBAL flat indices.

    s = HeightStrip.from_global(h, lo, hi, global_nx)          # or fill s.block yourself and call refresh()
    s.refresh(transport)                                       # halo rows from the neighbours (NCCL send/recv)
    d = s.direction(soil.d8); g = s.steepest(soil.d8); n = s.normal(scale)
"""
from __future__ import annotations

import torch

from . import ops, silt


class HeightStrip:
    """Rows [lo, hi) of a (global_nx, ny) float32 height map plus up to `halo` rows on each side."""

    def __init__(self, global_nx, ny, lo, hi, halo=2, device=None):
        if not 0 <= lo < hi <= global_nx or halo < 1:
            raise ValueError("HeightStrip: bad row range or halo")
        self.global_nx, self.ny, self.lo, self.hi, self.halo = int(global_nx), int(ny), int(lo), int(hi), int(halo)
        self.halo_lo, self.halo_hi = min(halo, lo), min(halo, global_nx - hi)
        self.row0 = lo - self.halo_lo
        self.rows = (hi - lo) + self.halo_lo + self.halo_hi
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.block = torch.zeros((self.rows, self.ny), dtype=torch.float32, device=self.device)

    @classmethod
    def from_global(cls, h, lo, hi, halo=2, device=None):
        """Cuts the strip (halo rows included) out of a whole map held on the host or on a device."""
        s = cls(h.shape[0], h.shape[1], lo, hi, halo, device)
        s.block.copy_(h[s.row0:s.row0 + s.rows].to(s.device))
        return s

    def own(self, t):
        return t[self.halo_lo:self.halo_lo + (self.hi - self.lo)]

    # ---- halo rows: contiguous row blocks, straight between the arrays
    def send_block(self, side, depth):
        o0, o1 = self.halo_lo, self.halo_lo + (self.hi - self.lo)
        if depth > self.hi - self.lo:
            raise ValueError("halo deeper than a neighbouring strip")
        return self.block[o0:o0 + depth] if side == 0 else self.block[o1 - depth:o1]

    def recv_block(self, side):
        o0, o1 = self.halo_lo, self.halo_lo + (self.hi - self.lo)
        return self.block[o0 - self.halo_lo:o0] if side == 0 else self.block[o1:o1 + self.halo_hi]

    def refresh(self, dist_transport):
        """Receives the halo rows from the neighbouring ranks (one batch of point-to-point ops)."""
        d, r, world, group = dist_transport.dist, dist_transport.rank, dist_transport.world, dist_transport.group
        ops_ = []
        if r > 0:
            ops_ += [d.P2POp(d.isend, self.send_block(0, self.halo), r - 1, group), d.P2POp(d.irecv, self.recv_block(0), r - 1, group)]
        if r + 1 < world:
            ops_ += [d.P2POp(d.isend, self.send_block(1, self.halo), r + 1, group), d.P2POp(d.irecv, self.recv_block(1), r + 1, group)]
        if ops_:
            for w in d.batch_isend_irecv(ops_):
                w.wait()

    # ---- the ops: the single-GPU kernel on the haloed block, own rows kept
    def _t(self):
        return silt.tensor(silt.float32, _data=self.block)

    def direction(self, e):
        return silt.tensor(silt.int32, _data=self.own(ops.direction(self._t(), e).data).contiguous())

    def steepest(self, e):
        """Receiver per owned cell as a flat index into the GLOBAL map (-1: none)."""
        g = self.own(ops.steepest(self._t(), e).data)
        g = torch.where(g >= 0, g + self.row0 * self.ny, g)
        return silt.tensor(silt.int32, _data=g.contiguous())

    def random_weighted(self, e, seed, offset, T):
        """Gibbs-random receiver per owned cell as a GLOBAL flat index: every cell draws from the stream of its global
        index (curand_init(seed, global n, offset)), so the strips decide what the undivided map decides."""
        g = self.own(ops.random_weighted(self._t(), e, seed, offset, T, first_cell=self.row0 * self.ny).data)
        g = torch.where(g >= 0, g + self.row0 * self.ny, g)
        return silt.tensor(silt.int32, _data=g.contiguous())

    def normal(self, scale=(1.0, 1.0, 1.0)):
        if self.halo < 2:
            raise ValueError("normal needs two halo rows")
        return silt.tensor(silt.float32, _data=self.own(ops.normal(self._t(), scale).data).contiguous())
