"""Flow accumulation over row strips (third row; new work - the reference is single-GPU).

Every rank holds rows [lo, hi) of a global (nx, ny) flow graph (receivers as GLOBAL flat indices, -1 = none; what
`flow_strips.HeightStrip.steepest` returns) and of the source map.  Upstream sums need what flows in from other
strips, and a river may cross strip boundaries several times, so:

  1. local pass: receivers outside the strip are cut; `accumulate` on the strip gives every cell its in-strip upstream
     sum.  Cells whose receiver was cut are the strip's EXITS (they lie on its first / last row);
  2. boundary summary: for every cell of the first / last row, the exit its in-strip flow path ends in (or none: a
     true root) - one short pointer walk per boundary cell (`soilb200_graph_terminal_i32`);
  3. one all-gather of those summaries (<= 2 ny records per rank): {exit cell, its receiver, its local sum} and
     {boundary cell, the exit it drains to}.  The exits of all strips form a forest - exit e hangs below the exit that
     the receiver of e drains to - which every rank solves redundantly: total(e) = local(e) + sum of total(children);
  4. second local pass with the totals of the foreign exits added to the source at the cells they flow into.

Cost: two local accumulations + a collective of a few hundred KiB.  Sums are formed in a different order than on one
GPU, so results agree to float rounding (exactly for integer-valued sources below 2^24).

`compute` / `terminal` / `gather` are injectable so that the protocol runs on CPU tensors in the gloo tests.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence

import torch

from . import _lib, silt


# ---- the two device ops -------------------------------------------------------------------------------------

def _accumulate_local(graph_local: torch.Tensor, source: torch.Tensor, e) -> torch.Tensor:
    from . import ops
    return ops.accumulate(silt.tensor(silt.int32, _data=graph_local), silt.tensor(silt.float32, _data=source), e).data


def _terminal(graph_local: torch.Tensor, starts: torch.Tensor) -> torch.Tensor:
    out = torch.empty_like(starts)
    if starts.numel():
        rows, ny = graph_local.shape
        s = torch.cuda.current_stream(graph_local.device).cuda_stream
        _lib.check(_lib.load().soilb200_graph_terminal_i32(graph_local.data_ptr(), starts.data_ptr(), out.data_ptr(),
                                                           starts.numel(), rows, ny, s), "graph_terminal")
    return out


# ---- per-strip summary ----------------------------------------------------------------------------------------

class _Summary:
    """What a strip tells the others: fixed-size (2 ny) arrays, unused slots carry index -1."""

    def __init__(self, ny: int, device):
        z = lambda dt, v: torch.full((2 * ny,), v, dtype=dt, device=device)  # noqa: E731
        self.exit_cell, self.exit_recv = z(torch.int64, -1), z(torch.int64, -1)
        self.exit_local = z(torch.float32, 0.0)
        self.b_cell, self.b_drain = z(torch.int64, -1), z(torch.int64, -1)

    def tensors(self):
        return [self.exit_cell, self.exit_recv, self.exit_local, self.b_cell, self.b_drain]


def _summarise(graph: torch.Tensor, lo: int, hi: int, nx: int, acc_local: torch.Tensor, graph_local: torch.Tensor,
               terminal: Callable) -> _Summary:
    rows, ny = graph.shape
    s = _Summary(ny, graph.device)
    brow = ([0] if lo > 0 else []) + ([rows - 1] if hi < nx and (rows > 1 or lo == 0) else [])
    if not brow:
        return s
    cols = torch.arange(ny, device=graph.device, dtype=torch.int64)
    local = torch.cat([r * ny + cols for r in brow])                      # local flat indices of the boundary cells
    k = local.numel()
    g = graph.reshape(-1)[local].to(torch.int64)
    is_exit = (g >= 0) & ((g < lo * ny) | (g >= hi * ny))
    every = ((graph >= 0) & ((graph < lo * ny) | (graph >= hi * ny))).sum()
    if int(every) != int(is_exit.sum()):
        raise ValueError("accumulate_strips: a receiver leaves the strip from an interior row (not a D4/D8 graph)")
    glob = local + lo * ny
    s.exit_cell[:k] = torch.where(is_exit, glob, torch.full_like(glob, -1))
    s.exit_recv[:k] = torch.where(is_exit, g, torch.full_like(g, -1))
    s.exit_local[:k] = torch.where(is_exit, acc_local.reshape(-1)[local], torch.zeros((), device=graph.device))
    # where does each boundary cell's in-strip path end, and is that end an exit?
    term = terminal(graph_local, local.to(torch.int32)).to(torch.int64)
    tg = graph.reshape(-1)[term].to(torch.int64)
    term_is_exit = (tg >= 0) & ((tg < lo * ny) | (tg >= hi * ny))
    s.b_cell[:k] = glob
    s.b_drain[:k] = torch.where(term_is_exit, term + lo * ny, torch.full_like(term, -1))
    return s


# ---- the replicated forest solve ------------------------------------------------------------------------------

def _solve_forest(summaries: Sequence[_Summary]):
    """-> (receiver cells, inflow they get): the totals of all exits, grouped by the cell they flow into."""
    cat = lambda i: torch.cat([s.tensors()[i] for s in summaries])  # noqa: E731
    e_cell, e_recv, e_local, b_cell, b_drain = (cat(i) for i in range(5))
    ev = e_cell >= 0
    e_cell, e_recv, e_local = e_cell[ev], e_recv[ev], e_local[ev]
    if e_cell.numel() == 0:
        return e_recv, e_local
    bv = b_cell >= 0
    b_cell, b_drain = b_cell[bv], b_drain[bv]
    b_cell, order = torch.sort(b_cell)
    b_drain = b_drain[order]
    # parent exit of e: the exit that the receiver of e drains to (the receiver is a boundary cell of its strip)
    pos = torch.searchsorted(b_cell, e_recv).clamp_(max=b_cell.numel() - 1)
    found = b_cell[pos] == e_recv
    parent_cell = torch.where(found, b_drain[pos], torch.full_like(e_recv, -1))
    e_sorted, e_order = torch.sort(e_cell)
    ppos = torch.searchsorted(e_sorted, parent_cell.clamp(min=0)).clamp_(max=e_sorted.numel() - 1)
    has_parent = (parent_cell >= 0) & (e_sorted[ppos] == parent_cell)
    parent = torch.where(has_parent, e_order[ppos], torch.full_like(ppos, -1))
    total = e_local.clone()
    child = has_parent.nonzero().squeeze(1)
    for _ in range(e_cell.numel() + 1):                                   # Jacobi sweeps: one per level of the forest
        nxt = e_local.clone()
        nxt.index_add_(0, parent[child], total[child])
        if torch.equal(nxt, total):
            break
        total = nxt
    else:
        raise RuntimeError("accumulate_strips: the boundary forest has a cycle (invalid flow graph)")
    return e_recv, total


# ---- drivers --------------------------------------------------------------------------------------------------

def _local_graph(graph: torch.Tensor, lo: int, hi: int) -> torch.Tensor:
    ny = graph.shape[1]
    own = (graph >= lo * ny) & (graph < hi * ny)
    return torch.where(own, graph - lo * ny, torch.full_like(graph, -1)).contiguous()


def _with_inflow(source: torch.Tensor, lo: int, hi: int, recv: torch.Tensor, total: torch.Tensor) -> torch.Tensor:
    ny = source.shape[1]
    mine = (recv >= lo * ny) & (recv < hi * ny)
    out = source.clone().reshape(-1)
    out.index_add_(0, recv[mine] - lo * ny, total[mine])
    return out.reshape(source.shape)


def accumulate_strips(graphs: List[torch.Tensor], sources: List[torch.Tensor], ranges, nx: int, e,
                      compute: Optional[Callable] = None, terminal: Optional[Callable] = None) -> List[torch.Tensor]:
    """All strips in one process (how the protocol is tested on one GPU).  graphs[r]: int32 (rows_r, ny) global
    receivers of strip r = rows ranges[r] = (lo, hi); sources[r]: float32 (rows_r, ny).  Returns the accumulation of
    every strip."""
    compute, terminal = compute or _accumulate_local, terminal or _terminal
    locals_ = [_local_graph(g, lo, hi) for g, (lo, hi) in zip(graphs, ranges)]
    first = [compute(gl, s.contiguous(), e) for gl, s in zip(locals_, sources)]
    summ = [_summarise(g, lo, hi, nx, a, gl, terminal) for g, (lo, hi), a, gl in zip(graphs, ranges, first, locals_)]
    recv, total = _solve_forest(summ)
    return [compute(gl, _with_inflow(s, lo, hi, recv, total), e) for gl, s, (lo, hi) in zip(locals_, sources, ranges)]


def accumulate_strip_dist(graph: torch.Tensor, source: torch.Tensor, lo: int, hi: int, nx: int, e, group=None,
                          compute: Optional[Callable] = None, terminal: Optional[Callable] = None) -> torch.Tensor:
    """One strip per rank (torch.distributed; NCCL on GPUs): the same four steps with one all-gather."""
    import torch.distributed as dist
    compute, terminal = compute or _accumulate_local, terminal or _terminal
    gl = _local_graph(graph, lo, hi)
    first = compute(gl, source.contiguous(), e)
    mine = _summarise(graph, lo, hi, nx, first, gl, terminal)
    world = dist.get_world_size(group)
    everyone = [_Summary(graph.shape[1], graph.device) for _ in range(world)]
    for i, t in enumerate(mine.tensors()):
        dist.all_gather([s.tensors()[i] for s in everyone], t, group=group)
    recv, total = _solve_forest(everyone)
    return compute(gl, _with_inflow(source, lo, hi, recv, total), e)
