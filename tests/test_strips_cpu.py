"""Host-side logic of the multi-GPU row-strip path, on CPU: partitioning and the neighbour exchange protocol
(torch.distributed, gloo backend, world_size 2 and 3)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from soillib_b200.strips import DistTransport, RECORD_FLOATS, partition_rows


def test_partition_rows():
    assert partition_rows(10, 1) == [(0, 10)]
    assert partition_rows(10, 3) == [(0, 4), (4, 7), (7, 10)]
    p = partition_rows(32768, 8)
    assert p[0] == (0, 4096) and p[-1] == (28672, 32768)
    assert all(a[1] == b[0] for a, b in zip(p, p[1:]))
    with pytest.raises(ValueError):
        partition_rows(2, 3)


def test_bind_host_to_gpu_without_a_gpu_changes_nothing():
    """bench.py's optional NUMA binding: no NVML device here, so it reports False and leaves the affinity alone."""
    from soillib_b200.strips import bind_host_to_gpu
    before = os.sched_getaffinity(0)
    if not torch.cuda.is_available():
        assert bind_host_to_gpu(0) is False
    assert os.sched_getaffinity(0) == before or torch.cuda.is_available()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, results):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        t = DistTransport()
        ny = 7
        # halo rows: every rank sends its first / last owned row, tagged with its rank
        first = torch.full((3 * ny,), float(10 * rank + 1))
        last = torch.full((3 * ny,), float(10 * rank + 2))
        lo, hi = t.exchange_rows(first, last)
        assert (lo is None) == (rank == 0) and (hi is None) == (rank == world - 1)
        if lo is not None:
            assert torch.equal(lo, torch.full((3 * ny,), float(10 * (rank - 1) + 2)))   # neighbour's LAST row
        if hi is not None:
            assert torch.equal(hi, torch.full((3 * ny,), float(10 * (rank + 1) + 1)))   # neighbour's FIRST row

        # particle records: rank r sends r+1 records down-rank and 2 records up-rank
        mine = torch.tensor([rank + 1 if rank > 0 else 0, 2 if rank + 1 < world else 0], dtype=torch.int32)
        counts = t.all_counts(mine)
        assert counts == [[r + 1 if r > 0 else 0, 2 if r + 1 < world else 0] for r in range(world)]
        out_lo = torch.full((8, RECORD_FLOATS), float(100 + rank))
        out_hi = torch.full((8, RECORD_FLOATS), float(200 + rank))
        box = t.exchange_records(out_lo, out_hi, counts, "cpu")
        want = []
        if rank > 0:
            want += [200.0 + rank - 1] * 2                    # what rank-1 sent up
        if rank + 1 < world:
            want += [100.0 + rank + 1] * (rank + 2)           # what rank+1 sent down
        got = [] if box is None else box[:, 0].tolist()
        assert sorted(got) == sorted(want)
        assert box is None or box.shape[1] == RECORD_FLOATS
        # a round with nothing in flight moves nothing
        zero = t.all_counts(torch.zeros(2, dtype=torch.int32))
        assert t.exchange_records(out_lo, out_hi, zero, "cpu") is None
        results[rank] = 1
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_neighbour_exchange_gloo(world):
    port = _free_port()
    results = mp.get_context("spawn").Manager().dict()
    mp.spawn(_worker, args=(world, port, results), nprocs=world, join=True)
    assert sorted(results.keys()) == list(range(world))


def _halo_worker(rank, world, port):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from soillib_b200.flow_strips import HeightStrip
        nx, ny = 23, 5
        full = torch.arange(nx * ny, dtype=torch.float32).reshape(nx, ny)
        lo, hi = partition_rows(nx, world)[rank]
        s = HeightStrip(nx, ny, lo, hi, halo=2, device=torch.device("cpu"))
        s.own(s.block).copy_(full[lo:hi])                    # every rank knows only its own rows
        s.refresh(DistTransport())
        assert torch.equal(s.block, full[s.row0:s.row0 + s.rows])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_height_strip_halo_refresh(world):
    mp.spawn(_halo_worker, args=(world, _free_port()), nprocs=world, join=True)


def _oracle_compute(graph_local, source, e):
    import oracle_api as orc
    return torch.from_numpy(orc.accumulate(graph_local.numpy(), source.numpy(), None, int(e), mode=1)[0])


def _python_terminal(graph_local, starts):
    g = graph_local.reshape(-1).tolist()
    out = []
    for c in starts.tolist():
        while g[c] >= 0 and g[c] != c:
            c = g[c]
        out.append(c)
    return torch.tensor(out, dtype=starts.dtype)


def _acc_worker(rank, world, port, graph, rain, want):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from soillib_b200.accumulate_strips import accumulate_strip_dist
        nx = graph.shape[0]
        lo, hi = partition_rows(nx, world)[rank]
        got = accumulate_strip_dist(graph[lo:hi].contiguous(), rain[lo:hi].contiguous(), lo, hi, nx, 1,
                                    compute=_oracle_compute, terminal=_python_terminal)
        assert torch.equal(got, want[lo:hi])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_accumulate_strips_protocol(world):
    """accumulate over strips (local pass, boundary forest over one all-gather, second pass) with the oracle as the
    per-rank compute: equal to the oracle on the undivided map."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import oracle_api as orc
    nx, ny = 90, 64
    h = orc.noise(nx, ny, 3.0)
    graph = torch.from_numpy(orc.steepest(h, 1))
    rain = torch.ones((nx, ny))
    want = torch.from_numpy(orc.accumulate(graph.numpy(), rain.numpy(), None, 1, mode=1)[0])
    assert want.max() > 100
    mp.spawn(_acc_worker, args=(world, _free_port(), graph, rain, want), nprocs=world, join=True)
