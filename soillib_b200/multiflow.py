"""Multi-flow accumulation: the K-sample loop of example/dem_multiflow.py:43-49 as one device-side call, and its
replica-parallel form over several GPUs (last row).

    mean = soil.multiflow(height, rain, soil.d8, K=512, T=10.0)            # float64 (nx, ny), on the GPU

Replicas: the K samples are independent, so R ranks (one process per GPU, each holding the whole DEM) take
contiguous sample ranges and one all-reduce (NCCL, summing float64 maps over NVLink) forms the mean - the only
collective of the path.  `sample_range` is the partition; `multiflow_replicas` the driver.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch

from . import _lib, silt
from .ops import _dims2, _edge, _gpu, _stream, _workspace


def multiflow(height: silt.tensor, source: silt.tensor, e, K: int, T: float, seed: int = 0, k_first: int = 0,
              k_count: Optional[int] = None, out: Optional[silt.tensor] = None) -> silt.tensor:
    """sum over k in [k_first, k_first+k_count) of accumulate(random_weighted(height, e, seed, k, T), source) / K.

    float64 result on the device; k_count defaults to all K samples (then the result is the script's mean, bit for
    bit).  `out` (float64, same shape) is added to instead of being overwritten when given."""
    h = _gpu(height, silt.float32, "multiflow")
    f = _gpu(source, silt.float32, "multiflow")
    nx, ny = _dims2(h, "multiflow")
    if tuple(f.shape[:2]) != (nx, ny):
        raise ValueError("multiflow: height and source shapes differ")
    if K <= 0:
        raise ValueError("multiflow: K must be positive")
    k_count = K - k_first if k_count is None else k_count
    if k_first < 0 or k_count < 0:
        raise ValueError("multiflow: negative sample range")
    lib = _lib.load()
    ek = _edge(e)
    nbytes = lib.soilb200_multiflow_ws(nx, ny, ek if ek in (0, 1) else 1)
    ws = _workspace(("mf", h.device, nx, ny), nbytes + 256, h.device)
    base = (ws.data_ptr() + 255) & ~255
    if out is None:
        acc = torch.empty((nx, ny), dtype=torch.float64, device=h.device)
        add = 0
    else:
        acc = out.data
        if acc.dtype != torch.float64 or tuple(acc.shape) != (nx, ny) or not acc.is_cuda:
            raise ValueError("multiflow: out must be a float64 GPU tensor of the map's shape")
        add = 1
    rc = lib.soilb200_multiflow_f64(h.data_ptr(), f.data_ptr(), acc.data_ptr(), nx, ny, ek, int(seed), int(k_first),
                                    int(k_count), int(K), float(T), add, base, nbytes, _stream())
    _lib.check(rc, "multiflow")
    return silt.tensor(silt.float64, _data=acc)


def sample_range(K: int, rank: int, world: int) -> Tuple[int, int]:
    """(k_first, k_count) of `rank`: contiguous ranges, the first K % world ranks take one sample more."""
    if world <= 0 or not 0 <= rank < world or K < 0:
        raise ValueError("sample_range: bad arguments")
    base, extra = divmod(K, world)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def multiflow_replicas(height, source, e, K: int, T: float, seed: int = 0, group=None,
                       compute: Optional[Callable] = None):
    """Every rank of `group` (torch.distributed; default group if None) holds the same DEM, computes its share of
    the K samples and takes part in one all-reduce; all ranks return the full mean.

    `compute(height, source, e, K, T, seed, k_first, k_count)` defaults to `multiflow`; it exists so that the
    partition + collective can be exercised on the gloo backend without a GPU (tests/test_multiflow_cpu.py)."""
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    k_first, k_count = sample_range(K, rank, world)
    part = (compute or multiflow)(height, source, e, K, T, seed, k_first, k_count)
    data = part.data if isinstance(part, silt.tensor) else part
    dist.all_reduce(data, op=dist.ReduceOp.SUM, group=group)
    return part
