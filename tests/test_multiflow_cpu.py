"""Replica-parallel multi-flow accumulation (soillib_b200/multiflow.py) on CPU: the sample partition, and the
partition + all-reduce protocol on the gloo backend (world_size 2 and 3) with the oracle as the per-rank compute."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from soillib_b200.multiflow import multiflow_replicas, sample_range


def test_sample_range_partitions_all_samples():
    for K in (0, 1, 7, 16, 512):
        for world in (1, 2, 3, 8):
            parts = [sample_range(K, r, world) for r in range(world)]
            assert parts[0][0] == 0 and sum(c for _, c in parts) == K
            assert all(a[0] + a[1] == b[0] for a, b in zip(parts, parts[1:]))
            assert max(c for _, c in parts) - min(c for _, c in parts) <= 1
    assert sample_range(512, 3, 8) == (192, 64)
    with pytest.raises(ValueError):
        sample_range(8, 2, 2)


def _oracle_partial(h, rain, e, K, T, seed, k_first, k_count):
    """dem_multiflow.py:43-49 on the CPU oracle for a sample range: float32 quotient, float64 sum."""
    import oracle_api as orc
    total = np.zeros(h.shape, np.float64)
    for k in range(k_first, k_first + k_count):
        g, _ = orc.random_weighted(h, e, seed, k, T)
        total += orc.accumulate(g, rain, None, e, mode=1)[0] / np.float32(K)
    return torch.from_numpy(total)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, h, rain, K, want):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        got = multiflow_replicas(h, rain, 1, K, 10.0, seed=0, compute=_oracle_partial)
        np.testing.assert_allclose(got.numpy(), want, rtol=1e-13, atol=0)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_replicas_sum_to_the_serial_mean(world):
    import oracle_api as orc
    h = orc.noise(24, 32, 3.0)
    rain = np.ones_like(h)
    K = 7
    want = _oracle_partial(h, rain, 1, K, 10.0, 0, 0, K).numpy()
    assert want.min() > 0 and abs(want.sum() / want.size - want.mean()) < 1e-9
    mp.spawn(_worker, args=(world, _free_port(), h, rain, K, want), nprocs=world, join=True)
