"""soilb200_accumulate_static_f32: accumulation for callers that may not allocate device memory.

Some hosts give this library nothing but pointers to their own maps: an application that bans cudaMalloc
from its processing loop because it locks and synchronises the whole device, or an embedding with no
allocator to take scratch from.  For them the pending-donor counters live in a fixed 16 MiB device array
inside the library instead of a caller workspace.  This file is named to be collected last: if the entry
point breaks, a run that stops at the first failure has already run every other test.

The check: through the C ABI, on the same graph and source, the static-scratch entry point must return
what the workspace entry point returns - on two calls in a row, the second of a different, smaller size so
that it runs over counters the first call left behind.  The static entry point always takes the global
dependency walk.  The workspace entry point takes the same walk when it is given 4 bytes of scratch per cell
or a decay map, and then the two results are identical bit for bit, real-valued sources included; given its
full workspace and no decay it sums inside shared-memory tiles first, which re-associates the float additions:
identical for integer-valued sources (unit rain), within 1e-5 relative otherwise - the tolerance of the other
accumulate tests.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_CELLS = 4 * 1024 * 1024   # SOILB200_ACCUMULATE_STATIC_MAX_CELLS in include/soilb200.h


def _both(lib, graph, src, decay, edge, walk_only=False):
    nx, ny = graph.shape
    ws = torch.empty(4 * nx * ny if walk_only else lib.soilb200_accumulate_ws(nx, ny, edge), dtype=torch.uint8, device="cuda")
    want = torch.full((nx, ny), -7.0, dtype=torch.float32, device="cuda")
    got = torch.full((nx, ny), -9.0, dtype=torch.float32, device="cuda")
    d = None if decay is None else decay.data_ptr()
    assert lib.soilb200_accumulate_f32(graph.data_ptr(), src.data_ptr(), d, want.data_ptr(), nx, ny, edge,
                                       ws.data_ptr(), ws.numel(), None) == 0
    assert lib.soilb200_accumulate_static_f32(graph.data_ptr(), src.data_ptr(), d, got.data_ptr(), nx, ny, edge,
                                              None) == 0
    torch.cuda.synchronize()
    return got, want


def test_static_scratch_accumulate_matches_workspace_accumulate(product_lib, oracle):
    lib = product_lib
    # first call: a 512 x 384 fBm DEM, D8 receivers, unit rain and a real-valued source with decay
    h1 = oracle.noise(512, 384, 3.0)
    g1 = torch.from_numpy(oracle.steepest(h1, 1)).cuda()
    ones = torch.ones((512, 384), dtype=torch.float32, device="cuda")
    got, want = _both(lib, g1, ones, None, 1)
    assert torch.equal(got, want)
    assert np.array_equal(got.cpu().numpy(), oracle.accumulate(oracle.steepest(h1, 1), np.ones_like(h1), None, 1, mode=1)[0])

    # second call, straight after: smaller ragged map, D4, decay - runs over the first call's leftovers
    h2 = oracle.noise(97, 61, 0.0)
    g2 = torch.from_numpy(oracle.steepest(h2, 0)).cuda()
    src = torch.from_numpy((np.abs(h2) + 0.25).astype(np.float32)).cuda()
    dec = torch.full((97, 61), 0.9, dtype=torch.float32, device="cuda")
    got2, want2 = _both(lib, g2, src, dec, 0)
    assert torch.equal(got2, want2)
    # a real-valued source without decay: bit-identical to the workspace call that takes the same walk (4 bytes of
    # scratch per cell), within 1e-5 relative of the tiled sums of the full workspace
    got2b, want2b = _both(lib, g2, src, None, 0, walk_only=True)
    assert torch.equal(got2b, want2b)
    _, tiled = _both(lib, g2, src, None, 0)
    torch.testing.assert_close(got2b, tiled, rtol=1e-5, atol=0)

    # and the first problem again: identical to its first answer
    got3, _ = _both(lib, g1, ones, None, 1)
    assert torch.equal(got3, got)


def test_static_scratch_accumulate_rejects_oversize_and_bad_arguments(product_lib):
    lib = product_lib
    from soillib_b200 import _lib
    nx, ny = 2048, 2049                                   # one row of cells beyond the fixed array
    assert nx * ny > MAX_CELLS
    g = torch.full((nx, ny), -1, dtype=torch.int32, device="cuda")
    s = torch.ones((nx, ny), dtype=torch.float32, device="cuda")
    o = torch.zeros((nx, ny), dtype=torch.float32, device="cuda")
    assert lib.soilb200_accumulate_static_f32(g.data_ptr(), s.data_ptr(), None, o.data_ptr(), nx, ny, 1, None) == _lib.E_LIMIT
    assert o.abs().sum().item() == 0                      # nothing was launched
    assert lib.soilb200_accumulate_static_f32(g.data_ptr(), s.data_ptr(), None, o.data_ptr(), 64, 64, 9, None) == _lib.E_EDGE
    assert lib.soilb200_accumulate_static_f32(None, s.data_ptr(), None, o.data_ptr(), 64, 64, 1, None) == _lib.E_ARG
    # exactly the maximum is accepted
    g4 = torch.full((2048, 2048), -1, dtype=torch.int32, device="cuda")
    assert lib.soilb200_accumulate_static_f32(g4.data_ptr(), s.data_ptr(), None, o.data_ptr(), 2048, 2048, 1, None) == 0
    torch.cuda.synchronize()
    assert torch.equal(o.ravel()[: 2048 * 2048], s.ravel()[: 2048 * 2048])   # no edges: every cell keeps its source
