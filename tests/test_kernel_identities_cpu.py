"""CPU checks of the arithmetic identities the CUDA sweeps rely on, restated in numpy (IEEE binary32, no contraction)
and compared bit for bit with the oracle:

* k_creep_sweep (soillib_b200/csrc/dense.cu): every pair of neighbours is evaluated once, creep_pair(c, q) ==
  -creep_pair(q, c), border cells contribute zeros (reference erosion.cu:654-708);
* k_steepest (soillib_b200/csrc/flow.cu): "strict > from 0, first maximum wins" (reference graph.cu:42-60) as the
  maximum of {0, slopes} followed by the two-level search for the lowest k whose slope equals it.

The GPU tests (-m gpu) check the kernels themselves; these pin the reformulations where no GPU is needed.
"""
import numpy as np
import pytest

f32 = np.float32


def _creep_pair(sc, sq, hc, hq, cd, sz):
    up = hq > hc
    ht, hb, st = np.where(up, hq, hc), np.where(up, hc, hq), np.where(up, sq, sc)
    with np.errstate(all="ignore"):
        tmax = f32(0.5) * ((ht - hb) - cd)
        give = np.fmax(f32(0), np.fmin(st * sz, tmax))
    return np.where(up, give, -give).astype(f32)


def _creep_shared_edges(delta, layers, scale, crit):
    nx, ny, _ = layers.shape
    sx, sy, sz = map(f32, scale)
    cdx, cdy = f32(crit) * sx, f32(crit) * sy
    h = ((layers[..., 0] + layers[..., 1]) * sz).astype(f32)
    sed = layers[..., 1]
    down = np.zeros((nx, ny), f32)            # edge (x,y)-(x+1,y) as seen from row x; zero below the last row
    if nx > 1:
        down[:-1] = _creep_pair(sed[:-1], sed[1:], h[:-1], h[1:], cdx, sz)
    along = np.zeros((nx, ny + 1), f32)       # edge (x,y-1)-(x,y) as seen from column y-1; zero at both borders
    if ny > 1:
        along[:, 1:ny] = _creep_pair(sed[:, :-1], sed[:, 1:], h[:, :-1], h[:, 1:], cdy, sz)
    up = np.zeros((nx, ny), f32)
    up[1:] = down[:-1]
    t = np.zeros((nx, ny), f32)
    t = t + down                              # the reference's order: x+1, x-1, y+1, y-1
    t = t + (-up)
    t = t + along[:, 1:]
    t = t + (-along[:, :-1])
    out = delta.copy()
    out[..., 1] = out[..., 1] + (f32(0.25) * t / sz).astype(f32)
    return out


@pytest.mark.parametrize("crit", [0.3, 0.0])
def test_creep_pairs_evaluated_once_match_the_oracle(oracle, crit):
    rng = np.random.default_rng(7)
    for nx, ny in ((64, 132), (300, 260), (5, 4), (1, 8), (67, 128)):
        layers = np.zeros((nx, ny, 2), f32)
        layers[..., 0] = oracle.noise(nx, ny, 3.0)
        layers[..., 1] = rng.random((nx, ny), dtype=f32) * f32(0.05)
        layers[nx // 3: nx // 3 + 4, : ny // 2] = (0.25, 0.01)   # a plateau: equal heights
        layers[nx // 2, ny // 2:, 1] = 0.0                        # a row without sediment
        scale = (20.0 / nx, 10.0 / ny, 4.0)
        delta = rng.random((nx, ny, 2), dtype=f32) * f32(1e-3)
        want = delta.copy()
        oracle.mass_creep(want, layers, scale, oracle.default_param(critSlopeSediment=crit))
        got = _creep_shared_edges(delta, layers, scale, crit)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (nx, ny, crit)


def _slopes_d8(h):
    """slopes towards the 8 neighbours in the reference's order (graph.hpp:35-44); NaN outside the map"""
    nx, ny = h.shape
    pad = np.full((nx + 2, ny + 2), np.nan, f32)
    pad[1:-1, 1:-1] = h
    shifts = [(-1, 0), (0, -1), (0, 1), (1, 0), (-1, -1), (-1, 1), (1, -1), (1, 1)]
    sl = []
    with np.errstate(all="ignore"):
        for k, (dx, dy) in enumerate(shifts):
            d = h - pad[1 + dx: 1 + dx + nx, 1 + dy: 1 + dy + ny]
            sl.append(d if k < 4 else (d / np.sqrt(f32(2.0))).astype(f32))
    return sl


def _direction_two_level(h):
    sl = _slopes_d8(h)
    zero = np.zeros_like(h)
    g0 = np.fmax(np.fmax(zero, sl[0]), sl[1])
    g1 = np.fmax(np.fmax(g0, sl[2]), sl[3])
    g2 = np.fmax(np.fmax(g1, sl[4]), sl[5])
    m = np.fmax(np.fmax(g2, sl[6]), sl[7])
    pick2 = lambda k: np.where(sl[k] == m, k, k + 1)
    win = pick2(6)
    win = np.where(g2 == m, pick2(4), win)
    win = np.where(g1 == m, pick2(2), win)
    win = np.where(g0 == m, pick2(0), win)
    return np.where(m > 0, win, -1).astype(np.int32)


def test_two_level_argmax_matches_the_oracle(oracle):
    rng = np.random.default_rng(3)
    cases = [oracle.noise(96, 80, 3.0), oracle.noise(33, 20, 1.0)]
    ties = rng.integers(0, 3, (64, 48)).astype(f32)                  # many exact ties, flats and pits
    cases.append(ties)
    diag = rng.integers(0, 4, (40, 40)).astype(f32) * np.sqrt(f32(2.0))  # diagonal and straight slopes that coincide
    cases.append(diag.astype(f32))
    holes = oracle.noise(50, 60, 2.0).copy()
    holes[rng.random(holes.shape) < 0.05] = np.nan                  # nodata cells
    cases.append(holes)
    for h in cases:
        assert np.array_equal(_direction_two_level(h), oracle.direction(h, 1)), h.shape
