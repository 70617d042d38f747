"""GPU parity of the flow-graph path (direction, steepest, random_weighted, accumulate, normal, noise).

Everything goes through the C ABI (soillib_b200 -> ctypes -> libsoilb200.so) and is compared with the
CPU oracle on the same seeded inputs; integer/index results must be bit-exact.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [(96, 80), (128, 256), (37, 53), (1, 1), (1, 9), (9, 1), (5, 4), (200, 132), (64, 1028)]


@pytest.fixture(scope="module")
def soil(product_lib):
    import soillib_b200 as soil
    return soil


def dev(a, soil):
    return soil.silt.tensor.from_numpy(np.ascontiguousarray(a)).gpu()


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("edge", [0, 1])
def test_direction_steepest_bit_exact(soil, oracle, shape, edge):
    h = oracle.noise(shape[0], shape[1], 3.0)
    e = soil.edge(edge)
    assert np.array_equal(soil.direction(dev(h, soil), e).numpy(), oracle.direction(h, edge))
    assert np.array_equal(soil.steepest(dev(h, soil), e).numpy(), oracle.steepest(h, edge))


def test_direction_edge_cases(soil, oracle):
    rng = np.random.default_rng(1)
    flat = np.zeros((16, 32), np.float32)
    assert (soil.direction(dev(flat, soil), soil.d8).numpy() == -1).all()
    q = rng.integers(0, 4, (64, 128)).astype(np.float32)       # quantised DEM: many exact ties
    assert np.array_equal(soil.direction(dev(q, soil), soil.d8).numpy(), oracle.direction(q, 1))
    assert np.array_equal(soil.steepest(dev(q, soil), soil.d4).numpy(), oracle.steepest(q, 0))
    n = oracle.noise(48, 64, 3.0)
    n[rng.random(n.shape) < 0.05] = np.nan                      # nodata cells never win, never crash
    assert np.array_equal(soil.direction(dev(n, soil), soil.d8).numpy(), oracle.direction(n, 1))
    near = (1.0 + rng.integers(0, 3, (64, 64)) * np.float32(2 ** -23)).astype(np.float32)  # 1-ulp differences
    assert np.array_equal(soil.direction(dev(near, soil), soil.d8).numpy(), oracle.direction(near, 1))


def test_direction_extreme_heights_take_the_exact_path(soil, oracle):
    """Slopes the three-instruction diagonal quotient cannot be trusted with (subnormal-range differences, +-inf,
    +-FLT_MAX nodata, overflowing differences) must still match the IEEE-division oracle bit for bit."""
    rng = np.random.default_rng(7)
    with np.errstate(all="ignore"):
        tiny = (rng.integers(0, 5, (64, 128)).astype(np.float32) * np.float32(3e-39))      # subnormal heights
        tiny32 = (rng.integers(0, 7, (64, 128)).astype(np.float32) * np.float32(7.3e-33))  # differences < 2.2e-32
        n = oracle.noise(64, 128, 3.0)
        inf = n.copy(); inf[rng.random(n.shape) < 0.03] = np.inf; inf[rng.random(n.shape) < 0.03] = -np.inf
        fmax = n.copy(); fmax[rng.random(n.shape) < 0.03] = np.float32(-3.4028235e38)
        fmax[rng.random(n.shape) < 0.03] = np.float32(3.4028235e38)
        mixed = n.copy(); mixed[rng.random(n.shape) < 0.02] = np.nan; mixed[rng.random(n.shape) < 0.02] = np.inf
        for h in (tiny, tiny32, inf, fmax, mixed):
            for e, ei in ((soil.d8, 1), (soil.d4, 0)):
                assert np.array_equal(soil.direction(dev(h, soil), e).numpy(), oracle.direction(h, ei))
                assert np.array_equal(soil.steepest(dev(h, soil), e).numpy(), oracle.steepest(h, ei))


@pytest.mark.parametrize("variant", ["staged", "tma"])
def test_tma_staged_sweep_matches(soil, oracle, tmp_path, variant):
    """The opt-in variants of the sweep that stage rows through shared memory - 1-D cp.async.bulk rows
    (SOILB200_STEEPEST=staged) and 2-D tensor-map boxes with NaN out-of-bound fill (SOILB200_STEEPEST=tma); the variable
    is read once per process, hence the subprocess - give the same bits, on ragged shapes and maps with NaN cells too."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, numpy as np; sys.path.insert(0, %r); sys.path.insert(0, %r + '/tests')\n"
        "import soillib_b200 as soil, oracle_api as orc\n"
        "rng = np.random.default_rng(3)\n"
        "for nx, ny in ((1, 4), (7, 12), (65, 516), (130, 1024), (300, 2052), (64, 512), (129, 768)):\n"
        "    h = orc.noise(nx, ny, 3.0)\n"
        "    h[rng.random(h.shape) < 0.01] = np.nan\n"
        "    d = soil.silt.tensor.from_numpy(h).gpu()\n"
        "    for e, ei in ((soil.d8, 1), (soil.d4, 0)):\n"
        "        assert np.array_equal(soil.direction(d, e).numpy(), orc.direction(h, ei)), (nx, ny, ei)\n"
        "        assert np.array_equal(soil.steepest(d, e).numpy(), orc.steepest(h, ei)), (nx, ny, ei)\n"
        "print('staged ok')\n" % (root, root))
    env = dict(os.environ, SOILB200_STEEPEST=variant)
    proc = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert proc.returncode == 0 and "staged ok" in proc.stdout, proc.stdout + proc.stderr


def test_fast_division_by_sqrt2_is_exact_for_every_float(soil):
    """The diagonal-slope division x / sqrtf(2) (graph.cu:56) is done with a multiply + two FMAs.  Over all 2^32 inputs:
    the guarded form agrees with the IEEE division bit for bit (NaN payloads aside); the unguarded form agrees from
    1e-30 to FLT_MAX, and below 1e-30 both stay below 1e-30 with the sign of x - what k_steepest relies on."""
    bad = torch.ones(1, dtype=torch.int64, device="cuda")
    assert soil._load_library().soilb200_selftest_div_sqrt2(bad.data_ptr(), None) == 0
    assert bad.item() == 0


def test_unaligned_view_uses_scalar_path(soil, oracle):
    h = oracle.noise(40, 64, 3.0)
    big = torch.zeros(40 * 64 + 1, dtype=torch.float32, device="cuda")
    big[1:] = torch.from_numpy(h).cuda().ravel()
    t = soil.silt.tensor(soil.silt.float32, _data=big[1:].view(40, 64))   # 4-byte aligned only
    assert np.array_equal(soil.direction(t, soil.d8).numpy(), oracle.direction(h, 1))


def test_bad_edge_raises(soil, oracle):
    with pytest.raises(ValueError):
        soil.direction(dev(oracle.noise(8, 8), soil), 5)


@pytest.mark.parametrize("shape", [(96, 80), (37, 53), (64, 128)])
def test_random_weighted_vs_oracle(soil, oracle, shape):
    h = oracle.noise(shape[0], shape[1], 3.0)
    lib = soil._load_library()
    hd = torch.from_numpy(h).cuda()
    for edge, seed, T in ((1, 0, 10.0), (0, 7, 2.5)):
        for k in range(3):
            want, amb = oracle.random_weighted(h, edge, seed, k, T)
            out = torch.empty(shape, dtype=torch.int32, device="cuda")
            rc = lib.soilb200_random_weighted_f32(hd.data_ptr(), out.data_ptr(), shape[0], shape[1], edge, seed, k,
                                                  T, None, 0, None)          # stateless: seeds like the reference
            assert rc == 0
            got = out.cpu().numpy()
            diff = got != want
            assert not (diff & ~amb).any()
            assert diff.mean() < 1e-3


def test_random_weighted_resume_is_bit_identical(soil, oracle):
    """Keeping the streams in scratch and advancing by one draw == re-seeding at offset k (graph.cu:100)."""
    h = oracle.noise(96, 80, 3.0)
    lib = soil._load_library()
    hd = torch.from_numpy(h).cuda()
    ws = torch.zeros(lib.soilb200_random_weighted_ws(96, 80), dtype=torch.uint8, device="cuda")
    for k in (4, 5, 6, 9, 10):      # 4: cold (seed), 5-6: resume, 9: gap -> re-seed, 10: resume
        a = torch.empty((96, 80), dtype=torch.int32, device="cuda")
        b = torch.empty_like(a)
        assert lib.soilb200_random_weighted_f32(hd.data_ptr(), a.data_ptr(), 96, 80, 1, 3, k, 10.0, ws.data_ptr(), ws.numel(), None) == 0
        assert lib.soilb200_random_weighted_f32(hd.data_ptr(), b.data_ptr(), 96, 80, 1, 3, k, 10.0, None, 0, None) == 0
        assert torch.equal(a, b)
    # the Python op caches the scratch and must agree too
    for k in range(3):
        b = torch.empty((96, 80), dtype=torch.int32, device="cuda")
        lib.soilb200_random_weighted_f32(hd.data_ptr(), b.data_ptr(), 96, 80, 1, 0, k, 10.0, None, 0, None)
        assert torch.equal(soil.random_weighted(dev(h, soil), soil.d8, 0, k, 10.0).data, b)


def test_seed_matches_curand_init(soil, oracle):
    r = soil.silt.tensor(soil.silt.rng, soil.silt.shape(300), soil.silt.gpu)
    soil.silt.seed(r, 11, 5)
    want = oracle.seed_states(300, seed=11, offset=5)
    assert np.array_equal(r.numpy()[:, :24], want[:, :24])


@pytest.mark.parametrize("shape", [(96, 80), (37, 53), (256, 192), (64, 64), (65, 129), (200, 63), (5, 2), (64, 2), (2, 5), (130, 3)])
def test_accumulate_vs_oracle(soil, oracle, shape):
    h = oracle.noise(shape[0], shape[1], 3.0)
    ones = np.ones_like(h)
    src = (np.abs(h) + 0.25).astype(np.float32)
    dec = (0.8 + 0.2 * np.abs(np.sin(np.arange(h.size, dtype=np.float32)))).reshape(h.shape).astype(np.float32)
    for edge in (1, 0):
        g = oracle.steepest(h, edge)
        e = soil.edge(edge)
        got = soil.accumulate(dev(g, soil), dev(ones, soil), e).numpy()
        assert np.array_equal(got, oracle.accumulate(g, ones, None, edge, mode=1)[0])
        assert np.array_equal(got, oracle.accumulate(g, ones, None, edge, mode=0)[0])    # the reference's schedule
        got = soil.accumulate(dev(g, soil), dev(src, soil), e).numpy()
        np.testing.assert_allclose(got, oracle.accumulate(g, src, None, edge, mode=1)[0], rtol=2e-6)  # tiles re-associate
        walk = torch.empty(shape, dtype=torch.float32, device="cuda")           # 4 B/cell scratch -> global walk:
        ws4 = torch.empty(4 * h.size, dtype=torch.uint8, device="cuda")         # the oracle's summation order, bit for bit
        assert soil._load_library().soilb200_accumulate_f32(dev(g, soil).ptr(), dev(src, soil).ptr(), None, walk.data_ptr(),
                                                            shape[0], shape[1], edge, ws4.data_ptr(), ws4.numel(), None) == 0
        assert np.array_equal(walk.cpu().numpy(), oracle.accumulate(g, src, None, edge, mode=1)[0])
        # with decay: the tiles multiply the edge weights along a path in another order than the walk (pointer jumping)
        # -> north_star's 1e-5 relative; the global walk (4 B/cell of scratch) keeps the oracle's order: 2e-6 (__powf)
        got = soil.accumulate_decay(dev(g, soil), dev(src, soil), dev(dec, soil), e).numpy()
        np.testing.assert_allclose(got, oracle.accumulate(g, src, dec, edge, mode=1)[0], rtol=1e-5)
        np.testing.assert_allclose(got, oracle.accumulate(g, src, dec, edge, mode=0)[0], rtol=1e-5)
        assert soil._load_library().soilb200_accumulate_f32(dev(g, soil).ptr(), dev(src, soil).ptr(), dev(dec, soil).ptr(),
                                                            walk.data_ptr(), shape[0], shape[1], edge, ws4.data_ptr(),
                                                            ws4.numel(), None) == 0
        np.testing.assert_allclose(walk.cpu().numpy(), oracle.accumulate(g, src, dec, edge, mode=1)[0], rtol=2e-6)
        # unit rain with decay 1: the tiled decay path must reproduce the exact counts
        one_dec = np.ones_like(h)
        got = soil.accumulate_decay(dev(g, soil), dev(ones, soil), dev(one_dec, soil), e).numpy()
        assert np.array_equal(got, oracle.accumulate(g, ones, None, edge, mode=1)[0])
    grw, _ = oracle.random_weighted(h, 1, 0, 1, 10.0)
    got = soil.accumulate(dev(grw, soil), dev(ones, soil), soil.d8).numpy()
    assert np.array_equal(got, oracle.accumulate(grw, ones, None, 1, mode=1)[0])


def test_accumulate_odd_graphs(soil, oracle):
    g = np.full((4, 4), -1, np.int32)
    g[0, 0] = 15
    g[0, 1] = 5
    got = soil.accumulate(dev(g, soil), dev(np.ones((4, 4), np.float32), soil), soil.d8).numpy()
    assert np.array_equal(got, oracle.accumulate(g, np.ones((4, 4), np.float32), None, 1, mode=1)[0])
    cyc = np.full((3, 3), -1, np.int32)
    cyc[0, 0], cyc[0, 1] = 1, 0           # a 2-cycle must not hang; cells keep their source
    got = soil.accumulate(dev(cyc, soil), dev(np.ones((3, 3), np.float32), soil), soil.d8).numpy()
    assert np.isfinite(got).all() and got[2, 2] == 1
    many = np.full((3, 3), 4, np.int32)   # 8 donors into the centre: slots 4..7 take the __powf decay
    many[1, 1] = -1
    src = np.arange(1, 10, dtype=np.float32).reshape(3, 3)
    dec = np.full((3, 3), 0.5, np.float32)
    got = soil.accumulate_decay(dev(many, soil), dev(src, soil), dev(dec, soil), soil.d8).numpy()
    np.testing.assert_allclose(got, oracle.accumulate(many, src, dec, 1, mode=1)[0], rtol=2e-6)


@pytest.mark.parametrize("shape", [(96, 80), (37, 53), (3, 3), (2, 7)])
def test_normal_bit_exact(soil, oracle, shape):
    h = oracle.noise(shape[0], shape[1], 3.0)
    scale = (20.0 / shape[0], 20.0 / shape[1], 4.0)
    assert np.array_equal(soil.normal(dev(h, soil), scale).numpy(), oracle.normal(h, scale))
    hn = h.copy()
    hn[0, 0] = np.nan
    assert np.array_equal(soil.normal(dev(hn, soil), scale).numpy(), oracle.normal(hn, scale))


def test_noise_host_and_device_bit_exact(soil, oracle):
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "fnl_noise.npz"))
    p = soil.noise_t()
    p.seed, p.ext = 3.0, (64.0, 48.0)
    host = soil.noise(soil.silt.shape(64, 48), p).numpy()
    devv = soil.noise(soil.silt.shape(64, 48), p, host=soil.silt.gpu).numpy()
    assert np.array_equal(host, g["fbm_64x48_z3"])          # the reference's FastNoiseLite
    assert np.array_equal(devv, g["fbm_64x48_z3"])
    p.ext = (1024.0, 1024.0)
    big = soil.noise(soil.silt.shape(1024, 1024), p, host=soil.silt.gpu).numpy()
    assert np.array_equal(big[::61, ::61], g["fbm_1024_z3_stride61"])


# ---- BASELINE-size properties (sizes the oracle cannot finish in seconds) ---------------------


def test_full_size_flow_properties(soil):
    nx = ny = 4096
    p = soil.noise_t()
    p.seed, p.ext = 3.0, (float(nx), float(ny))
    h = soil.noise(soil.silt.shape(nx, ny), p, host=soil.silt.gpu)
    s = soil.steepest(h, soil.d8).data
    d = soil.direction(h, soil.d8).data
    hv = h.data.ravel()
    ok = s >= 0
    assert torch.equal(ok, d >= 0)
    assert (hv[s[ok].long()] < h.data[ok]).all()                     # receivers strictly lower
    rain = soil.silt.tensor.from_torch(torch.ones((nx, ny), dtype=torch.float32, device="cuda"))
    acc = soil.accumulate(soil.silt.tensor.from_torch(s), rain, soil.d8).data
    assert acc[~ok].double().sum().item() == nx * ny                 # every cell drains to exactly one root
    assert acc.min().item() >= 1.0
    donors = torch.zeros(nx * ny, dtype=torch.float32, device="cuda")
    donors.index_add_(0, s[ok].long(), acc[ok])
    assert torch.equal(acc.ravel(), donors + 1.0)                    # local balance: acc = 1 + sum of donors
    rw = soil.random_weighted(h, soil.d8, 0, 0, 10.0).data
    okr = rw >= 0
    assert (hv[rw[okr].long()] < h.data[okr]).all()
