"""GPU parity of the gradient family, slope and the Gaussian blur through the C ABI:
bit-exact against the CPU oracle (the blur to 1e-6: its 33 weights come from __expf), bit-exact against the reference's
own kernels built with -fmad=false, within float tolerance against its stock build."""
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))

SHAPES = [(1, 1), (1, 9), (9, 1), (37, 53), (96, 80), (200, 260)]


@pytest.fixture(scope="module")
def soil(product_lib):
    import soillib_b200 as soil
    return soil


def dev(a, soil, dtype=None):
    return soil.silt.tensor.from_numpy(np.ascontiguousarray(a)).gpu()


def inputs(oracle, nx, ny):
    h = oracle.noise(nx, ny, 3.0)
    hn = h.copy()
    hn[(np.arange(nx * ny).reshape(nx, ny) * 7919) % 97 == 5] = np.nan
    v2 = np.stack([h, np.cos(7.0 * h)], axis=-1).astype(np.float32)
    return h, hn, v2, (20.0 / nx, 20.0 / ny)


@pytest.mark.parametrize("shape", SHAPES)
def test_gradient_negslope_laplacian_slope_bit_exact(soil, oracle, shape):
    h, hn, v2, scale = inputs(oracle, *shape)
    for a in (h, hn):
        assert np.array_equal(soil.gradient(dev(a, soil), scale).numpy(), oracle.gradient(a, scale), equal_nan=True)
        assert np.array_equal(soil.negslope(dev(a, soil), scale).numpy(), oracle.negslope(a, scale), equal_nan=True)
    assert np.array_equal(soil.laplacian(dev(h[..., None], soil), scale).numpy(), oracle.laplacian(h[..., None], scale))
    assert np.array_equal(soil.laplacian(dev(v2, soil), scale).numpy(), oracle.laplacian(v2, scale))
    for e, ei in ((soil.d8, 1), (soil.d4, 0)):
        flow = oracle.steepest(h, ei)
        assert np.array_equal(soil.slope(dev(h, soil), dev(flow, soil), scale).numpy(), oracle.slope(h, flow, scale))
    neg = (-scale[0], scale[1])                                         # one division per axis must survive a negative scale
    assert np.array_equal(soil.negslope(dev(hn, soil), neg).numpy(), oracle.negslope(hn, neg), equal_nan=True)
    assert np.array_equal(soil.gradient(dev(hn, soil), neg).numpy(), oracle.gradient(hn, neg), equal_nan=True)
    loops = np.arange(h.size, dtype=np.int32).reshape(h.shape)          # next == n -> 0 (graph.cu:285)
    assert (soil.slope(dev(h, soil), dev(loops, soil), scale).numpy() == 0).all()


@pytest.mark.parametrize("shape", [(1, 1), (9, 40), (37, 53), (96, 80)])
@pytest.mark.parametrize("sigma", [0.8, 2.5, 9.0])
def test_gaussian_blur_vs_oracle(soil, oracle, shape, sigma):
    h, _, v2, _ = inputs(oracle, *shape)
    for a in (h[..., None].copy(), v2):
        want, _ = oracle.gaussian_blur(a, sigma)
        t = dev(a, soil)
        got = soil.gaussian_blur(t, sigma)
        assert got is t                                                 # clobbers and returns its argument, filter.cu:82,90
        np.testing.assert_allclose(t.numpy(), want, rtol=2e-6, atol=1e-7)


def test_blur_of_a_constant_is_the_weight_sum(soil, oracle):
    """Weights are not renormalised and indices clamp: a constant map stays constant, scaled by
    (sum of the 33 weights)^2 - everywhere, borders included."""
    sigma = 4.0
    k = np.arange(-16, 17, dtype=np.float64)
    wsum = (np.exp(-0.5 * (k / sigma) ** 2) / (np.sqrt(2.0 * 3.14159265) * sigma)).sum()
    t = dev(np.full((40, 72, 1), 3.0, np.float32), soil)
    soil.gaussian_blur(t, sigma)
    np.testing.assert_allclose(t.numpy(), 3.0 * wsum * wsum, rtol=1e-5)


def test_matches_reference_kernels(soil, oracle):
    import ref_api as ref
    if not ref.available(nofma=True):
        pytest.skip("oracle/_ref not built")
    h, hn, v2, scale = inputs(oracle, 96, 80)
    hd, hnd, v2d = (torch.from_numpy(x).cuda() for x in (h, hn, v2))
    v1d = hd.unsqueeze(-1).contiguous()
    T = soil.silt.tensor.from_torch
    flow = soil.steepest(T(hd), soil.d8)
    ours = {
        "gradient": soil.gradient(T(hd), scale).data, "gradient_nan": soil.gradient(T(hnd), scale).data,
        "negslope": soil.negslope(T(hd), scale).data, "negslope_nan": soil.negslope(T(hnd), scale).data,
        "laplacian1": soil.laplacian(T(v1d), scale).data, "laplacian2": soil.laplacian(T(v2d), scale).data,
        "slope": soil.slope(T(hd), flow, scale).data,
        "blur1": soil.gaussian_blur(T(v1d.clone()), 2.5).data, "blur2": soil.gaussian_blur(T(v2d.clone()), 6.0).data,
    }
    for nofma in (True, False):
        theirs = {
            "gradient": ref.gradient(hd, scale, nofma), "gradient_nan": ref.gradient(hnd, scale, nofma),
            "negslope": ref.negslope(hd, scale, nofma), "negslope_nan": ref.negslope(hnd, scale, nofma),
            "laplacian1": ref.laplacian(v1d, scale, nofma), "laplacian2": ref.laplacian(v2d, scale, nofma),
            "slope": ref.slope(hd, flow.data, scale, nofma),
            "blur1": ref.gaussian_blur(v1d, 2.5, nofma), "blur2": ref.gaussian_blur(v2d, 6.0, nofma),
        }
        for k in ours:
            a, b = ours[k].cpu().numpy(), theirs[k].cpu().numpy()
            if nofma:       # same arithmetic, same intrinsics: every bit
                assert np.array_equal(a, b, equal_nan=True), k
            else:           # stock build contracts multiply-adds
                np.testing.assert_allclose(a, b, rtol=2e-5, atol=2e-5 * np.nanmax(np.abs(b)), equal_nan=True, err_msg=k)


def test_argument_errors(soil, oracle):
    h = dev(oracle.noise(16, 16, 3.0), soil)
    with pytest.raises(ValueError):
        soil.laplacian(h, (1.0, 1.0))                                   # needs (nx, ny, 1|2)
    with pytest.raises(ValueError):
        soil.gaussian_blur(dev(np.zeros((8, 8, 3), np.float32), soil), 1.0)
    with pytest.raises(ValueError):
        soil.gradient(h, (1.0, 1.0, 1.0))
    with pytest.raises(ValueError):
        soil.slope(h, dev(np.zeros((8, 8), np.int32), soil), (1.0, 1.0))
    with pytest.raises(ValueError):
        soil.negslope(soil.silt.tensor.from_numpy(np.zeros((8, 8), np.float32)), (1.0, 1.0))   # host tensor


def test_full_size_properties(soil):
    """8192^2: a planar ramp has the same gradient everywhere (one-sided at the borders), zero Laplacian inside,
    and its blur is the ramp scaled by the weight sum away from the borders."""
    n = 8192
    x = torch.arange(n, dtype=torch.float32, device="cuda")
    ramp = (0.5 * x[:, None] + 0.25 * x[None, :]).contiguous()           # exact in float32
    T = soil.silt.tensor.from_torch
    g = soil.gradient(T(ramp), (1.0, 1.0)).data
    assert bool((g[..., 0] == 0.5).all()) and bool((g[..., 1] == 0.25).all())
    s = soil.negslope(T(ramp), (1.0, 1.0)).data
    assert bool((s[1:, 1:] == np.float32(np.sqrt(np.float32(0.3125)))).all())
    lap = soil.laplacian(T(ramp.unsqueeze(-1).contiguous()), (1.0, 1.0)).data
    assert bool((lap[1:-1, 1:-1] == 0).all())
    flow = soil.steepest(T(ramp), soil.d8)
    sl = soil.slope(T(ramp), flow, (1.0, 1.0)).data
    assert bool((sl[1:, 1:] < 0).all()) and bool((sl[0, 0] == 0))
