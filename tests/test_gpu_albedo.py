"""GPU parity of the albedo generators (erosion.cu:759-919) through the C ABI: albedo_layer
and albedo_stratum bit-exact against the CPU oracle, albedo_discharge to 1e-6 (__expf); all three bit-exact against the
reference's own kernels built with -fmad=false and within float tolerance of its stock build."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [(1, 1), (5, 7), (37, 53), (96, 80)]          # cell counts not divisible by 4 take the scalar kernels
COL_A, COL_B, SHIFT = (0.8, 0.6, 0.4), (0.3, 0.35, 0.5), (0.1, 0.25, 0.6)


@pytest.fixture(scope="module")
def soil(product_lib):
    import soillib_b200 as soil
    return soil


def inputs(oracle, nx, ny):
    rng = np.random.default_rng(nx * 1000 + ny)
    h = oracle.noise(nx, ny, 3.0)
    layers = np.stack([h, 0.02 * rng.random((nx, ny), dtype=np.float32)], axis=-1).astype(np.float32)
    bedrock = rng.random((nx, ny, 3), dtype=np.float32)
    sediment = rng.random((nx, ny, 3), dtype=np.float32)
    uplift = np.clip(h, 0, 1).astype(np.float32)
    discharge = (rng.random((nx, ny), dtype=np.float32) * 4.0 - 1.0).astype(np.float32)     # some negative: clamped
    return layers, bedrock, sediment, uplift, discharge


def T(a, soil):
    return soil.silt.tensor.from_numpy(np.ascontiguousarray(a)).gpu()


def _param(soil, oracle):
    p = soil.param_t()
    C.memmove(C.byref(p._c), C.byref(oracle.example_param()), 112)
    return p


@pytest.mark.parametrize("shape", SHAPES)
def test_albedo_ops_vs_oracle(soil, oracle, shape):
    nx, ny = shape
    layers, bedrock, sediment, uplift, discharge = inputs(oracle, nx, ny)
    scale = (20.0 / nx, 20.0 / ny, 4.0)
    out = T(np.zeros_like(bedrock), soil)
    soil.albedo_layer(out, T(bedrock, soil), T(sediment, soil), T(layers, soil), 35.0, SHIFT)
    assert np.array_equal(out.numpy(), oracle.albedo_layer(bedrock, sediment, layers, 35.0, SHIFT))
    strat = T(np.zeros_like(bedrock), soil)
    soil.albedo_stratum(strat, T(uplift, soil), T(layers, soil), scale, _param(soil, oracle), COL_A, COL_B, 5000.0, 0.05)
    want = oracle.albedo_stratum(uplift, layers, scale, oracle.example_param(), COL_A, COL_B, 5000.0, 0.05)
    assert np.array_equal(strat.numpy(), want)
    if nx * ny > 100:
        assert 0.05 < (want[..., 0] == np.float32(COL_A[0])).mean() < 0.95          # both strata occur
    alb = T(bedrock, soil)
    soil.albedo_discharge(alb, T(discharge, soil), (0.1, 0.2, 0.9), 0.7, 0.8)
    np.testing.assert_allclose(alb.numpy(), oracle.albedo_discharge(bedrock, discharge, (0.1, 0.2, 0.9), 0.7, 0.8),
                               rtol=1e-6, atol=1e-7)


def test_albedo_ops_match_reference_kernels(soil, oracle):
    import ref_api as ref
    if not ref.available(nofma=True):
        pytest.skip("oracle/_ref not built")
    nx, ny = 96, 80
    layers, bedrock, sediment, uplift, discharge = inputs(oracle, nx, ny)
    scale = (20.0 / nx, 20.0 / ny, 4.0)
    d = {k: torch.from_numpy(v).cuda() for k, v in dict(layers=layers, bedrock=bedrock, sediment=sediment, uplift=uplift,
                                                        discharge=discharge).items()}
    W = soil.silt.tensor.from_torch
    ours_layer = torch.zeros_like(d["bedrock"])
    soil.albedo_layer(W(ours_layer), W(d["bedrock"]), W(d["sediment"]), W(d["layers"]), 35.0, SHIFT)
    ours_strat = torch.zeros_like(d["bedrock"])
    soil.albedo_stratum(W(ours_strat), W(d["uplift"]), W(d["layers"]), scale, _param(soil, oracle), COL_A, COL_B, 5000.0, 0.05)
    ours_dis = d["bedrock"].clone()
    soil.albedo_discharge(W(ours_dis), W(d["discharge"]), (0.1, 0.2, 0.9), 0.7, 0.8)
    for nofma in (True, False):
        r_layer = ref.albedo_layer(d["bedrock"], d["sediment"], d["layers"], 35.0, SHIFT, nofma)
        r_strat = ref.albedo_stratum(d["uplift"], d["layers"], scale, oracle.example_param(), COL_A, COL_B, 5000.0, 0.05, nofma)
        r_dis = ref.albedo_discharge(d["bedrock"], d["discharge"], (0.1, 0.2, 0.9), 0.7, 0.8, nofma)
        if nofma:
            assert torch.equal(ours_layer, r_layer) and torch.equal(ours_strat, r_strat) and torch.equal(ours_dis, r_dis)
        else:
            torch.testing.assert_close(ours_layer, r_layer, rtol=1e-6, atol=1e-7)
            assert (ours_strat != r_strat).float().mean().item() < 1e-3      # a contracted multiply-add can move a stratum edge
            torch.testing.assert_close(ours_dis, r_dis, rtol=1e-6, atol=1e-7)


def test_full_size(soil):
    """8192^2: zero sediment -> the bedrock colour; zero discharge leaves the albedo untouched."""
    n = 8192
    W = soil.silt.tensor.from_torch
    bedrock = torch.rand((n, n, 3), device="cuda")
    layers = torch.zeros((n, n, 2), device="cuda")
    out = torch.empty_like(bedrock)
    soil.albedo_layer(W(out), W(bedrock), W(torch.rand_like(bedrock)), W(layers), 10.0, (0, 0, 0))
    assert torch.equal(out, bedrock)
    soil.albedo_discharge(W(out), W(torch.zeros((n, n), device="cuda")), (0.0, 0.0, 1.0), 1.0, 1.0)
    assert torch.equal(out, bedrock)
