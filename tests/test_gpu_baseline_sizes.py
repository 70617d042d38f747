"""Parity at BASELINE.json's sizes against the REFERENCE'S OWN KERNELS running on the same GPU (oracle/_ref).

configs[0]: D8 direction + accumulation at 1024^2; the same ops at 8192^2 (the size the metric is quoted on);
configs[1]: one erosion step at 2048^2 with example/erosion_gpu.py's parameters (timeStep 1000, maxage 256, 8192
particles) - through the instantiation soil.erode() and bench.py run (k_pre on the side stream + k_fused_step with
albedo, ping-pong flux records, the pre scratch and the packed particle view), against the reference's unfused sequence built with the same
arithmetic (-fmad=false), and statistically against its stock build.

Tolerances (stated here, printed by the test):
  direction / steepest / accumulate of unit rain          bit-exact (graph.cu:73-91, 246-264, 578-584)
  one erosion step vs the -fmad=false build                 identical RNG states; every cell of `layers` within
                                                            ONE_STEP_ATOL absolute (+ 1e-5 relative); total mass to 1e-6;
                                                            mean and variance of the elevation change to 1e-4
  vs the stock build (FMA contraction on)                   >= 99 % of the cells within ONE_STEP_ATOL, the others within
                                                            FORKED_ATOL (a particle that takes another branch at a cell edge)
"""
import numpy as np
import pytest
import torch

import ref_api as ref

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")]

ONE_STEP_ATOL = 2e-6     # dimensionless map units (|h| ~ 0.5); flux sums differ by the order of their float atomics
FORKED_ATOL = 2e-2       # bound for cells on a forked particle path: one step's clamped transfer (erosion.cu:526-528)


@pytest.fixture(scope="module")
def soil(product_lib):
    import soillib_b200 as soil
    return soil


def S(soil, t):
    return soil.silt.tensor.from_torch(t)


def fbm(soil, n):
    npar = soil.noise_t()
    npar.seed, npar.ext = 3.0, (float(n), float(n))
    return soil.noise(soil.silt.shape(n, n), npar, host=soil.silt.gpu).data


@pytest.mark.parametrize("n", [1024, 8192])
def test_d8_flow_bit_exact_vs_reference_at_baseline_sizes(soil, n):
    h = fbm(soil, n)
    d_ref, d_our = ref.direction(h, 1), soil.direction(S(soil, h), soil.d8).data
    assert torch.equal(d_our, d_ref)
    del d_ref, d_our
    g_ref = ref.steepest(h, 1)
    g_our = soil.steepest(S(soil, h), soil.d8).data
    assert torch.equal(g_our, g_ref)
    del g_our
    ones = torch.ones_like(h)
    a_our = soil.accumulate(S(soil, g_ref), S(soil, ones), soil.d8).data
    a_ref = ref.accumulate(g_ref, ones, None, 1)
    torch.cuda.synchronize()
    # unit rain: every partial sum is an integer below 2^24, so both summation orders are exact
    assert float(a_ref.max()) < 2 ** 24
    assert torch.equal(a_our, a_ref), f"{(a_our != a_ref).sum().item()} cells differ"
    # the sum over the roots is the number of cells (nothing lost, nothing counted twice)
    roots = g_ref < 0
    assert float(a_our[roots].double().sum()) == float(n * n)


def test_accumulate_real_sources_vs_reference_1024(soil):
    """north_star: accumulation within 1e-5 relative of the reference (real-valued sources, with and without decay)."""
    n = 1024
    h = fbm(soil, n)
    g = ref.steepest(h, 1)
    src = h.abs() + 0.25
    torch.testing.assert_close(soil.accumulate(S(soil, g), S(soil, src), soil.d8).data, ref.accumulate(g, src, None, 1),
                               rtol=1e-5, atol=0)
    dec = torch.full_like(h, 0.97)
    torch.testing.assert_close(soil.accumulate_decay(S(soil, g), S(soil, src), S(soil, dec), soil.d8).data,
                               ref.accumulate(g, src, dec, 1), rtol=1e-5, atol=0)


def test_one_erosion_step_2048_vs_reference(soil, oracle):
    from test_gpu_erosion import _run_fused
    n, P = 2048, 8192
    op = oracle.example_param()                      # example/erosion_gpu.py:76-91
    h = fbm(soil, n).cpu().numpy()
    scale = (20.0 / n, 20.0 / n, 4.0)

    def make():
        return oracle.HostModel(h, scale, P, albedo=True, seed=0, uplift=np.clip(h, 0, 1))

    ours, msum, _ = _run_fused(soil, oracle, make(), op, 1, True, True, pingpong=True, pre=True, view=True)
    runs = {}
    for key, nofma in (("nofma", True), ("stock", False)):
        dm = ref.DeviceModel(make())
        dm.step(op, True, nofma=nofma)
        torch.cuda.synchronize()
        runs[key] = dm
    mine = ours["layers"].cpu().numpy()
    nf = runs["nofma"].get("layers")
    assert np.array_equal(ours["rngFluvial"].cpu().numpy()[:, :24], runs["nofma"].get("rngFluvial")[:, :24])
    assert np.array_equal(ours["rngDebris"].cpu().numpy()[:, :24], runs["nofma"].get("rngDebris")[:, :24])
    err = np.abs(mine - nf)
    print(f"2048^2, one step, vs reference built -fmad=false: max-abs layer error {err.max():.3e} "
          f"(stated tolerance {ONE_STEP_ATOL:g} + 1e-5 relative)")
    np.testing.assert_allclose(mine, nf, rtol=1e-5, atol=ONE_STEP_ATOL)
    np.testing.assert_allclose(msum.cpu().numpy(), nf.astype(np.float64).sum(axis=(0, 1)), rtol=1e-6)
    dz_m, dz_r = mine.sum(-1) - h, nf.sum(-1) - h
    np.testing.assert_allclose(dz_m.mean(), dz_r.mean(), rtol=1e-4, atol=1e-12)
    np.testing.assert_allclose(dz_m.var(), dz_r.var(), rtol=1e-4)
    for k, rtol, atol in (("waterHeight", 1e-4, 1e-9), ("momentum", 1e-4, 1e-6), ("debrisVel", 1e-4, 1e-6), ("albedoSurface", 1e-4, 1e-6)):
        a, b = ours[k].cpu().numpy(), runs["nofma"].get(k)
        fin = np.isfinite(b)
        assert np.array_equal(np.isfinite(a), fin), k
        bad = np.abs(a[fin] - b[fin]) > atol + rtol * np.abs(b[fin])
        assert bad.mean() <= 1e-5, (k, bad.mean())      # flux sums: atomic order; a handful of cells at most
    st = runs["stock"].get("layers")
    e2 = np.abs(mine.sum(-1) - st.sum(-1))
    out = e2 > ONE_STEP_ATOL
    print(f"vs the stock build: cells outside {ONE_STEP_ATOL:g}: {out.mean():.4%}, max-abs error over them "
          f"{e2[out].max() if out.any() else 0.0:.3e}")
    assert out.mean() <= 0.01
    assert e2.max() <= FORKED_ATOL
    np.testing.assert_allclose(msum.cpu().numpy().sum(), st.astype(np.float64).sum(), rtol=1e-6)
