"""The product against the REFERENCE'S OWN KERNELS on the same GPU (oracle/_ref, built by `make -C oracle ref`).

libsoilref.so is the reference's graph.cu / erosion.cu compiled unmodified for sm_100a with its stock
flags; libsoilref_nofma.so is the same source with -fmad=false.  Index results must be bit-exact against
both.  Erosion is compared bit-for-bit in trajectories against the -fmad=false build (same arithmetic as
ours) and statistically against the stock build, whose FMA contraction perturbs last bits of particle
positions; the tolerance is stated relative to the reference's own run-to-run (atomic order) noise, as
SURVEY.md section 8c prescribes.
"""
import ctypes as C

import numpy as np
import pytest
import torch

import ref_api as ref

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")]


@pytest.fixture(scope="module")
def soil(product_lib):
    import soillib_b200 as soil
    return soil


def S(soil, t):
    return soil.silt.tensor.from_torch(t)


@pytest.mark.parametrize("shape", [(96, 80), (37, 53), (512, 384)])
def test_graph_ops_bit_exact_vs_reference(soil, oracle, shape):
    h = torch.from_numpy(oracle.noise(shape[0], shape[1], 3.0)).cuda()
    for edge in (0, 1):
        e = soil.edge(edge)
        assert torch.equal(soil.direction(S(soil, h), e).data, ref.direction(h, edge))
        assert torch.equal(soil.steepest(S(soil, h), e).data, ref.steepest(h, edge))
        for k in range(3):
            assert torch.equal(soil.random_weighted(S(soil, h), e, 0, k, 10.0).data, ref.random_weighted(h, edge, 0, k, 10.0))
        g = ref.steepest(h, edge)
        ones = torch.ones_like(h)
        assert torch.equal(soil.accumulate(S(soil, g), S(soil, ones), e).data, ref.accumulate(g, ones, None, edge))
        src = h.abs() + 0.25
        dec = torch.full_like(h, 0.9)
        torch.testing.assert_close(soil.accumulate(S(soil, g), S(soil, src), e).data, ref.accumulate(g, src, None, edge), rtol=1e-5, atol=0)
        torch.testing.assert_close(soil.accumulate_decay(S(soil, g), S(soil, src), S(soil, dec), e).data,
                                   ref.accumulate(g, src, dec, edge), rtol=1e-5, atol=0)


def _fused(soil, oracle, host, op, steps, creep, pre=False):
    from test_gpu_erosion import _run_fused
    return _run_fused(soil, oracle, host, op, steps, creep, True, pre=bool(pre), view=pre == "view")


@pytest.mark.parametrize("preset", ["example", "default"])
@pytest.mark.parametrize("pre", [False, True, "view"], ids=["inline", "pre", "view"])
def test_erosion_vs_reference_kernels(soil, oracle, preset, pre):
    nx, ny, P, steps = 128, 96, 1024, 3
    op = oracle.example_param() if preset == "example" else oracle.default_param(maxage=128)
    h = oracle.noise(nx, ny, 3.0)
    scale = (20.0 / nx, 20.0 / ny, 4.0)

    def make():
        m = oracle.HostModel(h, scale, P, albedo=True, seed=0, uplift=np.clip(h, 0, 1))
        m.layers[..., 1] = 0.01
        return m

    ours, msum, _ = _fused(soil, oracle, make(), op, steps, True, pre=pre)
    runs = {}
    for key, nofma in (("stock", False), ("stock2", False), ("nofma", True)):
        dm = ref.DeviceModel(make())
        for _ in range(steps):
            dm.step(op, True, nofma=nofma)
        torch.cuda.synchronize()
        runs[key] = dm
    mine = ours["layers"].cpu().numpy()

    # same arithmetic (no FMA contraction): same particles, same cells; only the order of the float atomics differs.
    # That order is not reproducible (in either implementation), the sums it changes in the last bit feed the next
    # step's speeds, and once in a while one particle then crosses a cell edge differently: after three steps a
    # handful of cells (one short path) may differ, so the transported maps are required to agree in >= 99.5 % of
    # the cells rather than in all of them.
    nf = runs["nofma"]
    assert np.array_equal(ours["rngFluvial"].cpu().numpy()[:, :24], nf.get("rngFluvial")[:, :24])
    e_nf = np.abs(mine - nf.get("layers"))
    print(f"[{preset}/{pre if pre else 'inline'}] vs -fmad=false reference: max-abs layer error {e_nf.max():.3e}")
    np.testing.assert_allclose(mine, nf.get("layers"), rtol=1e-5, atol=2e-6)
    fin = np.isfinite(nf.get("waterHeight"))
    wh_o, wh_r = ours["waterHeight"].cpu().numpy()[fin], nf.get("waterHeight")[fin]
    assert (np.abs(wh_o - wh_r) <= 1e-4 * np.abs(wh_r) + 1e-9).mean() >= 0.995
    al_o, al_r = ours["albedoSurface"].cpu().numpy(), nf.get("albedoSurface")
    assert (np.abs(al_o - al_r) <= 1e-4 * np.abs(al_r) + 1e-6).mean() >= 0.995

    # stock flags (FMA contraction on): the reference itself is deterministic up to atomic order (run-to-run
    # spread `noise`), but contraction moves particle positions by an ulp and a few particles cross a cell edge
    # differently.  Stated tolerance: >= 97% of cells within max(2*noise, 2e-6) of the stock build's height,
    # total mass to 1e-4 (1e-6 holds against the same-arithmetic build above), mean / variance of the
    # elevation change within 2% / 5%.  The -fmad=false build of the SAME reference source differs from the
    # stock build by the same amount (checked below), i.e. this is the reference's own sensitivity.
    st, st2 = runs["stock"].get("layers"), runs["stock2"].get("layers")
    noise = np.abs(st - st2).max()
    tol = max(2.0 * noise, 2e-6)
    err = np.abs(mine.sum(-1) - st.sum(-1))
    err_ref = np.abs(nf.get("layers").sum(-1) - st.sum(-1))
    out = err > tol
    print(f"[{preset}/{pre if pre else 'inline'}] vs stock reference: reference run-to-run noise {noise:.3e}, cells outside "
          f"{tol:.3e}: {out.mean():.4%}, max-abs error over them {err[out].max() if out.any() else 0.0:.3e} "
          f"(the reference's own -fmad=false build: {(err_ref > tol).mean():.4%}, {err_ref.max():.3e})")
    assert (err <= tol).mean() >= 0.97
    assert (err <= tol).mean() >= (err_ref <= tol).mean() - 0.01
    assert err.max() <= max(2e-2, 2.0 * err_ref.max())      # stated max-abs bound for the forked cells (test_gpu_erosion.py)
    np.testing.assert_allclose(msum.cpu().numpy().sum(), st.astype(np.float64).sum(), rtol=1e-4)
    np.testing.assert_allclose(msum.cpu().numpy(), nf.get("layers").astype(np.float64).sum(axis=(0, 1)), rtol=1e-6)
    dz_m, dz_r = mine.sum(-1) - h, st.sum(-1) - h
    np.testing.assert_allclose(dz_m.mean(), dz_r.mean(), rtol=2e-2, atol=1e-9)
    np.testing.assert_allclose(dz_m.var(), dz_r.var(), rtol=5e-2)


def test_golden_fixture_matches_live_reference(soil, oracle):
    """tests/golden/ref_gpu.npz (committed) was produced by these same reference kernels."""
    import os
    path = os.path.join(os.path.dirname(__file__), "golden", "ref_gpu.npz")
    if not os.path.exists(path):
        pytest.skip("ref_gpu.npz not generated yet")
    g = np.load(path)
    h = torch.from_numpy(g["a_h"]).cuda()
    assert np.array_equal(ref.direction(h, 1).cpu().numpy(), g["a_direction_d8"])
    assert np.array_equal(soil.direction(S(soil, h), soil.d8).numpy(), g["a_direction_d8"])
    assert np.array_equal(soil.random_weighted(S(soil, h), soil.d8, 0, 2, 10.0).numpy(), g["a_rw_d8_k2"])
