"""The reference's example scripts on this library, end to end on the GPU: GeoTIFF -> multi-flow accumulation -> GeoTIFF."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_dem_multiflow_script(product_lib, oracle, tmp_path):
    import soillib_b200 as soil
    from soillib_b200 import silt
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import dem_multiflow

    nx, ny = 96, 128
    h = oracle.noise(nx, ny, 3.0)
    h[5, 7] = -32768.0
    src = soil.geotiff(silt.tensor.from_numpy(h))
    src.meta.scale, src.meta.coords = [30.0, 30.0, 1.0], [0.0, 0.0, 0.0, 4.5e5, 5.2e6, 0.0]
    src.meta.gdal_nodata = "-32768"
    dem, out = str(tmp_path / "dem.tif"), str(tmp_path / "acc.tif")
    src.write(dem)

    fused = dem_multiflow.main(dem, K=6, T=10.0, out=out)
    looped = dem_multiflow.main(dem, K=6, T=10.0, loop=True)
    assert fused.shape == (nx, ny) and np.array_equal(fused, looped)        # one device-side call == the script's loop
    assert fused.max() > 50 and fused.min() >= 0
    back = soil.geotiff(out)
    assert back.meta.scale == [30.0, 30.0, 1.0] and back.meta.coords[3:5] == [4.5e5, 5.2e6]
    assert np.array_equal(back.tensor.numpy(), fused.astype(np.float32))


def test_erosion_gpu_script(product_lib, tmp_path):
    """examples/erosion_gpu.py (the reference's example/erosion_gpu.py:9-113 on this library) at a small resolution."""
    import soillib_b200 as soil
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import erosion_gpu

    out = str(tmp_path / "height.tif")
    first = erosion_gpu.main(res=256, steps=0, quiet=True)
    got = erosion_gpu.main(res=256, steps=8, out=out, quiet=True)
    assert got.shape == (256, 256) and np.isfinite(got).all()
    assert np.abs(got - first).max() > 0                                     # the terrain was eroded
    assert np.abs(got - first).max() < 0.5                                   # ... and is still a terrain
    assert np.array_equal(soil.tiff(out).tensor.numpy(), got)


def test_erosion_gpu_multiscale_script(product_lib, tmp_path):
    """examples/erosion_gpu_multiscale.py (example/erosion_gpu_multiscale.py:28-171): 128^2 -> 256^2 -> 512^2.  At every
    scale-up the surface of the new model is the resampled surface of the old one: height (the base layer) and sediment
    are resized and assigned independently, in the script's order, and no sediment is counted twice; the transported
    maps travel along."""
    import torch
    import soillib_b200 as soil
    from soillib_b200 import silt
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import erosion_gpu_multiscale as ms

    seen = []

    def on_level(old, new, simres):
        want = silt.tensor(silt.float32, silt.shape(*simres), silt.gpu)
        silt.resize(want, old.surface, simres, (old.shape[0], old.shape[1]))
        torch.testing.assert_close(new.surface.data, want.data, rtol=0, atol=1e-6)
        torch.testing.assert_close(new.surface.data, new.height.data + new.sediment.data, rtol=0, atol=0)
        seen.append((tuple(int(v) for v in simres), float(new.sediment.data.double().sum())))

    out = str(tmp_path / "surface.tif")
    r = ms.main([([128, 128], 24), ([256, 256], 4), ([512, 512], 3)], out=out, quiet=True, on_level=on_level)
    assert [s for s, _ in seen] == [(128, 128), (256, 256), (512, 512)]
    assert seen[0][1] == 0.0 and seen[1][1] > 0.0                            # erosion at 128^2 left sediment that travels up
    for k in ("height", "sediment", "discharge", "surface"):
        a = r[k].numpy()
        assert a.size == 512 * 512 and np.isfinite(a).all(), k
    assert np.array_equal(r["surface"].numpy(), r["height"].numpy() + r["sediment"].numpy())
    assert r["discharge"].numpy().max() > 0
    back = soil.geotiff(out)
    assert np.array_equal(back.tensor.numpy().ravel(), r["surface"].numpy().ravel())
    assert back.meta.scale[:2] == [20.0 / 512, 20.0 / 512]
