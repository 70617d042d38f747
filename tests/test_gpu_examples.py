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
