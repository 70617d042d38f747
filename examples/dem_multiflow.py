#!/usr/bin/env python
"""The reference's example/dem_multiflow.py on this library: GeoTIFF in, multi-flow accumulation out.

    python examples/dem_multiflow.py dem.tif [--K 512] [--T 10.0] [--out accumulation.tif] [--loop]

The reference script (example/dem_multiflow.py:20-51) reads the DEM, draws K Gibbs-random flow graphs, accumulates unit
rain over each and averages on the host.  `--loop` runs exactly that loop through the same-named soil.* calls; the
default runs it as one device-side call (soil.multiflow), which gives the same mean bit for bit.  Plotting is left out;
the result is written as a GeoTIFF carrying the input's georeferencing.
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import soillib_b200 as soil  # noqa: E402
from soillib_b200 import silt  # noqa: E402


def main(data, K=512, T=10.0, out=None, loop=False):
    # Load GeoTIFF Data
    tiff = soil.geotiff(data)
    tensor = tiff.tensor.gpu()
    shape = tensor.shape
    res = (shape[0], shape[1])

    # Compute Accumulation
    rain = np.full(res, 1.0)
    rain = silt.tensor.from_numpy(rain.astype(np.float32)).gpu()

    t = soil.timer(soil.us)
    with t:
        if loop:
            multiflow = np.full(res, 0.0)
            for k in range(K):
                flow = soil.random_weighted(tensor, soil.d8, 0, k, T)
                accumulation = soil.accumulate(flow, rain, soil.d8)
                multiflow += accumulation.cpu().numpy() / float(K)
        else:
            multiflow = soil.multiflow(tensor, rain, soil.d8, K, T).numpy()
    print(f"Execution Time: {t.count} us")

    if out:
        result = soil.geotiff(silt.tensor.from_numpy(multiflow.astype(np.float32)))
        for name in ("scale", "coords", "params", "keydir", "gdal_ascii", "gdal_metadata"):
            setattr(result.meta, name, getattr(tiff.meta, name))
        result.write(out)
    return multiflow


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("data")
    ap.add_argument("--K", type=int, default=512)
    ap.add_argument("--T", type=float, default=10.0)
    ap.add_argument("--out", default=None)
    ap.add_argument("--loop", action="store_true")
    a = ap.parse_args()
    main(a.data, a.K, a.T, a.out, a.loop)
