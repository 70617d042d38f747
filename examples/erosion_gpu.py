#!/usr/bin/env python
"""The reference's example/erosion_gpu.py on this library: particle erosion of a synthetic fBm terrain.

    python examples/erosion_gpu.py [--res 256] [--steps 512] [--out height.tif]

The model set-up, the parameter block (including the names of the previous API that no longer exist in
soil::param_t - they are recorded, `samples` sets the particle count) and the stepping loop follow the reference
script line by line (example/erosion_gpu.py:9-106); `import silt` / `import soillib as soil` become the two imports
below.  Plotting (matplotlib relief view, :111-113) is left out: the final height map is returned, and written as a
TIFF when --out is given.
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import soillib_b200 as soil  # noqa: E402
from soillib_b200 import silt  # noqa: E402


def noise(shape, scale):
    noise_param = soil.noise_t()
    noise_param.ext = np.array([shape[0], shape[1]]) * scale
    noise_param.seed = 3
    tensor = soil.noise(shape, noise_param)
    soil.multiply(tensor, 1.0)
    return tensor.gpu()


def full(value, shape, dtype=silt.float32, host=silt.cpu):
    tensor = silt.tensor(dtype, shape, host)
    silt.set(tensor, value)
    return tensor


def parameters():
    """example/erosion_gpu.py:75-100"""
    param = soil.param_t()
    param.timeStep = 1000.0  # Geological Timestep [y]
    param.samples = 8192     # Number of Particle Samples
    param.maxage = 256       # Maximum Particle Lifetime
    param.lrate = 1          # Filter Learning Rate

    param.gravity = 9.81     # Specific Gravity [m/s^2]
    param.uplift = 0.01      # Uplift Rate [m/y]

    param.rainfall = 1.0             # Rainfall Rate [m/y]
    param.evapRate = 0.0005          # Evapotranspiration Rate [1/s]
    param.viscosity = 0.000001       # Water Viscosity [m^2/s]
    param.bedShear = 12.5            # Turbulent Shear Stress [Pa = kg/m/s^2]
    param.suspensionRate = 0.0008    # Fluvial Suspension Rate
    param.depositionRate = 0.00001   # Fluvial Deposition Rate
    param.fluvialExponent = 0.01     # Fluvial Power Exponent
    param.exitSlope = 0.025          # Slope Boundary Condition

    param.critSlope = 0.57                # Critical Slope
    param.debrisCreepRate = 0.0025        # Landslide Erosion Rate
    param.debrisSuspensionRate = 0.00025  # Debris Suspension Rate
    param.debrisDepositionRate = 0.0001   # Debris Deposition Rate
    param.debrisYieldStress = 2E6         # Yield Stress [Pa]
    param.debrisDensity = 2500.0          # Debris Density [kg/m^3]
    param.debrisViscosity = 0.004         # Debris Viscosity [m^2/s]
    param.debrisBedShear = 60 / 2500.0    # Debris Turbulent Shear Stress
    return param


def main(res=256, steps=512, out=None, quiet=False):
    # Simulation Resolution
    simres = np.array([res, res])          # Resolution [px]
    shape = silt.shape(*simres)            # Shape
    wscale = np.array([20.0, 20.0, 4.0])   # World Scale [km] (x, y, z)
    nscale = np.array([20.0, 20.0])        # Noise Feature Scale [km] (x, y)
    pscale = [wscale[0] / simres[0],       # Pixel Scale [km/px]
              wscale[1] / simres[1],
              wscale[2]]                   # Value Scale [km/unit]

    # Overall Model
    model = soil.map_t(shape, pscale)
    model.height = noise(shape, nscale / wscale[0:2])
    model.sediment = full(0.0, shape, dtype=silt.float32, host=silt.gpu)
    model.rainfall = full(1.0, shape, dtype=silt.float32, host=silt.gpu)
    model.uplift = full(0.0, shape, dtype=silt.float32, host=silt.gpu)

    # Transported Data
    data = soil.data_t(shape)
    data.discharge = full(0.0, shape, dtype=silt.float32, host=silt.gpu)
    data.mass = full(0.0, shape, dtype=silt.float32, host=silt.gpu)
    data.debris = full(0.0, shape, dtype=silt.float32, host=silt.gpu)
    data.momentum = full(0.0, silt.shape(*simres, 2), dtype=silt.float32, host=silt.gpu)
    data.debris_momentum = full(0.0, silt.shape(*simres, 2), dtype=silt.float32, host=silt.gpu)

    track = soil.data_t(shape)
    track.discharge = full(0.0, shape, dtype=silt.float32, host=silt.gpu)
    track.mass = full(0.0, shape, dtype=silt.float32, host=silt.gpu)
    track.debris = full(0.0, shape, dtype=silt.float32, host=silt.gpu)
    track.momentum = full(0.0, silt.shape(*simres, 2), dtype=silt.float32, host=silt.gpu)
    track.debris_momentum = full(0.0, silt.shape(*simres, 2), dtype=silt.float32, host=silt.gpu)

    # Construct Parameters
    param = parameters()

    timer = soil.timer()
    for i in range(steps):
        with timer:
            soil.erode(model, data, track, param, 1)
        if not quiet:
            print(f"Execution Time: {timer.count}ms")

    # the relief is the sum of the layers (example/erosion_gpu_multiscale.py:165-167: "we must also add the height of all
    # layers"): `height` is the base layer, `surface` = height + sediment
    height = model.surface.cpu()
    if out:
        soil.tiff(height).write(out)
    return height.numpy()


if __name__ == "__main__":
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--res", type=int, default=256)
    ap.add_argument("--steps", type=int, default=512)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    main(a.res, a.steps, a.out)
