#!/usr/bin/env python
"""The reference's example/erosion_gpu_multiscale.py on this library: erosion at increasing resolutions.

    python examples/erosion_gpu_multiscale.py [--to 1000] [--steps 2048,4,4] [--out surface.tif]

Multi-Scale Erosion Procedure (example/erosion_gpu_multiscale.py:18-24): construct the model at a resolution, erode,
construct a new model at an increased resolution from the resampled maps, repeat.  The physical scale attached to the
model (`pscale`) corrects the parameters, so that most of the work happens on small maps.  Set-up, parameter block
(the previous API's names included: recorded, `samples` sets the particle count), `scaleup` and the stepping loop follow
the reference script (lines 28-160); the level list is an argument here (the script's: 128^2 x 2048 steps, 256^2 x 4,
1000^2 x 4; BASELINE.json's configuration 3 continues to 8192^2).  Plotting and the zip archive are left out: the maps
are returned, and the surface is written as a GeoTIFF when --out is given.
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import soillib_b200 as soil  # noqa: E402

wscale = np.array([20.0, 20.0, 4.0])  # World Scale [km] (x, y, z)
nscale = np.array([20.0, 20.0])       # Noise Feature Scale [km] (x, y)


def pixel_scale(simres):
    return [wscale[0] / simres[0],    # Pixel Scale [km/px]
            wscale[1] / simres[1],
            wscale[2]]                # Value Scale [km/unit]


def parameters():
    """example/erosion_gpu_multiscale.py:72-98"""
    param = soil.param_t()
    param.timeStep = 10.0  # Geological Timestep [y]
    param.samples = 32768  # Number of Particle Samples
    param.maxage = 512     # Maximum Particle Lifetime
    param.lrate = 1.0      # Filter Learning Rate

    param.gravity = 9.81   # Specific Gravity [m/s^2]
    param.uplift = 0.01    # Uplift Rate [m/y]

    param.rainfall = 1.0              # Rainfall Rate [m/y]
    param.evapRate = 0.0005           # Evapotranspiration Rate [1/s]
    param.viscosity = 0.000001        # Water Viscosity [m^2/s]
    param.bedShear = 12.5             # Turbulent Shear Stress [Pa = kg/m/s^2]
    param.suspensionRate = 0.0000008  # Fluvial Suspension Rate
    param.depositionRate = 0.00001    # Fluvial Deposition Rate
    param.fluvialExponent = 0.01      # Fluvial Power Exponent
    param.exitSlope = 0.025           # Slope Boundary Condition

    param.critSlope = 0.57                # Critical Slope
    param.debrisCreepRate = 0.0025        # Landslide Erosion Rate
    param.debrisSuspensionRate = 0.00025  # Debris Suspension Rate
    param.debrisDepositionRate = 0.0001   # Debris Deposition Rate
    param.debrisYieldStress = 2E6         # Yield Stress [Pa]
    param.debrisDensity = 2500.0          # Debris Density [kg/m^3]
    param.debrisViscosity = 0.004         # Debris Viscosity [m^2/s]
    param.debrisBedShear = 60 / 2500.0    # Debris Turbulent Shear Stress
    return param


def scaleup(model, data, track, oldres, simres):
    """example/erosion_gpu_multiscale.py:102-140: every map of the model and of the transported data resampled to the
    next resolution (`soil.resize`; its body is in the absent `silt` - bilinear here, soillib_b200/silt.py)."""
    index = soil.index(simres)
    pscale = pixel_scale(simres)

    # Upscale Individual Buffers
    height = soil.buffer(soil.float32, index.elem(), soil.gpu)
    soil.resize(height, model.height, simres, oldres)

    sediment = soil.buffer(soil.float32, index.elem(), soil.gpu)
    soil.resize(sediment, model.sediment, simres, oldres)

    rainfall = soil.buffer(soil.float32, index.elem(), soil.gpu)
    soil.resize(rainfall, model.rainfall, simres, oldres)

    uplift = soil.buffer(soil.float32, index.elem(), soil.gpu)
    soil.resize(uplift, model.uplift, simres, oldres)

    model = soil.map_t(index, pscale)
    model.height = height
    model.sediment = sediment
    model.rainfall = rainfall
    model.uplift = uplift

    # Update Tracking
    newdata = soil.data_t(index.elem())
    newtrack = soil.data_t(index.elem())

    soil.resize(newdata.mass, data.mass, simres, oldres)
    soil.resize(newdata.discharge, data.discharge, simres, oldres)
    soil.resize(newdata.momentum, data.momentum, simres, oldres)
    soil.resize(newdata.debris_momentum, data.debris_momentum, simres, oldres)
    soil.resize(newdata.debris, data.debris, simres, oldres)

    # (the reference returns `newtrack, newdata` into `data, track`: the two swap roles at every level, which only
    # works there because both are re-filled; the resampled transported maps are the data here)
    return model, newdata, newtrack, index, simres, pscale


def default_levels(to=1000):
    """The script's list (128^2 x 2048 steps, 256^2 x 4, 1000^2 x 4), continued by doubling when `to` is larger."""
    levels = [([128, 128], 2048), ([256, 256], 4)]
    if to <= 1000:
        return levels + [([to, to], 4)] if to > 256 else levels[: 1 + (to >= 256)]
    res = 512
    while res < to:
        levels.append(([res, res], 4))
        res *= 2
    return levels + [([to, to], 4)]


def main(ksteps=None, out=None, quiet=False, on_level=None):
    simres = np.array([128, 128])         # Resolution [px]
    pscale = pixel_scale(simres)

    noise_param = soil.noise_t()
    noise_param.ext = simres * nscale / wscale[0:2]
    noise_param.seed = 3

    index = soil.index(simres)
    height = soil.noise(index, noise_param)
    soil.multiply(height, 1.0)

    sediment = soil.buffer(soil.float32, index.elem(), soil.gpu)
    sediment[:] = 0.0

    # Construct Model
    model = soil.map_t(index, pscale)
    model.height = height.gpu()
    model.sediment = sediment.gpu()

    model.rainfall = soil.buffer(soil.float32, index.elem(), soil.gpu)
    rain = model.rainfall
    soil.set(rain, 1.0)

    uplift = soil.noise(index, noise_param)
    soil.clamp(uplift, 0.0, 1.0)
    model.uplift = uplift.gpu()

    # Construct Data
    data = soil.data_t(index.elem())
    track = soil.data_t(index.elem())

    data.discharge[:] = 0.0
    data.momentum[:] = [0.0, 0.0]
    data.mass[:] = 0.0
    data.debris[:] = 0.0
    data.debris_momentum[:] = [0.0, 0.0]

    param = parameters()
    timer = soil.timer()

    ksteps = ksteps if ksteps is not None else default_levels()
    # Note: The first scale-up procedure here is redundant and can be removed.
    for nextres, steps in ksteps:
        before = model
        model, data, track, index, simres, pscale = scaleup(model, data, track, simres, nextres)
        if on_level is not None:
            on_level(before, model, simres)
        if not quiet:
            print(f"Simulating Resolution: {simres}")
        for i in range(steps):
            with timer:
                soil.erode(model, data, track, param, 1)
            if not quiet:
                print(f"Execution Time: {timer.count}ms")

    # Geotiff so that pixel and value scale are respected, and we must also add the height of all layers.
    result = {"height": model.height.cpu(), "sediment": model.sediment.cpu(), "discharge": data.discharge.cpu(),
              "surface": model.surface.cpu(), "index": index, "pscale": pscale}
    if out:
        img = soil.geotiff(result["surface"])
        img.meta.scale = [float(pscale[0]), float(pscale[1]), float(pscale[2])]
        img.write(out)
    return result


if __name__ == "__main__":
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--to", type=int, default=1000, help="final resolution (the script's list ends at 1000; 8192 = BASELINE configuration 3)")
    ap.add_argument("--steps", default=None, help="comma-separated steps per level, overriding the list's")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    levels = default_levels(a.to)
    if a.steps:
        st = [int(v) for v in a.steps.split(",")]
        levels = [(r, st[min(i, len(st) - 1)]) for i, (r, _) in enumerate(levels)]
    main(levels, a.out)
