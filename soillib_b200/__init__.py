"""soillib_b200 - B200-native erosion and flow-routing hot path behind soillib's `soil.*` names.

    import soillib_b200 as soil
    from soillib_b200 import silt

Reference surface: python/source/model.cpp:23-60,148-421, python/source/util.cpp:47-73.
Every op runs hand-written sm_100a kernels through the C ABI in include/soilb200.h
(libsoilb200.so); there is no CPU path, and importing without the built library raises
on first use.
"""
from . import silt
from ._lib import load as _load_library
from .ops import (edge, d4, d8, direction, steepest, flow, random_weighted, accumulate, accumulate_decay,
                  accumulation, normal, noise, noise_t, gradient, negslope, laplacian, slope, gaussian_blur,
                  solve_uniform)
from .erosion import (param_t, transport_fluvial, transport_debris, mass_transfer, mass_creep, layer_merge, albedo_layer, albedo_stratum, albedo_discharge, map_t,
                      data_t, erode)
from .multiflow import multiflow, multiflow_replicas, sample_range
from .io import tiff, geotiff, geotiff_meta
from .util import timer, duration, s, ms, us, ns

# the previous API kept the tensor helpers on `soil` itself
from .silt import tensor, shape, float32, float64, int32, cpu, gpu, set, add, multiply, clamp, clone, seed, resize  # noqa: A004
index = shape
buffer = tensor

__all__ = [
    "silt", "edge", "d4", "d8", "direction", "steepest", "flow", "random_weighted", "accumulate", "accumulate_decay",
    "accumulation", "normal", "noise", "noise_t", "gradient", "negslope", "laplacian", "slope", "gaussian_blur", "solve_uniform", "param_t", "transport_fluvial", "transport_debris", "mass_transfer",
    "mass_creep", "layer_merge", "albedo_layer", "albedo_stratum", "albedo_discharge", "map_t", "data_t", "erode", "multiflow", "multiflow_replicas", "sample_range", "tiff", "geotiff", "geotiff_meta", "timer", "duration", "s", "ms", "us", "ns",
]
