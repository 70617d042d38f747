"""The C-ABI library loads and exports every symbol include/soilb200.h declares (no GPU needed)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "soilb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(soilb200_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_hot_path():
    names = declared_symbols()
    for must in ("soilb200_direction_f32", "soilb200_steepest_f32", "soilb200_random_weighted_f32",
                 "soilb200_accumulate_f32", "soilb200_transport_fluvial", "soilb200_transport_debris",
                 "soilb200_mass_transfer", "soilb200_mass_creep", "soilb200_layer_merge", "soilb200_normal_f32",
                 "soilb200_erode_step", "soilb200_noise_f32"):
        assert must in names


def test_library_exports_every_declared_symbol(product_lib):
    from soillib_b200 import _lib
    lib = C.CDLL(_lib.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), f"{name} declared in soilb200.h but not exported"
    # and the Python binding table covers the same set
    assert sorted(_lib.SIGNATURES) == declared_symbols()


def test_param_layout_and_defaults(product_lib):
    from soillib_b200 import _lib
    import soillib_b200 as soil
    assert C.sizeof(_lib.Param) == 112  # soil::param_t, erosion.hpp:17-58
    p = soil.param_t()
    assert p.maxage == 512 and p.timeStep == 250.0 and abs(p.critSlopeBedrock - 0.57) < 1e-7
    assert abs(p.suspensionRateFluvial - 4.5e-8) < 1e-14 and p.force == (0.0, 0.0)
    p.samples = 8192  # legacy field of example/erosion_gpu.py:77
    assert p.legacy["samples"] == 8192
    assert product_lib.soilb200_rng_state_bytes() == 48


def test_argument_errors_without_gpu(product_lib):
    """Argument validation happens before any CUDA call, so it is testable on a CPU box."""
    from soillib_b200 import _lib
    assert product_lib.soilb200_direction_f32(None, None, 4, 4, 1, None) == _lib.E_ARG
    assert product_lib.soilb200_direction_f32(8, 8, 4, 4, 7, None) == _lib.E_EDGE
    assert product_lib.soilb200_accumulate_f32(8, 8, None, 8, 4, 4, 1, None, 0, None) == _lib.E_WORKSPACE
    assert product_lib.soilb200_accumulate_ws(1024, 1024, 1) == 12 * 1024 * 1024
    assert product_lib.soilb200_random_weighted_ws(16, 16) == 64 + 6 * 256 * 4


def test_python_surface_errors(product_lib):
    import numpy as np
    import soillib_b200 as soil
    from soillib_b200 import silt
    t = silt.tensor.from_numpy(np.zeros((8, 8), np.float32))
    with pytest.raises(ValueError):  # graph.cu:75-76 mismatch_host
        soil.steepest(t, soil.d8)
    with pytest.raises(ValueError):  # noise.hpp:44-45
        soil.noise(silt.shape(8), soil.noise_t())
    assert soil.flow is soil.steepest and soil.accumulation is soil.accumulate
