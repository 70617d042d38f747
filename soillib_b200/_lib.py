"""ctypes binding of the C ABI in include/soilb200.h (libsoilb200.so, built in-tree).

There is no fallback: if the library is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SOILB200_LIB") or os.path.join(HERE, "libsoilb200.so")  # env override: tuning builds only

E_ARG, E_EDGE, E_WORKSPACE, E_LIMIT = -1, -2, -3, -4


class Param(C.Structure):
    """soilb200_param_t == soil::param_t (reference erosion.hpp:17-58)."""

    _fields_ = [
        ("maxage", C.c_uint64),
        ("lrate", C.c_float), ("timeStep", C.c_float),
        ("exitSlope", C.c_float), ("uplift", C.c_float), ("rainfall", C.c_float), ("gravity", C.c_float),
        ("evapRate", C.c_float),
        ("frictionFactor", C.c_float), ("fluvialExponent", C.c_float),
        ("suspensionRateFluvial", C.c_float), ("depositionRateFluvial", C.c_float),
        ("suspensionRateDebris", C.c_float), ("depositionRateDebris", C.c_float), ("landslideRateDebris", C.c_float),
        ("critSlopeBedrock", C.c_float), ("critSlopeSediment", C.c_float), ("yieldStress", C.c_float),
        ("viscosityWater", C.c_float), ("bedShearWater", C.c_float), ("densityWater", C.c_float),
        ("viscosityDebris", C.c_float), ("bedShearDebris", C.c_float), ("densityDebris", C.c_float),
        ("force", C.c_float * 2),
    ]


class Model(C.Structure):
    """soilb200_model_t (include/soilb200.h)."""

    _fields_ = [
        ("nx", C.c_int32), ("ny", C.c_int32), ("scale", C.c_float * 3), ("with_creep", C.c_int32),
        ("layers", C.c_void_p), ("layers_next", C.c_void_p), ("height", C.c_void_p),
        ("rainfall", C.c_void_p), ("uplift", C.c_void_p),
        ("discharge", C.c_void_p), ("mass", C.c_void_p), ("momentum", C.c_void_p), ("debris", C.c_void_p),
        ("debris_momentum", C.c_void_p),
        ("albedo_bedrock", C.c_void_p), ("albedo_surface", C.c_void_p),
        ("rng_fluvial", C.c_void_p), ("rng_debris", C.c_void_p), ("n_particles", C.c_int64),
        ("flux", C.c_void_p), ("touched", C.c_void_p), ("flux_prev", C.c_void_p), ("touched_prev", C.c_void_p),
        ("mass_sum", C.c_void_p),
        ("global_nx", C.c_int32), ("global_row0", C.c_int32), ("own_row_lo", C.c_int32), ("own_row_hi", C.c_int32),
        ("outbox_lo", C.c_void_p), ("outbox_hi", C.c_void_p), ("outbox_counts", C.c_void_p),
        ("outbox_capacity", C.c_int32), ("phases", C.c_int32),
        ("peers", C.c_void_p), ("n_peers", C.c_int32), ("peer_self", C.c_int32),
        ("pre", C.c_void_p),
        ("spawn", C.c_void_p),
        ("dense_row_lo", C.c_int32), ("dense_row_hi", C.c_int32),
        ("view", C.c_void_p),
    ]


class Peer(C.Structure):
    """soilb200_peer_t (include/soilb200.h): one rank's rows and peer-mapped arrays, 64 bytes."""

    _fields_ = [
        ("own_row_lo", C.c_int32), ("own_row_hi", C.c_int32), ("global_row0", C.c_int32), ("rows", C.c_int32),
        ("height", C.c_void_p), ("discharge", C.c_void_p), ("momentum", C.c_void_p), ("debris_momentum", C.c_void_p),
        ("flux", C.c_void_p), ("touched", C.c_void_p), ("view", C.c_void_p), ("reserved", C.c_void_p),
    ]


assert C.sizeof(Peer) == 80


assert C.sizeof(Param) == 112

_P, _I, _I64, _U64, _F, _SZ = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_float, C.c_size_t
_F3 = C.POINTER(C.c_float)
_PP = C.POINTER(Param)

# every symbol include/soilb200.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "soilb200_param_default": (None, [_PP]),
    "soilb200_version": (C.c_char_p, []),
    "soilb200_device_count": (_I, []),
    "soilb200_rng_state_bytes": (_I, []),
    "soilb200_launch_count": (_U64, []),
    "soilb200_profile_enable": (_I, [_I]),
    "soilb200_profile_read": (_I, [C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int64), C.POINTER(C.c_uint64)]),
    "soilb200_selftest_div_sqrt2": (_I, [_P, _P]),
    "soilb200_selftest_div_const": (_I, [_F, _P, C.POINTER(C.c_int32), _P]),
    "soilb200_direction_f32": (_I, [_P, _P, _I, _I, _I, _P]),
    "soilb200_steepest_f32": (_I, [_P, _P, _I, _I, _I, _P]),
    "soilb200_seed_xorwow": (_I, [_P, _I64, _U64, _U64, _U64, _P]),
    "soilb200_random_weighted_ws": (_SZ, [_I, _I]),
    "soilb200_random_weighted_f32": (_I, [_P, _P, _I, _I, _I, _U64, _U64, _F, _P, _SZ, _P]),
    "soilb200_random_weighted_strip_f32": (_I, [_P, _P, _I, _I, _I, _U64, _U64, _F, _U64, _P, _SZ, _P]),
    "soilb200_accumulate_ws": (_SZ, [_I, _I, _I]),
    "soilb200_accumulate_decay_ws": (_SZ, [_I, _I, _I]),
    "soilb200_accumulate_f32": (_I, [_P, _P, _P, _P, _I, _I, _I, _P, _SZ, _P]),
    "soilb200_graph_terminal_i32": (_I, [_P, _P, _P, _I, _I, _I, _P]),
    "soilb200_accumulate_static_f32": (_I, [_P, _P, _P, _P, _I, _I, _I, _P]),
    "soilb200_gradient_f32": (_I, [_P, _P, _I, _I, _P, _P]),
    "soilb200_negslope_f32": (_I, [_P, _P, _I, _I, _P, _P]),
    "soilb200_laplacian_f32": (_I, [_P, _P, _I, _I, _I, _P, _P]),
    "soilb200_slope_f32": (_I, [_P, _P, _P, _I, _I, _P, _P]),
    "soilb200_gaussian_blur_f32": (_I, [_P, _P, _I, _I, _I, _F, _P]),
    "soilb200_solve_uniform_f32": (_I, [_P, _P, _P, _P, _I64, _P, _I, _I, _I, _P, _I64, _P]),
    "soilb200_albedo_layer": (_I, [_P, _P, _P, _P, _I, _I, _F, _P, _P]),
    "soilb200_albedo_stratum": (_I, [_P, _P, _P, _I, _I, _P, _PP, _P, _P, _F, _F, _P]),
    "soilb200_albedo_discharge": (_I, [_P, _P, _I, _I, _P, _F, _F, _P]),
    "soilb200_multiflow_ws": (_SZ, [_I, _I, _I]),
    "soilb200_multiflow_f64": (_I, [_P, _P, _P, _I, _I, _I, _U64, _U64, _U64, _U64, _F, _I, _P, _SZ, _P]),
    "soilb200_transport_fluvial": (_I, [_P] * 12 + [_I64, _I, _I, _F3, _PP, _P]),
    "soilb200_transport_debris": (_I, [_P] * 9 + [_I64, _I, _I, _F3, _PP, _P]),
    "soilb200_mass_transfer": (_I, [_P] * 12 + [_I, _I, _F3, _PP, _P]),
    "soilb200_mass_creep": (_I, [_P, _P, _I, _I, _F3, _PP, _P]),
    "soilb200_layer_merge": (_I, [_P, _P, _I, _I, _P]),
    "soilb200_normal_f32": (_I, [_P, _P, _I, _I, _F3, _P]),
    "soilb200_noise_f32": (_I, [_P, _I, _I, _F, _I, _F, _F, _F, _F, _F, _I, _P]),
    "soilb200_erode_ws": (_SZ, [_I, _I]),
    "soilb200_erode_flux_bytes": (_SZ, [_I, _I, _I]),
    "soilb200_erode_pre_bytes": (_SZ, [_I, _I]),
    "soilb200_erode_spawn_bytes": (_SZ, [_I64]),
    "soilb200_erode_view_bytes": (_SZ, [_I, _I]),
    "soilb200_erode_import": (_I, [C.POINTER(Model), _P]),
    "soilb200_erode_export": (_I, [C.POINTER(Model), _P]),
    "soilb200_erode_step": (_I, [C.POINTER(Model), _PP, _P]),
    "soilb200_enable_peer_access": (_I, [_I]),
    "soilb200_ipc_export": (_I, [_P, C.c_char_p, C.POINTER(C.c_uint64)]),
    "soilb200_ipc_open": (_I, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "soilb200_ipc_close": (_I, [_P]),
    "soilb200_erode_resume": (_I, [C.POINTER(Model), _PP, _P, _I, _P]),
}

_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(soillib_b200 has no CPU or pure-Python fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int, what: str) -> None:
    """Maps C-ABI status codes to the exceptions the reference's bindings raise."""
    if rc == 0:
        return
    if rc == E_EDGE:
        raise ValueError("invalid edge enumerator")  # std::invalid_argument, graph.cu:88
    if rc == E_ARG:
        raise ValueError(f"{what}: invalid argument (null tensor or non-positive size)")
    if rc == E_WORKSPACE:
        raise ValueError(f"{what}: workspace missing or too small")
    if rc == E_LIMIT:
        raise ValueError(f"{what}: size beyond a fixed limit of this library")
    raise RuntimeError(f"{what}: CUDA error {rc}")


def scale3(scale):
    s = [float(v) for v in scale]
    if len(s) == 2:
        s.append(1.0)
    if len(s) != 3:
        raise ValueError("scale must have 3 components")
    return (C.c_float * 3)(*s)
