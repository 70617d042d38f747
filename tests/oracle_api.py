"""ctypes access to the CPU oracle (oracle/liboracle.so) - test infrastructure only."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liboracle.so")


class Param(C.Structure):
    _fields_ = [
        ("maxage", C.c_uint64), ("lrate", C.c_float), ("timeStep", C.c_float), ("exitSlope", C.c_float),
        ("uplift", C.c_float), ("rainfall", C.c_float), ("gravity", C.c_float), ("evapRate", C.c_float),
        ("frictionFactor", C.c_float), ("fluvialExponent", C.c_float), ("suspensionRateFluvial", C.c_float),
        ("depositionRateFluvial", C.c_float), ("suspensionRateDebris", C.c_float), ("depositionRateDebris", C.c_float),
        ("landslideRateDebris", C.c_float), ("critSlopeBedrock", C.c_float), ("critSlopeSediment", C.c_float),
        ("yieldStress", C.c_float), ("viscosityWater", C.c_float), ("bedShearWater", C.c_float),
        ("densityWater", C.c_float), ("viscosityDebris", C.c_float), ("bedShearDebris", C.c_float),
        ("densityDebris", C.c_float), ("force", C.c_float * 2),
    ]


class Model(C.Structure):
    _fields_ = [("nx", C.c_int), ("ny", C.c_int), ("scale", C.c_float * 3)] + [
        (n, C.c_void_p) for n in (
            "layers", "height", "rainfall", "uplift", "waterHeight", "waterFlux", "mass", "massFlux", "momentum",
            "velocityFlux", "debris", "debrisFlux", "debrisVel", "debrisVelFlux", "albedoBedrock", "albedoSurface",
            "albedoFluxFluvial", "albedoFluxDebris", "delta", "rngFluvial", "rngDebris")
    ] + [("P", C.c_long)]


_lib = None


def build():
    if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(os.path.join(ORACLE_DIR, "soil_oracle.cu")):
        subprocess.run(["make", "-C", ORACLE_DIR, "oracle"], check=True, capture_output=True)
    return LIB


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB)
        _lib.orc_noise_sample.restype = C.c_float
        _lib.orc_noise_sample.argtypes = [C.c_float] * 4 + [C.c_int, C.c_float, C.c_float]
    return _lib


def vp(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def f3(scale):
    return (C.c_float * 3)(*[float(v) for v in scale])


def default_param(**kw) -> Param:
    """soil::param_t defaults (reference erosion.hpp:20-56)."""
    p = Param(512, 1.0, 250.0, 0.02, 0.001, 1.0, 9.81, 0.0002, 0.06, 2.0, 4.5e-8, 0.04, 0.001, 0.01, 0.003, 0.57, 0.3,
              0.001, 1e-6, 0.0075, 1.0, 0.0, 0.99, 2.0)
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def example_param(**kw) -> Param:
    """The fields of example/erosion_gpu.py:76-91 that still exist in soil::param_t."""
    base = dict(timeStep=1000.0, maxage=256, lrate=1.0, gravity=9.81, uplift=0.01, rainfall=1.0, evapRate=0.0005,
                fluvialExponent=0.01, exitSlope=0.025)
    base.update(kw)
    return default_param(**base)


def noise(nx, ny, seed=3.0, ext=None, frequency=1.0, octaves=8, gain=0.6, lacunarity=2.0):
    ext = ext or (float(nx), float(ny))
    out = np.zeros((nx, ny), np.float32)
    lib().orc_noise(vp(out), nx, ny, C.c_float(frequency), octaves, C.c_float(gain), C.c_float(lacunarity),
                    C.c_float(seed), C.c_float(ext[0]), C.c_float(ext[1]))
    return out


def direction(h, edge=1):
    out = np.zeros(h.shape, np.int32)
    lib().orc_direction(vp(np.ascontiguousarray(h)), vp(out), h.shape[0], h.shape[1], edge)
    return out


def steepest(h, edge=1):
    out = np.zeros(h.shape, np.int32)
    lib().orc_steepest(vp(np.ascontiguousarray(h)), vp(out), h.shape[0], h.shape[1], edge)
    return out


def random_weighted(h, edge, seed, offset, T, margin=1e-5):
    out = np.zeros(h.shape, np.int32)
    amb = np.zeros(h.shape, np.uint8)
    lib().orc_random_weighted(vp(np.ascontiguousarray(h)), vp(out), h.shape[0], h.shape[1], edge,
                              C.c_ulonglong(seed), C.c_ulonglong(offset), C.c_float(T), vp(amb), C.c_float(margin))
    return out, amb.astype(bool)


def accumulate(graph, src, decay=None, edge=1, mode=1, decay_scalar=1.0):
    out = np.zeros(src.shape, np.float32)
    rounds = C.c_int(0)
    lib().orc_accumulate(vp(np.ascontiguousarray(graph)), vp(np.ascontiguousarray(src)),
                         vp(None if decay is None else np.ascontiguousarray(decay)), C.c_float(decay_scalar), vp(out),
                         src.shape[0], src.shape[1], edge, mode, C.byref(rounds))
    return out, rounds.value


def seed_states(n, seed=0, offset=0, first=0):
    st = np.zeros((n, 48), np.uint8)
    lib().orc_seed_range(vp(st), C.c_long(first), C.c_long(n), C.c_ulonglong(seed), C.c_ulonglong(offset))
    return st


def uniform(states):
    out = np.zeros(states.shape[0], np.float32)
    lib().orc_uniform(vp(states), C.c_long(states.shape[0]), vp(out))
    return out


def normal(h, scale):
    out = np.zeros(h.shape + (3,), np.float32)
    lib().orc_normal(vp(np.ascontiguousarray(h)), vp(out), h.shape[0], h.shape[1], f3(scale))
    return out


def _f2(scale):
    return (C.c_float * 2)(*[float(v) for v in scale])


def gradient(h, scale):
    out = np.zeros(h.shape + (2,), np.float32)
    lib().orc_gradient(vp(np.ascontiguousarray(h)), vp(out), h.shape[0], h.shape[1], _f2(scale))
    return out


def negslope(h, scale):
    out = np.zeros(h.shape, np.float32)
    lib().orc_negslope(vp(np.ascontiguousarray(h)), vp(out), h.shape[0], h.shape[1], _f2(scale))
    return out


def laplacian(t, scale):
    t = np.ascontiguousarray(t)
    out = np.zeros_like(t)
    lib().orc_laplacian(vp(t), vp(out), t.shape[0], t.shape[1], t.shape[2], _f2(scale))
    return out


def slope(h, flow, scale):
    out = np.zeros(h.shape, np.float32)
    lib().orc_slope(vp(np.ascontiguousarray(h)), vp(np.ascontiguousarray(flow)), vp(out), h.shape[0], h.shape[1], _f2(scale))
    return out


def gaussian_blur(t, sigma):
    """returns (result, first-pass intermediate); the input array is left alone"""
    a = np.ascontiguousarray(t).copy()
    tmp = np.zeros_like(a)
    lib().orc_gaussian_blur(vp(a), vp(tmp), a.shape[0], a.shape[1], a.shape[2], C.c_float(sigma))
    return a, tmp


def solve_uniform(flow, source, decay, states, scale, count):
    """returns the normalised flux; `states` (n,48) uint8 is advanced in place"""
    flow, source, decay = (np.ascontiguousarray(a) for a in (flow, source, decay))
    flux = np.zeros_like(source)
    lib().orc_solve_uniform(vp(flow), vp(source), vp(decay), vp(states), C.c_long(states.shape[0]), vp(flux),
                            source.shape[0], source.shape[1], source.shape[2], _f2(scale), C.c_long(count))
    return flux


def albedo_layer(bedrock, sediment, layers, scale_sediment, shift):
    out = np.zeros_like(bedrock)
    lib().orc_albedo_layer(vp(out), vp(np.ascontiguousarray(bedrock)), vp(np.ascontiguousarray(sediment)),
                           vp(np.ascontiguousarray(layers)), C.c_long(layers.shape[0] * layers.shape[1]),
                           C.c_float(scale_sediment), f3(shift))
    return out


def albedo_stratum(uplift, layers, scale, param, color_a, color_b, age, freq):
    out = np.zeros(uplift.shape + (3,), np.float32)
    lib().orc_albedo_stratum(vp(out), vp(np.ascontiguousarray(uplift)), vp(np.ascontiguousarray(layers)),
                             C.c_long(uplift.size), f3(scale), C.byref(param), f3(color_a), f3(color_b),
                             C.c_float(age), C.c_float(freq))
    return out


def albedo_discharge(albedo, discharge, color, extinction, scale):
    out = np.ascontiguousarray(albedo).copy()
    lib().orc_albedo_discharge(vp(out), vp(np.ascontiguousarray(discharge)), C.c_long(discharge.size), f3(color),
                               C.c_float(extinction), C.c_float(scale))
    return out


def mass_creep(delta, layers, scale, param):
    lib().orc_mass_creep(vp(delta), vp(layers), layers.shape[0], layers.shape[1], f3(scale), C.byref(param))


def mass_transfer(delta, layers, uplift, mass, momentum, debris, scale, param, albedo=None):
    ab = albedo or {}
    lib().orc_mass_transfer(vp(delta), vp(layers), vp(uplift), None, vp(mass), vp(momentum), vp(debris), None,
                            vp(ab.get("bedrock")), vp(ab.get("fluvial")), vp(ab.get("debris")), vp(ab.get("surface")),
                            layers.shape[0], layers.shape[1], f3(scale), C.byref(param))


def transport_fluvial(f, rng, scale, param, particles=1, normalize=1):
    """f: dict of host arrays with the erosion.hpp:69-84 names."""
    nx, ny = f["layers"].shape[:2]
    lib().orc_transport_fluvial(vp(f["layers"]), vp(f["rainfall"]), vp(f["discharge"]), vp(f["discharge_track"]),
                                vp(f["mass"]), vp(f["mass_track"]), vp(f["momentum"]), vp(f["momentum_track"]),
                                vp(f["albedo_bedrock"]), vp(f["albedo_transport"]), vp(f["albedo_surface"]), vp(rng),
                                C.c_long(rng.shape[0]), nx, ny, f3(scale), C.byref(param), particles, normalize)


def transport_debris(f, rng, scale, param, particles=1, normalize=1):
    nx, ny = f["layers"].shape[:2]
    lib().orc_transport_debris(vp(f["layers"]), vp(f["velocity"]), vp(f["velocity_track"]), vp(f["mass"]),
                               vp(f["mass_track"]), vp(f["albedo_bedrock"]), vp(f["albedo_transport"]),
                               vp(f["albedo_surface"]), vp(rng), C.c_long(rng.shape[0]), nx, ny, f3(scale),
                               C.byref(param), particles, normalize)


class HostModel:
    """All maps of one erosion model as host arrays (field names of oracle model_t)."""

    MAPS1 = ("height", "rainfall", "uplift", "waterHeight", "waterFlux", "mass", "massFlux", "debris", "debrisFlux")
    MAPS2 = ("layers", "momentum", "velocityFlux", "debrisVel", "debrisVelFlux", "delta")
    MAPS3 = ("albedoBedrock", "albedoSurface", "albedoFluxFluvial", "albedoFluxDebris")

    def __init__(self, h, scale, P, albedo=True, seed=0, uplift=None):
        nx, ny = h.shape
        self.nx, self.ny, self.scale, self.P = nx, ny, tuple(scale), P
        for n in self.MAPS1:
            setattr(self, n, np.zeros((nx, ny), np.float32))
        for n in self.MAPS2:
            setattr(self, n, np.zeros((nx, ny, 2), np.float32))
        for n in self.MAPS3:
            setattr(self, n, np.full((nx, ny, 3), 0.5, np.float32) if albedo else None)
        if albedo:
            self.albedoBedrock[...] = (0.6, 0.5, 0.4)
            self.albedoSurface[...] = (0.3, 0.5, 0.2)
            self.albedoFluxFluvial[...] = 0
            self.albedoFluxDebris[...] = 0
        self.layers[..., 0] = h
        self.height[...] = h
        self.rainfall[...] = 1.0
        if uplift is not None:
            self.uplift[...] = uplift
        self.rngFluvial = seed_states(P, seed)
        self.rngDebris = seed_states(P, seed + 1)

    def c_model(self):
        m = Model()
        m.nx, m.ny, m.P = self.nx, self.ny, self.P
        m.scale[0], m.scale[1], m.scale[2] = self.scale
        for n, _ in Model._fields_[3:-1]:
            a = getattr(self, n)
            setattr(m, n, None if a is None else a.ctypes.data)
        return m

    def step(self, param, with_creep=True):
        m = self.c_model()
        lib().orc_erode_step(C.byref(m), C.byref(param), 1 if with_creep else 0)
