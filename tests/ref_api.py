"""ctypes access to the reference's own kernels (oracle/_ref/libsoilref*.so) - test infrastructure only.

The library is built by `make -C oracle ref` from the reference sources where
they lie under /root/reference, behind the silt/glm stand-in headers in
oracle/shim.  It needs a GPU; pointers are torch CUDA tensors.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")


class RefModel(C.Structure):
    _fields_ = [("nx", C.c_int), ("ny", C.c_int), ("scale", C.c_float * 3)] + [
        (n, C.c_void_p) for n in (
            "layers", "height", "rainfall", "uplift", "waterHeight", "waterFlux", "mass", "massFlux", "momentum",
            "velocityFlux", "debris", "debrisFlux", "debrisVel", "debrisVelFlux", "albedoBedrock", "albedoSurface",
            "albedoFluxFluvial", "albedoFluxDebris", "delta", "rngFluvial", "rngDebris")
    ] + [("P", C.c_long)]


_libs = {}


def available(nofma: bool = False) -> bool:
    return os.path.exists(os.path.join(REF_DIR, "libsoilref_nofma.so" if nofma else "libsoilref.so"))


def lib(nofma: bool = False):
    key = bool(nofma)
    if key not in _libs:
        _libs[key] = C.CDLL(os.path.join(REF_DIR, "libsoilref_nofma.so" if nofma else "libsoilref.so"))
    return _libs[key]


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _f3(scale):
    return (C.c_float * 3)(*[float(v) for v in scale])


def _ok(rc, what):
    if rc != 0:
        raise RuntimeError(f"reference {what}: CUDA error {rc}")


def _f2(scale):
    return (C.c_float * 2)(*[float(v) for v in scale])


def gradient(h, scale, nofma=False):
    out = torch.empty(tuple(h.shape) + (2,), dtype=torch.float32, device=h.device)
    _ok(lib(nofma).ref_gradient(_p(h), _p(out), h.shape[0], h.shape[1], _f2(scale)), "gradient")
    return out


def negslope(h, scale, nofma=False):
    out = torch.empty(h.shape, dtype=torch.float32, device=h.device)
    _ok(lib(nofma).ref_negslope(_p(h), _p(out), h.shape[0], h.shape[1], _f2(scale)), "negslope")
    return out


def laplacian(t, scale, nofma=False):
    out = torch.empty_like(t)
    _ok(lib(nofma).ref_laplacian(_p(t), _p(out), t.shape[0], t.shape[1], t.shape[2], _f2(scale)), "laplacian")
    return out


def slope(h, flow, scale, nofma=False):
    out = torch.empty(h.shape, dtype=torch.float32, device=h.device)
    _ok(lib(nofma).ref_slope(_p(h), _p(flow), _p(out), h.shape[0], h.shape[1], _f2(scale)), "slope")
    return out


def gaussian_blur(t, sigma, nofma=False):
    """blurs a copy (the reference overwrites its argument) and returns it"""
    a = t.clone()
    _ok(lib(nofma).ref_gaussian_blur(_p(a), a.shape[0], a.shape[1], a.shape[2], C.c_float(sigma)), "gaussian_blur")
    return a


def solve_uniform(flow, source, decay, states, scale, count, nofma=False):
    flux = torch.empty_like(source)
    _ok(lib(nofma).ref_solve_uniform(_p(flow), _p(source), _p(decay), _p(states), C.c_long(states.shape[0]), _p(flux),
                                     source.shape[0], source.shape[1], source.shape[2], _f2(scale), C.c_long(count)),
        "solve_uniform")
    return flux


def albedo_layer(bedrock, sediment, layers, scale_sediment, shift, nofma=False):
    out = torch.empty_like(bedrock)
    _ok(lib(nofma).ref_albedo_layer(_p(out), _p(bedrock), _p(sediment), _p(layers), layers.shape[0], layers.shape[1],
                                    C.c_float(scale_sediment), _f3(shift)), "albedo_layer")
    return out


def albedo_stratum(uplift, layers, scale, param, color_a, color_b, age, freq, nofma=False):
    out = torch.empty(tuple(uplift.shape) + (3,), dtype=torch.float32, device=uplift.device)
    _ok(lib(nofma).ref_albedo_stratum(_p(out), _p(uplift), _p(layers), uplift.shape[0], uplift.shape[1], _f3(scale),
                                      C.byref(param), _f3(color_a), _f3(color_b), C.c_float(age), C.c_float(freq)),
        "albedo_stratum")
    return out


def albedo_discharge(albedo, discharge, color, extinction, scale, nofma=False):
    out = albedo.clone()
    _ok(lib(nofma).ref_albedo_discharge(_p(out), _p(discharge), discharge.shape[0], discharge.shape[1], _f3(color),
                                        C.c_float(extinction), C.c_float(scale)), "albedo_discharge")
    return out


def direction(h, edge=1, nofma=False):
    out = torch.empty(h.shape, dtype=torch.int32, device=h.device)
    _ok(lib(nofma).ref_direction(_p(h), _p(out), h.shape[0], h.shape[1], edge), "direction")
    return out


def steepest(h, edge=1, nofma=False):
    out = torch.empty(h.shape, dtype=torch.int32, device=h.device)
    _ok(lib(nofma).ref_steepest(_p(h), _p(out), h.shape[0], h.shape[1], edge), "steepest")
    return out


def random_weighted(h, edge, seed, offset, T, nofma=False):
    out = torch.empty(h.shape, dtype=torch.int32, device=h.device)
    _ok(lib(nofma).ref_random_weighted(_p(h), _p(out), h.shape[0], h.shape[1], edge, C.c_ulonglong(seed),
                                       C.c_ulonglong(offset), C.c_float(T)), "random_weighted")
    return out


def accumulate(graph, src, decay=None, edge=1, nofma=False):
    out = torch.empty(src.shape, dtype=torch.float32, device=src.device)
    _ok(lib(nofma).ref_accumulate(_p(graph), _p(src), _p(decay), _p(out), src.shape[0], src.shape[1], edge), "accumulate")
    return out


def seed(states, seed_value=0, offset=0, nofma=False):
    _ok(lib(nofma).ref_seed(_p(states), C.c_long(states.shape[0]), C.c_ulonglong(seed_value), C.c_ulonglong(offset)), "seed")


def transport_fluvial(f, rng, scale, param, nofma=False):
    nx, ny = f["layers"].shape[:2]
    _ok(lib(nofma).ref_transport_fluvial(
        _p(f["layers"]), _p(f["rainfall"]), _p(f["discharge"]), _p(f["discharge_track"]), _p(f["mass"]),
        _p(f["mass_track"]), _p(f["momentum"]), _p(f["momentum_track"]), _p(f["albedo_bedrock"]),
        _p(f["albedo_transport"]), _p(f["albedo_surface"]), _p(rng), C.c_long(rng.shape[0]), nx, ny, _f3(scale),
        C.byref(param)), "transport_fluvial")


def transport_debris(f, rng, scale, param, nofma=False):
    nx, ny = f["layers"].shape[:2]
    _ok(lib(nofma).ref_transport_debris(
        _p(f["layers"]), _p(f["velocity"]), _p(f["velocity_track"]), _p(f["mass"]), _p(f["mass_track"]),
        _p(f["albedo_bedrock"]), _p(f["albedo_transport"]), _p(f["albedo_surface"]), _p(rng), C.c_long(rng.shape[0]),
        nx, ny, _f3(scale), C.byref(param)), "transport_debris")


def mass_transfer(delta, layers, uplift, discharge, mass, momentum, debris, momentum_debris, ab, af, ad, asf, scale,
                  param, nofma=False):
    nx, ny = layers.shape[:2]
    _ok(lib(nofma).ref_mass_transfer(_p(delta), _p(layers), _p(uplift), _p(discharge), _p(mass), _p(momentum),
                                     _p(debris), _p(momentum_debris), _p(ab), _p(af), _p(ad), _p(asf), nx, ny,
                                     _f3(scale), C.byref(param)), "mass_transfer")


def mass_creep(delta, layers, scale, param, nofma=False):
    nx, ny = layers.shape[:2]
    _ok(lib(nofma).ref_mass_creep(_p(delta), _p(layers), nx, ny, _f3(scale), C.byref(param)), "mass_creep")


def layer_merge(height, layers, nofma=False):
    _ok(lib(nofma).ref_layer_merge(_p(height), _p(layers), layers.shape[0], layers.shape[1]), "layer_merge")


class DeviceModel:
    """Device copy of tests/oracle_api.HostModel, steppable by the reference's unfused kernel sequence."""

    def __init__(self, host, device="cuda"):
        self.host = host
        self.t = {}
        names = host.MAPS1 + host.MAPS2 + host.MAPS3 + ("rngFluvial", "rngDebris")
        for n in names:
            a = getattr(host, n)
            self.t[n] = None if a is None else torch.from_numpy(a.copy()).to(device)
        if self.t["albedoSurface"] is None:  # the reference kernels always dereference albedo
            nx, ny = host.nx, host.ny
            for n in host.MAPS3:
                self.t[n] = torch.zeros((nx, ny, 3), dtype=torch.float32, device=device)
        self.m = RefModel()
        self.m.nx, self.m.ny, self.m.P = host.nx, host.ny, host.P
        self.m.scale[0], self.m.scale[1], self.m.scale[2] = host.scale
        for n, _ in RefModel._fields_[3:-1]:
            setattr(self.m, n, self.t[n].data_ptr())

    def step(self, param, with_creep=True, nofma=False):
        _ok(lib(nofma).ref_erode_step(C.byref(self.m), C.byref(param), 1 if with_creep else 0), "erode_step")

    def get(self, name):
        return self.t[name].cpu().numpy()
