"""Erosion ops and the `erode` driver (reference: python/source/model.cpp:23-60, 237-351).

`param_t` mirrors soil::param_t (erosion.hpp:17-58).  transport_* / mass_* /
layer_merge keep HEAD's in-place contracts.  `map_t`, `data_t` and `erode` are
the names of the previous soillib API that example/erosion_gpu.py still uses
(commented out at HEAD, model.cpp:62-143); `erode` runs the fused B200 step.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from . import _lib, silt
from .ops import _gpu, _stream, _dims2

_PARAM_FIELDS = [name for name, _ in _lib.Param._fields_]


class param_t:
    """soil.param_t: the 25 fields of erosion.hpp:17-58 with the reference defaults.

    Assigning a name that soil::param_t no longer has (the stale examples do:
    `samples`, `viscosity`, ...) is recorded in `.legacy` instead of raising, and
    `samples` is honoured by `erode` as the particle count.
    """

    def __init__(self):
        c = _lib.Param()
        _lib.load().soilb200_param_default(C.byref(c))
        object.__setattr__(self, "_c", c)
        object.__setattr__(self, "legacy", {})

    def __getattr__(self, name):
        if name in _PARAM_FIELDS:
            v = getattr(self._c, name)
            return (v[0], v[1]) if name == "force" else v
        if name in self.legacy:
            return self.legacy[name]
        raise AttributeError(name)

    def __setattr__(self, name, value):
        if name == "force":
            self._c.force[0], self._c.force[1] = float(value[0]), float(value[1])
        elif name == "maxage":
            self._c.maxage = int(value)
        elif name in _PARAM_FIELDS:
            setattr(self._c, name, float(value))
        else:
            self.legacy[name] = value

    def c(self):
        return C.byref(self._c)

    def copy(self):
        p = param_t()
        C.memmove(C.byref(p._c), C.byref(self._c), C.sizeof(_lib.Param))
        p.legacy.update(self.legacy)
        return p


def _f(t, what):
    return _gpu(t, silt.float32, what)


def _opt(t, what):
    return None if t is None else _f(t, what).data_ptr()


def transport_fluvial(layers, rainfall, discharge, discharge_track, mass, mass_track, momentum, momentum_track,
                      albedo_bedrock, albedo_transport, albedo_surface, rng, scale, param):
    """soil.transport_fluvial(...) - argument order of model.cpp:237-268 / erosion.hpp:69-84."""
    w = "transport_fluvial"
    L = _f(layers, w)
    nx, ny = _dims2(L, w)
    r = _gpu(rng, silt.rng, w)
    rc = _lib.load().soilb200_transport_fluvial(
        L.data_ptr(), _f(rainfall, w).data_ptr(), _f(discharge, w).data_ptr(), _f(discharge_track, w).data_ptr(),
        _f(mass, w).data_ptr(), _f(mass_track, w).data_ptr(), _f(momentum, w).data_ptr(),
        _f(momentum_track, w).data_ptr(), _opt(albedo_bedrock, w), _opt(albedo_transport, w), _opt(albedo_surface, w),
        r.data_ptr(), r.numel() // 48, nx, ny, _lib.scale3(scale), param.c(), _stream())
    _lib.check(rc, w)


def transport_debris(layers, velocity, velocity_track, mass, mass_track, albedo_bedrock, albedo_transport,
                     albedo_surface, rng, scale, param):
    """soil.transport_debris(...) - model.cpp:270-295 / erosion.hpp:86-98."""
    w = "transport_debris"
    L = _f(layers, w)
    nx, ny = _dims2(L, w)
    r = _gpu(rng, silt.rng, w)
    rc = _lib.load().soilb200_transport_debris(
        L.data_ptr(), _f(velocity, w).data_ptr(), _f(velocity_track, w).data_ptr(), _f(mass, w).data_ptr(),
        _f(mass_track, w).data_ptr(), _opt(albedo_bedrock, w), _opt(albedo_transport, w), _opt(albedo_surface, w),
        r.data_ptr(), r.numel() // 48, nx, ny, _lib.scale3(scale), param.c(), _stream())
    _lib.check(rc, w)


def mass_transfer(deltas, layers, uplift, discharge, mass, momentumFluvial, debris, momentumDebris, albedo_bedrock,
                  albedo_transport_fluvial, albedo_transport_debris, albedo_surface, scale, param):
    """soil.mass_transfer(...) - model.cpp:297-328 / erosion.hpp:104-119."""
    w = "mass_transfer"
    L = _f(layers, w)
    nx, ny = _dims2(L, w)
    rc = _lib.load().soilb200_mass_transfer(
        _f(deltas, w).data_ptr(), L.data_ptr(), _f(uplift, w).data_ptr(), _opt(discharge, w), _f(mass, w).data_ptr(),
        _f(momentumFluvial, w).data_ptr(), _f(debris, w).data_ptr(), _opt(momentumDebris, w), _opt(albedo_bedrock, w),
        _opt(albedo_transport_fluvial, w), _opt(albedo_transport_debris, w), _opt(albedo_surface, w), nx, ny,
        _lib.scale3(scale), param.c(), _stream())
    _lib.check(rc, w)


def mass_creep(delta, layers, scale, param):
    """soil.mass_creep(delta, layers, scale, param) - model.cpp:330-341."""
    L = _f(layers, "mass_creep")
    nx, ny = _dims2(L, "mass_creep")
    rc = _lib.load().soilb200_mass_creep(_f(delta, "mass_creep").data_ptr(), L.data_ptr(), nx, ny, _lib.scale3(scale),
                                         param.c(), _stream())
    _lib.check(rc, "mass_creep")


def _rgb(v, what):
    c = [float(x) for x in v]
    if len(c) != 3:
        raise ValueError(f"{what}: a colour has 3 components")
    return (_lib.C.c_float * 3)(*c)


def albedo_layer(albedo, albedoBedrock, albedoSediment, layers, scaleSediment, shiftSediment):
    """soil.albedo_layer(...) - model.cpp:353-369, erosion.cu:878-898."""
    A = _f(albedo, "albedo_layer")
    nx, ny = _dims2(A, "albedo_layer")
    rc = _lib.load().soilb200_albedo_layer(A.data_ptr(), _f(albedoBedrock, "albedo_layer").data_ptr(),
                                           _f(albedoSediment, "albedo_layer").data_ptr(),
                                           _f(layers, "albedo_layer").data_ptr(), nx, ny, float(scaleSediment),
                                           _rgb(shiftSediment, "albedo_layer"), _stream())
    _lib.check(rc, "albedo_layer")


def albedo_stratum(albedoBedrock, uplift, layers, scale, param, colorA, colorB, age, freq):
    """soil.albedo_stratum(...) - model.cpp:371-390, erosion.cu:828-855."""
    U = _f(uplift, "albedo_stratum")
    nx, ny = _dims2(U, "albedo_stratum")
    rc = _lib.load().soilb200_albedo_stratum(_f(albedoBedrock, "albedo_stratum").data_ptr(), U.data_ptr(),
                                             _f(layers, "albedo_stratum").data_ptr(), nx, ny, _lib.scale3(scale),
                                             param.c(), _rgb(colorA, "albedo_stratum"), _rgb(colorB, "albedo_stratum"),
                                             float(age), float(freq), _stream())
    _lib.check(rc, "albedo_stratum")


def albedo_discharge(albedo, discharge, colorDischarge, extinction, scale):
    """soil.albedo_discharge(...) - model.cpp:393-407, erosion.cu:900-917 (in place)."""
    A = _f(albedo, "albedo_discharge")
    nx, ny = _dims2(A, "albedo_discharge")
    rc = _lib.load().soilb200_albedo_discharge(A.data_ptr(), _f(discharge, "albedo_discharge").data_ptr(), nx, ny,
                                               _rgb(colorDischarge, "albedo_discharge"), float(extinction), float(scale),
                                               _stream())
    _lib.check(rc, "albedo_discharge")


def layer_merge(height, layers):
    """soil.layer_merge(height, layers) - model.cpp:343-351."""
    H = _f(height, "layer_merge")
    nx, ny = _dims2(H, "layer_merge")
    rc = _lib.load().soilb200_layer_merge(H.data_ptr(), _f(layers, "layer_merge").data_ptr(), nx, ny, _stream())
    _lib.check(rc, "layer_merge")


# ---- previous-API containers and the fused step ---------------------------------


def _zeros(dims, device):
    return torch.zeros(dims, dtype=torch.float32, device=device)


class data_t:
    """soil.data_t(shape): transported fields (model.cpp:101-140, commented out at HEAD): discharge, mass, debris
    (one float per cell), momentum, debris_momentum (two per cell).  As in the previous API the fields exist as soon as
    the object does (the scripts write `data.discharge[:] = 0.0`); here they are allocated, zeroed, on the GPU at
    first access - `track`, whose flux accumulators are private packed scratch in this library, therefore costs nothing.
    `shape` may be a 2-D shape or a flat cell count (`soil.data_t(index.elem())`)."""

    _FIELDS = {"discharge": 1, "mass": 1, "debris": 1, "momentum": 2, "debris_momentum": 2}

    def __init__(self, shp=None):
        self.shape = None if shp is None else (shp if isinstance(shp, silt.shape) else silt.shape(shp))
        self._t = {}
        self._on_access = None   # set by the fused step's state: its packed view may be ahead of these maps

    def _set(self, name, t):
        if self._on_access is not None:
            self._on_access()
        self._t[name] = t

    def _get(self, name):
        if self._on_access is not None:
            self._on_access()
        t = self._t.get(name)
        if t is None and self.shape is not None and torch.cuda.is_available():
            dims = tuple(self.shape) + ((2,) if self._FIELDS[name] == 2 else ())
            t = silt.tensor(silt.float32, _data=_zeros(dims, torch.device("cuda", torch.cuda.current_device())))
            self._t[name] = t
        return t


def _data_field(name):
    return property(lambda self: self._get(name), lambda self, t: self._set(name, t))


for _name in data_t._FIELDS:
    setattr(data_t, _name, _data_field(_name))


class map_t:
    """soil.map_t(shape, scale): terrain model (model.cpp:66-95, commented out at HEAD).

    Holds the two-layer terrain as HEAD's interleaved `layers` tensor {bedrock, sediment}, double-buffered for the
    fused step.  The previous API kept `height` and `sediment` as two independent tensors - the multiscale script
    resizes and assigns them one after the other and saves them with the note that the layers must be added
    (example/erosion_gpu_multiscale.py:110-124, 165-171) - so here

        height    is the base layer (bedrock), `bedrock` is the same map under HEAD's name,
        sediment  is the layer on top,
        surface   is their sum, the map the particles run on.

    The setters are independent of each other and of their order.  The getters return SNAPSHOTS (copies): writing into
    one does not change the model - assign it back.  `layers` is the live current buffer; the fused step ping-pongs
    between two buffers, so a `layers` tensor obtained before an `erode` call is stale after it.
    Optional albedo maps (HEAD's colour channels) are enabled with `enable_albedo`.
    """

    def __init__(self, shp, scale):
        self.shape = shp if isinstance(shp, silt.shape) else silt.shape(shp)
        self.scale = tuple(float(v) for v in scale)
        if not torch.cuda.is_available():
            raise RuntimeError("soil.map_t needs a CUDA device")
        dev = torch.device("cuda", torch.cuda.current_device())
        nx, ny = self.shape[0], self.shape[1]
        self._layers = [_zeros((nx, ny, 2), dev), _zeros((nx, ny, 2), dev)]
        self._cur = 0
        self._height = _zeros((nx, ny), dev)
        self._rainfall = torch.ones((nx, ny), dtype=torch.float32, device=dev)
        self._uplift = _zeros((nx, ny), dev)
        self._state = None  # lazily built by erode(): fluxes, rng, tags
        self._merged = False  # `_height` holds bedrock + sediment of the current layers
        self._edited = True   # the layers may have changed since the fused step last looked (view mode re-imports)
        self.albedo_bedrock = self.albedo_surface = None
        self.with_creep = True
        self.seed = 0

    def _map(self, t, what):
        return _gpu(t, silt.float32, what).reshape(self._height.shape)

    def _touch(self):
        self._merged, self._edited = False, True

    @property
    def layers(self):
        self._touch()  # the caller may edit the buffer in place
        return silt.tensor(silt.float32, _data=self._layers[self._cur])

    @property
    def height(self):
        return silt.tensor(silt.float32, _data=self._layers[self._cur][..., 0].contiguous())

    @height.setter
    def height(self, t):
        self._touch()
        self._layers[self._cur][..., 0] = self._map(t, "map_t.height")

    bedrock = height

    @property
    def sediment(self):
        return silt.tensor(silt.float32, _data=self._layers[self._cur][..., 1].contiguous())

    @sediment.setter
    def sediment(self, t):
        self._touch()
        self._layers[self._cur][..., 1] = self._map(t, "map_t.sediment")

    @property
    def surface(self):
        """bedrock + sediment (what layer_merge writes); a snapshot."""
        L = self._layers[self._cur]
        if not self._merged:
            st = self._state
            if st is not None and st.maps_stale and not self._edited:
                st.export()   # view mode: the merged heights live in the packed view
            else:
                torch.add(L[..., 0], L[..., 1], out=self._height)
            self._merged = True
        return silt.tensor(silt.float32, _data=self._height.clone())

    @property
    def rainfall(self):
        return silt.tensor(silt.float32, _data=self._rainfall)

    @rainfall.setter
    def rainfall(self, t):
        self._rainfall = self._map(t, "map_t.rainfall").contiguous().clone()

    @property
    def uplift(self):
        return silt.tensor(silt.float32, _data=self._uplift)

    @uplift.setter
    def uplift(self, t):
        self._uplift = self._map(t, "map_t.uplift").contiguous().clone()

    def enable_albedo(self, bedrock_rgb=(0.5, 0.5, 0.5), surface_rgb=(0.5, 0.5, 0.5)):
        nx, ny = self.shape[0], self.shape[1]
        dev = self._height.device
        self.albedo_bedrock = torch.tensor(bedrock_rgb, dtype=torch.float32, device=dev).repeat(nx, ny, 1).contiguous()
        self.albedo_surface = torch.tensor(surface_rgb, dtype=torch.float32, device=dev).repeat(nx, ny, 1).contiguous()
        if self._state is not None:
            self._state.export()
            self._state.data._on_access = None
        self._state = None


class _state:
    """Device-resident scratch of the fused step: packed flux records (x2, ping-pong), tags, RNG streams, mass sums."""

    def __init__(self, model: map_t, data: data_t, n_particles: int):
        lib = _lib.load()
        dev = model._height.device
        nx, ny = model.shape[0], model.shape[1]
        self.n_particles = int(n_particles)

        def field(obj, name, dims):
            t = getattr(obj, name, None)
            if t is None:   # a data_t built without a shape
                t = silt.tensor(silt.float32, _data=_zeros(dims, dev))
                setattr(obj, name, t)
            d = _gpu(t, silt.float32, name)
            if d.numel() != int(np.prod(dims)):
                raise ValueError(f"erode: {name} has the wrong size")
            return d

        self.discharge = field(data, "discharge", (nx, ny))
        self.mass = field(data, "mass", (nx, ny))
        self.debris = field(data, "debris", (nx, ny))
        self.momentum = field(data, "momentum", (nx, ny, 2))
        self.debris_momentum = field(data, "debris_momentum", (nx, ny, 2))
        self.albedo = model.albedo_surface is not None
        nflux = lib.soilb200_erode_flux_bytes(nx, ny, 1 if self.albedo else 0)
        # the dense pass only reads the current records and re-zeroes the previous step's with pure
        # stores (read-then-zero of the same lines is ~5x slower on B200), hence two sets
        self.flux = [torch.zeros(nflux, dtype=torch.uint8, device=dev) for _ in range(2)]
        self.tags = [torch.zeros(lib.soilb200_erode_ws(nx, ny), dtype=torch.uint8, device=dev) for _ in range(2)]
        self.cur = 0
        self.pre = torch.empty(lib.soilb200_erode_pre_bytes(nx, ny), dtype=torch.uint8, device=dev)
        self.rng_f = torch.empty((self.n_particles, 48), dtype=torch.uint8, device=dev)
        self.rng_d = torch.empty((self.n_particles, 48), dtype=torch.uint8, device=dev)
        s = _stream()
        _lib.check(lib.soilb200_seed_xorwow(self.rng_f.data_ptr(), self.n_particles, model.seed, 0, 0, s), "seed")
        _lib.check(lib.soilb200_seed_xorwow(self.rng_d.data_ptr(), self.n_particles, model.seed + 1, 0, 0, s), "seed")
        self.mass_sum = torch.zeros(2, dtype=torch.float64, device=dev)
        self.c = _lib.Model()
        # packed particle view (include/soilb200.h, soilb200_model_t.view): the dense pass writes it instead of the
        # maps height / discharge / momentum / debris / debris_momentum, which are brought up to date when somebody
        # looks at them (export) and folded back in when somebody may have written them (import)
        self.view = None
        if not os.environ.get("SOILB200_NO_VIEW"):
            self.view = torch.empty(lib.soilb200_erode_view_bytes(nx, ny), dtype=torch.uint8, device=dev)
        self.view_valid = False    # the view holds what the maps and the layers say
        self.maps_stale = False    # the view is ahead of the maps
        self.model, self.data = model, data
        data._on_access = self._data_accessed

    def _data_accessed(self):
        """Somebody asked for (or assigned) one of the transported maps: bring them up to date first; the caller may
        write into what it gets, so the view is rebuilt before the next step."""
        self.export()
        self.view_valid = False

    def export(self):
        if self.view is not None and self.maps_stale:
            self.fill(self.model, False)
            _lib.check(_lib.load().soilb200_erode_export(C.byref(self.c), _stream()), "erode export")
            self.maps_stale = False

    def fill(self, model: map_t, want_sum: bool):
        c, p = self.c, (lambda t: t.data_ptr())
        c.nx, c.ny = model.shape[0], model.shape[1]
        c.scale[0], c.scale[1], c.scale[2] = model.scale
        c.with_creep = 1 if model.with_creep else 0
        c.layers, c.layers_next = p(model._layers[model._cur]), p(model._layers[1 - model._cur])
        c.height, c.rainfall, c.uplift = p(model._height), p(model._rainfall), p(model._uplift)
        c.discharge, c.mass, c.momentum = p(self.discharge), p(self.mass), p(self.momentum)
        c.debris, c.debris_momentum = p(self.debris), p(self.debris_momentum)
        if self.albedo:
            c.albedo_bedrock, c.albedo_surface = p(model.albedo_bedrock), p(model.albedo_surface)
        else:
            c.albedo_bedrock = c.albedo_surface = None
        c.rng_fluvial, c.rng_debris, c.n_particles = p(self.rng_f), p(self.rng_d), self.n_particles
        c.flux, c.flux_prev = p(self.flux[self.cur]), p(self.flux[1 - self.cur])
        c.touched, c.touched_prev = p(self.tags[self.cur]), p(self.tags[1 - self.cur])
        c.mass_sum = p(self.mass_sum) if want_sum else None
        c.pre = p(self.pre)
        c.view = p(self.view) if self.view is not None else None


def erode(model: map_t, data: data_t, track: data_t, param: param_t, steps: int = 1, mass_sum: bool = False):
    """soil.erode(model, data, track, param, steps) (example/erosion_gpu.py:103-106).

    Each step = fluvial + debris particle transport, normalise, mass transfer,
    creep (if model.with_creep), apply, merge as one particle
    kernel and one fused dense kernel.  `track` is accepted for signature
    compatibility; the flux accumulators it used to hold are private packed
    scratch here.  Asynchronous on the current stream.  With mass_sum=True
    returns a 2-element float64 GPU tensor {sum bedrock, sum sediment} of the
    last step.
    """
    n_particles = int(param.legacy.get("samples", 8192))
    st = model._state
    if st is None or st.n_particles != n_particles or getattr(st, "_owner", None) is not data:
        if st is not None:      # the maps of the data it served are brought up to date before it goes
            st.export()
            st.data._on_access = None
        st = _state(model, data, n_particles)
        st._owner = data
        model._state = st
    lib = _lib.load()
    s = _stream()
    # the particles read the merged surface: make it current if the layers may have been edited since the last step
    # (map_t.layers / .height / .sediment handed out or assigned); the fused step itself leaves it current
    if st.view is not None:
        # view mode: the particles gather from the packed view; rebuild it if the layers or the transported maps may have
        # been edited since the last step
        if int(steps) > 0 and (not st.view_valid or model._edited):
            st.export()      # the transported maps first, if the view is ahead of them: the import reads them
            st.fill(model, False)
            _lib.check(lib.soilb200_erode_import(C.byref(st.c), s), "erode import")
            st.view_valid, st.maps_stale = True, False
    elif not model._merged:
        _lib.check(lib.soilb200_layer_merge(model._height.data_ptr(), model._layers[model._cur].data_ptr(),
                                            model.shape[0], model.shape[1], s), "layer_merge")
    if int(steps) > 0:
        model._edited = False
        model._merged = st.view is None     # the fused step keeps `_height` current only without the view
        st.maps_stale = st.view is not None
    for i in range(int(steps)):
        st.fill(model, mass_sum and i == int(steps) - 1)
        _lib.check(lib.soilb200_erode_step(C.byref(st.c), param.c(), s), "erode")
        model._cur = 1 - model._cur
        st.cur = 1 - st.cur
    return st.mass_sum if mass_sum else None
