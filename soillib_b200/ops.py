"""Host-side mirror of the reference's op bindings (python/source/model.cpp:157-421).

Same names, argument order and error behaviour as the nanobind lambdas; each
function validates its tensors, allocates the output the reference's op would
allocate, and forwards to one C-ABI entry point on the current torch stream.
"""
from __future__ import annotations

import ctypes as C
import enum

import torch

from . import _lib, silt


class edge(enum.IntEnum):
    """soil::edge_t (graph.hpp:11-14), exported into module scope like nb::enum_::export_values."""

    d4 = 0
    d8 = 1


d4, d8 = edge.d4, edge.d8


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _gpu(t: silt.tensor, dtype, what: str) -> torch.Tensor:
    """tensor.as<T>() plus the host check of graph.cu:75-76."""
    if not isinstance(t, silt.tensor):
        raise TypeError(f"{what}: expected a silt.tensor")
    if t.type is not dtype:
        raise ValueError(f"{what}: expected dtype {dtype.name}, got {t.type.name}")
    if not t.data.is_cuda:
        raise ValueError(f"{what}: tensor host mismatch (expected gpu, got cpu)")
    if not t.data.is_contiguous():
        t.data = t.data.contiguous()
    return t.data


def _edge(e) -> int:
    try:
        return int(e)
    except Exception as exc:  # noqa: BLE001
        raise ValueError("invalid edge enumerator") from exc


def _dims2(t: torch.Tensor, what: str):
    if t.dim() < 2:
        raise ValueError(f"{what}: expected a 2-D tensor")
    return int(t.shape[0]), int(t.shape[1])


# scratch the reference allocates inside its ops; cached per (op, device, shape) AND per stream: two calls of one op on
# different torch streams may overlap on the GPU and must not share counters
_scratch: dict = {}


def _workspace(key, nbytes: int, device, zero_head: bool = False) -> torch.Tensor:
    key = (key, torch.cuda.current_stream(device).cuda_stream if torch.device(device).type == "cuda" else 0)
    ws = _scratch.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 64), dtype=torch.uint8, device=device)
        if zero_head:
            ws[:64].zero_()
        _scratch[key] = ws
    return ws


def direction(height: silt.tensor, e) -> silt.tensor:
    """soil.direction(height, edge) -> int32 direction code per cell (graph.cu:246-264)."""
    h = _gpu(height, silt.float32, "direction")
    nx, ny = _dims2(h, "direction")
    out = torch.empty((nx, ny), dtype=torch.int32, device=h.device)
    _lib.check(_lib.load().soilb200_direction_f32(h.data_ptr(), out.data_ptr(), nx, ny, _edge(e), _stream()), "direction")
    return silt.tensor(silt.int32, _data=out)


def steepest(height: silt.tensor, e) -> silt.tensor:
    """soil.steepest(height, edge) -> int32 receiver index per cell (graph.cu:73-91)."""
    h = _gpu(height, silt.float32, "steepest")
    nx, ny = _dims2(h, "steepest")
    out = torch.empty((nx, ny), dtype=torch.int32, device=h.device)
    _lib.check(_lib.load().soilb200_steepest_f32(h.data_ptr(), out.data_ptr(), nx, ny, _edge(e), _stream()), "steepest")
    return silt.tensor(silt.int32, _data=out)


flow = steepest  # previous API name


def random_weighted(height: silt.tensor, e, seed: int, offset: int, T: float, first_cell: int = 0) -> silt.tensor:
    """soil.random_weighted(height, edge, seed, offset, T) (graph.cu:175-195).

    The per-cell generator state the reference re-creates on every call is kept
    in cached scratch, so a loop over offset = 0, 1, 2, ... (dem_multiflow.py:43-47)
    seeds once and then advances each stream by one draw - bit-identical.
    first_cell (not in the reference): the global index of the block's first cell when `height` is a row strip of a
    larger map, so that every cell draws from the stream the undivided map gives it.
    """
    h = _gpu(height, silt.float32, "random_weighted")
    nx, ny = _dims2(h, "random_weighted")
    lib = _lib.load()
    nbytes = lib.soilb200_random_weighted_ws(nx, ny)
    ws = _workspace(("rw", h.device, nx, ny, int(first_cell)), nbytes, h.device, zero_head=True)
    out = torch.empty((nx, ny), dtype=torch.int32, device=h.device)
    rc = lib.soilb200_random_weighted_strip_f32(h.data_ptr(), out.data_ptr(), nx, ny, _edge(e), int(seed), int(offset),
                                                float(T), int(first_cell), ws.data_ptr(), ws.numel(), _stream())
    _lib.check(rc, "random_weighted")
    return silt.tensor(silt.int32, _data=out)


def _accumulate(graph, field, decay, e, what):
    g = _gpu(graph, silt.int32, what)
    f = _gpu(field, silt.float32, what)
    nx, ny = _dims2(g, what)
    if tuple(f.shape[:2]) != (nx, ny):
        raise ValueError(f"{what}: graph and field shapes differ")
    d_ptr = None
    if decay is not None:
        d = _gpu(decay, silt.float32, what)
        if tuple(d.shape[:2]) != (nx, ny):
            raise ValueError(f"{what}: graph and decay shapes differ")
        d_ptr = d.data_ptr()
    lib = _lib.load()
    ek = _edge(e)
    size_of = lib.soilb200_accumulate_ws if decay is None else lib.soilb200_accumulate_decay_ws
    nbytes = size_of(nx, ny, ek if ek in (0, 1) else 1)
    ws = _workspace(("acc", g.device, nx, ny, decay is not None), nbytes, g.device)
    out = torch.empty((nx, ny), dtype=torch.float32, device=g.device)
    rc = lib.soilb200_accumulate_f32(g.data_ptr(), f.data_ptr(), d_ptr, out.data_ptr(), nx, ny, ek, ws.data_ptr(),
                                     ws.numel(), _stream())
    _lib.check(rc, what)
    return silt.tensor(silt.float32, _data=out)


def accumulate(graph: silt.tensor, field: silt.tensor, e) -> silt.tensor:
    """soil.accumulate(graph, field, edge) (graph.cu:578-584)."""
    return _accumulate(graph, field, None, e, "accumulate")


def accumulate_decay(graph: silt.tensor, field: silt.tensor, decay: silt.tensor, e) -> silt.tensor:
    """soil.accumulate_decay(graph, field, decay, edge) (graph.cu:586-593)."""
    return _accumulate(graph, field, decay, e, "accumulate_decay")


accumulation = accumulate  # previous API name


def normal(t: silt.tensor, scale=(1.0, 1.0, 1.0)) -> silt.tensor:
    """soil.normal(tensor, scale) -> (nx, ny, 3) float32 (op/normal.hpp:18-39)."""
    if t.shape.dim() != 2:
        raise ValueError("normal map can not be computed for non 2D-indexed buffers")  # normal.hpp:23
    h = _gpu(t, silt.float32, "normal")
    nx, ny = _dims2(h, "normal")
    out = torch.empty((nx, ny, 3), dtype=torch.float32, device=h.device)
    _lib.check(_lib.load().soilb200_normal_f32(h.data_ptr(), out.data_ptr(), nx, ny, _lib.scale3(scale), _stream()), "normal")
    return silt.tensor(silt.float32, _data=out)


def _scale2(scale):
    v = [float(x) for x in scale]
    if len(v) != 2:
        raise ValueError("scale must have 2 components")
    return (_lib.C.c_float * 2)(*v)


def gradient(t: silt.tensor, scale=(1.0, 1.0)) -> silt.tensor:
    """soil.gradient(tensor, scale) -> (nx, ny, 2) float32 (grad.cu:22-98)."""
    h = _gpu(t, silt.float32, "gradient")
    nx, ny = _dims2(h, "gradient")
    out = torch.empty((nx, ny, 2), dtype=torch.float32, device=h.device)
    _lib.check(_lib.load().soilb200_gradient_f32(h.data_ptr(), out.data_ptr(), nx, ny, _scale2(scale), _stream()), "gradient")
    return silt.tensor(silt.float32, _data=out)


def negslope(t: silt.tensor, scale=(1.0, 1.0)) -> silt.tensor:
    """soil.negslope(tensor, scale) -> (nx, ny) float32 (grad.cu:102-140)."""
    h = _gpu(t, silt.float32, "negslope")
    nx, ny = _dims2(h, "negslope")
    out = torch.empty((nx, ny), dtype=torch.float32, device=h.device)
    _lib.check(_lib.load().soilb200_negslope_f32(h.data_ptr(), out.data_ptr(), nx, ny, _scale2(scale), _stream()), "negslope")
    return silt.tensor(silt.float32, _data=out)


def _channels(h: torch.Tensor, what: str) -> int:
    if h.dim() != 3 or h.shape[2] not in (1, 2):
        raise ValueError(f"{what}: expected a (nx, ny, 1) or (nx, ny, 2) tensor")   # grad.cu:193-199, filter.cu:79-88
    return int(h.shape[2])


def laplacian(t: silt.tensor, scale=(1.0, 1.0)) -> silt.tensor:
    """soil.laplacian(tensor, scale): component-wise 9-point Laplacian of a (nx, ny, 1|2) tensor (grad.cu:137-203)."""
    h = _gpu(t, silt.float32, "laplacian")
    ch = _channels(h, "laplacian")
    nx, ny = int(h.shape[0]), int(h.shape[1])
    out = torch.empty_like(h)
    _lib.check(_lib.load().soilb200_laplacian_f32(h.data_ptr(), out.data_ptr(), nx, ny, ch, _scale2(scale), _stream()), "laplacian")
    return silt.tensor(silt.float32, _data=out)


def slope(t: silt.tensor, flow_graph: silt.tensor, scale=(1.0, 1.0)) -> silt.tensor:
    """soil.slope(tensor, flow, scale): slope along each cell's flow edge, 0 without a receiver (graph.cu:270-311)."""
    h = _gpu(t, silt.float32, "slope")
    g = _gpu(flow_graph, silt.int32, "slope")
    nx, ny = _dims2(h, "slope")
    if tuple(g.shape[:2]) != (nx, ny):
        raise ValueError("slope: tensor and flow shapes differ")
    out = torch.empty((nx, ny), dtype=torch.float32, device=h.device)
    _lib.check(_lib.load().soilb200_slope_f32(h.data_ptr(), g.data_ptr(), out.data_ptr(), nx, ny, _scale2(scale), _stream()), "slope")
    return silt.tensor(silt.float32, _data=out)


def gaussian_blur(t: silt.tensor, sigma: float) -> silt.tensor:
    """soil.gaussian_blur(tensor, sigma): like the reference's (filter.cu:82,90) it OVERWRITES its argument and returns it."""
    h = _gpu(t, silt.float32, "gaussian_blur")
    ch = _channels(h, "gaussian_blur")
    nx, ny = int(h.shape[0]), int(h.shape[1])
    tmp = torch.empty_like(h)
    _lib.check(_lib.load().soilb200_gaussian_blur_f32(h.data_ptr(), tmp.data_ptr(), nx, ny, ch, float(sigma), _stream()), "gaussian_blur")
    return t


def solve_uniform(flow_field: silt.tensor, source: silt.tensor, decay: silt.tensor, rng: silt.tensor, scale=(1.0, 1.0),
                  count: int = 1) -> silt.tensor:
    """soil.solve_uniform(flow, source, decay, rng, scale, count) (path.cu:169-217, model.cpp:209-226): Monte-Carlo
    estimate of the transported quantity; one sample per rng state, `count` = the number the estimate is divided by."""
    f = _gpu(flow_field, silt.float32, "solve_uniform")
    src = _gpu(source, silt.float32, "solve_uniform")
    dec = _gpu(decay, silt.float32, "solve_uniform")
    if rng.type is not silt.rng or not rng.data.is_cuda:
        raise ValueError("solve_uniform: rng must be a gpu tensor of generator states")
    if src.dim() != 3 or src.shape[2] not in (1, 2):
        raise ValueError("solve_uniform: source must be (nx, ny, 1) or (nx, ny, 2)")
    nx, ny, k = int(src.shape[0]), int(src.shape[1]), int(src.shape[2])
    if tuple(f.shape) != (nx, ny, 2) or tuple(dec.shape[:2]) != (nx, ny):
        raise ValueError("solve_uniform: flow must be (nx, ny, 2) and decay (nx, ny)")
    if count <= 0:
        raise ValueError("solve_uniform: count must be positive")
    out = torch.empty_like(src)
    rc = _lib.load().soilb200_solve_uniform_f32(f.data_ptr(), src.data_ptr(), dec.data_ptr(), rng.data.data_ptr(),
                                                rng.data.numel() // 48, out.data_ptr(), nx, ny, k, _scale2(scale),
                                                int(count), _stream())
    _lib.check(rc, "solve_uniform")
    return silt.tensor(silt.float32, _data=out)


class noise_t:
    """soil.noise_t (op/noise.hpp:14-40)."""

    def __init__(self):
        self.frequency = 1.0
        self.octaves = 8
        self.gain = 0.6
        self.lacunarity = 2.0
        self.seed = 0.0
        self.ext = (512.0, 512.0)


def noise(shp, param: noise_t, host=silt.cpu) -> silt.tensor:
    """soil.noise(shape, noise_t) -> float32 fBm map (op/noise.hpp:42-56).

    Like the reference it returns a CPU tensor; host=silt.gpu generates the same
    bits with one kernel directly in device memory.
    """
    shp = shp if isinstance(shp, silt.shape) else silt.shape(shp)
    if shp.dim() != 2:
        raise ValueError("can't extract a full noise buffer from a non-2D index")  # noise.hpp:44-45
    out = silt.tensor(silt.float32, shp, host)
    on_dev = 1 if host is silt.gpu else 0
    ext = [float(v) for v in param.ext]
    rc = _lib.load().soilb200_noise_f32(out.ptr(), shp[0], shp[1], float(param.frequency), int(param.octaves),
                                        float(param.gain), float(param.lacunarity), float(param.seed), ext[0], ext[1],
                                        on_dev, _stream() if on_dev else None)
    _lib.check(rc, "noise")
    return out
