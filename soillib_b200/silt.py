"""Stand-in for the `silt` tensor module that soillib's scripts import next to `soillib`.

The real `silt` (erosiv/silt) is a private submodule that is absent from the
reference tree; its surface is inferred from its call sites:
example/erosion_gpu.py:17-20,26,36, example/dem_multiflow.py:25,31,49,
example/dem_process.py:72-81.  Storage is a torch tensor (plumbing only).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib


class _dtype:
    def __init__(self, name, torch_dtype, itemsize):
        self.name, self.torch, self.itemsize = name, torch_dtype, itemsize

    def __repr__(self):
        return f"silt.{self.name}"


float32 = _dtype("float32", torch.float32, 4)
float64 = _dtype("float64", torch.float64, 8)
int32 = _dtype("int32", torch.int32, 4)
rng = _dtype("rng", torch.uint8, 48)  # curandStateXORWOW, App. A-4


class _host:
    def __init__(self, name):
        self.name = name

    def __repr__(self):
        return f"silt.{self.name}"


cpu = _host("cpu")
gpu = _host("gpu")


class shape:
    """silt.shape(*dims): row-major, flatten(x,y) = x*shape[1] + y (App. A-2)."""

    def __init__(self, *dims):
        if len(dims) == 1 and not np.isscalar(dims[0]):
            dims = tuple(dims[0])
        self.dims = tuple(int(d) for d in dims)

    def __getitem__(self, i):
        return self.dims[i]

    def __len__(self):
        return len(self.dims)

    def __iter__(self):
        return iter(self.dims)

    def __eq__(self, other):
        return tuple(self) == tuple(other)

    def elem(self):
        return int(np.prod(self.dims)) if self.dims else 0

    def dim(self):
        return len(self.dims)

    def flatten(self, pos):
        return int(pos[0]) * self.dims[1] + int(pos[1])

    def unflatten(self, n):
        return (int(n) // self.dims[1], int(n) % self.dims[1])

    def oob(self, pos):
        return any(p < 0 or p >= d for p, d in zip(pos, self.dims))

    def __repr__(self):
        return f"silt.shape{self.dims}"


def _device():
    if not torch.cuda.is_available():
        raise RuntimeError("silt.gpu requested but no CUDA device is available")
    return torch.device("cuda", torch.cuda.current_device())


class tensor:
    """silt.tensor(dtype, shape, host) - typed n-d array on the host or on the GPU."""

    def __init__(self, dtype, shp=None, host=cpu, _data=None):
        if _data is not None:
            self.type, self.data = dtype, _data
            return
        shp = shp if isinstance(shp, shape) else shape(shp)
        dims = tuple(shp) + ((48,) if dtype is rng else ())
        dev = _device() if host is gpu else torch.device("cpu")
        self.type = dtype
        self.data = torch.zeros(dims, dtype=dtype.torch, device=dev)

    # -- construction / conversion
    @staticmethod
    def from_numpy(arr):
        arr = np.ascontiguousarray(arr)
        for dt in (float32, float64, int32):
            if arr.dtype == np.dtype(dt.name):
                return tensor(dt, _data=torch.from_numpy(arr.copy()))
        raise ValueError(f"unsupported numpy dtype {arr.dtype}")

    @staticmethod
    def from_torch(t, dtype=None):
        if dtype is None:
            dtype = {torch.float32: float32, torch.float64: float64, torch.int32: int32}[t.dtype]
        return tensor(dtype, _data=t.contiguous())

    def numpy(self):
        return self.data.detach().cpu().numpy()

    def torch(self):
        return self.data

    def cpu(self):
        return tensor(self.type, _data=self.data.cpu())

    def gpu(self):
        return tensor(self.type, _data=self.data.to(_device()))

    # -- introspection
    @property
    def shape(self):
        dims = tuple(self.data.shape)
        return shape(*(dims[:-1] if self.type is rng else dims))

    @property
    def host(self):
        return gpu if self.data.is_cuda else cpu

    def elem(self):
        return self.shape.elem()

    def ptr(self):
        return self.data.data_ptr()

    def __getitem__(self, idx):
        return self.data[idx]

    def __setitem__(self, idx, value):
        self.data[idx] = torch.as_tensor(value, dtype=self.data.dtype, device=self.data.device)

    def __repr__(self):
        return f"silt.tensor({self.type.name}, {tuple(self.shape)}, {self.host.name})"


def set(t, value):
    """silt.set(tensor, scalar | tensor)"""
    if isinstance(value, tensor):
        t.data.copy_(value.data)
    else:
        t.data.fill_(value)


def add(a, b):
    a.data.add_(b.data if isinstance(b, tensor) else b)


def multiply(t, s):
    t.data.mul_(s.data if isinstance(s, tensor) else s)


def clamp(t, lo, hi):
    t.data.clamp_(lo, hi)


def clone(t):
    return tensor(t.type, _data=t.data.clone())


def resize(target, source, newres, oldres):
    """soil.resize(target, source, newres, oldres) of the multiscale script (example/erosion_gpu_multiscale.py:112-138):
    resample a map given at `oldres` = (nx, ny) into `target` at `newres`.  Both may be flat buffers of nx*ny cells or
    shaped maps, with or without interleaved trailing components (momentum: 2 per cell).

    The body lives in the absent `silt`; this is our stated definition (parity unpinned): bilinear interpolation between
    cell centres, edge cells continued outwards - torch's `interpolate(mode="bilinear", align_corners=False)`, plumbing
    between two erosion runs, not part of the hot path.  Equal resolutions copy."""
    ox, oy = (int(v) for v in oldres)
    nx, ny = (int(v) for v in newres)
    src, dst = source.data, target.data
    if src.numel() % (ox * oy) or dst.numel() % (nx * ny) or src.numel() // (ox * oy) != dst.numel() // (nx * ny):
        raise ValueError("silt.resize: tensor sizes do not match the resolutions")
    if not src.dtype.is_floating_point or dst.dtype != src.dtype:
        raise ValueError("silt.resize expects floating-point tensors of one type")
    ch = src.numel() // (ox * oy)
    planes = src.reshape(ox, oy, ch).permute(2, 0, 1).unsqueeze(0)          # (1, ch, ox, oy)
    if (ox, oy) != (nx, ny):
        planes = torch.nn.functional.interpolate(planes, size=(nx, ny), mode="bilinear", align_corners=False)
    dst.copy_(planes.squeeze(0).permute(1, 2, 0).reshape(dst.shape))


def seed(rng_tensor, seed_value, offset=0):
    """silt.seed(rng, seed, offset): curand_init(seed, i, offset, &rng[i]) (App. A-4)."""
    if rng_tensor.type is not rng:
        raise ValueError("silt.seed expects a tensor of type silt.rng")
    if not rng_tensor.data.is_cuda:
        raise ValueError("silt.seed expects a GPU tensor")
    lib = _lib.load()
    n = rng_tensor.data.numel() // 48
    rc = lib.soilb200_seed_xorwow(rng_tensor.ptr(), n, int(seed_value), 0, int(offset),
                                  torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "silt.seed")
